#!/bin/bash
# GPU call D (1 GPU): the ncu evidence for profiles/ -- launch list of the bench command and one `--set full` capture of the
# top kernels, each only after the same command has exited 0 without ncu (B200_PROFILING.md)
mkdir -p gpurun_out
T=${1:-r02}
CMD="python bench.py --steps 2 --warmup 3 --no_strong --no_cli --no_cpu_baseline"
timeout 600 $CMD > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${T}_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu1.log 2>&1
echo "launch list rc=$?"
# the last full build of the run: skip the launches of the warm-up steps (pattern-matched kernels only)
timeout 1500 ncu --set full --clock-control none --import-source on -k 'regex:k_dp_hsamples|k_bp_quad_fused|k_band_fill|k_quad_cells|k_bp_lattice|k_angle' -s 8 -c 8 \
    -o gpurun_out/${T}_full -f $CMD > gpurun_out/${T}_ncu2.log 2>&1
echo "full capture rc=$?"
grep -c "k_" gpurun_out/${T}_launches.csv; grep "Profiling" gpurun_out/${T}_ncu2.log | head -12; ls -la gpurun_out/${T}_full.ncu-rep
