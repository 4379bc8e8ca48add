#!/bin/bash
# GPU call G (1 GPU): whole -m gpu suite, bench N=1 + reference arm, then the ncu evidence (tools/gpu_call_d.sh)
mkdir -p gpurun_out
T=${1:-r02n}
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${T}_gpus.csv 2>&1
(timeout 1800 python -m pytest tests -m gpu -x -q -rs 2>&1 | tail -15) | tee gpurun_out/${T}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref_n1.json 2> gpurun_out/${T}_ref_n1.err; echo "ref rc=$?"
timeout 300 python tools/stepbench.py 4 100 2>&1 | tail -2 | tee gpurun_out/${T}_stepbench.log
python - <<PY
import json
d=json.load(open("gpurun_out/${T}_bench_n1.json"))
print("ms/step %.4f"%d["ms_per_step"], "value %.4g"%d["value"], "frac %.3f issue %.3f"%(d["roofline"]["frac"], d["roofline"]["frac_issue"]), "e2e", d["e2e"].get("ms"), d.get("cli"))
print(d["kernels_ms_per_step_single_stream_pass"])
PY
tools/gpu_call_d.sh ${T}
