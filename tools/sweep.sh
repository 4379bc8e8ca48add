#!/bin/bash
# scaling sweep on one multi-GPU box: weak (default workload per N) and strong (1/120 degree in 8 row-slabs)
mkdir -p gpurun_out
run() { # N extra-args tag
  N=$1; shift; TAG=$1; shift
  if [ "$N" = 1 ]; then python bench.py --gpus 1 "$@" > gpurun_out/sweep_${TAG}_n1.json 2> gpurun_out/sweep_${TAG}_n1.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N "$@" > gpurun_out/sweep_${TAG}_n$N.json 2> gpurun_out/sweep_${TAG}_n$N.err; fi
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/sweep_${TAG}_n$N.json"))
    print("${TAG} N=$N", d["config"]["workload"][:24], "ms/step %.3f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"], "e2e", (d.get("e2e") or {}).get("value"))
except Exception as e:
    print("${TAG} N=$N FAILED", e); print(open("gpurun_out/sweep_${TAG}_n$N.err").read()[-1500:])
PY
}
NG=$(nvidia-smi -L | wc -l)
for N in 1 2 4 8; do [ $N -le $NG ] && run $N weak --steps 20 --warmup 3 --no_cpu_baseline; done
for N in 1 2 4 8; do [ $N -le $NG ] && run $N r120 --resolution 120 --slabs 8 --steps 3 --warmup 3 --no_cpu_baseline; done
[ $NG -ge 2 ] && python -m pytest tests/test_gpu_multi_nccl.py -q 2>&1 | tail -2
