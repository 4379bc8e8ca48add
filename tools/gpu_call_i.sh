#!/bin/bash
# GPU call I (8 GPUs): the NCCL tests at world 8, bench N=8 and N=4 (weak line + halo_midas + the strong runs inside)
mkdir -p gpurun_out
T=${1:-r02w}
(timeout 600 python -m pytest tests/test_gpu_multi_nccl.py -q -rs -v 2>&1 | tail -12) > gpurun_out/${T}_pytest_nccl.log
tail -3 gpurun_out/${T}_pytest_nccl.log
for N in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_bench_n$N.json"))
    print("N=$N ms/step %.4f"%d["ms_per_step"], "value %.4g"%d["value"], "frac %.3f"%d["roofline"]["frac"], "e2e", d["e2e"]["ms"], d.get("halo_midas"), {k:(round(v["ms_per_step"],3), v["area_m2"]) for k,v in (d.get("strong") or {}).items()})
except Exception as e:
    print("N=$N FAILED", e); print(open("gpurun_out/${T}_bench_n$N.err").read()[-2500:])
PY
done
