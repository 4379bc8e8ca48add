#!/bin/bash
# GPU call A (2 GPUs): the whole -m gpu suite incl. the NCCL tests, bench N=1, bench N=2, the CPU reference arm
mkdir -p gpurun_out
T=${1:-r02a}
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${T}_gpus.csv 2>&1
(timeout 1500 python -m pytest tests -m gpu -x -q -rs 2>&1 | tail -40) > gpurun_out/${T}_pytest.log
(timeout 600 python -m pytest tests/test_gpu_multi_nccl.py -q -rs -v 2>&1 | tail -30) > gpurun_out/${T}_pytest_nccl.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_n1.json 2> gpurun_out/${T}_bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_ref_n1.json 2> gpurun_out/${T}_ref_n1.err
tail -5 gpurun_out/${T}_pytest.log; tail -3 gpurun_out/${T}_pytest_nccl.log
python - <<PY
import json
for n in ("n1","n2"):
    try:
        d=json.load(open("gpurun_out/${T}_bench_%s.json"%n))
        print(n, "ms/step %.4f"%d["ms_per_step"], "value %.4g"%d["value"], "frac %.3f issue %.3f"%(d["roofline"]["frac"], d["roofline"]["frac_issue"]), "e2e", d["e2e"]["ms"], d.get("halo_midas"), {k:(round(v["ms_per_step"],3), v["area_m2"]) for k,v in (d.get("strong") or {}).items()}, d.get("cli"))
    except Exception as e:
        print(n, "FAILED", e); print(open("gpurun_out/${T}_bench_%s.err"%n).read()[-2500:])
PY
