#!/bin/bash
# GPU call E (8 GPUs): topology, the NCCL tests at world 8, the device->host ceiling of the box (with / without NUMA-local
# pinned buffers) at 1/2/4/8 ranks, bench N=8 and N=4 (weak line + halo_midas + the strong runs inside)
mkdir -p gpurun_out
T=${1:-r02g}
(nvidia-smi topo -m; lscpu | grep -i -E "numa|socket|model name|^cpu\(s\)"; free -g | head -2) > gpurun_out/${T}_topology.txt 2>&1
(timeout 600 python -m pytest tests/test_gpu_multi_nccl.py -q -rs -v 2>&1 | tail -30) > gpurun_out/${T}_pytest_nccl.log
tail -3 gpurun_out/${T}_pytest_nccl.log
for N in 1 2 4 8; do for B in "" "--bind"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) tools/d2h_ceiling.py 350 $B 2>/dev/null | grep '^{'
done; done | tee gpurun_out/${T}_d2h_ceiling.jsonl
for N in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_bench_n$N.json"))
    print("N=$N ms/step %.4f"%d["ms_per_step"], "value %.4g"%d["value"], "frac %.3f"%d["roofline"]["frac"], "e2e", d["e2e"], d.get("halo_midas"), {k:(round(v["ms_per_step"],3), v["area_m2"]) for k,v in (d.get("strong") or {}).items()})
except Exception as e:
    print("N=$N FAILED", e); print(open("gpurun_out/${T}_bench_n$N.err").read()[-2500:])
PY
done
OGG_NUMA_BIND=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 10 --warmup 3 --no_strong > gpurun_out/${T}_bench_n8_nobind.json 2> gpurun_out/${T}_bench_n8_nobind.err
python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_n8_nobind.json')); print('N=8 no bind e2e', d['e2e'])"
