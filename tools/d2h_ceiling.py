#!/usr/bin/env python
"""What the box gives N ranks copying device -> pinned host memory at once (the floor under bench.py's e2e at N GPUs).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/d2h_ceiling.py [MB per rank] [--bind]

Every rank copies one contiguous buffer (default 350 MB, a rank's share of the weak-scaling grids) with a single
cudaMemcpyAsync, all ranks at once between barriers; also the kernel-store path (ogg_bswap64 writing straight into mapped
pinned memory).  Rank 0 prints one JSON line: per-rank and aggregate GB/s."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
mb = float(args[0]) if args else 350.0
bind = "--bind" in sys.argv
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
bound = None
if bind:   # sit on the GPU's NUMA node before the pinned buffers are allocated
    from ocean_model_grid_generator_b200 import device as _D
    bound = _D.bind_host_to_device(local)
n = int(mb * 1e6 / 8)
src = torch.rand(n, dtype=torch.float64, device="cuda")
dst = torch.empty(n, dtype=torch.float64, pin_memory=True)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps=8):
    fn()
    out = []
    for _ in range(reps):
        barrier()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        out.append(time.perf_counter() - t0)
        barrier()
    out.sort()
    return out[len(out) // 2]


t_copy = timed(lambda: dst.copy_(src, non_blocking=True))
from ocean_model_grid_generator_b200 import _cabi, device as D  # noqa: E402
D.ready()
dsti = torch.empty(n, dtype=torch.int64, pin_memory=True)
t_kern = timed(lambda: _cabi.call("ogg_bswap64", D.ptr(src), D.ptr(dsti), n, D.stream()))
res = torch.tensor([8 * n / t_copy / 1e9, 8 * n / t_kern / 1e9], dtype=torch.float64, device="cuda")
allr = [torch.zeros_like(res) for _ in range(world)]
if world > 1:
    dist.all_gather(allr, res)
else:
    allr = [res]
if rank == 0:
    per = [[round(float(v), 2) for v in t] for t in allr]
    print(json.dumps({"ranks": world, "mb_per_rank": mb, "cpus": len(os.sched_getaffinity(0)), "bound_to_gpu_numa_node": bool(bound),
                      "memcpy_gbs_per_rank": [p[0] for p in per], "memcpy_gbs_aggregate": round(sum(p[0] for p in per), 1),
                      "kernel_store_gbs_per_rank": [p[1] for p in per], "kernel_store_gbs_aggregate": round(sum(p[1] for p in per), 1)}))
if world > 1:
    dist.destroy_process_group()
