#!/bin/bash
# N = 2, 4, 8 of the weak and the 1/120-degree strong sweep + the NCCL test (run on an 8-GPU box; N = 1 runs on a 1-GPU box)
mkdir -p gpurun_out
run() { N=$1; shift; TAG=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N "$@" > gpurun_out/sweep_${TAG}_n$N.json 2> gpurun_out/sweep_${TAG}_n$N.err
  python -c "
import json
try:
    d = json.load(open('gpurun_out/sweep_${TAG}_n$N.json'))
    print('${TAG} N=$N', d['config']['workload'][:24], 'ms/step %.3f' % d['ms_per_step'], 'value %.4g' % d['value'], 'frac %.3f' % d['roofline']['frac'], 'e2e', (d.get('e2e') or {}).get('value'))
except Exception as e:
    print('${TAG} N=$N FAILED', e); print(open('gpurun_out/sweep_${TAG}_n$N.err').read()[-1500:])
"
}
for N in 2 4 8; do run $N weak --steps 20 --warmup 3 --no_cpu_baseline; done
for N in 8 4 2; do run $N r120 --resolution 120 --slabs 8 --steps 3 --warmup 3 --no_cpu_baseline; done
timeout 300 python -m pytest tests/test_gpu_multi_nccl.py -q 2>&1 | tail -2
