#!/bin/bash
# GPU call C (1 GPU): A/B timing of library variants (build/libogg_*.so): tools/gpu_call_c.sh TAG "kbench variants" "bandbench variants"
mkdir -p gpurun_out
T=${1:-r02d}
for v in $2; do
  [ -f build/libogg_$v.so ] && OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/kbench.py 4 20 2>&1 | tail -1
done | tee gpurun_out/${T}_kbench.log
for v in $3; do
  for r in 4 12; do OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/bandbench.py $r 20 2>&1 | tail -1; done
done | tee gpurun_out/${T}_bandbench.log
(timeout 900 python -m pytest tests/test_gpu_supergrid.py tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -3) | tee gpurun_out/${T}_pytest.log
timeout 300 python tools/stepbench.py 4 100 2>&1 | tail -2 | tee gpurun_out/${T}_stepbench.log
