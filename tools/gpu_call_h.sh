#!/bin/bash
# GPU call H (1 GPU): parity suite of the shipped build, then kbench + stepbench per library variant
#   tools/gpu_call_h.sh TAG "variants" [pytest: 1/0]
mkdir -p gpurun_out
T=${1:-r02p}
if [ "${3:-1}" = "1" ]; then
  (timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_supergrid.py -m gpu -q -x 2>&1 | tail -5) | tee gpurun_out/${T}_pytest.log
fi
for rep in 1 2; do
for v in $2; do
  [ -f build/libogg_$v.so ] && OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/kbench.py 4 20 2>&1 | tail -1
  [ -f build/libogg_$v.so ] && OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/stepbench.py 4 100 2>&1 | tail -1 | sed "s/^/$v /"
done
done | tee gpurun_out/${T}_kbench.log
