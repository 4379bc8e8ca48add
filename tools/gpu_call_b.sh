#!/bin/bash
# GPU call B (1 GPU): parity of the rewritten kernels, then A/B timing of library variants (build/libogg_*.so)
mkdir -p gpurun_out
T=${1:-r02b}; shift
VARIANTS=${@:-r01 new}
(timeout 1500 python -m pytest tests -m gpu -q -rs 2>&1 | tail -30) > gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
for v in $VARIANTS; do
  [ -f build/libogg_$v.so ] && OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/kbench.py 4 20 2>&1 | tail -1
done | tee gpurun_out/${T}_kbench.log
for v in $VARIANTS; do
  for r in 4 12; do OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/bandbench.py $r 20 2>&1 | tail -1; done
done | tee gpurun_out/${T}_bandbench.log
timeout 300 python tools/stepbench.py 4 100 2>&1 | tail -4 | tee gpurun_out/${T}_stepbench.log
