#!/bin/bash
# GPU call C (1 GPU): A/B timing of library variants (build/libogg_*.so) and a sweep of the band fill's rows per CTA
mkdir -p gpurun_out
T=${1:-r02d}; shift
VARIANTS=${@:-r01 new}
for v in $VARIANTS; do
  [ -f build/libogg_$v.so ] && OGG_B200_LIB=$PWD/build/libogg_$v.so timeout 300 python tools/kbench.py 4 20 2>&1 | tail -1
done | tee gpurun_out/${T}_kbench.log
for r in 4 12; do
  OGG_B200_LIB=$PWD/build/libogg_r01.so timeout 300 python tools/bandbench.py $r 20 2>&1 | tail -1
  timeout 300 python tools/bandbench.py $r 20 2>&1 | tail -1
  for rows in 4 5 6 8 10 12 15 16 20 24 30; do echo -n "rows=$rows "; OGG_BAND_ROWS=$rows timeout 300 python tools/bandbench.py $r 20 2>&1 | tail -1; done
done | tee gpurun_out/${T}_bandbench.log
(timeout 900 python -m pytest tests/test_gpu_supergrid.py -m gpu -q -x 2>&1 | tail -3) | tee gpurun_out/${T}_pytest.log
