#!/bin/bash
# Run tools/pcg_probe.py against every experiment build under build/exp (A/B of PCG kernel variants).
O=gpurun_out/${1:-variants}; mkdir -p $O
NX=${2:-2832}
if [ -z "$SKIP_TESTS" ]; then timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log; fi
for lib in build/exp/lib_*.so; do
  name=$(basename $lib .so)
  MPSB_LIB_PATH=$PWD/$lib timeout 300 python tools/pcg_probe.py $NX > $O/$name.log 2>&1
  echo "== $name"; grep "us/it" $O/$name.log | tail -1
done
tail -3 $O/pytest_gpu.log
