#!/bin/bash
# L2 residency experiment, second pass: persisting carve-out and stream policy, per-rank size of an 8-GPU run
O=gpurun_out/${1:-r2i}; mkdir -p $O
python - <<'PY' > $O/l2_limits.txt 2>&1
import torch
p = torch.cuda.get_device_properties(0)
print("L2", p.L2_cache_size)
import ctypes
rt = ctypes.CDLL("libcudart.so.12")
v = ctypes.c_int()
for name, attr in (("maxPersistingL2", 108), ("maxAccessPolicyWindow", 109), ("l2CacheSize", 38)):
    rt.cudaDeviceGetAttribute(ctypes.byref(v), attr, 0); print(name, v.value)
PY
cat $O/l2_limits.txt
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 300 python tools/pcg_probe.py ${NX:-1000} > $O/probe_$tag.log 2>&1
  echo "== $tag: $(grep 'us/it' $O/probe_$tag.log | tail -1 | cut -c1-220)"
}
run off X=1
for mb in 32 48 64; do run keep${mb} MPSB_L2_RESIDENT_MB=$mb; done
for mb in 48 80; do run keep${mb}_normal MPSB_L2_RESIDENT_MB=$mb MPSB_L2_STREAM_NORMAL=1; done
for mb in 48 80 104; do run keep${mb}_persist96 MPSB_L2_RESIDENT_MB=$mb MPSB_L2_PERSIST_MB=96; done
for mb in 48 80; do run keep${mb}_persist64 MPSB_L2_RESIDENT_MB=$mb MPSB_L2_PERSIST_MB=64; done
NX=708 run nx708_off X=1
NX=708 run nx708_keep48 MPSB_L2_RESIDENT_MB=48
NX=708 run nx708_keep80 MPSB_L2_RESIDENT_MB=80
NX=708 run nx708_keep80_persist96 MPSB_L2_RESIDENT_MB=80 MPSB_L2_PERSIST_MB=96
