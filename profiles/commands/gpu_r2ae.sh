#!/bin/bash
# where the per-iteration synchronisation time goes on 2 GPUs: default booking vs PCG_SPLIT_SYNC=1 (exchange after B booked
# under B, publication after C under C, so "sync" is the p.Ap exchange alone), at 2 M and 0.5 M rows per rank
O=gpurun_out/${1:-r2ae}; mkdir -p $O
for NX in 2832 1416; do for lib in mps_b200/libmps_b200.so build/exp/lib_split.so; do
  tag=$(basename $lib .so)_nx$NX
  MPSB_LIB_PATH=$PWD/$lib timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29620 \
      tools/pcg_probe_multi.py $NX > $O/probe_$tag.log 2>&1
  echo "== $tag"; grep "us/it" $O/probe_$tag.log
done; done
