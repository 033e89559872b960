#!/bin/bash
mkdir -p gpurun_out
for feats in none mo buffer geo; do
  echo "=== features $feats"
  SLAB_FEATURES=$feats SLAB_NSTEPS=1 SLAB_DIMS=64,64,32 MHH_SLAB_PATH=nccl timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/slab_parity.py 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|Starting\|Finished\|Loading" | tail -14
done
