#!/bin/bash
# GPU box with N GPUs: trail-block exchange A/B (peer stores vs packed ncclSend/ncclRecv): bash tools/shard_ab.sh tag N bins
tag=$1; N=$2; bins=${3:-6000}
mkdir -p gpurun_out
for p2p in 1 0; do
  FB_SHARD_P2P=$p2p FB_SHARD_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
     --master-port $((29700 + p2p)) tools/shard_check.py --big $bins > gpurun_out/${tag}_p2p${p2p}.log 2>&1
  echo "== p2p=$p2p rc=$?"; grep -E "SHARD CHECK|exchange of|per ADMM|big|rank 0 n=$bins" gpurun_out/${tag}_p2p${p2p}.log | tail -8
done
