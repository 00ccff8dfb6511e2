FB_SHARD_VERBOSE=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or two_gpus" 2>&1 | tail -6
for p2p in 1 0; do
FB_SHARD_P2P=$p2p timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2971$p2p tools/shard_check.py --big 6000 --iters 82 2>&1 | grep -E "BIG|SHARD CHECK"
done
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "matches_oracle or golden or c2" 2>&1 | tail -3
bash tools/variants_bench.sh r2r product
