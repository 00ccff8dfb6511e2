python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
bash tools/bench_ngpu.sh 2 2>&1 | grep -v "^ *[0-9]* \[fb shard" | tail -12
