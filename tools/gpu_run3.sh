python -m pytest tests/test_gpu_fused.py -m gpu -x -q 2>&1 | tail -2
python profiles/kernel_bench.py --iters 10 --only na --match C4 2>&1 | cut -c1-170
