python -m pytest tests/test_gpu_fused.py -m gpu -x -q 2>&1 | tail -3
python profiles/kernel_bench.py --iters 10 --only na --match C4 > gpurun_out/kernel_bench_na.jsonl 2>&1; echo "kbench rc=$?"; cut -c1-200 gpurun_out/kernel_bench_na.jsonl
ncu --set full --clock-control none --import-source on -k regex:norm_act -c 4 -o gpurun_out/ncu_normact -f python profiles/kernel_bench.py --iters 1 --only na --match "C4 GELU(LN(h1))+up N1536" > gpurun_out/ncu_na.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_na.log
