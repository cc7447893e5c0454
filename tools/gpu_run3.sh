python -m pytest tests/test_gpu_fused.py -m gpu -x -q 2>&1 | tail -12
python profiles/kernel_bench.py --iters 10 --only na --match C4 > gpurun_out/kernel_bench_na.jsonl 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench_na.jsonl
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json | cut -c1-300; tail -3 gpurun_out/bench_c4.err
ncu --set full --clock-control none --import-source on -k regex:norm_act_bwd -c 2 -o gpurun_out/ncu_normact_bwd -f python profiles/kernel_bench.py --iters 1 --only na --match "C4 GELU(LN(h1))+up N1536" > gpurun_out/ncu_na.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_na.log
