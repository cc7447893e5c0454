python -m pytest tests/test_gpu_fused.py -m gpu -x -q 2>&1 | tail -12
python profiles/kernel_bench.py --iters 10 --only na --match C4 > gpurun_out/kernel_bench_na.jsonl 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench_na.jsonl
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json | cut -c1-300; tail -3 gpurun_out/bench_c4.err
python tools/profile_step.py c4 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
