python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
python profiles/kernel_bench.py --iters 10 --only na > gpurun_out/kernel_bench_na.jsonl 2>&1; echo "kbench rc=$?"; cat gpurun_out/kernel_bench_na.jsonl
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"; cat gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
python bench.py --steps 10 --warmup 3 --batch 1024 --no-cpu-baseline > gpurun_out/bench_c4_b1024.json 2> gpurun_out/bench_c4_b1024.err; echo "bench1024 rc=$?"; cat gpurun_out/bench_c4_b1024.json; tail -3 gpurun_out/bench_c4_b1024.err
python tools/profile_step.py c4 > gpurun_out/step_profile_c4.txt 2>&1; echo "profile rc=$?"
