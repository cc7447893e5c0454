timeout 120 python tools/check_tc.py --bwd --time > gpurun_out/check_tc_bwd.txt 2>&1; echo "bwd rc=$?"; tail -9 gpurun_out/check_tc_bwd.txt
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python profiles/kernel_bench.py --iters 10 --only na --match "N1536" 2>&1 | cut -c1-200
python profiles/kernel_bench.py --iters 10 --only dwt 2>&1 | cut -c1-220
timeout 60 python tools/tc_timeline.py 1024 2>&1 | head -4
timeout 60 python tools/tc_timeline.py --bwd 1024 2>&1 | grep -E "setup|aux5|epiWG0  tile[01]"
