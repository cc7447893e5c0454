timeout 120 python tools/check_tc.py --time > gpurun_out/check_tc_fwd.txt 2>&1; echo "fwd rc=$?"; tail -12 gpurun_out/check_tc_fwd.txt
timeout 120 python tools/check_tc.py --bwd --time > gpurun_out/check_tc_bwd.txt 2>&1; echo "bwd rc=$?"; tail -12 gpurun_out/check_tc_bwd.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
