timeout 60 python tools/tc_timeline.py 1024 > gpurun_out/tl_fwd_1024.txt 2>&1; echo rc=$?
timeout 60 python tools/tc_timeline.py --bwd 1024 > gpurun_out/tl_bwd_1024.txt 2>&1; echo rc=$?
timeout 60 python tools/tc_timeline.py 256 > gpurun_out/tl_fwd_256.txt 2>&1; echo rc=$?
