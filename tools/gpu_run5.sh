timeout 60 python tools/tc_timeline.py 1024 2>&1 | grep -E "setup|CTA"
timeout 60 python tools/tc_timeline.py --bwd 1024 2>&1 | grep -E "setup|CTA"
timeout 60 python tools/tc_timeline.py --bwd 256 2>&1 | grep -E "setup|CTA"
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks.mem,power.draw --format=csv
