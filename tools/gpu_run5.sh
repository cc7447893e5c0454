SB200_EXTRA_NVCC_FLAGS="-DSB200_WALLCLOCK=1" python -m vit_spectre_experiments_b200.build --force > /dev/null 2>&1; echo "build rc=$?"
timeout 60 python tools/tc_timeline.py 1024 2>&1 | grep -E "setup|CTA"
timeout 60 python tools/tc_timeline.py --bwd 1024 2>&1 | grep -E "setup|CTA"
