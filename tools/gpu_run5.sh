timeout 120 python tools/tc_fixed_cost.py
