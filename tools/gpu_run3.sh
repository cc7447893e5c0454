python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 10 --warmup 3 --config c1 --no-cpu-baseline 2>&1 | cut -c1-200
