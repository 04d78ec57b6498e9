python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python profiles/profile_population.py --individuals 4
python bench.py --steps 1 --warmup 3 | cut -c1-1800
