# quick GPU check: parity tests, smoke, short bench of door and hammer
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python bench.py --steps 100 --warmup 10 --no-cpu-baseline --also hammer-v0:8192 2>&1 | tail -2 | cut -c1-1500
