# parity tests on the product library, then A/B bench of every variant in build_ab/
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
bash tools/gpu_ab.sh
