# parity tests on the product library, then A/B bench against everything in build_ab/
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
bash tools/gpu_ab.sh
