# parity tests + bench for one experiment library: bash tools/gpu_variant.sh build_ab/x.so
export ADROIT_B200_LIB=$PWD/$1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
unset ADROIT_B200_LIB
bash tools/gpu_ab.sh
