for lib in mj_envs_b200/libadroit_b200.so build_ab/*.so; do echo "== $lib"; ADROIT_B200_LIB=$PWD/$lib timeout 600 python tools/gpu_fp32_outliers.py 2>&1 | tail -4; done
