# exact A/B of the product library against build_ab/base.so (fingerprints must match), then a short bench of both
ADROIT_B200_LIB=$PWD/build_ab/base.so timeout 600 python tools/gpu_fingerprint.py 60 > gpurun_out/fp_base.txt 2>&1
timeout 600 python tools/gpu_fingerprint.py 60 > gpurun_out/fp_new.txt 2>&1
diff gpurun_out/fp_base.txt gpurun_out/fp_new.txt && echo FINGERPRINTS IDENTICAL
cat gpurun_out/fp_new.txt | tail -9
bash tools/gpu_ab.sh
