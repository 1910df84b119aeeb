#!/bin/bash
# On the GPU box: the allocation-free entry point's test alone in a fresh process, then the whole gpu suite (the file is
# collected last).  Logs under gpurun_out/noalloc_*.  (compute-sanitizer is closed on this pool: no memcheck leg.)
mkdir -p gpurun_out
O=gpurun_out/noalloc
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > ${O}_alone.log 2>&1; echo "alone rc=$?" > ${O}_rc.txt
timeout 900 python -m pytest tests -x -q -m gpu > ${O}_all.log 2>&1; echo "all rc=$?" >> ${O}_rc.txt
cat ${O}_rc.txt
tail -n 3 ${O}_alone.log
tail -n 3 ${O}_all.log
