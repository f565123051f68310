#!/bin/bash
# round 2, job ab: the allocation-free entry point: its test alone in a fresh process, then the whole -m gpu suite (it runs last)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/r02ab_noalloc_alone.log 2>&1; echo "alone rc=$?" >> $O/r02ab_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x > $O/r02ab_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02ab_pytest_gpu.log
tail -n 12 $O/r02ab_noalloc_alone.log; tail -n 5 $O/r02ab_pytest_gpu.log
