#!/bin/bash
# round 2, first GPU job: the whole -m gpu suite (new: drop-in binary, config-3/5 parity, NCCL one-rank
# communicator, bounded wait), the store-kernel sweep, bench N=1 and the reference arm at full size.
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
{ nvidia-smi -L; nproc; free -g | head -2; } > $O/r02a_env.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > $O/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02a_pytest_gpu.log
timeout 700 python tools/sweep_store.py > $O/r02a_sweep.jsonl 2> $O/r02a_sweep.err
timeout 900 python bench.py --steps 50 --warmup 5 > $O/r02a_bench.json 2> $O/r02a_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02a_bench_reference_arm.json 2> $O/r02a_ref.err
tail -n 15 $O/r02a_pytest_gpu.log; tail -n 3 $O/r02a_sweep.err $O/r02a_bench.err $O/r02a_ref.err; head -c 600 $O/r02a_bench.json
