#!/bin/bash
# round 2, second GPU job: forest-bin piece cut + level counters (parity suite), store-kernel sweep around the new
# defaults, bench N=1, ncu full capture of the store kernel, config-5 per-rank SA probes (8 / 16 / 64 chains).
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1500 python -m pytest tests -m gpu -q -x > $O/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02b_pytest_gpu.log
timeout 600 python tools/sweep_store.py --pf 3,5,7 --warps 96,128,192,256,384 --block 256,128 --shards 1,2,8 > $O/r02b_sweep.jsonl 2> $O/r02b_sweep.err
timeout 900 python bench.py --steps 100 --warmup 5 > $O/r02b_bench.json 2> $O/r02b_bench.err
# SA: what one rank of an 8-GPU run does (8 chains, 2 host threads per... all threads here), and the 1-GPU job
for spec in "8 8" "16 16" "64 16"; do set -- $spec
  echo "== chains $1 threads $2" >> $O/r02b_sa_probe.log
  LVB_GPU_PROFILE=2 OMP_NUM_THREADS=$2 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 >> $O/r02b_sa_probe.log 2>&1
done
TWS=192 LLS=0 BLKS=256 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fitch_store_pipe -c 3 -o $O/r02b_store_full -f python tools/time_fullpass.py > $O/r02b_ncu.log 2>&1
timeout 300 ncu -i $O/r02b_store_full.ncu-rep --page raw --csv > $O/r02b_ncu_full.csv 2>> $O/r02b_ncu.log
tail -n 8 $O/r02b_pytest_gpu.log; tail -n 3 $O/r02b_sweep.err $O/r02b_bench.err; head -c 400 $O/r02b_bench.json; tail -5 $O/r02b_ncu.log
