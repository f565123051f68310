#!/bin/bash
# round 2, third GPU job: level counters fixed, TMA streaming store kernel: parity suite, sweep, bench, ncu, SA probes
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1500 python -m pytest tests -m gpu -q > $O/r02c_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02c_pytest_gpu.log
timeout 600 python tools/sweep_store.py --stream 64,128,256,0 --pf 5 --warps 24,32,48,64,96,128,192,384 --block 128 --shards 1,2,8 > $O/r02c_sweep.jsonl 2> $O/r02c_sweep.err
timeout 900 python bench.py --steps 100 --warmup 5 > $O/r02c_bench.json 2> $O/r02c_bench.err
for spec in "8 8 1" "8 8 2" "8 8 4" "16 16 2" "64 16 4"; do set -- $spec
  echo "== chains $1 threads $2 groups $3" >> $O/r02c_sa_probe.log
  LVB_GPU_PROFILE=2 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep -v "Replicate" >> $O/r02c_sa_probe.log
done
TWS=192 LLS=0 BLKS=256 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fitch_stream -c 3 -o $O/r02c_store_full -f python tools/time_fullpass.py > $O/r02c_ncu.log 2>&1
timeout 300 ncu -i $O/r02c_store_full.ncu-rep --page raw --csv > $O/r02c_ncu_full.csv 2>> $O/r02c_ncu.log
tail -n 8 $O/r02c_pytest_gpu.log; tail -n 3 $O/r02c_sweep.err $O/r02c_bench.err; head -c 400 $O/r02c_bench.json; tail -5 $O/r02c_ncu.log
