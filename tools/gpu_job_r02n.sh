#!/bin/bash
# round 2, job n: evidence — ncu full captures (store, score, small-batch, NNI kernels), launch list of bench.py, SA probe
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "keep or routes" > $O/r02n_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02n_pytest_gpu.log
: > $O/r02n_sa_probe.log
for spec in "1 1 1" "8 8 8" "64 16 8"; do set -- $spec
  echo "== chains $1 threads $2 groups $3" >> $O/r02n_sa_probe.log
  LVB_GPU_PROFILE=1 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep "Wall\|All rep\|Host time\|ERROR" >> $O/r02n_sa_probe.log
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-sa > $O/r02n_bench_short.json 2> $O/r02n_bench_short.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02n_launches.csv python bench.py --steps 5 --warmup 3 --no-sa > $O/r02n_ncu_launches.log 2>&1
TWS=128 LLS=0 BLKS=256 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fitch_stream -c 2 -o $O/r02n_store_full -f python tools/time_fullpass.py > $O/r02n_ncu_store.log 2>&1
timeout 300 ncu -i $O/r02n_store_full.ncu-rep --page raw --csv > $O/r02n_ncu_store.csv 2>> $O/r02n_ncu_store.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fitch_score_kernel|fitch_pipe_kernel|nni_neighbourhood_kernel" -c 8 -o $O/r02n_score_full -f python tools/profile_score.py > $O/r02n_ncu_score.log 2>&1
timeout 300 ncu -i $O/r02n_score_full.ncu-rep --page raw --csv > $O/r02n_ncu_score.csv 2>> $O/r02n_ncu_score.log
rm -f $O/r02n_store_full.ncu-rep $O/r02n_score_full.ncu-rep
tail -n 3 $O/r02n_pytest_gpu.log; cat $O/r02n_sa_probe.log; tail -3 $O/r02n_ncu_store.log $O/r02n_ncu_score.log; wc -l $O/r02n_launches.csv
