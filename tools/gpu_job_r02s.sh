#!/bin/bash
# round 2, job s: the state to be judged — whole -m gpu suite, smoke, bench N=1 + reference arm, ncu capture of the store kernel,
# launch list, SA bench, time line
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1800 python -m pytest tests -m gpu -q > $O/r02s_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02s_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02s_smoke.log 2>&1
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02s_bench_reference_arm.json 2> $O/r02s_ref.err
timeout 900 python bench.py > $O/r02s_bench.json 2> $O/r02s_bench.err
timeout 300 python tools/trace_store.py 1 > $O/r02s_trace.log 2>&1
timeout 300 python tools/trace_store.py 8 >> $O/r02s_trace.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02s_launches.csv python bench.py --steps 5 --warmup 3 --no-sa > $O/r02s_ncu_launches.log 2>&1
TWS=96 LLS=0 BLKS=256 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fitch_stream -c 2 -o $O/r02s_store_full -f python tools/time_fullpass.py > $O/r02s_ncu_store.log 2>&1
timeout 300 ncu -i $O/r02s_store_full.ncu-rep --page raw --csv > $O/r02s_ncu_store.csv 2>> $O/r02s_ncu_store.log
rm -f $O/r02s_store_full.ncu-rep
timeout 900 python tools/sa_bench.py > $O/r02s_sa_bench.json 2> $O/r02s_sa_bench.err
tail -n 5 $O/r02s_pytest_gpu.log; cat $O/r02s_smoke.log | tail -2; head -c 700 $O/r02s_bench.json; echo; cat $O/r02s_trace.log | grep "shards\|level"
