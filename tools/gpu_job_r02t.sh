#!/bin/bash
# round 2, job t: chunk of 96 instructions (4 CTAs per SM at 4 stack slots), NNI kernel with adds on the FMA pipe, bench
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_hillclimb.py -m gpu -q -x > $O/r02t_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02t_pytest_gpu.log
timeout 300 python tools/trace_store.py 1 > $O/r02t_trace.log 2>&1
timeout 600 python tools/sweep_store.py --stream 256 --taper 1 --lag=-1 --warps 64,80,96,112,128,160 --shards 1 --reps 60 > $O/r02t_sweep.jsonl 2> $O/r02t_sweep.err
timeout 900 python bench.py --no-sa > $O/r02t_bench.json 2> $O/r02t_bench.err
tail -n 3 $O/r02t_pytest_gpu.log; grep "shards\|level\|flight" $O/r02t_trace.log | cut -c1-250; grep kernel_ms $O/r02t_sweep.jsonl | cut -c1-200; python - $O/r02t_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('ms/step', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_pass'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'])
for k in ('sa_trees_scored_per_s','nni_neighbourhood'):
    print(k, json.dumps(d.get(k))[:400])
PY
