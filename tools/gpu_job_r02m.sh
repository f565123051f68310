#!/bin/bash
# round 2, job m: score-and-keep (spare row bank per tree): parity suite, SA probes keep on/off, bench
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1800 python -m pytest tests -m gpu -q -x > $O/r02m_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02m_pytest_gpu.log
nproc > $O/r02m_sa_probe.log
for keep in 1 0; do
for spec in "1 1 1" "8 8 8" "16 16 16" "64 16 8"; do set -- $spec
  echo "== keep $keep chains $1 threads $2 groups $3" >> $O/r02m_sa_probe.log
  LVB_GPU_KEEP=$keep LVB_GPU_PROFILE=1 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep "Wall\|All rep\|Host time\|ERROR" >> $O/r02m_sa_probe.log
done; done
timeout 900 python bench.py --steps 200 --warmup 10 > $O/r02m_bench.json 2> $O/r02m_bench.err
tail -n 12 $O/r02m_pytest_gpu.log; cat $O/r02m_sa_probe.log; tail -3 $O/r02m_bench.err; python - $O/r02m_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('ms/step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'])
for k in ('sa_trees_scored_per_s','nni_neighbourhood','sa_replicates'):
    print(k, json.dumps(d.get(k))[:400])
PY
