#!/bin/bash
# round 2, job k: software-pipelined step of the cp.async kernels: parity suite, SA probes, bench
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1500 python -m pytest tests -m gpu -q > $O/r02k_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02k_pytest_gpu.log
nproc > $O/r02k_sa_probe.log
for spec in "8 8 8" "16 16 16" "64 16 4" "64 16 8"; do set -- $spec
  echo "== chains $1 threads $2 groups $3" >> $O/r02k_sa_probe.log
  LVB_GPU_PROFILE=1 timeout 300 tools/sa_probe.sh 1000 200000 0.02 $1 3000 -a 1 -p $2 -G $3 2>&1 | grep "Wall\|All rep\|Host time\|device_open" >> $O/r02k_sa_probe.log
done
timeout 900 python bench.py --steps 200 --warmup 10 > $O/r02k_bench.json 2> $O/r02k_bench.err
timeout 600 python tools/sa_bench.py > $O/r02k_sa_bench.json 2> $O/r02k_sa_bench.err
tail -n 4 $O/r02k_pytest_gpu.log; cat $O/r02k_sa_probe.log; python - $O/r02k_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('ms/step', d['ms_per_step'], 'kernel', d['per_rank'], 'e2e', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'])
for k in ('sa_trees_scored_per_s','roofline_score','nni_neighbourhood','sa_replicates'):
    print(k, json.dumps(d.get(k))[:500])
PY
tail -c 1500 $O/r02k_sa_bench.json
