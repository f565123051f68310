#!/bin/bash
# round 2, job x: the state to be judged, again (after the two-column score kernel, overlays in the bench, compile threshold)
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1800 python -m pytest tests -m gpu -q > $O/r02x_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02x_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02x_smoke.log 2>&1
timeout 900 python bench.py > $O/r02x_bench.json 2> $O/r02x_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02x_bench_reference_arm.json 2> $O/r02x_ref.err
tail -n 4 $O/r02x_pytest_gpu.log; tail -1 $O/r02x_smoke.log; python - $O/r02x_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('ms/step', d['ms_per_step'], 'kernel', d['roofline']['kernel_ms_per_pass'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'], d['clocks'])
for k in ('sa_trees_scored_per_s','nni_neighbourhood'):
    print(k, json.dumps(d.get(k))[:330])
print('sa', {k:v for k,v in d['sa_replicates'].items() if k in ('trees_scored_per_s','seconds','wall_seconds')})
PY
