#!/bin/bash
# round 2, job ae: final state on one GPU — whole -m gpu suite (keep/accept at config 3/5 shapes, no-alloc entry last), smoke,
# reference arm, bench
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 1800 python -m pytest tests -m gpu -x -q > $O/r02ae_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02ae_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02ae_smoke.log 2>&1
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $O/r02ae_bench_reference_arm.json 2> $O/r02ae_ref.err
timeout 900 python bench.py > $O/r02ae_bench.json 2> $O/r02ae_bench.err
LVB_GPU_PROFILE=2 timeout 300 python tools/profile_config1.py > $O/r02ae_config1_profile.log 2>&1
timeout 300 python tools/profile_config1.py >> $O/r02ae_config1_profile.log 2>&1
grep "Wall\|Host time\|kernel:" $O/r02ae_config1_profile.log
tail -n 4 $O/r02ae_pytest_gpu.log; tail -1 $O/r02ae_smoke.log; python - $O/r02ae_bench.json $O/r02ae_bench_reference_arm.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
print('ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'], d['roofline']['frac_on_fused_floor'], 'e2e', d['e2e']['ms_per_step'], d['clocks'])
print('ref', r['value'], 'ratio e2e', d['e2e']['value']/r['value'], 'device', d['value']/r['value'], d.get('length_match'))
print(json.dumps(d['sa_trees_scored_per_s'])[:260]); print({k:v for k,v in d['sa_replicates'].items() if k in ('trees_scored_per_s','seconds')})
PY
