#!/bin/bash
# round 2, job u: two-column score kernel A/B, host profile of the 512-candidate score call
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -x > $O/r02u_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02u_pytest_gpu.log
for v in 1 0; do echo "== LVB_GPU_SCORE2=$v" >> $O/r02u_score.log; LVB_GPU_SCORE2=$v LVB_GPU_PROFILE=1 timeout 300 python tools/profile_score.py >> $O/r02u_score.log 2>&1; done
LVB_GPU_SCORE2=1 timeout 600 python bench.py --no-sa --steps 50 > $O/r02u_bench.json 2> $O/r02u_bench.err
tail -n 3 $O/r02u_pytest_gpu.log; cat $O/r02u_score.log; python - $O/r02u_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(json.dumps(d.get('sa_trees_scored_per_s'))[:300]); print(d['roofline_score']['kernel_ms_per_batch'], d['nni_neighbourhood']['agrees_with_score_batch'])
PY
