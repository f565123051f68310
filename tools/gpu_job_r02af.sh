#!/bin/bash
# round 2, job af (2 GPUs): the driver's bench commands at N = 1 and N = 2 with the three-run SA sub-record
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python bench.py > $O/r02af_bench_n1.json 2> $O/r02af_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 200 --warmup 10 > $O/r02af_bench_n2.json 2> $O/r02af_bench_n2.err
for f in r02af_bench_n1 r02af_bench_n2; do tail -2 $O/$f.err | cut -c1-200; python - $O/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3))
sa=d['sa_replicates']; print('  sa', round(sa['trees_scored_per_s']), sa['seconds'], [r['search_s_all_runs'] for r in sa['per_rank']])
PY
done
