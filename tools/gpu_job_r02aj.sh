#!/bin/bash
# round 2, job aj (8 GPUs, final code of the round: three-run SA sub-record, 256 consumer threads at a quarter of the width): the driver's own N=8 command (site blocks + fused peer-memory exchange; replicas and SA sub-records),
# the same with NCCL for the all-reduce, N=4
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
nvidia-smi -L > $O/r02aj_env.txt; nproc >> $O/r02aj_env.txt; free -g | head -2 >> $O/r02aj_env.txt
run() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $n --steps 200 --warmup 10 "${@:3}" > $O/$2.json 2> $O/$2.err; }
run 8 29531 r02aj_bench_n8
run 4 29533 r02aj_bench_n4
run 2 29534 r02aj_bench_n2
timeout 600 python bench.py --steps 200 --warmup 10 > $O/r02aj_bench_n1.json 2> $O/r02aj_bench_n1.err
for f in r02aj_bench_n1 r02aj_bench_n2 r02aj_bench_n4 r02aj_bench_n8; do tail -2 $O/$f.err; python - $O/$f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],4), 'kernel', [round(r['kernel_ms'],4) for r in d['per_rank']], 'launches', d['gpu_launches'], 'e2e', round(d['e2e']['ms_per_step'],4))
    for k in ('replicas','sa_replicates'):
        v=d.get(k) or {}
        print(' ',k,{a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','trees_scored_per_s','rearrangements_per_s','seconds','wall_seconds','failed','refused')})
except Exception as e:
    print('no line', e)
PY
done
