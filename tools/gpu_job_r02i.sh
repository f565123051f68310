#!/bin/bash
# round 2, job i (2 GPUs): fused peer-memory exchange in the store kernel's finalize vs the separate all-reduce kernel vs NCCL;
# 2-rank tests; SA sub-record with the device-side DNAToBinary
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -m gpu -q > $O/r02i_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02i_pytest_gpu.log
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 200 --warmup 10 "${@:3}" > $O/$2.json 2> $O/$2.err; }
run 29521 r02i_bench_n2
LVB_GPU_P2P_FUSE=0 run 29522 r02i_bench_n2_unfused --no-extras
LVB_GPU_P2P=0 run 29523 r02i_bench_n2_nccl --no-extras
timeout 900 python bench.py --steps 200 --warmup 10 > $O/r02i_bench_n1.json 2> $O/r02i_bench_n1.err
tail -n 4 $O/r02i_pytest_gpu.log; tail -n 3 $O/r02i_bench_n2.err
for f in r02i_bench_n1 r02i_bench_n2 r02i_bench_n2_unfused r02i_bench_n2_nccl; do python - $O/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],4), 'kernel', [round(r['kernel_ms'],4) for r in d['per_rank']], 'launches', d['gpu_launches'], 'e2e', round(d['e2e']['ms_per_step'],4), 'frac', round(d['roofline']['frac'],3), 'sa', {k:(round(v,3) if isinstance(v,float) else v) for k,v in (d.get('sa_replicates') or {}).items() if k in ('trees_scored_per_s','seconds','wall_seconds')})
PY
done
