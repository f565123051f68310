#!/bin/bash
# round 2, job h (2 GPUs): the driver's own N=2 command (site blocks + all-reduce headline, replicas and SA sub-records),
# the -m gpu suite with two devices visible (2-rank NCCL / peer-memory tests), N=1 bench for the same box
R=${GRAFT_REPO_ROOT:-/root/repo}; O=$R/gpurun_out; mkdir -p $O; cd $R
nvidia-smi -L > $O/r02h_env.txt; nproc >> $O/r02h_env.txt
timeout 1500 python -m pytest tests -m gpu -q > $O/r02h_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r02h_pytest_gpu.log
timeout 900 python bench.py --steps 100 --warmup 5 > $O/r02h_bench_n1.json 2> $O/r02h_bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 5 > $O/r02h_bench_n2.json 2> $O/r02h_bench_n2.err
LVB_GPU_P2P=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 5 --no-extras > $O/r02h_bench_n2_nccl.json 2> $O/r02h_bench_n2_nccl.err
tail -n 6 $O/r02h_pytest_gpu.log; tail -n 4 $O/r02h_bench_n2.err; head -c 1500 $O/r02h_bench_n2.json; echo; head -c 600 $O/r02h_bench_n2_nccl.json
