#!/bin/bash
# The baselines table of BASELINE.md section 3b at N clients on one box: fedb200 (LeNet), N-nccl (LeNet), R-gpu (unmodified reference
# with one GPU per worker).  usage: tools/run_baselines.sh N        -> gpurun_out/base<N>/
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out/base$N; mkdir -p $O
echo "== fedb200 LeNet N=$N"; timeout 400 $TR --master-port 29701 bench.py --gpus $N --steps 20 --warmup 3 2> $O/bench.err | grep '^{' > $O/bench_N$N.json; cut -c1-330 $O/bench_N$N.json
echo "== N-nccl LeNet N=$N"; timeout 400 $TR --master-port 29702 baseline/nccl_fedavg.py --gpus $N --steps 10 --warmup 3 --model LeNet 2> $O/nccl.err | grep '^{' > $O/nccl_fedavg_N$N.json; cut -c1-300 $O/nccl_fedavg_N$N.json
echo "== R-gpu N=$N"; timeout 500 python bench.py --impl reference --ref-gpu-workers --gpus $N --steps 2 --warmup 1 2> $O/refgpu.err | grep '^{' > $O/ref_gpu_N$N.json; cut -c1-300 $O/ref_gpu_N$N.json
