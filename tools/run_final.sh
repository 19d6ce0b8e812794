#!/bin/bash
# Final bench lines of a round at N clients with the final code: LeNet (headline), ResNet-18 (config 3), MobileNet (the reference's
# default model) next to the N-nccl baseline for MobileNet, and the per-phase round profile.  usage: tools/run_final.sh N
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out/final$N; mkdir -p $O
for M in LeNet ResNet18 MobileNet; do
  echo "== fedb200 $M N=$N"; timeout 500 $TR --master-port 29801 bench.py --gpus $N --steps 20 --warmup 3 --model $M 2> $O/bench_$M.err | grep '^{' > $O/bench_${M}_N$N.json; cut -c1-330 $O/bench_${M}_N$N.json
done
echo "== N-nccl MobileNet N=$N"; timeout 500 $TR --master-port 29802 baseline/nccl_fedavg.py --gpus $N --steps 10 --warmup 3 --model MobileNet 2> $O/nccl.err | grep '^{' > $O/nccl_fedavg_mobilenet_N$N.json; cut -c1-300 $O/nccl_fedavg_mobilenet_N$N.json
echo "== round profile ResNet18"; timeout 400 $TR --master-port 29803 tools/prof_round.py --model ResNet18 --out $O/round_profile_resnet18_N$N.json > $O/prof_round.log 2>&1; tail -1 $O/prof_round.log | cut -c1-500
