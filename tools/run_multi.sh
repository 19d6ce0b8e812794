#!/bin/bash
# The multi-GPU evidence batch of a round on ONE box: control-plane failure phases, peer-plane checks, headline bench in both
# evaluation modes, ResNet-18 (config 3) with and without the fp8 codec, aggregation sweep, accuracy gate, N-nccl baselines,
# per-phase round profile.  EXTRA_EVAL=overlap adds the overlapped-evaluation bench line.  Logs and JSON go to gpurun_out/n<N>/.   usage: tools/run_multi.sh N [quick]
N=$1; Q=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out/n$N; mkdir -p $O
if [ "$Q" = "quick" ]; then ST="--steps 5 --warmup 3"; GATE="--rounds 6 --n-train 5120 --n-test 1000"; AGG="--sizes-mb 1 13 45"; PR="--n-train 5120 --n-test 1000"; else ST="--steps 20 --warmup 3"; GATE="--rounds 20 --n-train 25600 --n-test 5000"; AGG="--sizes-mb 1 13 45 256 359 1024"; PR=""; fi
echo "== grpc_gpu_check"; FEDB200_CHECK_WORLD=$N timeout 500 python tools/grpc_gpu_check.py > $O/grpc.log 2>&1; tail -12 $O/grpc.log | cut -c1-300
echo "== bench LeNet"; for m in inline $EXTRA_EVAL; do timeout 600 $TR --master-port 29602 bench.py --gpus $N $ST --eval $m > $O/bench_lenet_$m.json 2> $O/bench_lenet_$m.err; cut -c1-400 $O/bench_lenet_$m.json; done
echo "== bench ResNet18"; timeout 500 $TR --master-port 29603 bench.py --gpus $N $ST --model ResNet18 > $O/bench_resnet18_none.json 2> $O/bench_resnet18.err; cut -c1-400 $O/bench_resnet18_none.json
echo "== agg sweep"; timeout 500 $TR --master-port 29604 tools/bench_agg.py $AGG > $O/agg.log 2>&1; cp gpurun_out/agg_sweep_N$N.json $O/ 2>/dev/null; grep '"mb"' $O/agg.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print({k:r.get(k) for k in ('mb','p2p_ms','mcast_ms','nvls_ms','nccl_ms','nccl_avg_ms','p2p_frac_roofline','mcast_frac_roofline','deterministic_vs_nccl','best_vs_nccl','measured_copy_gbs')})"
echo "== codec gate"; timeout 500 $TR --master-port 29605 tools/codec_gate.py --models LeNet ResNet18 $GATE --out $O/codec_gate.json > $O/gate.log 2>&1; grep "PASS\|FAIL" $O/gate.log
if grep -q "\[ResNet18\].*PASS" $O/gate.log; then echo "== fp8 bench ResNet18 (licensed by the gate)"; timeout 600 $TR --master-port 29606 bench.py --gpus $N $ST --model ResNet18 --codec fp8 > $O/bench_resnet18_fp8.json 2>> $O/bench_resnet18.err; cut -c1-300 $O/bench_resnet18_fp8.json; else echo "== fp8 bench skipped: the accuracy gate did not pass"; fi
echo "== nccl baselines"; for M in ${NCCL_MODELS:-ResNet18}; do timeout 500 $TR --master-port 29607 baseline/nccl_fedavg.py --gpus $N $ST --model $M > $O/nccl_$M.json 2> $O/nccl_$M.err; cut -c1-300 $O/nccl_$M.json; done
echo "== round profile"; timeout 600 $TR --master-port 29608 tools/prof_round.py --model ResNet18 $PR --out $O/round_profile_resnet18.json > $O/prof_round.log 2>&1; tail -1 $O/prof_round.log | cut -c1-600
echo "== multigpu_check"; timeout 600 $TR --master-port 29601 tools/multigpu_check.py > $O/mg.log 2>&1; grep MULTIGPU_CHECK_OK $O/mg.log | cut -c1-400 || tail -5 $O/mg.log
