#!/bin/bash
# Multi-GPU evidence run (BASELINE.json configs #2-#5) on N GPUs of one box; every step under its own timeout.
#   gpurun --gpus N -- 'bash scripts/multi_gpu_suite.sh N [steps...]'      steps: peer sweep ours ddp vgg16 ddpvgg dense incep pipe main
N=${1:-2}; shift
STEPS=${@:-"peer sweep ours ddp vgg16 dense incep pipe"}
OUT=gpurun_out/suite_${N}gpu; mkdir -p $OUT
PORT=29700
run() {  # name timeout cmd...
  local name=$1 t=$2; shift 2
  PORT=$((PORT+1))
  local TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"
  local t0=$(date +%s)
  timeout -k 10 ${MPB_SUITE_TIMEOUT:-$t} $TR "$@" > $OUT/$name.json 2> $OUT/$name.err
  echo "== $name rc=$? $(( $(date +%s) - t0 ))s"; tail -n 1 $OUT/$name.json | cut -c1-300
}
K=${MPB_SUITE_STEPS:-10}; W=${MPB_SUITE_WARMUP:-3}
for s in $STEPS; do case $s in
  peer)   run peer 150 scripts/test_peer.py ;;
  sweep)  run sweep 240 scripts/peer_sweep.py ;;
  ours)   run ours_resnet34 150 bench.py --gpus $N --steps $K --warmup $W ;;
  wire)   MPB_GRAD_DTYPE=bf16 run ours_resnet34_bf16wire 150 bench.py --gpus $N --steps $K --warmup $W ;;
  ddp)    run ddp_resnet34 240 bench.py --impl ddp --gpus $N --steps $K --warmup $W ;;
  vgg16)  run ours_vgg16 200 bench.py --model vgg16 --gpus $N --steps $K --warmup $W ;;
  ddpvgg) run ddp_vgg16 240 bench.py --impl ddp --model vgg16 --gpus $N --steps $K --warmup $W ;;
  dense)  run ours_densenet 200 bench.py --model densenet --gpus $N --steps $K --warmup $W ;;
  ddpdense) run ddp_densenet 240 bench.py --impl ddp --model densenet --gpus $N --steps $K --warmup $W ;;
  incep)  run ours_inception 240 bench.py --model inception --gpus $N --steps $K --warmup $W ;;
  ddpincep) run ddp_inception 240 bench.py --impl ddp --model inception --gpus $N --steps $K --warmup $W ;;
  pipe)   run pipeline 240 bench.py --task pipeline --gpus $N --steps $K ;;
  main)   MPB_MODEL_NAME=resnet34 MPB_BATCH_SIZE=256 MPB_HEIGHT=224 MPB_WIDTH=224 MPB_NUM_CLASSES=1000 MPB_NUM_EPOCHS=2 \
          MPB_DEBUG=0 MPB_SYNTHETIC_SAMPLES=$((N*256*40)) MPB_VALIDATE=0 MPB_USE_PRETRAINED=0 MPB_LOG_DIR=$OUT/logs \
          MPB_CHECKPOINT_DIR=/tmp/mpb_ck/ MPB_MODELS_DIR=/tmp/mpb_m/ run main_resnet34 240 main.py ;;
esac; done
