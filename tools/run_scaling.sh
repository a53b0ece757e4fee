#!/bin/bash
# usage: tools/run_scaling.sh N   -- benches of all four models + the allreduce sweep on N GPUs, results in gpurun_out/
N=${1:-2}
mkdir -p gpurun_out
PORT=29520
launch() {   # launch <timeout> <args...>
  local to=$1; shift
  if [ "$N" = 1 ]; then timeout $to python "$@"
  else timeout $to python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@"; fi
}
MODELS=${MODELS:-"mnist_cnn nyiso_lstm nyiso_gru stacked_lstm"}
SWEEP_MB=${SWEEP_MB:-64}
TAG=""; [ "$SWEEP_MB" != 64 ] && TAG="_le${SWEEP_MB}mb"
for m in $MODELS; do
  PORT=$((PORT+1))
  steps=300; [ $m = stacked_lstm ] && steps=12
  launch 300 bench.py --gpus $N --model $m --steps $steps --warmup 10 > gpurun_out/bench_n${N}_$m.log 2>&1
  grep '^{' gpurun_out/bench_n${N}_$m.log > gpurun_out/bench_n${N}_$m.json
  cut -c1-330 gpurun_out/bench_n${N}_$m.json
done
if [ "$N" != 1 ] && [ -z "$SKIP_SWEEP" ]; then
  PORT=$((PORT+1))
  launch 400 -m distributeddeeplearningspark_b200.bench.allreduce_sweep --max-mb $SWEEP_MB --graph --out gpurun_out/sweep${TAG}_n$N.json > gpurun_out/sweep${TAG}_n$N.log 2>&1
  tail -12 gpurun_out/sweep${TAG}_n$N.log | cut -c1-400
fi
