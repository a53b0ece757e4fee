#!/bin/bash
# usage: tools/run_n8.sh [N] [full]   -- the multi-GPU evidence run (default 8 GPUs): multi-process tests, default bench line,
# MNIST step timeline, the two examples through spark.run; with `full` also the allreduce sweep to 256 MB with NCCL side
# by side, the dp_step phase trace and the graph-captured stand-in.  Everything lands in gpurun_out/.
N=${1:-8}
FULL=${2:-}
mkdir -p gpurun_out
run() { local to=$1 port=$2; shift 2; timeout $to python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; }
echo "== pytest multigpu ($N GPUs)"
timeout 500 python -m pytest tests/test_gpu_comm.py -m gpu -q -k "multi_gpu or silent_peer" 2>&1 | tail -15 | tee gpurun_out/pytest_multigpu_n$N.log
python - <<PY
import json
try:
    for o in json.load(open('gpurun_out/multigpu_checks.json')):
        print({k: o[k] for k in ('rank', 'mc', 'weights_equal', 'fused_matches', 'sharded_matches', 'sharded_err', 'cnn_weights_equal', 'cnn_modes', 'resume_ok', 'resume_err', 'll_mc_push_mismatches') if k in o})
except Exception as e:
    print('no multigpu_checks.json:', e)
PY
echo "== bench (default line)"
DDLS_BENCH_WATCHDOG_S=200 run 250 29541 bench.py --gpus $N --steps 300 --warmup 20 2>gpurun_out/bench_n$N.err | grep '^{' > gpurun_out/bench_n$N.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_n$N.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches_per_step')}, d['e2e'], d['config']['dp_step'])
s=d['secondary']; print({k:s[k] for k in ('value','ms_per_step','vs_published')}, s['e2e'], s['config']['dp_step'])
print(d.get('allreduce_us')); print(d['clocks'])
PY
echo "== timeline mnist"
run 150 29543 tools/step_timeline.py mnist_cnn 2>/dev/null | grep -v '^\*\|OMP' | cut -c1-150
echo "== examples through spark.run"
timeout 300 python examples/ddl_mnist.py --num-workers $N 2>&1 | grep -v Warning | tail -12 | tee gpurun_out/example_mnist_n$N.log
W=$N; [ $N -gt 4 ] && W=4
timeout 400 python examples/ddl_nyiso.py --num-workers $W --hours 11566 --report gpurun_out/ddl_nyiso_run.md 2>&1 | grep -i "parameter updates\|training time\|training loop\|MAPE\|run record\|Error\|error" | tee gpurun_out/example_nyiso_n$W.log
if [ -n "$FULL" ]; then
echo "== sweep"
run 400 29542 -m distributeddeeplearningspark_b200.bench.allreduce_sweep --max-mb ${SWEEP_MB:-256} --graph --out gpurun_out/sweep_n$N.json 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['bytes'], {k:round(v,1) for k,v in r.items() if k.endswith('_us')}, r['best'], round(r['speedup_vs_nccl'],2))"
echo "== dp_step trace"
SHADOW=1 BLOCKS=0,16 run 120 29546 tools/dp_step_trace.py 2>&1 | grep '^blocks'
echo "== stand-in (graph-captured)"
for m in mnist_cnn nyiso_lstm; do run 200 29547 bench.py --gpus $N --impl standin --model $m --steps 300 --warmup 20 2>gpurun_out/standin_n${N}_$m.err | grep '^{' > gpurun_out/standin_n${N}_$m.json; cut -c1-420 gpurun_out/standin_n${N}_$m.json; done
fi
