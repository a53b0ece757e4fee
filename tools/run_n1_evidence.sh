#!/bin/bash
# usage: tools/run_n1_evidence.sh  -- single-GPU evidence: warm per-kernel times with clocks, GEMM bench (L2 flushed),
# step timelines, graph-captured stand-in, full-size stacked-LSTM drift test and bench.  Results in gpurun_out/.
mkdir -p gpurun_out
for m in mnist_cnn nyiso_lstm nyiso_gru; do timeout 100 python tools/kernel_times.py $m 2>&1 | tail -3; done
timeout 150 python tools/step_timeline.py mnist_cnn 2>/dev/null | tail -4
timeout 100 python tools/step_timeline.py nyiso_lstm 2>/dev/null | tail -3
timeout 300 python tools/gemm_bench.py 2>&1 | cut -c1-260
for m in mnist_cnn nyiso_lstm nyiso_gru stacked_lstm; do
  s=300; [ $m = stacked_lstm ] && s=12
  timeout 200 python bench.py --impl standin --model $m --steps $s --warmup 10 2>gpurun_out/standin_n1_$m.err | grep '^{' > gpurun_out/standin_n1_$m.json; cut -c1-160 gpurun_out/standin_n1_$m.json
done
timeout 300 python -m pytest tests/test_gpu_recurrent.py -m gpu -x -q -k "benchmark_size" 2>&1 | tail -4; cat gpurun_out/stacked_lstm_drift.json
for m in nyiso_gru stacked_lstm; do
  s=300; [ $m = stacked_lstm ] && s=12
  timeout 300 python bench.py --model $m --steps $s --warmup 10 2>gpurun_out/bench_n1_$m.err | grep '^{' > gpurun_out/bench_n1_$m.json; cut -c1-200 gpurun_out/bench_n1_$m.json
done
