set -x
for m in mnist_cnn nyiso_lstm nyiso_gru; do
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --model $m --steps 300 --warmup 20 2>&1 | grep '^{' > gpurun_out/bench_n8_$m.json
cat gpurun_out/bench_n8_$m.json
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 -m distributeddeeplearningspark_b200.bench.allreduce_sweep --max-mb 64 --graph --out gpurun_out/sweep_n8.json 2>&1 | tail -20
