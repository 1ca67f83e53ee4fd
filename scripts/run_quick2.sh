N=$1
T="timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
echo "== fused criteo"; DFM_BENCH_PHASES=1 $T 29542 bench.py --gpus $N --steps 3000 --warmup 100 --no-pipeline --no-check 2>gpurun_out/q.err | cut -c1-330; grep -E "phases" gpurun_out/q.err | head -2 | cut -c1-420
echo "== fused large";  DFM_BENCH_PHASES=1 $T 29543 bench.py --gpus $N --config large --steps 500 --warmup 30 --no-pipeline --no-check 2>gpurun_out/q2.err | cut -c1-330; grep -E "phases" gpurun_out/q2.err | head -2 | cut -c1-420
