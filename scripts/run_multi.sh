# usage: run_multi.sh N  -- the multi-GPU evidence run: self-check, headline bench with phases, large config, both baseline arms, sweep
N=$1
T="timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
mkdir -p gpurun_out
echo "== fused criteo";  DFM_BENCH_PHASES=1 $T 29542 bench.py --gpus $N --steps 3000 --warmup 100 2>gpurun_out/fused_n$N.err | tee gpurun_out/fused_n$N.json | cut -c1-1800; grep -E "phases" gpurun_out/fused_n$N.err | head -3 | cut -c1-420
echo "== fused large";   DFM_BENCH_PHASES=1 $T 29543 bench.py --gpus $N --config large --steps 500 --warmup 30 --no-pipeline --no-check 2>gpurun_out/large_n$N.err | tee gpurun_out/large_n$N.json | cut -c1-400; grep -E "phases" gpurun_out/large_n$N.err | head -2 | cut -c1-420
echo "== eager";         $T 29544 bench.py --gpus $N --impl eager --steps 50 --warmup 5 --no-pipeline 2>/dev/null | tee gpurun_out/eager_n$N.json | cut -c1-420
echo "== sweep";         $T 29546 bench.py --gpus $N --config sweep 2>/dev/null | tee gpurun_out/sweep_n$N.json | cut -c1-200
