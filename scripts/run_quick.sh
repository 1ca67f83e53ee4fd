# usage: run_quick.sh N [steps] -- headline bench on N GPUs with the collective kernels' phase stamps (no self-check, no pipeline)
N=$1; STEPS=${2:-3000}
T="timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
DFM_BENCH_PHASES=1 $T 29542 bench.py --gpus $N --steps $STEPS --warmup 100 --no-pipeline --no-check 2>gpurun_out/q.err | cut -c1-330; grep -E "phases" gpurun_out/q.err | tr '[' '\n' | grep rank | sort | head -3 | cut -c1-330
