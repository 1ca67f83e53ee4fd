# usage: run_variants.sh N "VAR=1 VAR2=2" "..."   -- bench.py under torchrun on N GPUs, once per environment variant
N=$1; shift
B="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
for v in "$@"; do
  echo "== variant [$v]"
  env $v DFM_BENCH_PHASES=1 $B 29542 bench.py --gpus $N --steps 2000 --warmup 50 2>&1 | grep -E "^\{|phases|Error|symmetric" | sed -e 's/"config".*"e2e"/"e2e"/' | cut -c1-700
done
