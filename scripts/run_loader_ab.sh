#!/bin/bash
# Old-vs-new host data layer on a GPU box, with the CPU budget one rank has in an 8-rank job (4 of 32 cores):
#   1. the loader alone (build/abtest/lb: dfio_loader_* over a page-cached file, best of 6), old and new library
#   2. bench.py's `pipeline` figure (file -> loader -> Trainer.train_batches -> loss) pinned to 4 cores, new then old
# build/abtest/ holds the previous commit's libdfm_io.so and the harness (scratch, not in git).
set -u
cd "$(dirname "$0")/.."
P=deepfm-tensorflow-distributed-training-on-amazon-sagemaker_b200/_native
OUT=gpurun_out
mkdir -p $OUT
cp $P/libdfm_io.so /tmp/dfm_io_new.so
python - <<'PY'
import sys
sys.path.insert(0, ".")
import deepfm_b200
from deepfm_b200.data.tfrecord import generate_synthetic
generate_synthetic("/tmp/lb.tfrecords", 524288, 39, 117581, n_dense=13, seed=1, pos_rate=0.2553, zipf_s=1.05)
PY
{
  nproc; grep -m1 "model name" /proc/cpuinfo
  for cores in 0-3 0-15; do
    w=4; [ $cores = 0-15 ] && w=16
    for rep in 1 2; do
      echo "cores $cores workers $w old:"; taskset -c $cores timeout 60 build/abtest/lb build/abtest/libdfm_io_old.so /tmp/lb.tfrecords $w 6
      echo "cores $cores workers $w new:"; taskset -c $cores timeout 60 build/abtest/lb /tmp/dfm_io_new.so /tmp/lb.tfrecords $w 6
    done
  done
  echo "stream mode, cores 0-3, 3 workers old/new:"
  STREAM=1 taskset -c 0-3 timeout 60 build/abtest/lb build/abtest/libdfm_io_old.so /tmp/lb.tfrecords 3 4
  STREAM=1 taskset -c 0-3 timeout 60 build/abtest/lb /tmp/dfm_io_new.so /tmp/lb.tfrecords 3 4
} > $OUT/loader_ab.txt 2>&1
taskset -c 0-3 timeout 100 python bench.py --steps 1000 --warmup 50 > $OUT/bench_4cores_new.json 2> $OUT/bench_4cores_new.err
cp build/abtest/libdfm_io_old.so $P/libdfm_io.so
taskset -c 0-3 timeout 100 python bench.py --steps 1000 --warmup 50 > $OUT/bench_4cores_old.json 2> $OUT/bench_4cores_old.err
cp /tmp/dfm_io_new.so $P/libdfm_io.so
cat $OUT/loader_ab.txt
for f in new old; do python -c "
import json,sys
d=json.loads(open('$OUT/bench_4cores_$f.json').read().strip().splitlines()[-1])
print('$f', 'value', round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'pipeline', round(d['pipeline']['value']/1e6,2), 'threads', d['pipeline']['parser_threads_per_rank'], 'loader_records_per_sec', round(d.get('loader_records_per_sec',0)/1e6,2))
"; done
