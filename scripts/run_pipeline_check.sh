#!/bin/bash
# bench.py's pipeline figure with the CPU budget of one rank of an 8-rank job (4 cores) and with the whole box,
# after the train_batches tests.  Prints value / e2e / pipeline / CPU cores used per thread group.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 80 python -m pytest tests/test_engine_gpu.py -q -x -k "train_batches or pipelined or partial_batch" 2>&1 | tail -2
taskset -c 0-3 timeout 60 python bench.py --steps 1000 --warmup 50 > gpurun_out/bench_pipe_4cores.json 2> gpurun_out/bench_pipe_4cores.err; echo rc=$?
timeout 60 python bench.py --steps 2000 --warmup 50 > gpurun_out/bench_pipe_all.json 2> gpurun_out/bench_pipe_all.err; echo rc=$?
python - <<'PY'
import json
for f in ("4cores", "all"):
    d = json.loads(open("gpurun_out/bench_pipe_%s.json" % f).read().strip().splitlines()[-1])
    print(f, round(d["value"] / 1e6, 2), round(d["e2e"]["value"] / 1e6, 2), round(d["pipeline"]["value"] / 1e6, 2),
          d["pipeline"]["rank0_cpu_cores_used_by_thread"], d["final_loss"])
PY
