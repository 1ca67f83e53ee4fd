"""Fused NVLink collectives vs the restated NCCL path on real GPUs (skipped on boxes with fewer than 2; the same
protocol runs on ONE GPU in tests/test_multi_rank_emu_gpu.py, and bench.py repeats the check before its timed region
at every N > 1)."""
import json
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_dp_check(world: int, opt: str, port: int, **extra_env):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", **extra_env)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "gpu_dp_check.py"), opt]
    p = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    results = [json.loads(m) for m in re.findall(r'\{"rank": \d+.*?"ok": (?:true|false)\}', p.stdout)]
    return p, results


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("opt,env", [("GD", {}), ("Adam", {}), ("GD", {"DFM_NO_MULTICAST": "1"}),
                                     ("GD", {"DFM_COMM_PROVIDER": "ipc"})])
def test_fused_collectives_two_ranks(opt, env):
    """Replicas stay bit-identical and the update matches eager fp32 + NCCL (scripts/gpu_dp_check.py), with NVLS
    multicast, with peer loops on symmetric memory, and on plain CUDA IPC mappings."""
    p, results = _run_dp_check(2, opt, 29631, **env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert len(results) == 2 and all(r["ok"] for r in results)
    for r in results:
        assert all(v for k, v in r.items() if k.startswith("identical:"))
        assert r.get("stress_2000_steps_identical", True)      # delayed-rank stress of the flag protocol
        if opt == "GD":
            assert r["worst_rel_update_err"] < 0.05


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("ke", ["16", "64"])
def test_fused_collectives_other_embedding_sizes(ke):
    """Ke/4 lanes serve one record in the exchange kernel: 4 and 16 lanes per record, and a 400-400 MLP."""
    p, results = _run_dp_check(2, "GD", 29633, DFM_DP_KE=ke, DFM_DP_LAYERS="400,400")
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert len(results) == 2 and all(r["ok"] for r in results)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs 4 GPUs")
def test_fused_collectives_four_ranks():
    p, results = _run_dp_check(4, "GD", 29634)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert len(results) == 4 and all(r["ok"] for r in results)


@pytest.mark.skipif(torch.cuda.device_count() < 8, reason="needs 8 GPUs")
def test_fused_collectives_eight_ranks():
    p, results = _run_dp_check(8, "GD", 29632)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert len(results) == 8 and all(r["ok"] for r in results)
