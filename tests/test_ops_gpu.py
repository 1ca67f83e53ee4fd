"""Standalone ops of libdfm_cuda.so vs PyTorch references."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,V", [(1, 10), (2048 * 39, 117581), (5000, 300), (100000, 2000000)])
def test_unique_slots_matches_torch_unique(n, V):
    from deepfm_b200.ops.unique import UniqueSlots
    g = torch.Generator(device="cuda").manual_seed(n)
    ids = torch.randint(0, V, (n,), device="cuda", generator=g, dtype=torch.int32)
    if n > 100:
        ids[::3] = 7                                   # a hot id, like the numeric Criteo fields
    op = UniqueSlots(V)
    for _ in range(2):                                 # the workspace is left clean for the next call
        uniq, pos, count = op(ids)
        torch.cuda.synchronize()
        c = int(count.item())
        ref = torch.unique(ids)
        assert c == ref.numel()
        assert torch.equal(torch.sort(uniq[:c]).values, ref)
        assert torch.equal(uniq[pos.long()], ids)
    hdr = op.workspace.numel() - 4 * V                 # the id -> slot map follows a small header
    assert int(op.workspace[hdr:].view(torch.int32).max().item()) == -1
