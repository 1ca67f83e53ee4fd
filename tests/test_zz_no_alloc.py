"""dfm_unique_slots_static -- the entry point for callers that cannot allocate GPU memory (a plug-in
inside another framework's process, whose allocator owns the device; a loop that cannot afford
cudaMalloc's device-wide synchronisation).  It must give what dfm_unique_slots gives, using only the
scratch array built into the library, and keep doing so on repeated calls.

Called straight through the C ABI with ctypes, as such a caller would.  This file sorts last so a
``pytest -x`` run has already been through every other test when it gets here."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from deepfm_b200 import _native
    lib = _native.cuda()
    lib.dfm_last_error.restype = C.c_char_p
    lib.dfm_unique_slots_workspace_bytes.restype = C.c_longlong
    lib.dfm_unique_slots_workspace_bytes.argtypes = [C.c_longlong]
    lib.dfm_unique_slots_init.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p]
    lib.dfm_unique_slots.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p]
    lib.dfm_unique_slots_static_max_v.restype = C.c_longlong
    lib.dfm_unique_slots_static.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p]
    return lib


def _outputs(n):
    return (torch.full((n,), -7, dtype=torch.int32, device="cuda"), torch.full((n,), -7, dtype=torch.int32, device="cuda"),
            torch.full((1,), -7, dtype=torch.int32, device="cuda"))


def test_static_scratch_variant_matches_workspace_variant_twice_in_a_row():
    lib = _lib()
    V, n = 117581, 2048 * 39
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(11)
    batches = []
    for k in range(2):                                   # two different batches, one call each
        ids = torch.randint(0, V, (n,), device="cuda", generator=g, dtype=torch.int32)
        ids[k::5] = 3 + k                                # hot ids
        batches.append(ids)

    # reference: the existing entry point with a caller-owned workspace
    ws = torch.empty(lib.dfm_unique_slots_workspace_bytes(V), dtype=torch.uint8, device="cuda")
    assert lib.dfm_unique_slots_init(ws.data_ptr(), V, stream) == 0
    ref = []
    for ids in batches:
        u, p, c = _outputs(n)
        rc = lib.dfm_unique_slots(ids.data_ptr(), n, V, ws.data_ptr(), u.data_ptr(), p.data_ptr(), c.data_ptr(), stream)
        assert rc == 0, lib.dfm_last_error()
        ref.append((u, p, c))

    # the variant under test: no workspace argument, nothing allocated inside
    got = []
    for ids in batches:
        u, p, c = _outputs(n)
        rc = lib.dfm_unique_slots_static(ids.data_ptr(), n, V, u.data_ptr(), p.data_ptr(), c.data_ptr(), stream)
        assert rc == 0, lib.dfm_last_error()
        got.append((u, p, c))
    torch.cuda.synchronize()

    for ids, (ur, pr, cr), (ug, pg, cg) in zip(batches, ref, got):
        c = int(cr.item())
        assert int(cg.item()) == c == torch.unique(ids).numel()
        # slot numbering depends on which thread claims first, so compare what is order-free ...
        assert torch.equal(torch.sort(ug[:c]).values, torch.sort(ur[:c]).values)
        # ... and that each numbering is self-consistent: the slot of a position names that position's id
        assert torch.equal(ug[pg.long()], ids) and torch.equal(ur[pr.long()], ids)


def test_static_scratch_variant_rejects_oversized_id_space():
    lib = _lib()
    max_v = lib.dfm_unique_slots_static_max_v()
    assert max_v == 1 << 20
    ids = torch.zeros(8, dtype=torch.int32, device="cuda")
    u, p, c = _outputs(8)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.dfm_unique_slots_static(ids.data_ptr(), 8, max_v + 1, u.data_ptr(), p.data_ptr(), c.data_ptr(), stream)
    assert rc != 0
    msg = lib.dfm_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(max_v) in msg
    torch.cuda.synchronize()
    assert int(c.item()) == -7                           # nothing ran


def test_out_of_range_ids_stay_inside_the_scratch():
    """An id outside [0, V) -- a corrupt record, a too-small feature_size -- is treated as the nearest valid id by both
    entry points, so it can never index outside the id -> slot map (for the built-in scratch that would mean
    scribbling over whatever the module placed next to the array)."""
    lib = _lib()
    V, n = 1000, 4096
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(5)
    ids = torch.randint(0, V, (n,), device="cuda", generator=g, dtype=torch.int32)
    ids[7], ids[8], ids[9] = -3, V, 2 ** 31 - 1
    clamped = ids.clamp(0, V - 1)
    ws = torch.empty(lib.dfm_unique_slots_workspace_bytes(V), dtype=torch.uint8, device="cuda")
    assert lib.dfm_unique_slots_init(ws.data_ptr(), V, stream) == 0
    u1, p1, c1 = _outputs(n)
    assert lib.dfm_unique_slots(ids.data_ptr(), n, V, ws.data_ptr(), u1.data_ptr(), p1.data_ptr(), c1.data_ptr(), stream) == 0
    u2, p2, c2 = _outputs(n)
    assert lib.dfm_unique_slots_static(ids.data_ptr(), n, V, u2.data_ptr(), p2.data_ptr(), c2.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    for u, p, c in ((u1, p1, c1), (u2, p2, c2)):
        k = int(c.item())
        assert k == torch.unique(clamped).numel()
        assert torch.equal(u[p.long()], clamped)
    # the caller-owned map is back to -1 everywhere: nothing was written past an edge and left behind
    assert int(ws[ws.numel() - 4 * V:].view(torch.int32).max().item()) == -1
