"""Warm per-kernel timings of the fused training step (CUDA events, serialized) at the bench config."""
import ctypes as C, json, sys
import torch
sys.path.insert(0, ".")
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.engine import FusedEngine
from deepfm_b200 import _native

B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
V = int(sys.argv[2]) if len(sys.argv) > 2 else 117581
cfg = DeepFMConfig(39, V, 32, (256, 128, 64), (0.5, 0.5, 0.5), l2_reg=1e-4)
eng = FusedEngine(cfg, B, optimizer="Adam", lr=5e-4)
eng.load_from_module(DeepFM(cfg, seed=1))
g = torch.Generator(device="cuda").manual_seed(0)
# zipf-ish categorical ids + 13 constant numeric ids
u = torch.rand(B, 39, device="cuda", generator=g)
ids = (13 + ((V - 600) ** u)).long().clamp(14, V - 1)
ids[:, :13] = torch.arange(1, 14, device="cuda")
vals = torch.rand(B, 39, device="cuda", generator=g); vals[:, 13:] = 1.0
labels = (torch.rand(B, device="cuda", generator=g) < 0.25).float()
ids32 = ids.int().contiguous()
eng.set_state(B, train=True)
out = (C.c_float * 3)()
lib = eng._lib
lib.dfm_engine_profile.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
_native.check(lib.dfm_engine_profile(eng._h, ids32.data_ptr(), vals.data_ptr(), labels.data_ptr(), 50, out), "profile")
res = {"forward_us": round(out[0] * 1000, 2), "backward_update_us": round(out[1] * 1000, 2),
       "step_us": round(out[2] * 1000, 2)}
# the same step replayed as a CUDA graph (what training uses)
eng.capture(B)
slab, dids, dvals, dlab = eng.device_slab(0)
dids.copy_(ids32); dvals.copy_(vals); dlab.copy_(labels)
for _ in range(10): eng.replay(0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200): eng.replay(0)
e1.record(); torch.cuda.synchronize()
res["graph_step_us"] = round(e0.elapsed_time(e1) * 1000 / 200, 2)
res["unique_ids"] = int(ids.unique().numel())
res["batch"] = B
res["V"] = V
res["launches"] = eng.launches_per_step()
print(json.dumps(res))
