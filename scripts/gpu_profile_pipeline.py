"""Where the main thread's time goes in the file -> loader -> Trainer.train_batches pipeline: cProfile over ~1.5 s of
steps on cuda:0 (bench.py's model and batch), top functions by own time.  Host-side only: the GPU is not the subject."""
import cProfile
import io
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import deepfm_b200  # noqa: E402,F401
from deepfm_b200.data.loader import BatchLoader  # noqa: E402
from deepfm_b200.data.tfrecord import generate_synthetic  # noqa: E402
from deepfm_b200.models.deepfm import DeepFMConfig  # noqa: E402
from deepfm_b200.runtime.trainer import Trainer, TrainerOptions  # noqa: E402

B, F, V = 2048, 39, 117581
path = "/tmp/dfm_prof_pipeline.tfrecords"
if not os.path.exists(path):
    generate_synthetic(path, B * 128, F, V, n_dense=13, seed=1, pos_rate=0.2553, zipf_s=1.05)
dev = torch.device("cuda", 0)
cfg = DeepFMConfig(field_size=F, feature_size=V, embedding_size=32, deep_layers=(256, 128, 64), dropout_keep=(0.5,) * 3, l2_reg=1e-4)
trainer = Trainer(cfg, TrainerOptions(batch_size=B, optimizer="Adam", learning_rate=5e-4, backend="fused", seed=1, cuda_graph=True), device=dev)
loader = BatchLoader([path], F, B, num_epochs=100000, num_workers=max(2, len(os.sched_getaffinity(0))), num_slots=8,
                     ids_dtype=torch.int32, verify_crc=True)


def run(seconds):
    n, t0 = 0, time.time()
    gen = trainer.train_batches(loader)
    for _ in gen:
        n += 1
        if n % 64 == 0:
            trainer.last_loss()
            if time.time() - t0 >= seconds:
                break
    gen.close()
    return n, time.time() - t0


run(0.5)
torch.cuda.synchronize()
n, dt = run(1.5)
print(f"unprofiled: {n / dt:.0f} batches/s, {dt / n * 1e6:.1f} us per batch")
pr = cProfile.Profile()
pr.enable()
n, dt = run(1.5)
pr.disable()
print(f"profiled:   {n / dt:.0f} batches/s, {dt / n * 1e6:.1f} us per batch (n = {n})")
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
print("\n".join(l for l in s.getvalue().splitlines() if l.strip())[:6000])
loader.close()
