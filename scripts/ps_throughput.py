"""BASELINE.json config 4 on this host: async parameter-server mode, 2 PS + 4 workers over gloo, batch 1024,
MLP 128-64-32, F 39, V 117,581 (the NB-PS notebook's hyper-parameters).  Prints examples/s over all workers."""
import os, socket, sys, time
import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepfm_b200.models.deepfm import DeepFMConfig
from deepfm_b200.parallel import ps as P

N_PS, N_W, B, STEPS = 2, 4, 1024, int(os.environ.get("DFM_PS_STEPS", "30"))
CFG = dict(field_size=39, feature_size=117581, embedding_size=32, deep_layers=[128, 64, 32], dropout_keep=[0.5] * 3, l2_reg=1e-4)


def proc(rank, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(max(1, (os.cpu_count() or 8) // (N_PS + N_W)))
    dist.init_process_group("gloo", rank=rank, world_size=N_PS + N_W)
    cfg = DeepFMConfig(**CFG)
    if rank < N_PS:
        P.ParameterServer(cfg, rank, N_PS, N_W, "Adam", 5e-4).serve()
    else:
        w = rank - N_PS
        client = P.PSClient(cfg, N_PS)
        worker = P.PSWorker(cfg, client, seed=7)
        if w == 0:
            for n, t in worker.model.tf_variables().items():
                client.set(n, t)
            client.broadcast_simple(P.OP_INIT_DONE, aux=0)
        else:
            while not client.ask(P.OP_IS_INIT):
                time.sleep(0.01)
        rng = np.random.default_rng(100 + w)
        data = []
        for _ in range(4):
            ids = rng.zipf(1.3, (B, 39)).clip(1, 117580); ids[:, :13] = np.arange(1, 14)
            data.append((torch.from_numpy(ids), torch.from_numpy(rng.random((B, 39)).astype(np.float32)),
                         torch.from_numpy((rng.random(B) < 0.25).astype(np.float32))))
        for i in range(3):
            worker.train_batch(*data[i % 4])
        t0 = time.time()
        for i in range(STEPS):
            worker.train_batch(*data[i % 4])
        q.put((w, STEPS * B / (time.time() - t0)))
        client.broadcast_simple(P.OP_DONE)
    dist.destroy_process_group()


if __name__ == "__main__":
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=proc, args=(r, port, q)) for r in range(N_PS + N_W)]
    [p.start() for p in ps]
    rates = [q.get(timeout=600) for _ in range(N_W)]
    [p.join() for p in ps]
    print({"workers": N_W, "ps": N_PS, "batch": B, "steps_per_worker": STEPS, "examples_per_s_total": round(sum(r for _, r in rates)),
           "per_worker": [round(r) for _, r in sorted(rates)], "host_cores": os.cpu_count()})
