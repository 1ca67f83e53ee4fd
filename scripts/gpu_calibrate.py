"""What do trivially small kernels cost on this box?  (calibration for the latency-bound step kernels)"""
import torch, json
dev = "cuda"
def timeit(fn, iters=200, graph=True):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(5): fn()
        if graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                for _ in range(20): fn()
            g.replay(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(s)
            for _ in range(iters // 20): g.replay()
            e1.record(s); torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1000 / (iters // 20 * 20)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(s)
        for _ in range(iters): fn()
        e1.record(s); torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1000 / iters
res = {}
V, K = 117581, 32
tab = torch.randn(V, K, device=dev); m = torch.zeros_like(tab); v = torch.zeros_like(tab)
ids = torch.randint(0, V, (16850,), device=dev)
out = torch.empty(16850, K, device=dev)
res["gather_16850x128B_us"] = timeit(lambda: torch.index_select(tab, 0, ids, out=out))
ids80 = torch.randint(0, V, (2048 * 39,), device=dev); out80 = torch.empty(2048 * 39, K, device=dev)
res["gather_79872x128B_us"] = timeit(lambda: torch.index_select(tab, 0, ids80, out=out80))
p = torch.randn(361000, device=dev); g = torch.randn(361000, device=dev)
res["axpy_361k_us"] = timeit(lambda: p.add_(g, alpha=0.1))
big = torch.randn(64 << 20, device=dev); big2 = torch.empty_like(big)
res["copy_256MB_us"] = timeit(lambda: big2.copy_(big), iters=40)
res["copy_256MB_GBs"] = 2 * big.numel() * 4 / res["copy_256MB_us"] / 1e3
src = torch.randn(16850, K, device=dev)
res["scatter_rows_index_copy_us"] = timeit(lambda: tab.index_copy_(0, ids, src))
res["index_add_16850_us"] = timeit(lambda: tab.index_add_(0, ids, src))
res["index_add_79872_us"] = timeit(lambda: tab.index_add_(0, ids80, out80))
tiny = torch.zeros(1, device=dev)
res["tiny_kernel_us"] = timeit(lambda: tiny.add_(1.0))
a = torch.randn(2048, 1248, device=dev, dtype=torch.bfloat16); w = torch.randn(1248, 256, device=dev, dtype=torch.bfloat16)
res["cublas_2048x1248x256_us"] = timeit(lambda: a @ w)
res["nograph_tiny_kernel_us"] = timeit(lambda: tiny.add_(1.0), graph=False)
print(json.dumps(res, indent=1))
