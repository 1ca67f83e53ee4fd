import torch, json
dev="cuda"
def timeit(fn, iters=200):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(5): fn()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(20): fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(s)
        for _ in range(iters // 20): g.replay()
        e1.record(s); torch.cuda.synchronize()
        return round(e0.elapsed_time(e1) * 1000 / (iters // 20 * 20), 2)
res={}
n=16850
for V in (8192, 117581, 2000000):
    ids=torch.randperm(V, device=dev)[:n].contiguous() if V>=n else torch.randint(0,V,(n,),device=dev)
    for W in (32, 96, 128):
        tab=torch.randn(V,W,device=dev); out=torch.empty(n,W,device=dev); src=torch.randn(n,W,device=dev)
        res[f"gather_V{V}_W{W}"]=timeit(lambda: torch.index_select(tab,0,ids,out=out))
        res[f"scatter_V{V}_W{W}"]=timeit(lambda: tab.index_copy_(0,ids,src))
    v1=torch.randn(V,device=dev); o1=torch.empty(n,device=dev)
    res[f"gather_scalar_V{V}"]=timeit(lambda: torch.index_select(v1,0,ids,out=o1))
    res[f"scatter_scalar_V{V}"]=timeit(lambda: v1.index_copy_(0,ids,o1))
# embedding-style gather via F.embedding (different kernel)
tab=torch.randn(117581,32,device=dev); ids80=torch.randint(0,117581,(2048,39),device=dev)
res["F_embedding_80k"]=timeit(lambda: torch.nn.functional.embedding(ids80,tab))
res["take_along_80k"]=timeit(lambda: tab[ids80])
print(json.dumps(res,indent=1))
