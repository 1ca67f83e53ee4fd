"""Stage-by-stage comparison of the fused engine against the fp32 PyTorch oracle.
Prints one line per intermediate so a single GPU run localises a broken kernel."""
import sys
import torch
sys.path.insert(0, ".")
from deepfm_b200.models.deepfm import DeepFM, DeepFMConfig
from deepfm_b200.runtime.engine import FusedEngine


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item(), (a - b).abs().max().item()


def run(B, layers, Ke=32, F=39, V=5000, opt="GD", lr=1.0, l2=0.0, mode="lazy", tag=""):
    torch.manual_seed(0)
    cfg = DeepFMConfig(field_size=F, feature_size=V, embedding_size=Ke, deep_layers=layers,
                       dropout_keep=[1.0] * len(layers), l2_reg=l2)
    m = DeepFM(cfg, seed=1).cuda()
    with torch.no_grad():
        m.fm_v.mul_(30); m.fm_w.mul_(30)          # make the FM terms matter
        for b in m.mlp_b: b.uniform_(-0.1, 0.1)
        m.out_b.fill_(0.05); m.fm_bias.fill_(-0.3)
    eng = FusedEngine(cfg, batch_max=B, optimizer=opt, lr=lr, table_update_mode=mode)
    eng.load_from_module(m)
    g = torch.Generator(device="cuda").manual_seed(3)
    ids = torch.randint(1, V, (B, F), device="cuda", generator=g, dtype=torch.int64)
    ids[:, :13] = torch.arange(1, 14, device="cuda")          # Criteo-like hot numeric ids
    ids[0, 14] = ids[0, 15]                                    # intra-row duplicate
    vals = torch.rand(B, F, device="cuda", generator=g)
    vals[:, 13:] = 1.0
    labels = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    ids32 = ids.int().contiguous()

    print(f"=== {tag} B={B} layers={layers} Ke={Ke} opt={opt} mode={mode}")
    # ---------------- forward ----------------
    prob = eng.forward(ids32, vals, labels).clone()
    torch.cuda.synchronize()
    with torch.no_grad():
        y_w, y_v, e = m.fm_terms(ids, vals)
        h = e.reshape(B, -1)
        print("e        ", rel(eng.buffer("act", 0, "bf16").view(-1, (eng.K[0] + 1 + 15) // 16 * 16)[:B, :eng.K[0]], h))
        print("fm_part  ", rel(eng.buffer("fm_part")[:B], m.fm_bias + y_w + y_v))
        print("s_sum    ", rel(eng.buffer("s_sum").view(-1, Ke)[:B], e.sum(1)))
        hs = [h]
        for l in range(len(layers)):
            h = torch.relu(h.bfloat16().float() @ m.mlp_w[l].bfloat16().float() + m.mlp_b[l])
            hs.append(h)
            pitch = (eng.K[l + 1] + 1 + 15) // 16 * 16
            got = eng.buffer("act", l + 1, "bf16").view(-1, pitch)[:B, :layers[l]]
            print(f"act[{l+1}]   ", rel(got, h))
        logits = m(ids, vals)
        print("logit    ", rel(eng.buffer("logit")[:B], logits))
        print("prob     ", rel(prob, torch.sigmoid(logits)))
        st = eng.read_state()
        ref_loss = torch.nn.functional.binary_cross_entropy_with_logits(logits, labels, reduction="sum").item()
        print("loss_sum ", st["loss_sum"], ref_loss)

    # ---------------- one GD step == gradients ----------------
    before = {k: v.clone() for k, v in eng.variables().items()}
    eng.set_state(B, train=True, step=0)
    eng.train_step(ids32, vals, labels, set_state=False)
    torch.cuda.synchronize()
    st = eng.read_state()
    print("state after step:", st, "unique ref", ids.unique().numel())
    m.zero_grad()
    def rt(x): return x + (x.bfloat16().float() - x).detach()
    y_w, y_v, e = m.fm_terms(ids, vals)
    h = rt(e.reshape(B, -1))
    for l in range(len(layers)):
        h = rt(torch.relu(h @ rt(m.mlp_w[l]) + m.mlp_b[l]))
    loss = m.loss(m.fm_bias + y_w + y_v + (h @ m.out_w).view(-1) + m.out_b, labels)
    loss.backward()
    after = eng.variables()
    # dz buffers
    print("dlogit   ", rel(eng.buffer("dlogit")[:B], (torch.sigmoid(logits) - labels) / B))
    if opt == "GD":
        for (name, p) in m.tf_variables().items():
            got = (before[name] - after[name]) / lr
            ref = p.grad if p.grad is not None else torch.zeros_like(p)
            if name in ("fm_w", "fm_v") and l2 > 0 and mode == "lazy":
                # lazy mode applies L2 to touched rows only
                touched = torch.zeros(V, dtype=torch.bool, device="cuda"); touched[ids.unique()] = True
                mask = touched if name == "fm_w" else touched[:, None]
                ref = torch.where(mask, ref, torch.zeros_like(ref))
            print(f"grad {name:32s}", rel(got, ref))
    print("pending: n_unique", st["n_unique"], "G max", eng.buffer("G").abs().max().item(),
          "slot_of max", eng.buffer("slot_of", dtype="i4").max().item())
    eng.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    cases = {
        "a": lambda: run(256, [256, 128, 64], tag="a"),
        "b": lambda: run(300, [256, 128, 64], l2=1e-3, tag="b partial tile + l2"),
        "c": lambda: run(2048, [256, 128, 64], tag="c full"),
        "d": lambda: run(512, [128, 64, 32], tag="d notebook layers"),
        "e": lambda: run(384, [400, 400, 400], Ke=64, V=3000, tag="e config3 shape"),
        "f": lambda: run(200, [64], tag="f one layer"),
        "g": lambda: run(256, [256, 128, 64], l2=1e-3, mode="dense_exact", tag="g dense_exact"),
        "h": lambda: run(256, [100, 50], tag="h padded widths"),
    }
    for k, fn in cases.items():
        if which in ("all", k):
            fn()
