"""Diagnostic: per-iteration difference between two eager runs, and between an eager and a graph-replayed run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import p2p_pytorch_b200 as P
from oracle import pix2pix_oracle as orc

DEV = "cuda"
prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
P.set_precision(prec)


def run(graphed, iters=8):
    g = P.ExpandNetwork().to(DEV)
    d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True).to(DEV)
    g.load_state_dict(orc.init_g_state(123))
    d.load_state_dict(orc.init_d_state(124))
    tr = P.Pix2PixTrainer(g, d, cuda_graph=graphed, graph_warmup=1)
    rec = []
    for it in range(iters):
        a, b, c = orc.synthetic_batch(2, 32, 32, 50 + it)
        out = tr.step(a.to(DEV), b.to(DEV), c.to(DEV))
        rec.append({k: v.detach().float().cpu().clone() for k, v in out.items()})
    return rec


e1, e2, gr = run(False), run(False), run(True)
for name, x, y in (("eager/eager", e1, e2), ("eager/graph", e1, gr)):
    for it in range(len(x)):
        dl = {k: abs(x[it][k].item() - y[it][k].item()) / abs(x[it][k].item()) for k in ("loss_d", "loss_g")}
        df = (x[it]["fake_b"] - y[it]["fake_b"]).abs().max().item()
        print(f"{prec} {name} it {it}: loss_d {dl['loss_d']:.2e} loss_g {dl['loss_g']:.2e} fake_b max {df:.2e}")
