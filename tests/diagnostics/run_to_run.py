"""Run-to-run reproducibility of the trainer (same init, same batches, twice): losses per iteration.
    MODE=graph|eager python tests/diagnostics/run_to_run.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import p2p_pytorch_b200 as P
from oracle import pix2pix_oracle as orc

mode = os.environ.get("MODE", "graph")
P.set_precision(os.environ.get("PREC", "fp32"))
batches = [tuple(t.cuda() for t in orc.synthetic_batch(2, 32, 32, 70 + i)) for i in range(6)]
runs = []
for r in range(3):
    g = P.ExpandNetwork().cuda(); d = P.MultiscaleDiscriminator(6, 64, getIntermFeat=True).cuda()
    g.load_state_dict(orc.init_g_state(123)); d.load_state_dict(orc.init_d_state(124))
    tr = P.Pix2PixTrainer(g, d, cuda_graph=(mode == "graph"), graph_warmup=1, two_lanes=os.environ.get("LANES", "2") == "2")
    rec = []
    for bt in batches:
        o = tr.step(*bt)
        rec.append((o["loss_d"].item(), o["loss_g"].item()))
    runs.append(rec)
for i in range(len(batches)):
    print(i, " | ".join(f"{r[i][0]:.6f} {r[i][1]:.6f}" for r in runs))
