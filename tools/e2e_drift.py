#!/usr/bin/env python
"""Does the host-bound build get faster with use on a fresh box?  N builds of cfg3 with pinned host posteriors, wall time each."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ipk_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
wl = synth.WORKLOADS["cfg3_16s_10k"]
thr, lthr = capi.score_threshold(wl.omega, wl.sigma, wl.k)
h = torch.empty((2 * wl.n_branches, wl.width, wl.sigma), dtype=torch.float32, pin_memory=True)
_, nb = synth.posteriors(wl.n_taxa, wl.width, wl.sigma, seed=42, out=h.numpy())
ctx = capi.Context(0)
inp, keep = capi.make_input(h.numpy(), nb, wl.width, wl.sigma, wl.bits, wl.k, lthr, thr, wl.n_tree_nodes, capi.FILTER_MIF0)
out = []
for i in range(n):
    t0 = time.perf_counter()
    res = ctx.build_raw(inp, result_on_device=False, indexed=True, branch16=True)
    out.append((time.perf_counter() - t0) * 1e3)
print("e2e ms per build:", " ".join(f"{x:.0f}" for x in out))
