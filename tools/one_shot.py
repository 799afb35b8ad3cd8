#!/usr/bin/env python
"""The ipk-b200 executable, one shot and cold, on a synthetic workload: posteriors to /dev/shm, then the process from start to exit.
  python tools/one_shot.py [WORKLOAD] [extra ipk-b200 arguments ...]      (environment is passed through: IPKGPU_TRACE=1 ...)"""
import os, shutil, subprocess, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ipk_b200 import hostlib, synth

wl = synth.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else "cfg3_16s_10k"]
extra = [a for a in sys.argv[1:] if a.startswith("-") or a[0].isdigit()]
d = tempfile.mkdtemp(prefix="ipkb200_oneshot_", dir="/dev/shm")
try:
    logp, node_branch = synth.posteriors(wl.n_taxa, wl.width, wl.sigma, seed=42)
    logp.tofile(os.path.join(d, "logp.f32"))
    with open(os.path.join(d, "node_map.tsv"), "w") as f:
        f.write("".join(f"n{i}\t{int(b)}\n" for i, b in enumerate(node_branch)))
    argv = [hostlib.EXE_PATH, "--logp-f32", os.path.join(d, "logp.f32"), "--width", str(wl.width), "--node-map", os.path.join(d, "node_map.tsv"),
            "--states", "dna" if wl.sigma == 4 else "aa", "-k", str(wl.k), "--omega", str(wl.omega), "--mif0" if wl.filter == "mif0" else "--random",
            "--node-count", str(wl.n_tree_nodes), "-o", os.path.join(d, "DB.ipk")] + extra
    for rep in range(2):
        t0 = time.perf_counter()
        r = subprocess.run(argv, capture_output=True, text=True)
        print(f"run {rep}: wall {time.perf_counter() - t0:.3f} s rc={r.returncode}")
        print(r.stderr.strip())
finally:
    shutil.rmtree(d, ignore_errors=True)
