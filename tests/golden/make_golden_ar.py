"""Fixtures for the posterior loader: small raxml-ng style ancestralProbs files and what the REFERENCE's csv reader
(oracle/_ref/ref_csv_{dna,aa}, i.e. /root/reference's csv.h instantiated as in ipk/src/ar.cpp:200-259) makes of them.
Run in the build container (needs oracle/_ref):   python tests/golden/make_golden_ar.py"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_host import random_probs, write_tsv, REF_CSV       # noqa: E402

for sigma, name in ((4, "dna"), (20, "aa")):
    rng = np.random.default_rng(100 + sigma)
    tsv = os.path.join(HERE, f"ar_{name}.tsv")
    write_tsv(tsv, random_probs(rng, 3, 25, sigma), ["Node1", "Node2", "Node3"], sigma, rng, tricky=True)
    subprocess.run([REF_CSV[sigma], tsv, os.path.join(HERE, f"ar_{name}.f32")], check=True)
    print(tsv, os.path.getsize(tsv))
