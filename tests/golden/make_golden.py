"""Generates tests/golden/*.npz from the REFERENCE'S OWN OBJECT CODE (oracle/_ref/libipkref_{dna,aa}.so =
/root/reference/ipk/src/pk_compute.cpp + window.cpp compiled unmodified, see oracle/Makefile, oracle/ref_driver.cpp).

Run in the build container (needs /root/reference to have built oracle/_ref):   python tests/golden/make_golden.py
The fixtures carry their inputs, so the tests that read them need neither /root/reference nor oracle/_ref.

Per case:  logp [n_nodes, W, sigma] f32, parameters, and for every node the reference's emission as
  counts[n_nodes, n_windows] u32, sha256 of the canonical (node, window, key, score-bits) stream, and -- for cases with
  < 30k records -- the records themselves.  "build" cases add the merged database (ref_driver.cpp's restated glue around
  the reference DCLA): entries by (key, branch) and MIF0 filter values.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oraclelib as O  # noqa: E402


def thresholds(omega, sigma, k):
    t = np.float32(np.float32(omega) / np.float32(sigma)) ** np.float32(k)
    # float powf / log10f, as include/ipkgpu.h: ipkgpu_score_threshold (values are stored in the fixture)
    import ctypes as C
    libm = C.CDLL("libm.so.6")
    libm.powf.argtypes = [C.c_float, C.c_float]; libm.powf.restype = C.c_float
    libm.log10f.argtypes = [C.c_float]; libm.log10f.restype = C.c_float
    t = libm.powf(C.c_float(omega / sigma if False else np.float32(omega) / np.float32(sigma)), C.c_float(k))
    return float(t), float(libm.log10f(t))


def posteriors(rng, n_nodes, W, sigma, peaked, uniform_share=0.02, zero_share=0.3):
    p = rng.dirichlet(np.ones(sigma) * 0.3, size=(n_nodes, W))
    sharp = rng.random((n_nodes, W)) < peaked
    top = rng.integers(0, sigma, size=(n_nodes, W))
    eps = 10.0 ** (-rng.uniform(1, 4, size=(n_nodes, W)))
    q = p.copy()
    np.put_along_axis(q, top[..., None], 0.0, axis=2)
    q = q / np.maximum(q.sum(axis=2, keepdims=True), 1e-300) * eps[..., None]
    np.put_along_axis(q, top[..., None], (1.0 - eps)[..., None], axis=2)
    p = np.where(sharp[..., None], q, p)
    p = np.round(p, 9)                                   # raxml-ng --precision 9 (ar.cpp:670): exact zeros appear
    hard = rng.random((n_nodes, W)) < zero_share * peaked    # some columns with a certain state: p = (1, 0, 0, ...)
    onehot = np.zeros_like(p); np.put_along_axis(onehot, top[..., None], 1.0, axis=2)
    p = np.where((hard & sharp)[..., None], onehot, p)
    uni = rng.random((n_nodes, W)) < uniform_share       # exactly uniform columns: scores sit AT the leaf thresholds
    p = np.where(uni[..., None], 1.0 / sigma, p)
    with np.errstate(divide="ignore"):
        return np.log10(p.astype(np.float32)).astype(np.float32)


def digest(node, win, key, score):
    node, win, key, score = O.canonical_windows(node, win, key, score)
    h = hashlib.sha256()
    h.update(node.astype("<u4").tobytes()); h.update(win.astype("<u4").tobytes())
    h.update(key.astype("<u8").tobytes()); h.update(score.view(np.uint32).astype("<u4").tobytes())
    return h.hexdigest(), (node, win, key, score)


def enum_case(name, seed, sigma, bits, k, omega, W, n_nodes, peaked, **kw):
    rng = np.random.default_rng(seed)
    logp = posteriors(rng, n_nodes, W, sigma, peaked, **kw)
    thr, lthr = thresholds(omega, sigma, k)
    nwin = max(W - k + 1, 0)
    counts = np.zeros((n_nodes, nwin), np.uint32)
    N, Wn, K, S = [], [], [], []
    for n in range(n_nodes):
        w, key, sc = O.ref_enumerate_node(logp[n], sigma, k, lthr)
        np.add.at(counts[n], w, 1)
        N.append(np.full(len(w), n, np.uint32)); Wn.append(w); K.append(key); S.append(sc)
    node, win, key, score = map(np.concatenate, (N, Wn, K, S))
    sha, (node, win, key, score) = digest(node, win, key, score)
    # k-mers a root-only test in exact arithmetic would accept but the reference's fp32 look-ahead drops
    out = dict(logp=logp, sigma=sigma, bits=bits, k=k, omega=omega, W=W, threshold=np.float32(thr), log_threshold=np.float32(lthr),
               counts=counts, sha256=sha, n_records=len(key))
    if len(key) < 30000:
        out.update(node=node, window=win, key=key, score=score)
    np.savez_compressed(os.path.join(HERE, f"enum_{name}.npz"), **out)
    print(f"enum_{name}: {len(key)} records, {sha[:12]}")


def build_case(name, seed, sigma, bits, k, omega, W, n_branches, peaked):
    rng = np.random.default_rng(seed)
    n_nodes = 2 * n_branches
    logp = posteriors(rng, n_nodes, W, sigma, peaked)
    # X0 / X1 of one branch share most of their columns, so that the per-branch max actually merges duplicates
    for b in range(n_branches):
        keep = rng.random(W) < 0.8
        logp[2 * b + 1][keep] = logp[2 * b][keep]
    # ghost nodes in tree-iteration order are not grouped by branch: interleave
    order = rng.permutation(n_nodes)
    node_branch = (order // 2 + 3).astype(np.uint32)          # branch ids need not start at 0
    logp = np.ascontiguousarray(logp[order])
    thr, lthr = thresholds(omega, sigma, k)
    T = n_branches + 2
    db = O.ref_build(logp, node_branch, W, sigma, k, lthr, thr, T, 0, 1)
    ek, eb, es = O.db_entries(db)
    o = np.argsort(db.keys, kind="stable")
    np.savez_compressed(os.path.join(HERE, f"build_{name}.npz"), logp=logp, node_branch=node_branch, sigma=sigma, bits=bits, k=k,
                        omega=omega, W=W, threshold=np.float32(thr), log_threshold=np.float32(lthr), total_num_groups=T,
                        n_emitted=db.n_emitted, entry_key=ek, entry_branch=eb, entry_score=es,
                        kmer_key=db.keys[o], kmer_fv=db.fv[o], ordered_keys=db.keys, offsets=db.offsets, branch=db.branch)
    print(f"build_{name}: {db.n_kmers} k-mers, {db.n_entries} entries, {db.n_emitted} emitted")


if __name__ == "__main__":
    assert O.have_ref(4) and O.have_ref(20), "oracle/_ref is not built (needs /root/reference): make -C oracle"
    # (1) per-window emission; odd / even k, W in {k, k+1, ...}
    enum_case("dna_k4", 1, 4, 2, 4, 1.5, 120, 3, 0.5)
    enum_case("dna_k7_o2", 2, 4, 2, 7, 2.0, 200, 3, 0.6)
    enum_case("dna_k8", 3, 4, 2, 8, 1.5, 200, 2, 0.7)
    enum_case("dna_k10", 4, 4, 2, 10, 1.5, 200, 2, 0.85)
    enum_case("dna_k12", 5, 4, 2, 12, 1.5, 150, 2, 0.9)
    enum_case("dna_k14_o2", 6, 4, 2, 14, 2.0, 150, 2, 0.9)
    enum_case("dna_k5_Weqk", 7, 4, 2, 5, 1.5, 5, 4, 0.5)
    enum_case("dna_k5_Wk1", 8, 4, 2, 5, 1.5, 6, 4, 0.5)
    enum_case("aa_k3", 9, 20, 5, 3, 1.5, 60, 2, 0.7)
    enum_case("aa_k4_o10", 10, 20, 5, 4, 10.0, 150, 3, 0.7)
    enum_case("aa_k6", 11, 20, 5, 6, 1.5, 40, 1, 0.95)
    # (2) long alignment: the fp32 whole-alignment prefix sums make the look-ahead bound lossy
    enum_case("dna_k6_W9700", 123, 4, 2, 6, 1.5, 9700, 1, 0.7)
    enum_case("dna_k7_W20000", 124, 4, 2, 7, 1.5, 20000, 1, 0.8)
    # (3)/(4) many exact zeros (-inf) and many exactly uniform columns
    enum_case("dna_k8_zeros", 12, 4, 2, 8, 1.5, 150, 2, 0.9, zero_share=0.9)
    enum_case("dna_k6_uniform", 13, 4, 2, 6, 1.5, 150, 2, 0.5, uniform_share=0.5)
    enum_case("dna_k6_alluniform", 14, 4, 2, 6, 1.0, 60, 1, 0.0, uniform_share=1.0)
    # (5) group max over X0/X1 + MIF0
    build_case("dna_k8", 21, 4, 2, 8, 1.5, 120, 6, 0.8)
    build_case("dna_k10", 22, 4, 2, 10, 1.5, 100, 5, 0.9)
    build_case("aa_k4", 23, 20, 5, 4, 10.0, 80, 4, 0.8)
