"""CPU tests (-m "not gpu"): the oracle (oracle/ipk_oracle.c, kind "port") against the golden fixtures produced by the
reference's own object code (tests/golden/make_golden.py), against its own brute-force predicate, and -- when
oracle/_ref is present -- against the reference object code live."""
import glob
import hashlib
import os

import numpy as np
import pytest

import oraclelib as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ENUM_FIXTURES = sorted(glob.glob(os.path.join(GOLDEN, "enum_*.npz")))
BUILD_FIXTURES = sorted(glob.glob(os.path.join(GOLDEN, "build_*.npz")))


def _digest(node, win, key, score):
    node, win, key, score = O.canonical_windows(node, win, key, score)
    h = hashlib.sha256()
    h.update(node.astype("<u4").tobytes()); h.update(win.astype("<u4").tobytes())
    h.update(key.astype("<u8").tobytes()); h.update(score.view(np.uint32).astype("<u4").tobytes())
    return h.hexdigest(), (node, win, key, score)


def _oracle_emission(f, fn=O.enumerate_node):
    logp, sigma, bits, k = f["logp"], int(f["sigma"]), int(f["bits"]), int(f["k"])
    lthr = float(f["log_threshold"])
    N, Wn, K, S = [], [], [], []
    for n in range(logp.shape[0]):
        w, key, sc = fn(logp[n], sigma, bits, k, lthr)
        N.append(np.full(len(w), n, np.uint32)); Wn.append(w); K.append(key); S.append(sc)
    return tuple(map(np.concatenate, (N, Wn, K, S)))


def test_fixtures_present():
    assert len(ENUM_FIXTURES) >= 16 and len(BUILD_FIXTURES) >= 3


@pytest.mark.parametrize("path", ENUM_FIXTURES, ids=[os.path.basename(p)[:-4] for p in ENUM_FIXTURES])
def test_port_enumeration_matches_reference_fixture(path):
    f = np.load(path)
    node, win, key, score = _oracle_emission(f)
    sha, (node, win, key, score) = _digest(node, win, key, score)
    assert len(key) == int(f["n_records"])
    counts = np.zeros_like(f["counts"])
    np.add.at(counts, (node, win), 1)
    assert np.array_equal(counts, f["counts"])
    assert sha == str(f["sha256"])
    if "key" in f.files:
        assert np.array_equal(key, f["key"]) and np.array_equal(score.view(np.uint32), f["score"].view(np.uint32))


@pytest.mark.parametrize("name", ["dna_k4", "dna_k5_Weqk", "dna_k5_Wk1", "dna_k6_uniform", "dna_k6_alluniform", "aa_k3", "dna_k7_o2"])
def test_port_predicate_equals_dc(name):
    """DCLA's output == 'every split-tree node's fp32 partial sum > its fp32 threshold'"""
    f = np.load(os.path.join(GOLDEN, f"enum_{name}.npz"))
    a = _digest(*_oracle_emission(f, O.enumerate_node_predicate))[0]
    assert a == str(f["sha256"])


def test_long_alignment_lookahead_only_ever_drops():
    """the fp32 whole-alignment prefix sums make the reference's look-ahead bound lossy on long
    alignments -- it may DROP k-mers whose balanced fp32 score clears the root threshold, never add one.  The emitted
    set (fixture = reference object code, checked above) must be a subset of the root-only set, window by window."""
    f = np.load(os.path.join(GOLDEN, "enum_dna_k6_W9700.npz"))
    M, k, lthr = f["logp"][0], int(f["k"]), np.float32(f["log_threshold"])
    W = M.shape[0]
    w, key, sc = O.enumerate_node(M, 4, 2, k, float(lthr))
    cnt = np.bincount(w, minlength=W - k + 1)
    # balanced fp32 sums 6 -> 3|3 -> 1|2 for every window and every 6-mer
    T3 = (M[:-2, :, None, None] + (M[1:-1, None, :, None] + M[2:, None, None, :])).reshape(W - 2, 64)
    for s0 in range(0, W - k + 1, 1024):
        s1 = min(s0 + 1024, W - k + 1)
        S = (T3[s0:s1, :, None] + T3[s0 + 3:s1 + 3, None, :]).reshape(s1 - s0, 4096)
        assert np.all((S > lthr).sum(1) >= cnt[s0:s1])
    ws, ks = w[:2000].astype(np.int64), key[:2000].astype(np.int64)
    S = T3[ws, ks >> 6] + T3[ws + 3, ks & 63]
    assert np.array_equal(S.astype(np.float32).view(np.uint32), sc[:2000].view(np.uint32))


@pytest.mark.parametrize("path", BUILD_FIXTURES, ids=[os.path.basename(p)[:-4] for p in BUILD_FIXTURES])
def test_port_build_matches_reference_fixture(path):
    f = np.load(path)
    db = O.build(f["logp"], f["node_branch"], int(f["W"]), int(f["sigma"]), int(f["bits"]), int(f["k"]), float(f["log_threshold"]),
                 float(f["threshold"]), int(f["total_num_groups"]), 0, 2)
    assert db.n_emitted == int(f["n_emitted"])
    ek, eb, es = O.db_entries(db)
    assert np.array_equal(ek, f["entry_key"]) and np.array_equal(eb, f["entry_branch"])
    assert np.array_equal(es.view(np.uint32), f["entry_score"].view(np.uint32))
    o = np.argsort(db.keys, kind="stable")
    assert np.array_equal(db.keys[o], f["kmer_key"])
    np.testing.assert_allclose(db.fv[o], f["kmer_fv"], rtol=1e-6, atol=0)
    assert np.array_equal(db.keys, f["ordered_keys"])
    assert np.array_equal(db.offsets, f["offsets"]) and np.array_equal(db.branch, f["branch"])


def test_port_threads_do_not_change_the_database():
    f = np.load(os.path.join(GOLDEN, "build_dna_k8.npz"))
    args = (f["logp"], f["node_branch"], int(f["W"]), 4, 2, 8, float(f["log_threshold"]), float(f["threshold"]), int(f["total_num_groups"]), 0)
    a, b = O.build(*args, 1), O.build(*args, 4)
    for x in ("keys", "fv", "offsets", "branch", "score"):
        assert np.array_equal(getattr(a, x), getattr(b, x))


def test_mif0_known_answers():
    """hand-computable cases of filter.cpp:60-119"""
    T, thr = 10, 0.01
    # a k-mer present on every branch with the same score: uniform distribution => H = log2 T => fv = 0
    assert abs(O.mif0(np.full(T, np.log10(0.5), np.float32), T, thr)) < 1e-6
    # one entry of score s: S = s + (T-1) thr;  fv = S * (H - log2 T)
    s = np.float32(10.0) ** np.float32(-0.25)
    ls = np.float32(-0.25)
    s = float(np.float32(min(10.0 ** float(ls), 1.0)))
    thr32 = float(np.float32(thr))
    S = s + (T - 1) * thr32
    h = lambda x: -x * np.log2(x)
    want = S * ((T - 1) * h(thr32 / S) + h(s / S) - np.log2(T))
    assert abs(O.mif0(np.array([ls], np.float32), T, thr) - want) < 1e-6 * abs(want)
    # scores above 1 are clamped (log > 0 cannot occur, but the clamp is in the reference)
    assert O.mif0(np.array([0.5], np.float32), T, thr) == O.mif0(np.array([0.0], np.float32), T, thr)


def test_random_fv_is_libstdcxx_minstd():
    fv = O.random_fv(np.arange(4))
    # std::default_random_engine(42) + uniform_real_distribution<double>(0,1), libstdc++: keys 0..3 take draws (1,2), (3,4), ...
    x = 42; out = []
    for _ in range(4):
        x = x * 16807 % 2147483647; g1 = x - 1
        x = x * 16807 % 2147483647; g2 = x - 1
        out.append(np.float32((g1 + g2 * 2147483646.0) / (2147483646.0 ** 2)))
    assert np.array_equal(fv, np.array(out, np.float32))
    assert np.all((fv >= 0) & (fv < 1))
    # a function of the key alone (key shards of a multi-GPU build draw what one GPU draws), periodic in 2^31-2
    big = np.array([3, 2**40 + 12345, (2**31 - 2) + 3], np.uint64)
    got = O.random_fv(big)
    assert got[0] == fv[3] == got[2] and 0 <= got[1] < 1


@pytest.mark.skipif(not (O.have_ref(4) and O.have_ref(20)), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("sigma,bits,k,omega,W", [(4, 2, 9, 1.5, 300), (4, 2, 11, 1.5, 200), (4, 2, 13, 2.0, 150), (20, 5, 5, 1.5, 40),
                                                   (4, 2, 3, 1.2, 64), (4, 2, 2, 1.5, 30), (4, 2, 1, 1.5, 30)])
def test_port_equals_reference_object_code_live(sigma, bits, k, omega, W):
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden as G
    rng = np.random.default_rng(k * 1000 + W)
    logp = G.posteriors(rng, 2, W, sigma, 0.85)
    thr, lthr = G.thresholds(omega, sigma, k)
    for n in range(2):
        w1, k1, s1 = O.enumerate_node(logp[n], sigma, bits, k, lthr)
        w2, k2, s2 = O.ref_enumerate_node(logp[n], sigma, k, lthr)
        o1, o2 = np.lexsort((k1, w1)), np.lexsort((k2, w2))
        assert np.array_equal(w1[o1], w2[o2]) and np.array_equal(k1[o1], k2[o2])
        assert np.array_equal(s1[o1].view(np.uint32), s2[o2].view(np.uint32))
        assert np.array_equal(O.best_prefix(logp[n], sigma).view(np.uint32), O.ref_best_prefix(logp[n], sigma).view(np.uint32))
