"""-m gpu: the CUDA path (through the C ABI, libipkgpu.so) against the oracle on the same seeded inputs."""
import numpy as np
import pytest

import oraclelib as O
from parity import assert_db_equal

pytestmark = pytest.mark.gpu


def _thr(omega, sigma, k):
    from ipk_b200 import capi as libipkgpu
    return libipkgpu.score_threshold(omega, sigma, k)


def _synth(n_taxa, W, sigma, seed=42, params=None):
    from ipk_b200 import synth
    return synth.posteriors(n_taxa, W, sigma, seed=seed, params=params)


def _rand_posteriors(rng, n_nodes, W, sigma, peaked=0.7, zeros=True):
    """independent columns; a share of exact zeros (-inf) and exactly uniform columns"""
    p = rng.dirichlet(np.ones(sigma) * 0.3, size=(n_nodes, W))
    sharp = rng.random((n_nodes, W)) < peaked
    top = rng.integers(0, sigma, size=(n_nodes, W))
    eps = 10.0 ** (-rng.uniform(1, 4, size=(n_nodes, W)))
    q = p * eps[..., None]
    np.put_along_axis(q, top[..., None], 0.0, axis=2)
    q = q / np.maximum(q.sum(axis=2, keepdims=True), 1e-300) * eps[..., None]
    np.put_along_axis(q, top[..., None], (1.0 - eps)[..., None], axis=2)
    p = np.where(sharp[..., None], q, p)
    p = np.round(p, 9)
    if zeros:
        uni = rng.random((n_nodes, W)) < 0.02
        p = np.where(uni[..., None], 1.0 / sigma, p)
    with np.errstate(divide="ignore"):
        return np.log10(p.astype(np.float32)).astype(np.float32)


@pytest.mark.parametrize("sigma,bits,W", [(4, 2, 257), (20, 5, 131), (4, 2, 9700)])
def test_best_prefix_bitexact(ctx, sigma, bits, W):
    rng = np.random.default_rng(1)
    logp = _rand_posteriors(rng, 5, W, sigma)
    nb = np.arange(5, dtype=np.uint32)
    B = ctx.best_prefix(logp, nb, W, sigma, bits, 3)
    for n in range(5):
        want = O.best_prefix(logp[n], sigma)
        assert np.array_equal(B[n].view(np.uint32), want.view(np.uint32))


ENUM_CASES = [
    # sigma, bits, k, omega, W, n_nodes
    (4, 2, 1, 1.5, 40, 3), (4, 2, 2, 1.5, 40, 3), (4, 2, 3, 1.5, 50, 3), (4, 2, 5, 1.5, 140, 4), (4, 2, 7, 2.0, 300, 4),
    (4, 2, 8, 1.5, 300, 4), (4, 2, 10, 1.5, 400, 3), (4, 2, 12, 1.5, 300, 2), (4, 2, 14, 2.0, 300, 2), (4, 2, 10, 1.5, 10, 2),
    (4, 2, 10, 1.5, 9, 2), (4, 2, 6, 1.5, 129 + 5, 2), (4, 2, 6, 1.5, 128 + 5, 2), (20, 5, 3, 1.5, 90, 3), (20, 5, 4, 10.0, 150, 3),
    (20, 5, 6, 1.5, 60, 2),
]


# enumeration strategies (include/ipkgpu.h: ipkgpu_set_enumeration): the list-join kernel with its default pool, the
# depth-first kernel alone, and the join kernel with a pool so small that many tiles are deferred to the depth-first one
ENUM_MODES = {"join": (0, 0), "dfs": (1, 0), "join_tiny_pool": (0, 1)}


@pytest.fixture(params=list(ENUM_MODES))
def mode(request, ctx):
    m, pool = ENUM_MODES[request.param]
    ctx.set_enumeration(m, pool)
    yield request.param
    ctx.set_enumeration(0, 0)


@pytest.mark.parametrize("sigma,bits,k,omega,W,n_nodes", ENUM_CASES)
def test_enumerate_per_window(ctx, mode, sigma, bits, k, omega, W, n_nodes):
    """every (node, window, k-mer, score) the reference's DCLA emits, bit-exact scores"""
    rng = np.random.default_rng(100 * k + sigma + W)
    logp = _rand_posteriors(rng, n_nodes, W, sigma, peaked=0.85 if k > 8 or sigma == 20 else 0.6)
    thr, lthr = _thr(omega, sigma, k)
    nb = np.arange(n_nodes, dtype=np.uint32)
    wn, ww, wk, ws = [], [], [], []
    for n in range(n_nodes):
        w, key, sc = O.enumerate_node(logp[n], sigma, bits, k, lthr)
        wn.append(np.full(len(w), n, np.uint32)); ww.append(w); wk.append(key); ws.append(sc)
    wn, ww, wk, ws = O.canonical_windows(np.concatenate(wn), np.concatenate(ww), np.concatenate(wk), np.concatenate(ws))
    cap = max(len(wk), 1)
    n_total, gn, gw, gk, gs = ctx.enumerate(logp, nb, W, sigma, bits, k, lthr, cap)
    assert n_total == len(wk)
    gn, gw, gk, gs = O.canonical_windows(gn, gw, gk, gs)
    assert np.array_equal(gn, wn) and np.array_equal(gw, ww) and np.array_equal(gk, wk)
    assert np.array_equal(gs.view(np.uint32), ws.view(np.uint32))


def test_enumerate_cap_overflow_counts_everything(ctx):
    rng = np.random.default_rng(5)
    logp = _rand_posteriors(rng, 3, 200, 4, peaked=0.5)
    thr, lthr = _thr(1.5, 4, 8)
    nb = np.arange(3, dtype=np.uint32)
    n_total, *_ = ctx.enumerate(logp, nb, 200, 4, 2, 8, lthr, 1 << 20)
    n2, gn, gw, gk, gs = ctx.enumerate(logp, nb, 200, 4, 2, 8, lthr, 100)
    assert n2 == n_total and len(gk) == 100


BUILD_CASES = [
    # name, n_taxa, W, sigma, bits, k, omega
    ("dna_k7_o2", 24, 220, 4, 2, 7, 2.0), ("dna_k8", 24, 220, 4, 2, 8, 1.5), ("dna_k10", 16, 200, 4, 2, 10, 1.5),
    ("dna_k12", 8, 150, 4, 2, 12, 1.5), ("aa_k4_o10", 12, 120, 20, 5, 4, 10.0), ("aa_k6", 4, 60, 20, 5, 6, 1.5),
    ("dna_k14_o2", 6, 400, 4, 2, 14, 2.0),
]


@pytest.mark.parametrize("name,n_taxa,W,sigma,bits,k,omega", BUILD_CASES)
@pytest.mark.parametrize("filt", [0, 1])
def test_build_matches_oracle(ctx, mode, name, n_taxa, W, sigma, bits, k, omega, filt):
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    want = O.build(logp, nb, W, sigma, bits, k, lthr, thr, T, filt, 8)
    got = ctx.build(logp, nb, W, sigma, bits, k, lthr, thr, T, filt)
    assert got.stats["n_emitted"] == want.n_emitted
    assert_db_equal(got, want, exact_fv=(filt == 1))
    if mode == "dfs":
        assert got.stats["n_deferred_tiles"] == got.stats["n_tiles"] > 0
    elif mode == "join" and sigma == 4:
        assert got.stats["n_deferred_tiles"] < got.stats["n_tiles"]       # the fast path really ran


@pytest.mark.parametrize("name,n_taxa,W,sigma,bits,k,omega", [BUILD_CASES[1], BUILD_CASES[4]])
@pytest.mark.parametrize("result_mode", ["host", "device_indexed"])
def test_result_digest_is_the_numpy_digest_of_the_database(ctx, name, n_taxa, W, sigma, bits, k, omega, result_mode):
    """ipkgpu_result_digest (digest.cuh) == the same definition in numpy (parity.db_digest) over the database the build
    handed back, and its entry / k-mer parts == the digest of the ORACLE's database: the order-independent check bench.py
    uses where two builds never share a memory (device-resident vs host-bound, 1 GPU vs G shards)"""
    from ipk_b200 import capi as libipkgpu
    from parity import db_digest
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    got = ctx.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    if result_mode == "device_indexed":
        inp, keep = libipkgpu.make_input(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
        res = ctx.build_raw(inp, result_on_device=True, indexed=True)
        assert int(res.n_entries) == got.n_entries
    d = ctx.result_digest(lthr, 1e-5)
    assert d == db_digest(got, lthr, 1e-5)
    want = db_digest(O.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0, 4), lthr, 1e-5)
    for f in ("entries_sum", "entries_xor", "kmers_sum", "kmers_xor", "n_boundary_above", "n_kmers", "n_entries"):
        assert d[f] == want[f], f
    # one flipped score bit must show
    got.score.view(np.uint32)[got.n_entries // 2] ^= 1
    assert db_digest(got, lthr, 1e-5)["entries_xor"] != d["entries_xor"]


def _threshold_columns(rng, n_nodes, W, k, omega):
    """posteriors in which many k-mers score within rounding of the threshold: runs of columns whose two best states have
    exactly p = omega / 4 each (a k-mer over k such columns scores k log10(omega / 4) = log10(threshold) up to fp32 rounding),
    between ordinary peaked columns"""
    logp = _rand_posteriors(rng, n_nodes, W, 4, peaked=0.8, zeros=False)
    a = np.float32(np.log10(np.float32(omega / 4.0)))
    b = np.float32(np.log10(np.float32((1.0 - 2 * omega / 4.0) / 2.0)))
    for n in range(n_nodes):
        for start in rng.integers(0, W - k - 3, 3):
            for c in range(start, start + k + 2):
                perm = rng.permutation(4)
                logp[n, c, perm[:2]] = a
                logp[n, c, perm[2:]] = b
    return logp


@pytest.mark.parametrize("k", [5, 8])
def test_boundary_list_matches_oracle(ctx, k):
    """ipkgpu_boundary_list: the entries whose presence or score flips when the threshold moves by less than the band (the
    boundary cases BASELINE.json sets aside), against the same set taken from two ORACLE builds at threshold -/+ band"""
    from ipk_b200 import capi as libipkgpu
    from parity import by_key
    rng = np.random.default_rng(70 + k)
    n_taxa, W, omega, band = 5, 90, 1.5, 1e-5
    n_nodes = 2 * (2 * n_taxa - 2)
    logp = _threshold_columns(rng, n_nodes, W, k, omega)
    nb = np.repeat(np.arange(n_nodes // 2, dtype=np.uint32), 2)
    thr, lthr = _thr(omega, 4, k)
    T = 2 * n_taxa - 1
    inp, keep = libipkgpu.make_input(logp, nb, W, 4, 2, k, lthr, thr, T, 0)
    n, kmer, branch, s_lo, s_hi = ctx.boundary_list(inp, band)
    loose = O.build(logp, nb, W, 4, 2, k, np.float32(lthr) - np.float32(band), thr, T, 0, 4)
    tight = O.build(logp, nb, W, 4, 2, k, np.float32(lthr) + np.float32(band), thr, T, 0, 4)
    _, _, _, lk, lb, ls = by_key(loose)
    _, _, _, tk, tb, ts = by_key(tight)
    have = {(int(a), int(b_)): np.float32(c) for a, b_, c in zip(tk, tb, ts)}
    want = sorted((int(a), int(b_), np.float32(c), have.get((int(a), int(b_)))) for a, b_, c in zip(lk, lb, ls)
                  if have.get((int(a), int(b_))) is None or have[(int(a), int(b_))].view(np.uint32) != np.float32(c).view(np.uint32))
    assert n == len(want) and n > 20, (n, len(want))          # the fixture does straddle the threshold
    assert [(int(a), int(b_)) for a, b_ in zip(kmer, branch)] == [(w[0], w[1]) for w in want]
    assert np.array_equal(s_lo.view(np.uint32), np.array([w[2] for w in want], np.float32).view(np.uint32))
    absent = np.array([w[3] is None for w in want])
    assert np.array_equal(np.isnan(s_hi), absent) and absent.any()
    assert np.array_equal(s_hi[~absent].view(np.uint32), np.array([w[3] for w in want if w[3] is not None], np.float32).view(np.uint32))
    # every listed score lies in the band around the threshold, on the emitted side of the loose build
    assert np.all(np.abs(s_lo[absent] - np.float32(lthr)) <= np.float32(2 * band))
    # the build itself is untouched by the analysis
    got = ctx.build(logp, nb, W, 4, 2, k, lthr, thr, T, 0)
    assert_db_equal(got, O.build(logp, nb, W, 4, 2, k, lthr, thr, T, 0, 4))
    assert ctx.result_digest(lthr, band)["n_boundary_above"] >= int(absent.sum()) - n      # (a lower bound; just exercises the counter)


def _positions_oracle(logp, nb, sigma, bits, k, lthr):
    """ipk::put under KEEP_POSITIONS (ipk/src/branch_group.cpp:72-86; db_builder.cpp:642-690) restated over the oracle's
    per-window emission: a group's ghost nodes one after the other, windows ascending, `if (it->second.score < kmer.score)`
    -- the FIRST window that reaches the maximum keeps its position.  -> {(key, branch): (score bits, position)}"""
    out = {}
    order = []
    for b in nb:
        if int(b) not in order:
            order.append(int(b))
    for b in order:
        best = {}
        for n in np.where(nb == b)[0]:
            w, key, sc = O.enumerate_node(logp[n], sigma, bits, k, lthr, cap=1 << 20)
            o = np.argsort(w, kind="stable")
            for wi, ki, si in zip(w[o], key[o], sc[o]):
                cur = best.get(int(ki))
                if cur is None or cur[0] < si:
                    best[int(ki)] = (np.float32(si), int(wi))
        for ki, (si, wi) in best.items():
            out[(ki, b)] = (int(np.float32(si).view(np.uint32)), wi)
    return out


@pytest.mark.parametrize("case", ["dna_ties", "aa"])
@pytest.mark.parametrize("indexed", [False, True])
def test_keep_positions_build(case, indexed):
    """ipkgpu_set_keep_positions (the reference's ipk-aa-pos variant): same database as the plain build, and every entry's
    position is the first window (in processing order) that reached its score"""
    from ipk_b200 import capi as libipkgpu
    if case == "dna_ties":
        sigma, bits, k, omega, W, n_taxa = 4, 2, 5, 1.5, 90, 5
        rng = np.random.default_rng(12)
        n_nodes = 2 * (2 * n_taxa - 2)
        logp = _threshold_columns(rng, n_nodes, W, k, omega)        # runs of identical column values: equal scores at several windows
        nb = np.repeat(np.arange(n_nodes // 2, dtype=np.uint32), 2)
        nb[[2, 3]] = nb[[3, 2]]                                      # ... and interleaved groups
    else:
        sigma, bits, k, omega, W, n_taxa = 20, 5, 3, 1.5, 50, 4
        logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    want_db = O.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0, 4)
    want = _positions_oracle(logp, nb, sigma, bits, k, lthr)
    assert len(want) == want_db.n_entries
    with libipkgpu.Context(0) as c:
        c.set_keep_positions(True)
        inp, keep = libipkgpu.make_input(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
        res = c.build_raw(inp, result_on_device=False, indexed=indexed)
        pos = c.result_positions(int(res.n_entries))
        got = libipkgpu.Context.to_db(res)
        assert pos is not None and len(pos) == got.n_entries
        if indexed:      # to_db gathered the entries into filter-value order; do the same to the positions
            nk = int(res.n_kmers)
            start = libipkgpu._view(res.offsets, nk, np.uint64).astype(np.int64)
            cnt = libipkgpu._view(res.counts, nk, np.uint32).astype(np.int64)
            pos = pos[np.repeat(start - got.offsets[:-1].astype(np.int64), cnt) + np.arange(int(cnt.sum()))]
        assert_db_equal(got, want_db)
        keys = np.repeat(got.keys, np.diff(got.offsets.astype(np.int64)))
        have = {(int(a), int(b)): (int(s), int(p)) for a, b, s, p in zip(keys, got.branch, got.score.view(np.uint32), pos)}
        assert have == want
        if case == "dna_ties":
            # the fixture does contain ties that a "last maximum" or "any maximum" rule would resolve differently
            assert sum(1 for v in want.values() if v[1] > 0) > 10
        c.set_keep_positions(False)
        plain = c.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
        assert c.result_positions(plain.n_entries) is None
        assert_db_equal(plain, want_db)


def test_build_interleaved_groups_and_singletons(ctx):
    """ghost nodes of one branch need not be adjacent; a branch may have one ghost only (--inner-only/--outer-only)"""
    logp, nb = _synth(10, 150, 4)
    rng = np.random.default_rng(3)
    perm = rng.permutation(len(nb))[: len(nb) - 5]
    logp2 = np.ascontiguousarray(logp[perm]); nb2 = nb[perm]
    thr, lthr = _thr(1.5, 4, 8)
    want = O.build(logp2, nb2, 150, 4, 2, 8, lthr, thr, 19, 0, 4)
    got = ctx.build(logp2, nb2, 150, 4, 2, 8, lthr, thr, 19, 0)
    assert_db_equal(got, want)


def test_build_empty_and_short(ctx):
    thr, lthr = _thr(1.5, 4, 8)
    z = ctx.build(np.zeros((0, 10, 4), np.float32), np.zeros(0, np.uint32), 10, 4, 2, 8, lthr, thr, 1, 0)
    assert z.n_kmers == 0 and z.n_entries == 0
    logp, nb = _synth(4, 7, 4)      # W < k: no windows
    z = ctx.build(logp, nb, 7, 4, 2, 8, lthr, thr, 7, 0)
    assert z.n_kmers == 0 and z.stats["n_windows"] == 0
    # nothing clears an impossible threshold
    logp, nb = _synth(4, 50, 4)
    z = ctx.build(logp, nb, 50, 4, 2, 8, 0.5, 3.0, 7, 0)
    assert z.n_kmers == 0 and z.stats["n_emitted"] == 0


def test_build_batched_equals_single(ctx):
    """a tiny workspace forces the overflow -> count -> batch path; the database must not change"""
    from ipk_b200 import capi as libipkgpu
    logp, nb = _synth(24, 200, 4)
    thr, lthr = _thr(1.5, 4, 8)
    want = ctx.build(logp, nb, 200, 4, 2, 8, lthr, thr, 47, 0)
    c2 = libipkgpu.Context(0)
    try:
        c2.set_workspace_limit(24 * 2 * (1 << 16))      # batch capacity = 65536 records
        got = c2.build(logp, nb, 200, 4, 2, 8, lthr, thr, 47, 0)
        assert got.stats["enum_launches"] > 2
        assert_db_equal(got, want, exact_fv=True)
    finally:
        c2.close()


@pytest.mark.parametrize("raw", [True, False])
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("sigma,bits,k,n_taxa,W", [(4, 2, 8, 20, 200), (20, 5, 4, 8, 90)])
def test_sharded_stages_equal_single(world, sigma, bits, k, n_taxa, W, raw):
    """The multi-GPU path, its ranks emulated one after the other on ONE device (B200_PROFILING.md: never several
    waiting ranks on one GPU): stage 1 per rank -> partition by k-mer owner -> the all-to-all done with device copies
    -> stage 2 per rank -> G-way merge.  Entry-for-entry equal to the single-GPU build."""
    import torch
    from types import SimpleNamespace
    from ipk_b200 import capi as libipkgpu, dist as ipkdist
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    dev = torch.device("cuda", 0)
    with libipkgpu.Context(0) as c1:
        want = c1.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    parts = ipkdist.split_branches(2 * n_taxa - 2, world)
    group_branch = ipkdist.local_groups(nb)
    group_bits = max(int(np.ceil(np.log2(len(group_branch)))), 0)
    ctxs = [libipkgpu.Context(0) for _ in range(world)]
    try:
        send = []
        for r, (b0, b1) in enumerate(parts):
            sl = slice(2 * b0, 2 * b1)
            inp, keep = libipkgpu.make_input(np.ascontiguousarray(logp[sl]), nb[sl], W, sigma, bits, k, lthr, thr, T, 0)
            n_unique, st1 = ctxs[r].stage1(inp, b0, group_bits, raw)
            assert n_unique == st1["n_emitted"] if raw else n_unique <= st1["n_emitted"]
            keys = torch.empty(max(n_unique, 1), dtype=torch.int64, device=dev)
            scores = torch.empty(max(n_unique, 1), dtype=torch.float32, device=dev)
            counts = ctxs[r].stage1_partition(world, keys, scores).astype(np.int64)
            assert counts.sum() == n_unique
            send.append((keys, scores, np.concatenate(([0], np.cumsum(counts)))))
        shards = []
        for r in range(world):
            rk = torch.cat([send[s][0][send[s][2][r]:send[s][2][r + 1]] for s in range(world)])      # source-rank order
            rs = torch.cat([send[s][1][send[s][2][r]:send[s][2][r + 1]] for s in range(world)])
            n = int(rk.numel())
            if n == 0:
                rk = torch.empty(1, dtype=torch.int64, device=dev); rs = torch.empty(1, dtype=torch.float32, device=dev)
            torch.cuda.synchronize()
            res = ctxs[r].stage2(rk, rs, n, group_bits, group_branch, k * bits, thr, T, 0, not raw, False)
            shards.append(libipkgpu.Context.to_db(res))
        assert sum(s.n_entries for s in shards) == want.n_entries
        merged = ipkdist.merge_shards(shards)
        got = SimpleNamespace(keys=merged[0], fv=merged[1], offsets=merged[2], branch=merged[3], score=merged[4],
                              n_kmers=len(merged[0]), n_entries=len(merged[3]))
        assert_db_equal(got, want, exact_fv=True)
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("world", [2, 3, 5])
@pytest.mark.parametrize("k,n_taxa,W,enum_mode", [(8, 20, 200, 0), (10, 12, 150, 0), (4, 16, 120, 0), (8, 20, 200, 1)])
@pytest.mark.parametrize("streamed,filt", [(False, 0), (True, 0), (False, 1)])
def test_packed_exchange_equals_single(world, k, n_taxa, W, enum_mode, streamed, filt, monkeypatch):
    """The exchange as the MSD pass of the merge sort (ipkgpu_stage1_bucket_counts / ipkgpu_exchange_plan / recv_packed /
    stage1_scatter_packed / stage2_packed), ranks emulated as contexts of ONE process on one device so that the "peer"
    buffers are plain device pointers (the IPC mapping is exercised by bench.py --gpus N, whose database digest must equal
    the single-GPU one).  Entry-for-entry equal to the single-GPU build, filter values bit for bit -- also under the random
    filter, whose draw is a function of the k-mer.  enum_mode 1 (depth-first kernel: no runs) takes the unordered path."""
    from types import SimpleNamespace
    from ipk_b200 import capi as libipkgpu, dist as ipkdist, hostlib
    if streamed:
        monkeypatch.setenv("IPKGPU_PART_MIN", "1")
        monkeypatch.setenv("IPKGPU_PARTS", "3")
    sigma, bits, omega = 4, 2, 1.5
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    with libipkgpu.Context(0) as c1:
        want = c1.build(logp, nb, W, sigma, bits, k, lthr, thr, T, filt)
    parts = ipkdist.split_branches(2 * n_taxa - 2, world)
    group_branch = ipkdist.local_groups(nb)
    group_bits = max(int(np.ceil(np.log2(len(group_branch)))), 0)
    assert ipkdist.packed_eligible(k * bits, group_bits)
    ctxs = [libipkgpu.Context(0) for _ in range(world)]
    try:
        matrix = np.zeros((world, 256), np.uint64)
        for r, (b0, b1) in enumerate(parts):
            ctxs[r].set_enumeration(enum_mode, 0)
            sl = slice(2 * b0, 2 * b1)
            inp, keep = libipkgpu.make_input(np.ascontiguousarray(logp[sl]), nb[sl], W, sigma, bits, k, lthr, thr, T, filt)
            n1, _ = ctxs[r].stage1(inp, b0, group_bits, True)
            matrix[r] = ctxs[r].stage1_bucket_counts()
            assert int(matrix[r].sum()) == n1
        plans = [libipkgpu.exchange_plan(world, r, matrix) for r in range(world)]
        assert all(np.array_equal(plans[0][0], p[0]) for p in plans) and np.all(np.diff(plans[0][0].astype(int)) >= 0)     # owners: the same contiguous ranges everywhere
        assert sum(p[3] for p in plans) == int(matrix.sum())
        bufs = [ctxs[r].recv_packed(max(plans[r][3], 1)) for r in range(world)]
        ordered = all([ctxs[r].stage1_scatter_packed([b[0] for b in bufs], plans[r][0], plans[r][1]) for r in range(world)])
        assert ordered == (enum_mode == 0)
        shards = []
        for r in range(world):
            res = ctxs[r].stage2_packed(plans[r][3], plans[r][2], group_bits, ordered, group_branch, k * bits, thr, T, filt, False,
                                        indexed=streamed, branch16=streamed)
            assert bool(res.branch16) == (streamed and plans[r][3] > 0)
            if ordered and plans[r][3]:
                assert res.stats.sort_passes == (k * bits - 8 + 7) // 8           # the owner sorts the k-mer bits below the bucket byte, nothing else
            shards.append(libipkgpu.Context.to_db(res))
            d = ctxs[r].result_digest(lthr)
            assert d["n_entries"] == shards[-1].n_entries
        assert sum(s.n_entries for s in shards) == want.n_entries
        # key RANGES: every k-mer of shard r lies below every k-mer of shard r + 1
        tops = [(int(s.keys.min()), int(s.keys.max())) for s in shards if s.n_kmers]
        assert all(a[1] < b[0] for a, b in zip(tops, tops[1:]))
        merged = hostlib.merge_shards(shards)
        got = SimpleNamespace(keys=merged[0], fv=merged[1], offsets=merged[2], branch=merged[3], score=merged[4],
                              n_kmers=len(merged[0]), n_entries=len(merged[3]))
        assert_db_equal(got, want, exact_fv=True)
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("raw", [True, False])
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("sigma,bits,k,n_taxa,W", [(4, 2, 8, 20, 200), (20, 5, 4, 8, 90)])
@pytest.mark.parametrize("streamed", [False, True])
def test_peer_scatter_equals_single(world, sigma, bits, k, n_taxa, W, raw, streamed, monkeypatch):
    """The fused partition + exchange (ipkgpu_stage1_owner_counts / recv_buffers / stage1_scatter), ranks emulated as
    contexts of ONE process on one device, so the "peer" buffers are plain device pointers (the IPC mapping itself is
    exercised by bench.py --gpus 2).  Entry-for-entry equal to the single-GPU build.  streamed: the owner's stage 2 hands
    back an indexed host shard with 16-bit branch ids, sorted / merged / copied in k-mer ranges (forced at this size)."""
    from types import SimpleNamespace
    from ipk_b200 import capi as libipkgpu, dist as ipkdist
    if streamed and not raw:
        pytest.skip("k-mer ranges apply to the single sort of raw records")
    if streamed:
        monkeypatch.setenv("IPKGPU_PART_MIN", "1")
        monkeypatch.setenv("IPKGPU_PARTS", "3")
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    with libipkgpu.Context(0) as c1:
        want = c1.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    parts = ipkdist.split_branches(2 * n_taxa - 2, world)
    group_branch = ipkdist.local_groups(nb)
    group_bits = max(int(np.ceil(np.log2(len(group_branch)))), 0)
    ctxs = [libipkgpu.Context(0) for _ in range(world)]
    try:
        matrix = np.zeros((world, world), np.int64)
        for r, (b0, b1) in enumerate(parts):
            sl = slice(2 * b0, 2 * b1)
            inp, keep = libipkgpu.make_input(np.ascontiguousarray(logp[sl]), nb[sl], W, sigma, bits, k, lthr, thr, T, 0)
            n1, _ = ctxs[r].stage1(inp, b0, group_bits, raw)
            matrix[r] = ctxs[r].stage1_owner_counts(world).astype(np.int64)
            assert matrix[r].sum() == n1
        bufs = [ctxs[r].recv_buffers(max(int(matrix[:, r].sum()), 1)) for r in range(world)]
        assert all(len(b[2]) == 64 and len(b[3]) == 64 and b[2] != b[3] for b in bufs)
        for r in range(world):
            ctxs[r].stage1_scatter([b[0] for b in bufs], [b[1] for b in bufs], [int(matrix[:r, d].sum()) for d in range(world)])
        shards = []
        for r in range(world):
            res = ctxs[r].stage2(bufs[r][0], bufs[r][1], int(matrix[:, r].sum()), group_bits, group_branch, k * bits, thr, T, 0, not raw, False,
                                 indexed=streamed, branch16=streamed)
            assert bool(res.branch16) == streamed
            shards.append(libipkgpu.Context.to_db(res))
        assert sum(s.n_entries for s in shards) == want.n_entries
        assert all(s.n_entries > 0 for s in shards)
        merged = ipkdist.merge_shards(shards)
        got = SimpleNamespace(keys=merged[0], fv=merged[1], offsets=merged[2], branch=merged[3], score=merged[4],
                              n_kmers=len(merged[0]), n_entries=len(merged[3]))
        assert_db_equal(got, want, exact_fv=True)
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("pool_cap", [0, 700])
@pytest.mark.parametrize("sigma,bits,k,n_taxa,W", [(4, 2, 10, 12, 400), (20, 5, 4, 6, 150)])
def test_ordered_emission_equals_default(sigma, bits, k, n_taxa, W, pool_cap, monkeypatch):
    """IPKGPU_ORDERED=1: the join kernel's first level emits in tile (= branch-group) order through a decoupled look-back,
    the main stream is sorted on the k-mer bits only, and the tiles that level deferred (forced here by a small list
    pool) arrive through a side stream that the max-merge kernel merges in.  Same database, bit for bit."""
    from ipk_b200 import capi as libipkgpu
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    with libipkgpu.Context(0) as c1:
        want = c1.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    monkeypatch.setenv("IPKGPU_ORDERED", "1")
    with libipkgpu.Context(0) as c2:
        if pool_cap:
            c2.set_enumeration(0, pool_cap)
        got = c2.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
        if pool_cap and sigma == 4:
            assert got.stats["n_deferred_tiles"] > 0, "the small pool should have deferred some tiles to the side stream"
        assert got.stats["sort_passes"] <= want.stats["sort_passes"]
    assert want.n_entries > 1000
    assert_db_equal(got, want, exact_fv=True)


@pytest.mark.parametrize("shuffled", [False, True])
@pytest.mark.parametrize("sigma,bits,k,n_taxa,W", [(4, 2, 8, 20, 300), (20, 5, 3, 6, 120)])
def test_chunked_upload_equals_single_copy(sigma, bits, k, n_taxa, W, shuffled, monkeypatch):
    """Host posteriors above IPKGPU_H2D_CHUNK_MIN bytes go up in node chunks on the copy stream, with the best-prefix and
    enumeration kernels of a chunk running under the next upload (forced here at test size).  Same database; nodes that do
    not arrive grouped by branch keep the single copy."""
    from ipk_b200 import capi as libipkgpu
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    if shuffled:
        perm = np.random.default_rng(5).permutation(len(nb))
        logp, nb = np.ascontiguousarray(logp[perm]), np.ascontiguousarray(nb[perm])
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    with libipkgpu.Context(0) as c:
        want = c.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    monkeypatch.setenv("IPKGPU_H2D_CHUNK_MIN", "1")
    monkeypatch.setenv("IPKGPU_H2D_CHUNKS", "3")
    with libipkgpu.Context(0) as c:
        for _ in range(2):
            got = c.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
            if shuffled:
                assert got.stats["kernel_launches"] == want.stats["kernel_launches"]
            else:
                assert got.stats["kernel_launches"] > want.stats["kernel_launches"]
            assert got.stats["n_emitted"] == want.stats["n_emitted"]
            assert_db_equal(got, want, exact_fv=True)
    assert want.n_entries > 500


def test_pinned_loader_buffer_chunked_upload(monkeypatch):
    """ipkgpu_host_alloc: posteriors loaded into the library's page-locked buffer upload asynchronously in node chunks
    (forced at this size); same database as from a pageable numpy array, and the buffer can be reused and freed."""
    from ipk_b200 import capi as libipkgpu
    sigma, bits, k, n_taxa, W = 4, 2, 8, 24, 300
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(1.5, sigma, k)
    T = 2 * n_taxa - 1
    with libipkgpu.Context(0) as c:
        want = c.build(logp, nb, W, sigma, bits, k, lthr, thr, T, 0)
    monkeypatch.setenv("IPKGPU_H2D_CHUNK_MIN", "1")
    with libipkgpu.PinnedArray(logp.shape) as pin, libipkgpu.Context(0) as c:
        for seed_shift in (0, 1):
            pin.array[...] = logp if seed_shift == 0 else np.roll(logp, 1, axis=0)
            nbs = nb if seed_shift == 0 else np.roll(nb, 1)
            got = c.build(pin.array, nbs, W, sigma, bits, k, lthr, thr, T, 0)
            if seed_shift == 0:
                assert got.stats["kernel_launches"] > want.stats["kernel_launches"]          # chunked
                assert_db_equal(got, want, exact_fv=True)
            else:
                assert got.n_entries == want.n_entries                    # same nodes in another order: same database size
    assert want.n_entries > 500


@pytest.mark.parametrize("parts", [2, 5, 300])
@pytest.mark.parametrize("sigma,bits,k,n_taxa,W,filt", [(4, 2, 10, 12, 400, 0), (4, 2, 4, 9, 60, 0), (4, 2, 6, 40, 90, 0), (20, 5, 4, 6, 150, 1)])
def test_streamed_parts_equal_single_sort(sigma, bits, k, n_taxa, W, filt, parts, monkeypatch):
    """Host-bound indexed results above IPKGPU_PART_MIN records are sorted and max-merged in k-mer ranges (one MSD radix pass,
    then per range the remaining LSD passes + merge appended to the CSR), each range's entries copied to the host while the
    next is sorted.  Forced here at test sizes; same database as the single sort, bit for bit, with and without 16-bit
    branch ids, with more requested parts than there are digit ranges, and with keys of 13 .. 25 bits (sort units of 1 .. 8
    digit ranges)."""
    from ipk_b200 import capi as libipkgpu
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    inp, keep = libipkgpu.make_input(logp, nb, W, sigma, bits, k, lthr, thr, T, filt)
    with libipkgpu.Context(0) as c:
        want = libipkgpu.Context.to_db(c.build_raw(inp, False, indexed=True))
    monkeypatch.setenv("IPKGPU_PART_MIN", "1")
    monkeypatch.setenv("IPKGPU_PARTS", str(parts))
    with libipkgpu.Context(0) as c:
        for b16 in (False, True, False):                 # the third build reuses every buffer of the first two
            res = c.build_raw(inp, False, indexed=True, branch16=b16)
            assert bool(res.branch16) == b16
            got = libipkgpu.Context.to_db(res)
            assert got.stats["kernel_launches"] > want.stats["kernel_launches"]        # it did go through the parts
            assert got.stats["sort_passes"] == want.stats["sort_passes"]
            assert_db_equal(got, want, exact_fv=True)
        # contiguous and device-resident results keep the single sort
        res = c.build_raw(inp, True, indexed=True)
        assert res.stats.kernel_launches == want.stats["kernel_launches"]
    assert want.n_entries > 500


@pytest.mark.parametrize("sigma,bits,k,n_taxa,W,filt", [(4, 2, 8, 20, 300, 0), (20, 5, 3, 6, 120, 1)])
def test_indexed_result_equals_contiguous(sigma, bits, k, n_taxa, W, filt):
    """IPKGPU_RESULT_HOST_INDEXED / _DEVICE_INDEXED: no gather kernel, the entries stay in k-mer key order and every k-mer
    (in filter-value order) points at its own -- walking it yields the contiguous result, bit for bit."""
    import torch
    from ipk_b200 import capi as libipkgpu
    omega = 1.5 if sigma == 4 else 10.0
    logp, nb = _synth(n_taxa, W, sigma)
    thr, lthr = _thr(omega, sigma, k)
    T = 2 * n_taxa - 1
    inp, keep = libipkgpu.make_input(logp, nb, W, sigma, bits, k, lthr, thr, T, filt)
    with libipkgpu.Context(0) as c:
        want = libipkgpu.Context.to_db(c.build_raw(inp, False))
        res = c.build_raw(inp, False, indexed=True)
        assert res.counts and not res.on_device
        got = libipkgpu.Context.to_db(res)
        launches = got.stats["kernel_launches"]
        assert launches < want.stats["kernel_launches"]                    # the scan and the gather are gone
        assert_db_equal(got, want, exact_fv=True)
        # 16-bit branch ids (2 of the 8 bytes per entry that cross PCIe)
        res = c.build_raw(inp, False, indexed=True, branch16=True)
        assert res.branch16 and not res.branch
        assert res.stats.d2h_bytes < got.stats["d2h_bytes"]
        assert_db_equal(libipkgpu.Context.to_db(res), want, exact_fv=True)
        # the device-resident flavour: same index arrays
        res = c.build_raw(inp, True, indexed=True)
        assert res.on_device and res.counts
        nk = int(res.n_kmers)
        torch.cuda.synchronize()
        import ctypes as C
        cnt = np.empty(nk, np.uint32); off = np.empty(nk + 1, np.uint64)
        cudart = C.CDLL("libcudart.so")
        assert cudart.cudaMemcpy(C.c_void_p(cnt.ctypes.data), C.c_void_p(res.counts), C.c_size_t(nk * 4), 2) == 0
        assert cudart.cudaMemcpy(C.c_void_p(off.ctypes.data), C.c_void_p(res.offsets), C.c_size_t((nk + 1) * 8), 2) == 0
        assert np.array_equal(cnt, np.diff(want.offsets).astype(np.uint32))
        assert int(off[-1]) == want.n_entries and int(cnt.sum()) == want.n_entries
    assert want.n_entries > 1000


def test_branch16_falls_back_when_ids_do_not_fit():
    """IPKGPU_RESULT_BRANCH16 is a request: with a branch id >= 65,536 the result keeps 32-bit ids"""
    from ipk_b200 import capi as libipkgpu
    logp, nb = _synth(6, 80, 4)
    nb = (nb.astype(np.uint64) + 70000).astype(np.uint32)
    thr, lthr = _thr(1.5, 4, 8)
    inp, keep = libipkgpu.make_input(logp, nb, 80, 4, 2, 8, lthr, thr, 11, 0)
    with libipkgpu.Context(0) as c:
        want = libipkgpu.Context.to_db(c.build_raw(inp, False))
        res = c.build_raw(inp, False, indexed=True, branch16=True)
        assert not res.branch16 and res.branch and res.counts
        got = libipkgpu.Context.to_db(res)
    assert int(got.branch.min()) >= 70000
    assert_db_equal(got, want, exact_fv=True)


def test_invalid_arguments(ctx):
    from ipk_b200.capi import IpkGpuError
    logp, nb = _synth(4, 50, 4)
    with pytest.raises(IpkGpuError):
        ctx.build(logp, nb, 50, 4, 2, 33, -1.0, 0.1, 7, 0)        # k*bits > 63
    with pytest.raises(IpkGpuError):
        ctx.build(logp, nb, 50, 4, 1, 8, -1.0, 0.1, 7, 0)         # 1 bit cannot encode 4 states
    with pytest.raises(IpkGpuError):
        ctx.build(logp, nb, 50, 4, 2, 8, -1.0, 0.1, 7, 7)         # unknown filter


import glob
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "enum_*.npz"))), ids=lambda p: os.path.basename(p)[:-4])
def test_enumerate_matches_reference_fixture(ctx, mode, path):
    """the committed fixtures were produced by the reference's own object code (tests/golden/make_golden.py)"""
    import hashlib
    f = np.load(path)
    logp, sigma, bits, k, W = f["logp"], int(f["sigma"]), int(f["bits"]), int(f["k"]), int(f["W"])
    nb = np.arange(logp.shape[0], dtype=np.uint32)
    n_total, gn, gw, gk, gs = ctx.enumerate(logp, nb, W, sigma, bits, k, float(f["log_threshold"]), max(int(f["n_records"]), 1))
    assert n_total == int(f["n_records"])
    gn, gw, gk, gs = O.canonical_windows(gn, gw, gk, gs)
    h = hashlib.sha256()
    h.update(gn.astype("<u4").tobytes()); h.update(gw.astype("<u4").tobytes())
    h.update(gk.astype("<u8").tobytes()); h.update(gs.view(np.uint32).astype("<u4").tobytes())
    assert h.hexdigest() == str(f["sha256"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "build_*.npz"))), ids=lambda p: os.path.basename(p)[:-4])
def test_build_matches_reference_fixture(ctx, mode, path):
    f = np.load(path)
    got = ctx.build(f["logp"], f["node_branch"], int(f["W"]), int(f["sigma"]), int(f["bits"]), int(f["k"]), float(f["log_threshold"]),
                    float(f["threshold"]), int(f["total_num_groups"]), 0)
    assert got.stats["n_emitted"] == int(f["n_emitted"])
    ek, eb, es = O.db_entries(got)
    assert np.array_equal(ek, f["entry_key"]) and np.array_equal(eb, f["entry_branch"])
    assert np.array_equal(es.view(np.uint32), f["entry_score"].view(np.uint32))
    o = np.argsort(got.keys, kind="stable")
    np.testing.assert_allclose(got.fv[o], f["kmer_fv"], rtol=2e-5, atol=1e-12)
    assert np.array_equal(got.offsets[-1:], f["offsets"][-1:])


@pytest.mark.parametrize("name", ["cfg2_dna_1k", "cfg3_16s_10k", "cfg5_viral", "cfg4_aa_2k", "cfg1_d652_k7"])
def test_workload_sample_matches_oracle(ctx, mode, name):
    """a few branch groups of each BASELINE workload at its real width (calibrated, sharply peaked posteriors)"""
    from ipk_b200 import synth
    w = synth.WORKLOADS[name]
    nbr = 3 if w.width < 5000 else 1
    logp, nb = synth.posteriors(w.n_taxa, w.width, w.sigma, seed=42, branch_begin=50, branch_end=50 + nbr)
    thr, lthr = _thr(w.omega, w.sigma, w.k)
    want = O.build(logp, nb, w.width, w.sigma, w.bits, w.k, lthr, thr, w.n_tree_nodes, 0, 8)
    got = ctx.build(logp, nb, w.width, w.sigma, w.bits, w.k, lthr, thr, w.n_tree_nodes, 0)
    assert got.stats["n_emitted"] == want.n_emitted
    assert_db_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("key_ranges", [True, False])
@pytest.mark.parametrize("n_shards", [1, 3, 8])
def test_device_merge_order_equals_host_merge_order(ctx, key_ranges, n_shards):
    """ipkgpu_merge_order (one stable device sort of the filter values) against ipkhost_merge_order (the N-way merge of
    merge_stage2, db_builder.cpp:392-458, on the host cores): same (shard, index) per merged k-mer, ties and signed zeros included"""
    from types import SimpleNamespace
    from ipk_b200 import hostlib
    rng = np.random.default_rng(11 + n_shards)
    per = 5000
    keys_all = rng.permutation(1 << 22)[: n_shards * per].astype(np.uint64)
    if key_ranges:
        keys_all = np.sort(keys_all)                      # shard s = the s-th contiguous key range (k-mer-range owners)
    shards = []
    for s in range(n_shards):
        keys = keys_all[s * per:(s + 1) * per]
        fv = rng.normal(size=per).astype(np.float32)
        fv[rng.integers(0, per, per // 8)] = np.float32(0.125)       # equal filter values inside and across shards
        fv[:4] = [-0.0, 0.0, -1.5, np.float32(-3.0e38)]
        img = fv.view(np.uint32).astype(np.int64)
        o = np.lexsort((keys, np.where(img >> 31, img ^ 0xffffffff, img ^ 0x80000000)))     # the order the device leaves a shard in
        keys, fv = keys[o], fv[o]
        cnt = np.ones(per, np.int64)
        offsets = np.concatenate(([0], np.cumsum(cnt))).astype(np.uint64)
        shards.append(SimpleNamespace(keys=keys, fv=fv, offsets=offsets, counts=None, branch=np.zeros(per, np.uint32), score=np.zeros(per, np.float32), branch16=None))
    want_s, want_i = hostlib.merge_order(shards)
    got_s, got_i = ctx.merge_order(shards, key_ranges, 22)
    got_s, got_i = got_s.copy(), got_i.copy()            # (views of the context's staging, overwritten by the next call)
    assert np.array_equal(got_s, want_s) and np.array_equal(got_i, want_i)
    if not key_ranges and n_shards > 1:
        # and the promise matters: without it the tie order of interleaved key sets needs the keys
        bad_s, bad_i = ctx.merge_order(shards, True, 22)
        assert not (np.array_equal(bad_s, want_s) and np.array_equal(bad_i, want_i))
