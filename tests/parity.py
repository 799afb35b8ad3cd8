"""Comparison of two phylo-k-mer databases (CSR form) at the bar BASELINE.json sets:
(k-mer, branch) set bit-exact, scores bit-exact here (fp32 max of identically computed fp32 sums; the north star
allows 1e-5), entry order within a k-mer = branch-group processing order, filter values within FV_RTOL, and the
k-mer order = ascending filter value."""
import numpy as np

FV_RTOL = 2e-5      # fp64 pow/log2 of two libms, rounded to fp32
FV_ATOL = 1e-12


def by_key(db):
    """(key, branch, score) per entry with k-mers in ascending key order, entries of a k-mer in stored order"""
    order = np.argsort(db.keys, kind="stable")
    offs = db.offsets.astype(np.int64)
    counts = np.diff(offs)
    starts = offs[:-1][order]
    cnt = counts[order]
    total = int(cnt.sum())
    # gather index: for each k-mer in key order, its entry range
    idx = np.repeat(starts - np.concatenate(([0], np.cumsum(cnt)[:-1])), cnt) + np.arange(total)
    return db.keys[order], db.fv[order], cnt, np.repeat(db.keys[order], cnt), db.branch[idx], db.score[idx]


def assert_db_equal(got, want, check_fv=True, exact_fv=False):
    assert got.n_kmers == want.n_kmers, (got.n_kmers, want.n_kmers)
    assert got.n_entries == want.n_entries, (got.n_entries, want.n_entries)
    assert int(got.offsets[0]) == 0 and int(got.offsets[-1]) == got.n_entries
    gk, gfv, gcnt, gek, gb, gs = by_key(got)
    wk, wfv, wcnt, wek, wb, ws = by_key(want)
    assert np.array_equal(gk, wk), "k-mer sets differ"
    assert len(np.unique(gk)) == len(gk), "duplicate k-mers"
    assert np.array_equal(gcnt, wcnt), "entries per k-mer differ"
    assert np.array_equal(gb, wb), "branch lists differ (set or order)"
    assert np.array_equal(gs.view(np.uint32), ws.view(np.uint32)), "scores differ bitwise"
    if check_fv:
        if exact_fv:
            assert np.array_equal(gfv.view(np.uint32), wfv.view(np.uint32))
        else:
            np.testing.assert_allclose(gfv, wfv, rtol=FV_RTOL, atol=FV_ATOL)
        # serialization order: ascending filter value, ties by ascending key
        fv = got.fv
        assert np.all(fv[1:] >= fv[:-1]), "k-mers are not in ascending filter-value order"
        tie = fv[1:] == fv[:-1]
        assert np.all(got.keys[1:][tie] > got.keys[:-1][tie]), "ties are not in ascending key order"


# ---- the digest of ipk_b200/csrc/digest.cuh in numpy (uint64 arithmetic wraps) ---------------------------------------
def _mix64(x):
    x = x.astype(np.uint64)
    x ^= x >> np.uint64(30); x *= np.uint64(0xbf58476d1ce4e5b9)
    x ^= x >> np.uint64(27); x *= np.uint64(0x94d049bb133111eb)
    x ^= x >> np.uint64(31)
    return x


def db_digest(db, log_threshold=0.0, band=1e-5):
    """the dict Context.result_digest returns, computed from a host database (any k-mer order)"""
    with np.errstate(over="ignore"):
        hk = _mix64(db.keys.astype(np.uint64) + np.uint64(0x9e3779b97f4a7c15))
        cnt = np.diff(db.offsets.astype(np.uint64))
        he = _mix64(np.repeat(hk, cnt.astype(np.int64)) ^ ((db.branch.astype(np.uint64) << np.uint64(32)) | db.score.view(np.uint32).astype(np.uint64)))
        hc = _mix64(hk ^ (cnt + np.uint64(0x632be59bd9b4e019)))
        hf = _mix64(hk ^ (db.fv.view(np.uint32).astype(np.uint64) | np.uint64(0x5bd1e99500000000)))
        xor = lambda a: int(np.bitwise_xor.reduce(a)) if a.size else 0
        add = lambda a: int(a.sum(dtype=np.uint64)) if a.size else 0
        hi = np.float32(log_threshold) + np.float32(band)
        return {"entries_sum": add(he), "entries_xor": xor(he), "kmers_sum": add(hc), "kmers_xor": xor(hc), "fv_sum": add(hf), "fv_xor": xor(hf),
                "n_boundary_above": int((db.score < hi).sum()), "n_kmers": db.n_kmers, "n_entries": db.n_entries}
