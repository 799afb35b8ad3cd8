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
