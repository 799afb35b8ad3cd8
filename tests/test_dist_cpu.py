"""world_size-2/3 `gloo` tests (CPU) of the multi-GPU host logic in ipk_b200/dist.py.

The device work behind `ctx.stage1 / stage1_partition / stage2` is replaced by a numpy stand-in built on the oracle, so
what is tested here is exactly the plumbing a rank runs around the C ABI: branch-group split, the group table
exchange, the counts all-gather + all-to-all-v of records, and the G-way merge of the shards.  Invariant:
the database at G in {2, 3} is entry-for-entry identical to G = 1.
"""
from __future__ import annotations

import os
import socket
import sys
from types import SimpleNamespace

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

N_TAXA, W, SIGMA, BITS, K, OMEGA = 12, 120, 4, 2, 8, 1.5


def mix64(x: np.ndarray) -> np.ndarray:
    """same mixer as ipk_b200/csrc/radix_sort.cuh: mix64 (owner rank of a k-mer = (mix64(kmer) >> 32) * G >> 32)"""
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


class OracleStages:
    """numpy stand-in for ipk_b200.capi.Context's stage methods (CPU; oracle arithmetic)"""

    def __init__(self):
        self.records = None

    def stage1(self, inp, group_base, group_bits, raw=False):
        import oraclelib as O
        from ipk_b200.dist import local_groups
        db = O.build(inp.logp, inp.node_branch, inp.W, inp.sigma, inp.bits, inp.k, inp.lthr, inp.thr, inp.total, 1, 1)
        groups = local_groups(inp.node_branch)
        gidx = {int(b): group_base + i for i, b in enumerate(groups)}
        cnt = np.diff(db.offsets.astype(np.int64))
        kmer = np.repeat(db.keys, cnt)
        grp = np.array([gidx[int(b)] for b in db.branch], dtype=np.uint64)
        o = np.lexsort((grp, kmer))
        ckey = (kmer[o] << np.uint64(group_bits)) | grp[o]
        score = db.score[o].copy()
        if raw:       # as emitted: any order, with duplicates of a (k-mer, group) that stage 2 must merge by maximum
            rng = np.random.default_rng(5 + group_base)
            dup = rng.integers(0, len(ckey), size=len(ckey) // 7) if len(ckey) else np.zeros(0, np.int64)
            ckey = np.concatenate([ckey, ckey[dup]])
            score = np.concatenate([score, score[dup] - np.float32(0.25)])
            o2 = rng.permutation(len(ckey))
            ckey, score = ckey[o2], score[o2]
        self.records = (ckey, score, group_bits)
        return len(ckey), dict(n_emitted=db.n_emitted)

    def stage1_partition(self, n_parts, d_keys, d_scores):
        ckey, score, gb = self.records
        owner = (((mix64(ckey >> np.uint64(gb)) >> np.uint64(32)) * np.uint64(n_parts)) >> np.uint64(32)).astype(np.int64)
        o = np.argsort(owner, kind="stable")
        n = len(ckey)
        d_keys[:n] = torch.from_numpy(ckey[o].view(np.int64).copy())
        d_scores[:n] = torch.from_numpy(score[o].copy())
        return np.bincount(owner, minlength=n_parts).astype(np.uint64)

    def stage2(self, d_keys, d_scores, n, group_bits, group_branch, key_bits, threshold, total_num_groups, filter=0,
               chunks_group_ordered=True, result_on_device=False):
        import oraclelib as O
        ckey = d_keys[:n].numpy().view(np.uint64)
        score = d_scores[:n].numpy()
        kmer = ckey >> np.uint64(group_bits)
        grp = (ckey & np.uint64((1 << group_bits) - 1)).astype(np.int64)
        if chunks_group_ordered:
            o = np.argsort(kmer, kind="stable")        # chunks arrive in ascending group ranges: stable on the k-mer alone
            assert np.all(np.diff(grp[o])[np.diff(kmer[o].astype(np.int64)) == 0] > 0), "chunks were not group ordered"
        else:
            o = np.lexsort((-score.astype(np.float64), grp, kmer))
        kmer, grp, score = kmer[o], grp[o], score[o]
        if not chunks_group_ordered and len(kmer):      # per-branch max: the first of every (k-mer, group) run
            head = np.ones(len(kmer), bool)
            head[1:] = (kmer[1:] != kmer[:-1]) | (grp[1:] != grp[:-1])
            kmer, grp, score = kmer[head], grp[head], score[head]
        keys, first = np.unique(kmer, return_index=True)
        offsets = np.append(first, len(kmer)).astype(np.uint64)
        fv = np.array([O.mif0(score[int(offsets[i]):int(offsets[i + 1])], total_num_groups, threshold) for i in range(len(keys))],
                      dtype=np.float32)
        order = np.lexsort((keys, fv))
        cnt = np.diff(offsets.astype(np.int64))[order]
        starts = offsets[:-1].astype(np.int64)[order]
        idx = np.concatenate([np.arange(s, s + c) for s, c in zip(starts, cnt)]) if len(order) else np.zeros(0, np.int64)
        out_off = np.zeros(len(order) + 1, np.uint64)
        out_off[1:] = np.cumsum(cnt)
        return SimpleNamespace(keys=keys[order], fv=fv[order], offsets=out_off, branch=np.asarray(group_branch)[grp[idx]].astype(np.uint32),
                               score=score[idx], n_kmers=len(order), n_entries=len(idx))


def _inputs():
    from ipk_b200 import synth
    return synth.posteriors(N_TAXA, W, SIGMA, seed=11)


def _thresholds():
    thr = np.float32(np.float32(OMEGA / SIGMA) ** np.float32(K))
    return float(thr), float(np.log10(thr, dtype=np.float32))


def _worker(rank, world, port, out_dir, raw):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ipk_b200 import dist as ipkdist
        logp, nb = _inputs()
        thr, lthr = _thresholds()
        n_branches = 2 * N_TAXA - 2
        b0, b1 = ipkdist.split_branches(n_branches, world)[rank]
        sl = slice(2 * b0, 2 * b1)                      # synth emits X0, X1 of branch b at rows 2b, 2b+1
        inp = SimpleNamespace(logp=logp[sl], node_branch=nb[sl], W=W, sigma=SIGMA, bits=BITS, k=K, lthr=lthr, thr=thr, total=2 * N_TAXA - 1)
        table = ipkdist.exchange_group_table(nb[sl], torch.device("cpu"))
        assert table.group_base == b0 and len(table.group_branch) == n_branches
        res, info = ipkdist.sharded_build(OracleStages(), inp, nb[sl], K * BITS, thr, 2 * N_TAXA - 1, 0, torch.device("cpu"), table=table, raw=raw)
        assert info["n_received"] > res.n_entries if raw else info["n_received"] == res.n_entries
        shards = [None] * world
        dist.all_gather_object(shards, dict(keys=res.keys, fv=res.fv, offsets=res.offsets, branch=res.branch, score=res.score))
        if rank == 0:
            merged = ipkdist.merge_shards([SimpleNamespace(**s) for s in shards])
            np.savez(os.path.join(out_dir, "merged.npz"), keys=merged[0], fv=merged[1], offsets=merged[2], branch=merged[3], score=merged[4],
                     shard_sizes=np.array([len(s["keys"]) for s in shards]))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("raw", [True, False])
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_build_equals_single(world, raw, tmp_path):
    import oraclelib as O
    from parity import assert_db_equal
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), raw), nprocs=world, join=True)
    z = np.load(os.path.join(str(tmp_path), "merged.npz"))
    got = SimpleNamespace(keys=z["keys"], fv=z["fv"], offsets=z["offsets"], branch=z["branch"], score=z["score"],
                          n_kmers=len(z["keys"]), n_entries=len(z["branch"]))
    logp, nb = _inputs()
    thr, lthr = _thresholds()
    want = O.build(logp, nb, W, SIGMA, BITS, K, lthr, thr, 2 * N_TAXA - 1, 0, 1)
    assert want.n_entries > 1000
    assert all(n > 0 for n in z["shard_sizes"]), "every rank should own part of the key space"
    assert_db_equal(got, want)


def test_split_branches_covers_everything():
    from ipk_b200.dist import split_branches
    for n in (0, 1, 7, 19998):
        for g in (1, 2, 3, 8):
            parts = split_branches(n, g)
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1


def test_merge_shards_orders_by_fv_then_key():
    from ipk_b200.dist import merge_shards
    a = SimpleNamespace(keys=np.array([5, 1], np.uint64), fv=np.array([-2.0, 0.5], np.float32), offsets=np.array([0, 2, 3], np.uint64),
                        branch=np.array([1, 2, 3], np.uint32), score=np.array([-1, -2, -3], np.float32))
    b = SimpleNamespace(keys=np.array([4, 9], np.uint64), fv=np.array([-2.0, 0.1], np.float32), offsets=np.array([0, 1, 2], np.uint64),
                        branch=np.array([7, 8], np.uint32), score=np.array([-7, -8], np.float32))
    e = SimpleNamespace(keys=np.zeros(0, np.uint64), fv=np.zeros(0, np.float32), offsets=np.zeros(1, np.uint64),
                        branch=np.zeros(0, np.uint32), score=np.zeros(0, np.float32))
    keys, fv, offs, br, sc = merge_shards([a, e, b])
    assert keys.tolist() == [4, 5, 9, 1]
    assert offs.tolist() == [0, 1, 3, 4, 5]
    assert br.tolist() == [7, 1, 2, 8, 3]
    assert sc.tolist() == [-7, -1, -2, -8, -3]
