"""ctypes bindings of libipkhost.so (ipk_b200/host: posterior loader + .ipk file) -- test / tooling harness of the
C++ host side; the product path is the ipk-b200 executable built from the same sources."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libipkhost.so")
EXE_PATH = os.path.join(_HERE, "host", "ipk-b200")


class HostDb(C.Structure):
    _fields_ = [("sequence_type", C.c_char_p), ("tree", C.c_char_p), ("n_index", C.c_uint64),
                ("index_num_nodes", C.POINTER(C.c_uint64)), ("index_subtree_branch_length", C.POINTER(C.c_double)),
                ("kmer_size", C.c_uint64), ("omega", C.c_float), ("n_kmers", C.c_uint64), ("n_entries", C.c_uint64),
                ("keys", C.POINTER(C.c_uint64)), ("fv", C.POINTER(C.c_float)), ("offsets", C.POINTER(C.c_uint64)),
                ("counts", C.POINTER(C.c_uint32)), ("branch", C.POINTER(C.c_uint32)), ("score", C.POINTER(C.c_float)),
                ("branch16", C.POINTER(C.c_uint16)), ("position", C.POINTER(C.c_uint32))]


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        lib = C.CDLL(LIB_PATH)
        P = C.POINTER
        lib.ipkhost_ar_open.argtypes = [C.c_char_p, C.c_uint32, C.c_int, P(C.c_void_p)]
        lib.ipkhost_ar_close.argtypes = [C.c_void_p]
        lib.ipkhost_ar_close.restype = None
        lib.ipkhost_last_error.restype = C.c_char_p
        lib.ipkhost_ar_num_nodes.argtypes = [C.c_void_p]
        lib.ipkhost_ar_num_nodes.restype = C.c_uint32
        lib.ipkhost_ar_width.argtypes = [C.c_void_p]
        lib.ipkhost_ar_width.restype = C.c_uint32
        lib.ipkhost_ar_label.argtypes = [C.c_void_p, C.c_uint32]
        lib.ipkhost_ar_label.restype = C.c_char_p
        lib.ipkhost_ar_read_nodes.argtypes = [C.c_void_p, P(C.c_char_p), C.c_uint32, C.c_void_p]
        lib.ipkhost_parse_float.argtypes = [C.c_char_p, P(C.c_float)]
        lib.ipkhost_write_ipk.argtypes = [C.c_char_p, P(HostDb)]
        lib.ipkhost_read_ipk.argtypes = [C.c_char_p, P(C.c_void_p), P(HostDb)]
        lib.ipkhost_free_ipk.argtypes = [C.c_void_p]
        lib.ipkhost_free_ipk.restype = None
        lib.ipkhost_write_ipk_mt.argtypes = [C.c_char_p, P(HostDb), C.c_int]
        lib.ipkhost_merge_order.argtypes = [C.c_uint32, P(HostDb), C.c_int, C.c_void_p, C.c_void_p]
        lib.ipkhost_write_ipk_sharded.argtypes = [C.c_char_p, P(HostDb), C.c_uint32, P(HostDb), C.c_void_p, C.c_void_p, C.c_int]
        lib.ipkhost_digest.argtypes = [C.c_uint32, P(HostDb), C.c_int, P(C.c_uint64)]
        lib.ipkhost_plan_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, P(C.c_void_p)]
        lib.ipkhost_plan_free.argtypes = [C.c_void_p]
        lib.ipkhost_plan_free.restype = None
        for name, res in (("num_ghosts", C.c_uint32), ("num_groups", C.c_uint32), ("node_count", C.c_uint32), ("node_branch", P(C.c_uint32)),
                          ("index_num_nodes", P(C.c_uint64)), ("index_lengths", P(C.c_double)), ("newick", C.c_char_p), ("extended_newick", C.c_char_p)):
            fn = getattr(lib, "ipkhost_plan_" + name)
            fn.argtypes = [C.c_void_p]
            fn.restype = res
        for name in ("ext_label", "ar_label"):
            fn = getattr(lib, "ipkhost_plan_" + name)
            fn.argtypes = [C.c_void_p, C.c_uint32]
            fn.restype = C.c_char_p
        _lib = lib
    return _lib


class HostError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"ipkhost error {code}: {message}")
        self.code = code


def parse_float(text: str) -> float:
    v = C.c_float(0)
    rc = load().ipkhost_parse_float(text.encode(), C.byref(v))
    if rc != 0:
        raise HostError(rc, load().ipkhost_last_error().decode())
    return v.value


def read_ar(path: str, sigma: int, labels: list[str] | None = None, threads: int = 0):
    """-> (labels, logp [n, width, sigma] f32) for the given node labels (default: every node, file order)"""
    lib = load()
    h = C.c_void_p(0)
    rc = lib.ipkhost_ar_open(path.encode(), sigma, threads, C.byref(h))
    if rc != 0:
        raise HostError(rc, lib.ipkhost_last_error().decode())
    try:
        if labels is None:
            labels = [lib.ipkhost_ar_label(h, i).decode() for i in range(lib.ipkhost_ar_num_nodes(h))]
        width = lib.ipkhost_ar_width(h)
        out = np.empty((len(labels), width, sigma), np.float32)
        arr = (C.c_char_p * len(labels))(*[s.encode() for s in labels])
        rc = lib.ipkhost_ar_read_nodes(h, arr, len(labels), out.ctypes.data)
        if rc != 0:
            raise HostError(rc, lib.ipkhost_last_error().decode())
        return labels, out
    finally:
        lib.ipkhost_ar_close(h)


def plan(newick: str, ar_tree: str | None = None, ghosts: str = "both", use_unrooted: bool = False) -> dict:
    """the build plan of ipkhost.h: ghost nodes in processing order (extended / AR labels, branch of each), tree index, trees"""
    lib = load()
    h = C.c_void_p(0)
    rc = lib.ipkhost_plan_create(newick.encode(), ar_tree.encode() if ar_tree is not None else None,
                                 {"both": 0, "inner-only": 1, "outer-only": 2}[ghosts], 1 if use_unrooted else 0, C.byref(h))
    if rc != 0:
        raise HostError(rc, lib.ipkhost_last_error().decode())
    try:
        n, nc = lib.ipkhost_plan_num_ghosts(h), lib.ipkhost_plan_node_count(h)
        return dict(n_groups=lib.ipkhost_plan_num_groups(h), node_count=nc,
                    ext_labels=[lib.ipkhost_plan_ext_label(h, i).decode() for i in range(n)],
                    ar_labels=[lib.ipkhost_plan_ar_label(h, i).decode() for i in range(n)],
                    node_branch=np.ctypeslib.as_array(lib.ipkhost_plan_node_branch(h), (n,)).copy() if n else np.zeros(0, np.uint32),
                    index_num_nodes=np.ctypeslib.as_array(lib.ipkhost_plan_index_num_nodes(h), (nc,)).copy(),
                    index_len=np.ctypeslib.as_array(lib.ipkhost_plan_index_lengths(h), (nc,)).copy(),
                    newick=lib.ipkhost_plan_newick(h).decode(), extended_newick=lib.ipkhost_plan_extended_newick(h).decode())
    finally:
        lib.ipkhost_plan_free(h)


def write_ipk(path: str, keys, fv, offsets, branch, score, kmer_size: int, omega: float, sequence_type: str = "DNA",
              tree: str = "", index_num_nodes=None, index_len=None, threads: int = 0, position=None) -> None:
    keys = np.ascontiguousarray(keys, np.uint64); fv = np.ascontiguousarray(fv, np.float32)
    pos = np.ascontiguousarray(position, np.uint32) if position is not None else None
    offsets = np.ascontiguousarray(offsets, np.uint64)
    branch = np.ascontiguousarray(branch, np.uint32); score = np.ascontiguousarray(score, np.float32)
    inn = np.ascontiguousarray(index_num_nodes if index_num_nodes is not None else [], np.uint64)
    il = np.ascontiguousarray(index_len if index_len is not None else [], np.float64)
    db = HostDb(sequence_type.encode(), tree.encode(), len(inn), inn.ctypes.data_as(C.POINTER(C.c_uint64)),
                il.ctypes.data_as(C.POINTER(C.c_double)), kmer_size, omega, len(keys), len(branch),
                keys.ctypes.data_as(C.POINTER(C.c_uint64)), fv.ctypes.data_as(C.POINTER(C.c_float)),
                offsets.ctypes.data_as(C.POINTER(C.c_uint64)), None, branch.ctypes.data_as(C.POINTER(C.c_uint32)),
                score.ctypes.data_as(C.POINTER(C.c_float)), None, pos.ctypes.data_as(C.POINTER(C.c_uint32)) if pos is not None else None)
    rc = load().ipkhost_write_ipk_mt(path.encode(), C.byref(db), threads)
    if rc != 0:
        raise HostError(rc, "cannot write " + path)


def read_ipk(path: str) -> dict:
    lib = load()
    h = C.c_void_p(0)
    v = HostDb()
    rc = lib.ipkhost_read_ipk(path.encode(), C.byref(h), C.byref(v))
    if rc != 0:
        raise HostError(rc, "cannot read " + path)
    try:
        nk, ne, ni = int(v.n_kmers), int(v.n_entries), int(v.n_index)
        arr = lambda p, n: np.ctypeslib.as_array(p, (n,)).copy() if n else np.zeros(0, np.ctypeslib.as_array(p, (1,)).dtype if p else np.uint8)
        return dict(sequence_type=v.sequence_type.decode(), tree=v.tree.decode(), kmer_size=int(v.kmer_size), omega=float(v.omega),
                    index_num_nodes=arr(v.index_num_nodes, ni), index_len=arr(v.index_subtree_branch_length, ni),
                    keys=arr(v.keys, nk), fv=arr(v.fv, nk), offsets=np.ctypeslib.as_array(v.offsets, (nk + 1,)).copy(),
                    branch=arr(v.branch, ne), score=arr(v.score, ne), position=arr(v.position, ne) if v.position else None)
    finally:
        lib.ipkhost_free_ipk(h)


def _host_db(d, keep: list) -> HostDb:
    """ipkhost_db view of an object with keys / fv / offsets / branch (or branch16) / score (+ counts for indexed shards)"""
    def arr(name, dtype):
        a = getattr(d, name, None)
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype)
        keep.append(a)
        return a
    keys, fv, offsets = arr("keys", np.uint64), arr("fv", np.float32), arr("offsets", np.uint64)
    counts, branch, b16, score = arr("counts", np.uint32), arr("branch", np.uint32), arr("branch16", np.uint16), arr("score", np.float32)
    ptr = lambda a, t: a.ctypes.data_as(C.POINTER(t)) if a is not None else None
    n_entries = len(score) if score is not None else 0
    return HostDb(b"", b"", 0, None, None, 0, 0.0, len(keys), n_entries, ptr(keys, C.c_uint64), ptr(fv, C.c_float), ptr(offsets, C.c_uint64),
                  ptr(counts, C.c_uint32), ptr(branch, C.c_uint32), ptr(score, C.c_float), ptr(b16, C.c_uint16), None)


def merge_order(shards, threads: int = 0):
    """N-way merge of shard heads by (filter value, key) (merge_stage2, db_builder.cpp:392-458): -> (shard[u8], index[u32]) per merged k-mer"""
    keep: list = []
    views = (HostDb * len(shards))(*[_host_db(d, keep) for d in shards])
    total = sum(len(d.keys) for d in shards)
    out_shard, out_index = np.zeros(total, np.uint8), np.zeros(total, np.uint32)
    rc = load().ipkhost_merge_order(len(shards), views, threads, out_shard.ctypes.data, out_index.ctypes.data)
    if rc != 0:
        raise HostError(rc, "ipkhost_merge_order")
    return out_shard, out_index


def merge_shards(shards, threads: int = 0):
    """the merged database as five arrays (keys, fv, offsets, branch, score): what a reader of the one file would see"""
    sh, ix = merge_order(shards, threads)
    total = len(sh)
    keys, fv, cnt = np.zeros(total, np.uint64), np.zeros(total, np.float32), np.zeros(total, np.int64)
    starts = np.zeros(total, np.int64)
    bases = np.concatenate(([0], np.cumsum([len(d.score) for d in shards]))).astype(np.int64)
    for s, d in enumerate(shards):
        m = sh == s
        i = ix[m].astype(np.int64)
        keys[m] = np.asarray(d.keys)[i]; fv[m] = np.asarray(d.fv)[i]
        off = np.asarray(d.offsets).astype(np.int64)
        c = np.asarray(d.counts).astype(np.int64)[i] if getattr(d, "counts", None) is not None else off[i + 1] - off[i]
        cnt[m] = c; starts[m] = off[i] + bases[s]
    offsets = np.zeros(total + 1, np.uint64)
    offsets[1:] = np.cumsum(cnt)
    branch_all = np.concatenate([np.asarray(d.branch if getattr(d, "branch", None) is not None else d.branch16).astype(np.uint32) for d in shards]) if shards else np.zeros(0, np.uint32)
    score_all = np.concatenate([np.asarray(d.score, np.float32) for d in shards]) if shards else np.zeros(0, np.float32)
    idx = np.repeat(starts - offsets[:-1].astype(np.int64), cnt) + np.arange(int(cnt.sum()), dtype=np.int64)
    return keys, fv, offsets, branch_all[idx], score_all[idx]


def write_ipk_sharded(path: str, shards, kmer_size: int, omega: float, sequence_type: str = "DNA", tree: str = "",
                      index_num_nodes=None, index_len=None, threads: int = 0) -> None:
    keep: list = []
    views = (HostDb * len(shards))(*[_host_db(d, keep) for d in shards])
    sh, ix = merge_order(shards, threads)
    inn = np.ascontiguousarray(index_num_nodes if index_num_nodes is not None else [], np.uint64)
    il = np.ascontiguousarray(index_len if index_len is not None else [], np.float64)
    header = HostDb(sequence_type.encode(), tree.encode(), len(inn), inn.ctypes.data_as(C.POINTER(C.c_uint64)), il.ctypes.data_as(C.POINTER(C.c_double)),
                    kmer_size, omega, 0, 0, None, None, None, None, None, None, None, None)
    rc = load().ipkhost_write_ipk_sharded(path.encode(), C.byref(header), len(shards), views, sh.ctypes.data, ix.ctypes.data, threads)
    if rc != 0:
        raise HostError(rc, "cannot write " + path)


def digest(shards, threads: int = 0) -> dict:
    """ipkhost_digest: the device's database digest (ipkgpu_result_digest) recomputed from host arrays"""
    keep: list = []
    views = (HostDb * len(shards))(*[_host_db(d, keep) for d in shards])
    out = (C.c_uint64 * 6)()
    rc = load().ipkhost_digest(len(shards), views, threads, out)
    if rc != 0:
        raise HostError(rc, "ipkhost_digest")
    names = ("entries_sum", "entries_xor", "kmers_sum", "kmers_xor", "fv_sum", "fv_xor")
    return {n: int(v) for n, v in zip(names, out)}
