"""ctypes bindings of the CHECKERS: oracle/libipk_oracle.so (plain-C restatement, kind "port") and
oracle/_ref/libipkref_{dna,aa}.so (the reference's own pk_compute.cpp + window.cpp object code, kind "reference").

Test infrastructure only -- imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  Nothing under ipk_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libipk_oracle.so")
REF_SO = {4: os.path.join(ROOT, "oracle", "_ref", "libipkref_dna.so"),
          20: os.path.join(ROOT, "oracle", "_ref", "libipkref_aa.so")}


@dataclass
class DB:
    keys: np.ndarray
    fv: np.ndarray
    offsets: np.ndarray
    branch: np.ndarray
    score: np.ndarray
    n_emitted: int = 0
    ms_stage1: float = 0.0
    ms_stage2: float = 0.0

    @property
    def n_kmers(self):
        return int(self.keys.shape[0])

    @property
    def n_entries(self):
        return int(self.branch.shape[0])


class _IpkoDb(C.Structure):
    _fields_ = [("n_kmers", C.c_uint64), ("n_entries", C.c_uint64), ("n_emitted", C.c_uint64),
                ("keys", C.POINTER(C.c_uint64)), ("fv", C.POINTER(C.c_float)), ("offsets", C.POINTER(C.c_uint64)),
                ("branch", C.POINTER(C.c_uint32)), ("score", C.POINTER(C.c_float)),
                ("ms_stage1", C.c_double), ("ms_stage2", C.c_double)]


_port = None
_ref = {}


def port() -> C.CDLL:
    global _port
    if _port is None:
        lib = C.CDLL(ORACLE_SO)
        lib.ipko_best_prefix.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.ipko_best_prefix.restype = None
        for name in ("ipko_enumerate", "ipko_enumerate_predicate"):
            f = getattr(lib, name)
            f.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p, C.c_void_p,
                          C.c_void_p, C.c_int64]
            f.restype = C.c_int64
        lib.ipko_mif0.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_float]
        lib.ipko_mif0.restype = C.c_float
        lib.ipko_random_fv.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p]
        lib.ipko_random_fv.restype = None
        lib.ipko_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                   C.c_float, C.c_float, C.c_uint32, C.c_int, C.c_int]
        lib.ipko_build.restype = C.POINTER(_IpkoDb)
        lib.ipko_db_free.argtypes = [C.POINTER(_IpkoDb)]
        lib.ipko_db_free.restype = None
        _port = lib
    return _port


def have_ref(sigma: int) -> bool:
    return sigma in REF_SO and os.path.exists(REF_SO[sigma])


def ref(sigma: int) -> C.CDLL:
    if sigma not in _ref:
        lib = C.CDLL(REF_SO[sigma])
        lib.ipkref_sigma.restype = C.c_int
        assert lib.ipkref_sigma() == sigma
        lib.ipkref_best_prefix.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        lib.ipkref_best_prefix.restype = None
        lib.ipkref_enumerate.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int64]
        lib.ipkref_enumerate.restype = C.c_int64
        lib.ipkref_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_float,
                                     C.c_uint32, C.c_int, C.c_int]
        lib.ipkref_build.restype = C.c_void_p
        for n in ("ipkref_db_num_kmers", "ipkref_db_num_entries", "ipkref_db_num_emitted"):
            getattr(lib, n).argtypes = [C.c_void_p]
            getattr(lib, n).restype = C.c_uint64
        lib.ipkref_db_ms.argtypes = [C.c_void_p, C.c_int]
        lib.ipkref_db_ms.restype = C.c_double
        lib.ipkref_db_copy.argtypes = [C.c_void_p] * 6
        lib.ipkref_db_copy.restype = None
        lib.ipkref_db_free.argtypes = [C.c_void_p]
        lib.ipkref_db_free.restype = None
        _ref[sigma] = lib
    return _ref[sigma]


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


# ---- port -----------------------------------------------------------------------------------------------------------
def best_prefix(logp: np.ndarray, sigma: int) -> np.ndarray:
    logp = _f32(logp).reshape(-1, sigma)
    W = logp.shape[0]
    B = np.empty(W + 1, np.float32)
    port().ipko_best_prefix(logp.ctypes.data, W, sigma, B.ctypes.data)
    return B


def _enum(fn, logp, sigma, bits, k, eps, cap):
    logp = _f32(logp).reshape(-1, sigma)
    W = logp.shape[0]
    while True:
        wpos = np.empty(cap, np.uint32)
        keys = np.empty(cap, np.uint64)
        scores = np.empty(cap, np.float32)
        n = fn(logp.ctypes.data, W, sigma, bits, k, eps, wpos.ctypes.data, keys.ctypes.data, scores.ctypes.data, cap)
        if n <= cap:
            return wpos[:n], keys[:n], scores[:n]
        cap = int(n)


def enumerate_node(logp, sigma, bits, k, eps, cap=1 << 16):
    """one matrix [W, sigma] -> (window, key, score) of every emitted k-mer, reference emission order"""
    return _enum(port().ipko_enumerate, logp, sigma, bits, k, eps, cap)


def enumerate_node_predicate(logp, sigma, bits, k, eps, cap=1 << 16):
    return _enum(port().ipko_enumerate_predicate, logp, sigma, bits, k, eps, cap)


def mif0(log_scores, total_num_groups: int, threshold: float) -> float:
    s = _f32(log_scores)
    return float(port().ipko_mif0(s.ctypes.data, s.shape[0], total_num_groups, threshold))


def random_fv(keys) -> np.ndarray:
    """the canonical random filter value of every k-mer key (a function of the key alone)"""
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    fv = np.empty(len(keys), np.float32)
    port().ipko_random_fv(len(keys), keys.ctypes.data, fv.ctypes.data)
    return fv


def build(logp, node_branch, W, sigma, bits, k, log_threshold, threshold, total_num_groups, filter=0, threads=1) -> DB:
    logp = _f32(logp)
    nb = np.ascontiguousarray(node_branch, dtype=np.uint32)
    p = port().ipko_build(logp.ctypes.data, nb.ctypes.data, nb.shape[0], W, sigma, bits, k, log_threshold, threshold,
                          total_num_groups, filter, threads)
    d = p.contents
    nk, ne = int(d.n_kmers), int(d.n_entries)
    out = DB(np.ctypeslib.as_array(d.keys, (nk + 1,))[:nk].copy(), np.ctypeslib.as_array(d.fv, (nk + 1,))[:nk].copy(),
             np.ctypeslib.as_array(d.offsets, (nk + 1,)).copy(), np.ctypeslib.as_array(d.branch, (ne + 1,))[:ne].copy(),
             np.ctypeslib.as_array(d.score, (ne + 1,))[:ne].copy(), int(d.n_emitted), d.ms_stage1, d.ms_stage2)
    port().ipko_db_free(p)
    return out


# ---- reference object code ------------------------------------------------------------------------------------------
def ref_best_prefix(logp, sigma):
    logp = _f32(logp).reshape(-1, sigma)
    W = logp.shape[0]
    B = np.empty(W + 1, np.float32)
    ref(sigma).ipkref_best_prefix(logp.ctypes.data, W, B.ctypes.data)
    return B


def ref_enumerate_node(logp, sigma, k, eps, cap=1 << 16):
    logp = _f32(logp).reshape(-1, sigma)
    W = logp.shape[0]
    lib = ref(sigma)
    while True:
        wpos = np.empty(cap, np.uint32)
        keys = np.empty(cap, np.uint64)
        scores = np.empty(cap, np.float32)
        n = lib.ipkref_enumerate(logp.ctypes.data, W, k, eps, wpos.ctypes.data, keys.ctypes.data, scores.ctypes.data, cap)
        if n <= cap:
            return wpos[:n], keys[:n], scores[:n]
        cap = int(n)


def ref_build(logp, node_branch, W, sigma, k, log_threshold, threshold, total_num_groups, filter=0, threads=1) -> DB:
    logp = _f32(logp)
    nb = np.ascontiguousarray(node_branch, dtype=np.uint32)
    lib = ref(sigma)
    h = lib.ipkref_build(logp.ctypes.data, nb.ctypes.data, nb.shape[0], W, k, log_threshold, threshold, total_num_groups,
                         filter, threads)
    nk, ne = int(lib.ipkref_db_num_kmers(h)), int(lib.ipkref_db_num_entries(h))
    keys = np.empty(nk, np.uint64); fv = np.empty(nk, np.float32); offs = np.empty(nk + 1, np.uint64)
    branch = np.empty(ne, np.uint32); score = np.empty(ne, np.float32)
    lib.ipkref_db_copy(h, keys.ctypes.data, fv.ctypes.data, offs.ctypes.data, branch.ctypes.data, score.ctypes.data)
    out = DB(keys, fv, offs, branch, score, int(lib.ipkref_db_num_emitted(h)), lib.ipkref_db_ms(h, 1), lib.ipkref_db_ms(h, 2))
    lib.ipkref_db_free(h)
    return out


# ---- comparison -----------------------------------------------------------------------------------------------------
def canonical_windows(node, win, key, score):
    """sort unmerged emissions by (node, window, key) so that two emission orders can be compared"""
    o = np.lexsort((key, win, node))
    return node[o], win[o], key[o], score[o]


def db_entries(db) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """(key, branch, score) per entry, sorted by (key, branch): the set BASELINE.json's parity is defined on"""
    counts = np.diff(db.offsets.astype(np.int64))
    k = np.repeat(db.keys, counts)
    o = np.lexsort((db.branch, k))
    return k[o], db.branch[o], db.score[o]
