"""ctypes binding of the C ABI in include/ipkgpu.h (libipkgpu.so, sm_100a).

This is plumbing only: every entry point below forwards to the CUDA library.  There is no CPU fallback -- if the
shared library is missing or no B200 is visible the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("IPKGPU_LIB") or os.path.join(_HERE, "libipkgpu.so")      # IPKGPU_LIB: A/B builds while tuning

FILTER_MIF0 = 0
FILTER_RANDOM = 1

ERRORS = {0: "OK", -1: "EINVAL", -2: "ENODEV", -3: "ENOMEM", -4: "ECUDA"}


class IpkGpuError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libipkgpu {ERRORS.get(code, code)}: {message}")
        self.code = code


class Input(C.Structure):
    _fields_ = [
        ("logp", C.c_void_p),
        ("node_branch", C.c_void_p),
        ("n_nodes", C.c_uint32),
        ("width", C.c_uint32),
        ("sigma", C.c_uint32),
        ("bits", C.c_uint32),
        ("k", C.c_uint32),
        ("log_threshold", C.c_float),
        ("threshold", C.c_float),
        ("total_num_groups", C.c_uint32),
        ("filter", C.c_int32),
        ("logp_on_device", C.c_int32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("n_windows", C.c_uint64),
        ("n_emitted", C.c_uint64),
        ("n_entries", C.c_uint64),
        ("n_kmers", C.c_uint64),
        ("n_groups", C.c_uint32),
        ("sort_passes", C.c_uint32),
        ("enum_launches", C.c_uint32),
        ("kernel_launches", C.c_uint32),
        ("ms_h2d", C.c_float),
        ("ms_prefix", C.c_float),
        ("ms_enumerate", C.c_float),
        ("ms_sort", C.c_float),
        ("ms_merge", C.c_float),
        ("ms_filter", C.c_float),
        ("ms_order", C.c_float),
        ("ms_d2h", C.c_float),
        ("ms_total", C.c_float),
        ("h2d_bytes", C.c_uint64),
        ("d2h_bytes", C.c_uint64),
        ("n_tiles", C.c_uint64),
        ("n_deferred_tiles", C.c_uint64),
    ]

    def as_dict(self) -> dict:
        return {n: getattr(self, n) for n, _ in self._fields_}


class Result(C.Structure):
    _fields_ = [
        ("n_kmers", C.c_uint64),
        ("n_entries", C.c_uint64),
        ("keys", C.c_void_p),
        ("fv", C.c_void_p),
        ("offsets", C.c_void_p),
        ("branch", C.c_void_p),
        ("score", C.c_void_p),
        ("on_device", C.c_int32),
        ("counts", C.c_void_p),
        ("branch16", C.c_void_p),
        ("stats", Stats),
    ]


class Digest(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("entries_sum", "entries_xor", "kmers_sum", "kmers_xor", "fv_sum", "fv_xor",
                                           "n_boundary_above", "n_kmers", "n_entries")]

    def as_dict(self) -> dict:
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


_lib = None


def load() -> C.CDLL:
    """dlopen libipkgpu.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). ipk_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.ipkgpu_create.argtypes = [P(C.c_void_p), C.c_int, C.c_void_p]
    lib.ipkgpu_create.restype = C.c_int
    lib.ipkgpu_destroy.argtypes = [C.c_void_p]
    lib.ipkgpu_destroy.restype = None
    lib.ipkgpu_last_error.argtypes = [C.c_void_p]
    lib.ipkgpu_last_error.restype = C.c_char_p
    lib.ipkgpu_stream.argtypes = [C.c_void_p]
    lib.ipkgpu_stream.restype = C.c_void_p
    lib.ipkgpu_host_alloc.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
    lib.ipkgpu_host_alloc.restype = C.c_int
    lib.ipkgpu_host_free.argtypes = [C.c_void_p]
    lib.ipkgpu_host_free.restype = None
    lib.ipkgpu_set_workspace_limit.argtypes = [C.c_void_p, C.c_uint64]
    lib.ipkgpu_set_workspace_limit.restype = C.c_int
    lib.ipkgpu_set_enumeration.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
    lib.ipkgpu_set_enumeration.restype = C.c_int
    lib.ipkgpu_kernel_profile.argtypes = [C.c_void_p, C.c_int]
    lib.ipkgpu_kernel_profile.restype = C.c_int
    lib.ipkgpu_kernel_profile_get.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p, C.c_size_t, P(C.c_uint32), P(C.c_double), P(C.c_uint64)]
    lib.ipkgpu_kernel_profile_get.restype = C.c_int
    lib.ipkgpu_build.argtypes = [C.c_void_p, P(Input), C.c_int, P(Result)]
    lib.ipkgpu_build.restype = C.c_int
    lib.ipkgpu_result_digest.argtypes = [C.c_void_p, C.c_float, C.c_float, P(Digest)]
    lib.ipkgpu_result_digest.restype = C.c_int
    lib.ipkgpu_set_keep_positions.argtypes = [C.c_void_p, C.c_int]
    lib.ipkgpu_set_keep_positions.restype = C.c_int
    lib.ipkgpu_result_positions.argtypes = [C.c_void_p, P(C.c_void_p), P(C.c_int32)]
    lib.ipkgpu_result_positions.restype = C.c_int
    lib.ipkgpu_boundary_list.argtypes = [C.c_void_p, P(Input), C.c_float, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, P(C.c_uint64)]
    lib.ipkgpu_boundary_list.restype = C.c_int
    lib.ipkgpu_free_result.argtypes = [C.c_void_p, P(Result)]
    lib.ipkgpu_free_result.restype = None
    lib.ipkgpu_best_prefix.argtypes = [C.c_void_p, P(Input), C.c_void_p]
    lib.ipkgpu_best_prefix.restype = C.c_int
    lib.ipkgpu_enumerate.argtypes = [C.c_void_p, P(Input), C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     P(C.c_uint64)]
    lib.ipkgpu_enumerate.restype = C.c_int
    lib.ipkgpu_stage1.argtypes = [C.c_void_p, P(Input), C.c_uint32, C.c_uint32, P(C.c_uint64), P(Stats)]
    lib.ipkgpu_stage1.restype = C.c_int
    lib.ipkgpu_stage1_raw.argtypes = [C.c_void_p, P(Input), C.c_uint32, C.c_uint32, P(C.c_uint64), P(Stats)]
    lib.ipkgpu_stage1_raw.restype = C.c_int
    lib.ipkgpu_stage1_partition.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, P(C.c_uint64)]
    lib.ipkgpu_stage1_partition.restype = C.c_int
    lib.ipkgpu_stage1_owner_counts.argtypes = [C.c_void_p, C.c_uint32, P(C.c_uint64)]
    lib.ipkgpu_stage1_owner_counts.restype = C.c_int
    lib.ipkgpu_recv_buffers.argtypes = [C.c_void_p, C.c_uint64, P(C.c_void_p), P(C.c_void_p), C.c_void_p, C.c_void_p]
    lib.ipkgpu_recv_buffers.restype = C.c_int
    lib.ipkgpu_open_peer.argtypes = [C.c_void_p, C.c_void_p, P(C.c_void_p)]
    lib.ipkgpu_open_peer.restype = C.c_int
    lib.ipkgpu_stage1_scatter.argtypes = [C.c_void_p, C.c_uint32, P(C.c_void_p), P(C.c_void_p), P(C.c_uint64)]
    lib.ipkgpu_stage1_scatter.restype = C.c_int
    lib.ipkgpu_stage2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32,
                                  C.c_uint32, C.c_float, C.c_uint32, C.c_int32, C.c_int32, C.c_int, P(Result)]
    lib.ipkgpu_stage2.restype = C.c_int
    lib.ipkgpu_enable_peer.argtypes = [C.c_void_p, C.c_int]
    lib.ipkgpu_enable_peer.restype = C.c_int
    lib.ipkgpu_stage1_bucket_counts.argtypes = [C.c_void_p, P(C.c_uint64)]
    lib.ipkgpu_stage1_bucket_counts.restype = C.c_int
    lib.ipkgpu_exchange_plan.argtypes = [C.c_uint32, C.c_uint32, P(C.c_uint64), P(C.c_uint8), P(C.c_uint64), P(C.c_uint64), P(C.c_uint64)]
    lib.ipkgpu_exchange_plan.restype = C.c_int
    lib.ipkgpu_recv_packed.argtypes = [C.c_void_p, C.c_uint64, P(C.c_void_p), C.c_void_p]
    lib.ipkgpu_recv_packed.restype = C.c_int
    lib.ipkgpu_open_peer_slot.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, P(C.c_void_p)]
    lib.ipkgpu_open_peer_slot.restype = C.c_int
    lib.ipkgpu_stage1_scatter_packed.argtypes = [C.c_void_p, C.c_uint32, P(C.c_void_p), P(C.c_uint8), P(C.c_uint64), P(C.c_int32)]
    lib.ipkgpu_stage1_scatter_packed.restype = C.c_int
    lib.ipkgpu_stage2_packed.argtypes = [C.c_void_p, C.c_uint64, P(C.c_uint64), C.c_uint32, C.c_int32, C.c_void_p, C.c_uint32, C.c_uint32,
                                         C.c_float, C.c_uint32, C.c_int32, C.c_int, P(Result)]
    lib.ipkgpu_stage2_packed.restype = C.c_int
    lib.ipkgpu_set_result_arena.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]
    lib.ipkgpu_set_result_arena.restype = C.c_int
    lib.ipkgpu_result_arena_layout.argtypes = [C.c_uint64, C.c_uint64, P(C.c_size_t)]
    lib.ipkgpu_result_arena_layout.restype = C.c_size_t
    lib.ipkgpu_score_threshold.argtypes = [C.c_float, C.c_uint32, C.c_uint32, P(C.c_float), P(C.c_float)]
    lib.ipkgpu_score_threshold.restype = None
    lib.ipkgpu_merge_order.argtypes = [C.c_void_p, C.c_uint32, P(C.c_void_p), P(C.c_void_p), P(C.c_uint64), C.c_int32, C.c_uint32, C.c_void_p, C.c_void_p]
    lib.ipkgpu_merge_order.restype = C.c_int
    lib.ipkgpu_best_prefix_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
    lib.ipkgpu_best_prefix_noalloc.restype = C.c_int
    lib.ipkgpu_noalloc_last_error.argtypes = []
    lib.ipkgpu_noalloc_last_error.restype = C.c_char_p
    _lib = lib
    return lib


def result_arena_layout(cap_entries: int, cap_kmers: int) -> tuple[int, list[int]]:
    """(bytes, byte offsets of branch / score / keys / fv / offsets / counts) of a result arena (ipkgpu_result_arena_layout)"""
    off = (C.c_size_t * 6)()
    total = load().ipkgpu_result_arena_layout(cap_entries, cap_kmers, off)
    return int(total), [int(o) for o in off]


NOALLOC_MAX_COLUMNS = 1 << 22          # IPKGPU_NOALLOC_MAX_COLUMNS


def best_prefix_noalloc(device: int, stream: int, d_logp: int, n_nodes: int, width: int, sigma: int, d_B: int) -> None:
    """ipkgpu_best_prefix_noalloc: DEVICE addresses in and out, work queued on `stream` (0 = default stream); allocates nothing"""
    lib = load()
    rc = lib.ipkgpu_best_prefix_noalloc(device, stream or None, d_logp or None, n_nodes, width, sigma, d_B or None)
    if rc != 0:
        raise IpkGpuError(rc, (lib.ipkgpu_noalloc_last_error() or b"").decode())


def exchange_plan(n_parts: int, rank: int, counts: np.ndarray):
    """ipkgpu_exchange_plan: counts [n_parts, 256] -> (owner[256] u8, base[256] u64, bucket_start[257] u64, n_recv)"""
    counts = np.ascontiguousarray(counts, np.uint64)
    assert counts.shape == (n_parts, 256)
    owner = np.zeros(256, np.uint8); base = np.zeros(256, np.uint64); start = np.zeros(257, np.uint64)
    n_recv = C.c_uint64(0)
    P = C.POINTER
    rc = load().ipkgpu_exchange_plan(n_parts, rank, counts.ctypes.data_as(P(C.c_uint64)), owner.ctypes.data_as(P(C.c_uint8)),
                                     base.ctypes.data_as(P(C.c_uint64)), start.ctypes.data_as(P(C.c_uint64)), C.byref(n_recv))
    if rc != 0:
        raise IpkGpuError(rc, "ipkgpu_exchange_plan: bad arguments or a shard above 2^32-1 records")
    return owner, base, start, int(n_recv.value)


def score_threshold(omega: float, sigma: int, k: int) -> tuple[float, float]:
    """(threshold, log10 threshold) as the library's host helper computes them (float powf / log10f)."""
    t, lt = C.c_float(0), C.c_float(0)
    load().ipkgpu_score_threshold(omega, sigma, k, C.byref(t), C.byref(lt))
    return float(t.value), float(lt.value)


class PinnedArray:
    """float32 numpy array in page-locked memory from ipkgpu_host_alloc (include/ipkgpu.h): what a loader fills so that
    ipkgpu_build can upload in node chunks under its first kernels.  Freed by close() / the context manager."""

    def __init__(self, shape, device: int = 0):
        self._lib = load()
        n = int(np.prod(shape))
        p = C.c_void_p(0)
        rc = self._lib.ipkgpu_host_alloc(device, n * 4, C.byref(p))
        if rc != 0 or not p.value:
            raise IpkGpuError(rc, "ipkgpu_host_alloc failed")
        self._p = p
        self.array = np.ctypeslib.as_array((C.c_float * max(n, 1)).from_address(p.value))[:n].reshape(shape)

    def close(self):
        if self._p:
            self.array = None
            self._lib.ipkgpu_host_free(self._p)
            self._p = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def _ptr(a) -> int:
    """address of a numpy array / torch tensor / raw int"""
    if a is None:
        return 0
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    raise TypeError(type(a))


@dataclass
class PhyloKmerDB:
    """What db_builder.cpp:323-332 hands to save_phylo_kmer, k-mer by k-mer, as CSR arrays (host numpy copies)."""
    keys: np.ndarray       # u64 [n_kmers], ascending filter value (ties: ascending key)
    fv: np.ndarray         # f32 [n_kmers]
    offsets: np.ndarray    # u64 [n_kmers+1]
    branch: np.ndarray     # u32 [n_entries]
    score: np.ndarray      # f32 [n_entries]
    stats: dict = field(default_factory=dict)

    @property
    def n_kmers(self) -> int:
        return int(self.keys.shape[0])

    @property
    def n_entries(self) -> int:
        return int(self.branch.shape[0])


def _view(addr, n, dtype):
    if n == 0 or not addr:
        return np.zeros(0, dtype=dtype)
    nbytes = n * np.dtype(dtype).itemsize
    buf = (C.c_char * nbytes).from_address(addr)
    return np.frombuffer(buf, dtype=dtype, count=n)


def make_input(logp, node_branch: np.ndarray, width: int, sigma: int, bits: int, k: int, log_threshold: float,
               threshold: float, total_num_groups: int, filter: int = FILTER_MIF0) -> tuple[Input, tuple]:
    """logp: numpy [n_nodes, W, sigma] float32 (host) or a CUDA torch tensor of that shape."""
    node_branch = np.ascontiguousarray(node_branch, dtype=np.uint32)
    n_nodes = int(node_branch.shape[0])
    on_device = 0
    if isinstance(logp, np.ndarray):
        if logp.dtype != np.float32 or not logp.flags.c_contiguous:
            raise ValueError("logp must be C-contiguous float32")
        if logp.size != n_nodes * width * sigma:
            raise ValueError("logp has the wrong number of elements")
    else:
        import torch
        if logp.dtype != torch.float32 or not logp.is_contiguous():
            raise ValueError("logp must be a contiguous float32 tensor")
        if logp.numel() != n_nodes * width * sigma:
            raise ValueError("logp has the wrong number of elements")
        on_device = 1 if logp.is_cuda else 0
    inp = Input(_ptr(logp), _ptr(node_branch), n_nodes, width, sigma, bits, k, log_threshold, threshold,
                total_num_groups, filter, on_device)
    return inp, (logp, node_branch)      # keep-alive


class Context:
    """One ipkgpu_ctx: one CUDA device, one stream."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self._lib = load()
        self._h = C.c_void_p()
        rc = self._lib.ipkgpu_create(C.byref(self._h), device, C.c_void_p(stream or 0))
        if rc != 0:
            raise IpkGpuError(rc, (self._lib.ipkgpu_last_error(None) or b"").decode())
        self.device = device

    def close(self):
        if self._h:
            self._free_merge_order_staging()
            self._lib.ipkgpu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise IpkGpuError(rc, (self._lib.ipkgpu_last_error(self._h) or b"").decode())

    @property
    def stream(self) -> int:
        return int(self._lib.ipkgpu_stream(self._h) or 0)

    def set_workspace_limit(self, nbytes: int):
        self._check(self._lib.ipkgpu_set_workspace_limit(self._h, nbytes))

    def set_enumeration(self, mode: int = 0, pool_cap: int = 0):
        """0: list-join kernel + depth-first kernel on deferred tiles; 1: depth-first only.  pool_cap 0 = default."""
        self._check(self._lib.ipkgpu_set_enumeration(self._h, mode, pool_cap))

    def kernel_profile(self, enable: bool = True):
        """start (and reset) / stop per-kernel CUDA-event timing"""
        self._check(self._lib.ipkgpu_kernel_profile(self._h, 1 if enable else 0))

    def kernel_profile_rows(self) -> list[dict]:
        rows, i = [], 0
        name = C.create_string_buffer(64)
        n, ms, b = C.c_uint32(0), C.c_double(0), C.c_uint64(0)
        while self._lib.ipkgpu_kernel_profile_get(self._h, i, name, 64, C.byref(n), C.byref(ms), C.byref(b)) == 0:
            rows.append(dict(kernel=name.value.decode(), launches=int(n.value), ms=float(ms.value), bytes=int(b.value)))
            i += 1
        return rows

    # ---- whole path ------------------------------------------------------------------------------------------------
    def build_raw(self, inp: Input, result_on_device: bool | int = False, indexed: bool = False, branch16: bool = False) -> Result:
        """result_on_device: bool, or one of the RESULT_* codes; indexed: leave the entries in k-mer key order (no gather);
        branch16: host indexed results carry 16-bit branch ids when they all fit"""
        res = Result()
        mode = (int(result_on_device) & 7) | (2 if indexed else 0) | (4 if branch16 else 0)
        self._check(self._lib.ipkgpu_build(self._h, C.byref(inp), mode, C.byref(res)))
        return res

    def result_digest(self, log_threshold: float = 0.0, band: float = 1e-5) -> dict:
        """order-independent digest of the database the last build / stage2 of this context left on the device"""
        d = Digest()
        self._check(self._lib.ipkgpu_result_digest(self._h, log_threshold, band, C.byref(d)))
        return d.as_dict()

    def set_keep_positions(self, on: bool = True) -> None:
        """KEEP_POSITIONS builds (ipk-aa-pos): every entry also carries the alignment position of the first window that reached its score"""
        self._check(self._lib.ipkgpu_set_keep_positions(self._h, 1 if on else 0))

    def result_positions(self, n_entries: int):
        """positions parallel to the last HOST result's branch / score arrays (numpy copy), or None when the build kept none"""
        p = C.c_void_p(0)
        dev = C.c_int32(0)
        self._check(self._lib.ipkgpu_result_positions(self._h, C.byref(p), C.byref(dev)))
        if not p.value or dev.value:
            return None
        return _view(p.value, n_entries, np.uint32).copy()

    def boundary_list(self, inp: Input, band: float = 1e-5, cap: int = 1 << 20):
        """(k-mer, branch) entries whose presence or score flips when the threshold moves by less than `band`:
        -> (n_total, kmer[u64], branch[u32], score at threshold - band, score at threshold + band or NaN), sorted by (k-mer, branch)"""
        kmer, branch = np.empty(cap, np.uint64), np.empty(cap, np.uint32)
        lo, hi = np.empty(cap, np.float32), np.empty(cap, np.float32)
        n = C.c_uint64(0)
        self._check(self._lib.ipkgpu_boundary_list(self._h, C.byref(inp), band, cap, kmer.ctypes.data, branch.ctypes.data, lo.ctypes.data,
                                                   hi.ctypes.data, C.byref(n)))
        m = min(int(n.value), cap)
        o = np.lexsort((branch[:m], kmer[:m]))
        return int(n.value), kmer[:m][o], branch[:m][o], lo[:m][o], hi[:m][o]

    @staticmethod
    def to_db(res: Result) -> PhyloKmerDB:
        if res.on_device:
            raise ValueError("result lives on the device")
        nk, ne = int(res.n_kmers), int(res.n_entries)
        branch = _view(res.branch16, ne, np.uint16).astype(np.uint32) if res.branch16 else _view(res.branch, ne, np.uint32)
        if res.counts:
            # indexed result -> contiguous CSR in filter-value order (what the serializer's walk produces)
            start = _view(res.offsets, nk, np.uint64).astype(np.int64)
            cnt = _view(res.counts, nk, np.uint32).astype(np.int64)
            offsets = np.zeros(nk + 1, np.uint64)
            offsets[1:] = np.cumsum(cnt)
            idx = np.repeat(start - offsets[:-1].astype(np.int64), cnt) + np.arange(int(cnt.sum()), dtype=np.int64)
            return PhyloKmerDB(_view(res.keys, nk, np.uint64).copy(), _view(res.fv, nk, np.float32).copy(), offsets,
                               branch[idx], _view(res.score, ne, np.float32)[idx], res.stats.as_dict())
        return PhyloKmerDB(_view(res.keys, nk, np.uint64).copy(), _view(res.fv, nk, np.float32).copy(),
                           _view(res.offsets, nk + 1, np.uint64).copy() if nk or res.offsets else np.zeros(1, np.uint64),
                           _view(res.branch, ne, np.uint32).copy(), _view(res.score, ne, np.float32).copy(),
                           res.stats.as_dict())

    def build(self, logp, node_branch, width, sigma, bits, k, log_threshold, threshold, total_num_groups,
              filter=FILTER_MIF0) -> PhyloKmerDB:
        inp, keep = make_input(logp, node_branch, width, sigma, bits, k, log_threshold, threshold, total_num_groups, filter)
        res = self.build_raw(inp, False)
        del keep
        return self.to_db(res)

    # ---- stage-level -----------------------------------------------------------------------------------------------
    def best_prefix(self, logp, node_branch, width, sigma, bits, k) -> np.ndarray:
        inp, keep = make_input(logp, node_branch, width, sigma, bits, k, 0.0, 0.0, 0)
        out = np.empty((inp.n_nodes, width + 1), dtype=np.float32)
        self._check(self._lib.ipkgpu_best_prefix(self._h, C.byref(inp), out.ctypes.data))
        del keep
        return out

    def enumerate(self, logp, node_branch, width, sigma, bits, k, log_threshold, cap: int):
        """-> (n_total, node[u32], window[u32], key[u64], score[f32]) with min(n_total, cap) records, any order"""
        inp, keep = make_input(logp, node_branch, width, sigma, bits, k, log_threshold, 0.0, 0)
        node = np.empty(cap, np.uint32)
        win = np.empty(cap, np.uint32)
        key = np.empty(cap, np.uint64)
        score = np.empty(cap, np.float32)
        n = C.c_uint64(0)
        self._check(self._lib.ipkgpu_enumerate(self._h, C.byref(inp), cap, node.ctypes.data, win.ctypes.data,
                                               key.ctypes.data, score.ctypes.data, C.byref(n)))
        del keep
        m = min(int(n.value), cap)
        return int(n.value), node[:m], win[:m], key[:m], score[:m]

    def stage1(self, inp: Input, group_base: int, group_bits: int, raw: bool = False) -> tuple[int, dict]:
        """raw: leave the records as emitted (no local sort / per-branch max): they are sorted once, after the exchange"""
        n = C.c_uint64(0)
        st = Stats()
        fn = self._lib.ipkgpu_stage1_raw if raw else self._lib.ipkgpu_stage1
        self._check(fn(self._h, C.byref(inp), group_base, group_bits, C.byref(n), C.byref(st)))
        return int(n.value), st.as_dict()

    def stage1_partition(self, n_parts: int, d_keys, d_scores) -> np.ndarray:
        counts = np.zeros(n_parts, np.uint64)
        self._check(self._lib.ipkgpu_stage1_partition(self._h, n_parts, _ptr(d_keys), _ptr(d_scores),
                                                      counts.ctypes.data_as(C.POINTER(C.c_uint64))))
        return counts

    # ---- fused partition + exchange over peer memory (include/ipkgpu.h) ----
    def stage1_owner_counts(self, n_parts: int) -> np.ndarray:
        counts = np.zeros(n_parts, np.uint64)
        self._check(self._lib.ipkgpu_stage1_owner_counts(self._h, n_parts, counts.ctypes.data_as(C.POINTER(C.c_uint64))))
        return counts

    def recv_buffers(self, n_records: int) -> tuple[int, int, bytes, bytes]:
        """-> (device address of the key buffer, of the score buffer, their CUDA IPC handles)"""
        k, s = C.c_void_p(0), C.c_void_p(0)
        hk, hs = C.create_string_buffer(64), C.create_string_buffer(64)
        self._check(self._lib.ipkgpu_recv_buffers(self._h, n_records, C.byref(k), C.byref(s), hk, hs))
        return int(k.value or 0), int(s.value or 0), hk.raw, hs.raw

    def open_peer(self, handle: bytes) -> int:
        p = C.c_void_p(0)
        self._check(self._lib.ipkgpu_open_peer(self._h, C.create_string_buffer(handle, 64), C.byref(p)))
        return int(p.value or 0)

    def stage1_scatter(self, peer_keys: list[int], peer_scores: list[int], base_offsets) -> None:
        n = len(peer_keys)
        pk = (C.c_void_p * n)(*peer_keys)
        ps = (C.c_void_p * n)(*peer_scores)
        base = (C.c_uint64 * n)(*[int(b) for b in base_offsets])
        self._check(self._lib.ipkgpu_stage1_scatter(self._h, n, pk, ps, base))

    # ---- the exchange on packed records: the MSD pass of the merge sort stores into the owners' buffers (include/ipkgpu.h) ----
    def stage1_bucket_counts(self) -> np.ndarray:
        counts = np.zeros(256, np.uint64)
        self._check(self._lib.ipkgpu_stage1_bucket_counts(self._h, counts.ctypes.data_as(C.POINTER(C.c_uint64))))
        return counts

    def recv_packed(self, n_records: int) -> tuple[int, bytes]:
        """-> (device address of this rank's receive buffer, its CUDA IPC handle)"""
        p = C.c_void_p(0)
        h = C.create_string_buffer(64)
        self._check(self._lib.ipkgpu_recv_packed(self._h, n_records, C.byref(p), h))
        return int(p.value or 0), h.raw

    def open_peer_slot(self, slot: int, handle: bytes) -> int:
        p = C.c_void_p(0)
        self._check(self._lib.ipkgpu_open_peer_slot(self._h, slot, C.create_string_buffer(handle, 64), C.byref(p)))
        return int(p.value or 0)

    def stage1_scatter_packed(self, peer_recs: list[int], owner: np.ndarray, base: np.ndarray) -> bool:
        """-> whether this rank's records went out in branch-group order"""
        n = len(peer_recs)
        pr = (C.c_void_p * n)(*peer_recs)
        owner = np.ascontiguousarray(owner, np.uint8); base = np.ascontiguousarray(base, np.uint64)
        ordered = C.c_int32(0)
        self._check(self._lib.ipkgpu_stage1_scatter_packed(self._h, n, pr, owner.ctypes.data_as(C.POINTER(C.c_uint8)),
                                                           base.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(ordered)))
        return bool(ordered.value)

    def stage2_packed(self, n: int, bucket_start: np.ndarray, group_bits: int, group_ordered: bool, group_branch: np.ndarray, key_bits: int,
                      threshold: float, total_num_groups: int, filter: int = FILTER_MIF0, result_on_device: bool | int = False,
                      indexed: bool = False, branch16: bool = False) -> Result:
        group_branch = np.ascontiguousarray(group_branch, dtype=np.uint32)
        bucket_start = np.ascontiguousarray(bucket_start, np.uint64)
        res = Result()
        mode = (int(result_on_device) & 7) | (2 if indexed else 0) | (4 if branch16 else 0)
        self._check(self._lib.ipkgpu_stage2_packed(self._h, n, bucket_start.ctypes.data_as(C.POINTER(C.c_uint64)), group_bits, 1 if group_ordered else 0,
                                                   group_branch.ctypes.data, int(group_branch.shape[0]), key_bits, threshold, total_num_groups, filter,
                                                   mode, C.byref(res)))
        return res

    def merge_order(self, shards, key_ranges: bool, key_bits: int = 64):
        """ipkgpu_merge_order over shard objects with .keys / .fv host arrays (numpy): -> (shard[u8], index[u32]) per merged k-mer"""
        n = len(shards)
        fvs = [np.ascontiguousarray(d.fv, np.float32) for d in shards]
        kys = [np.ascontiguousarray(d.keys, np.uint64) for d in shards]
        cnt = (C.c_uint64 * n)(*[len(f) for f in fvs])
        pf = (C.c_void_p * n)(*[f.ctypes.data for f in fvs])
        pk = (C.c_void_p * n)(*[k.ctypes.data for k in kys])
        total = sum(len(f) for f in fvs)
        # the order comes back into page-locked memory kept with the context (views are valid until the next call / close())
        need = total * 5 + 64
        if getattr(self, "_mo_cap", 0) < need:
            self._free_merge_order_staging()
            p = C.c_void_p(0)
            rc = self._lib.ipkgpu_host_alloc(self.device, need + need // 8, C.byref(p))
            if rc != 0 or not p.value:
                raise IpkGpuError(rc, "ipkgpu_host_alloc failed")
            self._mo_pin, self._mo_cap = p, need + need // 8
        base = self._mo_pin.value
        out_index = np.ctypeslib.as_array((C.c_uint32 * max(total, 1)).from_address(base))[:total]
        out_shard = np.ctypeslib.as_array((C.c_uint8 * max(total, 1)).from_address(base + 4 * total))[:total]
        self._check(self._lib.ipkgpu_merge_order(self._h, n, pf, pk, cnt, 1 if key_ranges else 0, key_bits, out_shard.ctypes.data, out_index.ctypes.data))
        return out_shard, out_index

    def _free_merge_order_staging(self):
        if getattr(self, "_mo_pin", None):
            self._lib.ipkgpu_host_free(self._mo_pin)
        self._mo_pin, self._mo_cap = None, 0

    def set_result_arena(self, base, cap_entries: int, cap_kmers: int) -> None:
        """host results go to the caller's page-locked arena (e.g. shared memory another process maps); base None resets"""
        self._check(self._lib.ipkgpu_set_result_arena(self._h, _ptr(base), cap_entries, cap_kmers))

    def stage2(self, d_keys, d_scores, n: int, group_bits: int, group_branch: np.ndarray, key_bits: int, threshold: float,
               total_num_groups: int, filter: int = FILTER_MIF0, chunks_group_ordered: bool = True,
               result_on_device: bool | int = False, indexed: bool = False, branch16: bool = False) -> Result:
        group_branch = np.ascontiguousarray(group_branch, dtype=np.uint32)
        res = Result()
        self._check(self._lib.ipkgpu_stage2(self._h, _ptr(d_keys), _ptr(d_scores), n, group_bits, group_branch.ctypes.data,
                                            int(group_branch.shape[0]), key_bits, threshold, total_num_groups, filter,
                                            1 if chunks_group_ordered else 0, (int(result_on_device) & 7) | (2 if indexed else 0) | (4 if branch16 else 0), C.byref(res)))
        return res
