"""Synthetic trees and ghost-node posterior matrices of the BASELINE.json shapes (host tooling, csrc/synth.c).

Shared by tests/ and both arms of bench.py; not part of the timed path.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libipksynth.so")


class SynthParams(C.Structure):
    _fields_ = [("p_peak_x0", C.c_float), ("p_peak_x1", C.c_float), ("mut_mean", C.c_float), ("x1_mut", C.c_float),
                ("diffuse_lo", C.c_float), ("diffuse_hi", C.c_float), ("peak_exp_lo", C.c_float), ("peak_exp_hi", C.c_float)]


# calibrated so that stored entries per window land in the 1-20 range the reference's golden DB sizes imply
# (D652 golden at k=7, omega=2.0 => ~1.2 stored/window, this generator: 1.2): almost all columns are
# sharply peaked (p_max = 1 - 10^-U(2.5,7)), a few percent are diffuse.  Measured on 16-branch samples:
# cfg2 9.9, cfg3 23.7, cfg4 136, cfg5 14.7 stored entries/window.
DEFAULT_PARAMS = dict(p_peak_x0=0.98, p_peak_x1=0.95, mut_mean=0.04, x1_mut=0.05, diffuse_lo=0.30, diffuse_hi=0.90,
                      peak_exp_lo=2.5, peak_exp_hi=7.0)

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        lib = C.CDLL(LIB_PATH)
        lib.ipk_synth_tree.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p]
        lib.ipk_synth_tree.restype = C.c_int
        lib.ipk_synth_posteriors.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(SynthParams),
                                             C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        lib.ipk_synth_posteriors.restype = C.c_int
        lib.ipk_synth_score_threshold.argtypes = [C.c_float, C.c_uint32, C.c_uint32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        lib.ipk_synth_score_threshold.restype = None
        _lib = lib
    return _lib


@dataclass(frozen=True)
class Workload:
    """One BASELINE.json configuration."""
    name: str
    n_taxa: int
    width: int
    k: int
    omega: float
    sigma: int
    bits: int
    filter: str = "mif0"

    @property
    def n_branches(self) -> int:       # non-root nodes of a rooted binary tree = branch groups
        return 2 * self.n_taxa - 2

    @property
    def n_tree_nodes(self) -> int:     # original_tree.get_node_count() -> total_num_groups of the MIF0 filter
        return 2 * self.n_taxa - 1

    @property
    def n_ghost_nodes(self) -> int:
        return 2 * self.n_branches

    @property
    def n_windows(self) -> int:
        return self.n_ghost_nodes * (self.width - self.k + 1)


# SURVEY.md §8 table.  cfg1 stands in for tests/data D652 (LFS stub); both the script's and BASELINE.json's parameters.
WORKLOADS = {
    "cfg1_d652_k7": Workload("cfg1_d652_k7", 652, 1700, 7, 2.0, 4, 2),
    "cfg1_d652_k8": Workload("cfg1_d652_k8", 652, 1700, 8, 1.5, 4, 2),
    "cfg1_d140_aa_k4": Workload("cfg1_d140_aa_k4", 140, 600, 4, 10.0, 20, 5),
    "cfg2_dna_1k": Workload("cfg2_dna_1k", 1000, 1500, 10, 1.5, 4, 2),
    "cfg3_16s_10k": Workload("cfg3_16s_10k", 10000, 1600, 12, 1.5, 4, 2),
    "cfg4_aa_2k": Workload("cfg4_aa_2k", 2000, 800, 6, 1.5, 20, 5),
    "cfg5_viral": Workload("cfg5_viral", 500, 9700, 14, 2.0, 4, 2),
}


def score_threshold(omega: float, sigma: int, k: int) -> tuple[float, float]:
    """(threshold, log10 threshold): the same float powf / log10f as ipkgpu_score_threshold, without loading libipkgpu.so"""
    t, lt = C.c_float(0), C.c_float(0)
    load().ipk_synth_score_threshold(omega, sigma, k, C.byref(t), C.byref(lt))
    return float(t.value), float(lt.value)


def tree(n_taxa: int, seed: int = 42):
    """-> (parent[u32, post-order ids; root = last, parent UINT32_MAX], branch_len[f32])"""
    n_nodes = 2 * n_taxa - 1
    parent = np.empty(n_nodes, np.uint32)
    blen = np.empty(n_nodes, np.float32)
    if load().ipk_synth_tree(n_taxa, seed, parent.ctypes.data, blen.ctypes.data) != 0:
        raise ValueError("ipk_synth_tree failed")
    return parent, blen


def posteriors(n_taxa: int, width: int, sigma: int, seed: int = 42, branch_begin: int = 0, branch_end: int | None = None,
               params: dict | None = None, out: np.ndarray | None = None):
    """Ghost matrices X0, X1 of branches [branch_begin, branch_end): (logp [2*nb, W, sigma] f32 log10, node_branch [2*nb] u32)."""
    if branch_end is None:
        branch_end = 2 * n_taxa - 2
    nb = branch_end - branch_begin
    prm = SynthParams(**{**DEFAULT_PARAMS, **(params or {})})
    if out is None:
        out = np.empty((2 * nb, width, sigma), np.float32)
    assert out.dtype == np.float32 and out.size == 2 * nb * width * sigma
    node_branch = np.empty(2 * nb, np.uint32)
    rc = load().ipk_synth_posteriors(n_taxa, width, sigma, seed, C.byref(prm), branch_begin, branch_end, out.ctypes.data,
                                     node_branch.ctypes.data)
    if rc != 0:
        raise ValueError("ipk_synth_posteriors failed")
    return out, node_branch
