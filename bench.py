#!/usr/bin/env python
"""bench.py -- phylo-k-mer database build throughput on B200 (BASELINE.json metric) and the reference CPU arm.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU; strong scaling: the SAME database)

A step = one complete build of the workload's database: posteriors -> enumerate -> per-branch max-merge -> regroup by
k-mer -> MIF0 filter values -> order by filter value (db_builder::run stages 1+2, ipk/src/db_builder.cpp:182-290).
  value  phylo-k-mers/s = stored (k-mer, branch) entries / device time, posteriors already resident in HBM, result left
         in HBM.
  e2e    same metric through the C ABI with HOST buffers: pinned posteriors -> H2D -> build -> D2H of the database.
Prints ONE JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "phylo-k-mers/s"
NVLINK_GBS = 900.0          # NVLink 5 per direction per GPU (BASELINE.json north_star)
DEFAULT_WORKLOAD = "cfg3_16s_10k"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("IPK_BENCH_WORKLOAD", DEFAULT_WORKLOAD))
    ap.add_argument("--cpu-sample-branches", type=int, default=0, help="branch groups in the CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--no-file", action="store_true", help="skip the serialization leg (posteriors in RAM -> DB.ipk on /dev/shm, and the one-shot ipk-b200 run)")
    ap.add_argument("--write-digest", action="store_true", help="record this run's database digest in profiles/digests.json")
    return ap.parse_args()


def peaks():
    """roofline denominator: measured copy bandwidth (MEASURED_PEAKS.json), else the profiling recipe's fallback"""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md, clocks line)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.thread.join(timeout=2)
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); power.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own object code (oracle/_ref, kind "reference") or the plain-C restatement (kind "port").
# The only place outside tests/ and smoke() where oracle/ is executed -- as the thing timed, never on the GPU path.
# ---------------------------------------------------------------------------------------------------------------------
def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_build_sample(wl, seed, b0, b1, threads, keep_db=False):
    """one CPU build over branch groups [b0, b1) of the workload; returns (n_entries, n_emitted, seconds, kind[, db])"""
    import oraclelib as O
    from ipk_b200 import synth                      # libipksynth.so only: the CPU arm maps no GPU library
    logp, nb = synth.posteriors(wl.n_taxa, wl.width, wl.sigma, seed=seed, branch_begin=b0, branch_end=b1)
    thr, lthr = synth.score_threshold(wl.omega, wl.sigma, wl.k)
    filt = 0 if wl.filter == "mif0" else 1
    t = time.perf_counter()
    if O.have_ref(wl.sigma):
        db, kind = O.ref_build(logp, nb, wl.width, wl.sigma, wl.k, lthr, thr, wl.n_tree_nodes, filt, threads), "reference"
    else:
        db, kind = O.build(logp, nb, wl.width, wl.sigma, wl.bits, wl.k, lthr, thr, wl.n_tree_nodes, filt, threads), "port"
    dt = time.perf_counter() - t
    if keep_db:
        return db.n_entries, db.n_emitted, dt, kind, db
    return db.n_entries, db.n_emitted, dt, kind


def cpu_stage3(db, wl):
    """the CPU arm's serialization (db_builder.cpp:297-332: one save_phylo_kmer per k-mer, a single thread) of its sample
    database, timed the way the B200 arm's file leg is: records -> file closed on /dev/shm.  Reported beside the build time,
    not added to it (the build metric of both arms ends before serialization)."""
    import tempfile
    from ipk_b200 import hostlib
    base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
    path = os.path.join(base, f"ipkb200_cpu_stage3_{os.getpid()}.ipk")
    try:
        t = time.perf_counter()
        hostlib.write_ipk(path, db.keys, db.fv, db.offsets, db.branch, db.score, wl.k, wl.omega, "DNA" if wl.sigma == 4 else "Proteins", threads=1)
        dt = time.perf_counter() - t
        size = os.path.getsize(path)
        return {"seconds": round(dt, 3), "file_bytes": size, "gbs": round(size / dt * 1e-9, 2), "threads": 1}
    finally:
        if os.path.exists(path):
            os.remove(path)


def cpu_sample_size(wl, seed, threads, target_s, forced=0):
    """branch groups in the bounded sample: probe with one group per thread, scale to ~target_s seconds"""
    if forced:
        return min(forced, wl.n_branches)
    probe = min(max(threads, 4), wl.n_branches)
    _, _, dt, _ = cpu_build_sample(wl, seed, 0, probe, threads)
    n = int(probe * max(target_s / max(dt, 1e-3), 1.0))
    n = max(probe, n - n % max(threads, 1))
    return min(n, wl.n_branches)


def cpu_baseline(wl, seed, target_s=15.0, forced=0, gpu_check=None):
    """gpu_check(nb, cpu_db) -> str: the B200 arm builds the same slice and compares the two databases entry for entry"""
    threads = host_cores()
    nb = cpu_sample_size(wl, seed, threads, target_s, forced)
    n_entries, n_emitted, dt, kind, db = cpu_build_sample(wl, seed, 0, nb, threads, keep_db=True)
    out = {"value": n_entries / dt, "unit": METRIC, "cores": threads, "kind": kind,
           "sample": f"branch groups [0,{nb}) of {wl.n_branches} ({2 * nb} ghost matrices, {2 * nb * (wl.width - wl.k + 1)} windows): "
                     f"{n_entries} entries, {n_emitted} emitted in {dt:.2f} s; OpenMP over branch groups "
                     f"(the loop db_builder.cpp:602-605 once marked parallel), in-RAM stages 1+2, no serialization"}
    out["stage3"] = cpu_stage3(db, wl)
    if gpu_check is not None:
        out["parity"] = gpu_check(nb, db)
    return out


def workload_config(wl, n_gpus, seed=42):
    """identical in both arms (the driver compares the two dicts)"""
    return {"seed": seed, "l2": "B200 arm: flushed between steps (256 MiB write)",
            "timing": "B200 arm: CUDA events on the launching stream, per step, summed, max over ranks; CPU arm: wall clock",
            "result": "B200 arm: indexed CSR (ipkgpu.h IPKGPU_RESULT_*_INDEXED): k-mers in filter-value order, each pointing at its entries",
            "workload": wl.name, "n_taxa": wl.n_taxa, "alignment_columns": wl.width, "k": wl.k, "omega": wl.omega,
            "states": wl.sigma, "filter": wl.filter, "branch_groups": wl.n_branches, "ghost_matrices": wl.n_ghost_nodes,
            "windows": wl.n_windows, "partition": f"branch groups over {n_gpus} GPU(s)" + (", the first pass of the merge sort stores 8-byte packed records into the k-mer-range owners' buffers over NVLink peer memory"
                                                                   if n_gpus > 1 and os.environ.get("IPK_EXCHANGE", "peer") == "peer" else
                                                                   ", NCCL all-to-all by k-mer hash" if n_gpus > 1 else "")}


def run_reference(args):
    from ipk_b200 import synth
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wl = synth.WORKLOADS[args.workload]
    threads = host_cores()
    per_step = max(3.0, min(20.0, 150.0 / max(args.steps + args.warmup, 1)))
    nb = cpu_sample_size(wl, args.seed, threads, per_step, args.cpu_sample_branches)
    for _ in range(args.warmup):
        cpu_build_sample(wl, args.seed, 0, nb, threads)
    tot_t, tot_e, kind, n_emitted = 0.0, 0, "port", 0
    for _ in range(args.steps):
        n_entries, n_emitted, dt, kind = cpu_build_sample(wl, args.seed, 0, nb, threads)
        tot_t += dt; tot_e += n_entries
    value = tot_e / tot_t
    sample = (f"each step = branch groups [0,{nb}) of {wl.n_branches} ({2 * nb * (wl.width - wl.k + 1)} windows, {tot_e // max(args.steps, 1)} entries, "
              f"{n_emitted} emitted); reference pk_compute.cpp/window.cpp object code + restated put/unsafe_insert/MIF0, "
              f"OpenMP over branch groups on {threads} threads")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(args.steps, 1), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(wl, args.gpus, args.seed),
            "cpu_baseline": {"value": value, "unit": METRIC, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def file_leg(ctx, inp_host, h_logp, node_branch, wl, n_entries, n_kmers):
    """BASELINE's metric is "phylo-k-mers/s & .ipk build time": (1) posteriors in pinned RAM -> ipkgpu_build -> the multi-threaded
    writer -> DB.ipk closed on /dev/shm, in this process (warm); (2) the ipk-b200 executable, one shot and cold: raw float32
    posteriors from /dev/shm -> DB.ipk on /dev/shm, wall clock of the whole process (context creation, first-touch allocations
    and page-locking of 9.5 GB included).  Outside the timed legs; files are removed afterwards."""
    import ctypes as C
    import shutil
    import tempfile
    from ipk_b200 import capi, hostlib
    base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
    need = 20 * n_kmers + 8 * n_entries + h_logp.numel() * 4 + (1 << 28)
    if shutil.disk_usage(base).free < need:
        return {"skipped": f"{base} has less than {need >> 30} GiB free"}
    d = tempfile.mkdtemp(prefix="ipkb200_bench_", dir=base)
    out = {}
    try:
        lib = hostlib.load()
        path = os.path.join(d, "DB.ipk")
        t0 = time.perf_counter()
        res = ctx.build_raw(inp_host, result_on_device=False, indexed=True, branch16=True)
        t1 = time.perf_counter()
        db = hostlib.HostDb(b"DNA" if wl.sigma == 4 else b"Proteins", b"", 0, None, None, wl.k, wl.omega, res.n_kmers, res.n_entries,
                            C.cast(res.keys, C.POINTER(C.c_uint64)), C.cast(res.fv, C.POINTER(C.c_float)), C.cast(res.offsets, C.POINTER(C.c_uint64)),
                            C.cast(res.counts, C.POINTER(C.c_uint32)), C.cast(res.branch, C.POINTER(C.c_uint32)), C.cast(res.score, C.POINTER(C.c_float)),
                            C.cast(res.branch16, C.POINTER(C.c_uint16)), None)
        rc = lib.ipkhost_write_ipk_mt(path.encode(), C.byref(db), 0)
        t2 = time.perf_counter()
        if rc != 0:
            raise RuntimeError("ipkhost_write_ipk_mt failed")
        size = os.path.getsize(path)
        out.update(build_ms=round((t1 - t0) * 1e3, 1), write_ms=round((t2 - t1) * 1e3, 1), total_ms=round((t2 - t0) * 1e3, 1), file_bytes=size,
                   writer_gbs=round(size / (t2 - t1) * 1e-9, 2), threads=host_cores(),
                   what="pinned posteriors -> ipkgpu_build (indexed host result, 16-bit branches) -> ipkhost_write_ipk_mt -> file closed on " + base)
        os.remove(path)
        # one shot, cold
        logp_path, map_path = os.path.join(d, "logp.f32"), os.path.join(d, "node_map.tsv")
        h_logp.numpy().tofile(logp_path)
        with open(map_path, "w") as f:
            f.write("".join(f"n{i}\t{int(b)}\n" for i, b in enumerate(node_branch)))
        argv = [hostlib.EXE_PATH, "--logp-f32", logp_path, "--width", str(wl.width), "--node-map", map_path, "--states", "dna" if wl.sigma == 4 else "aa",
                "-k", str(wl.k), "--omega", str(wl.omega), "--mif0" if wl.filter == "mif0" else "--random", "--node-count", str(wl.n_tree_nodes), "-o", path]
        t0 = time.perf_counter()
        r = subprocess.run(argv, capture_output=True, text=True)
        t1 = time.perf_counter()
        out["one_shot"] = {"wall_s": round(t1 - t0, 3), "returncode": r.returncode, "file_bytes": os.path.getsize(path) if os.path.exists(path) else 0,
                           "log": [l_ for l_ in r.stderr.splitlines() if l_.startswith(("Loaded", "Built", "Wrote", "Exception"))],
                           "what": "ipk-b200 --logp-f32 ... -o DB.ipk: process start to exit, posteriors and database on " + base}
        # the same executable over several devices (ranks = threads, shards merged while serializing): IPK_ONE_SHOT_DEVICES=0,1
        devs = os.environ.get("IPK_ONE_SHOT_DEVICES")
        if devs and r.returncode == 0:
            path2 = os.path.join(d, "DB_multi.ipk")
            argv2 = argv[:-1] + [path2, "--devices", devs]
            t0 = time.perf_counter()
            r2 = subprocess.run(argv2, capture_output=True, text=True)
            t1 = time.perf_counter()
            same = r2.returncode == 0 and subprocess.run(["cmp", "-s", path, path2]).returncode == 0
            out["one_shot_devices"] = {"devices": devs, "wall_s": round(t1 - t0, 3), "returncode": r2.returncode, "byte_equal_to_one_device": same,
                                       "log": [l_ for l_ in r2.stderr.splitlines() if l_.startswith(("Loaded", "Built", "Wrote", "Exception"))]}
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    g.build(only_missing=True)
    from ipk_b200 import capi, synth
    from ipk_b200 import dist as ipkdist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- ipk_b200 has no CPU path (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_bound = world > 1 and os.environ.get("IPK_BIND", "1") != "0" and ipkdist.bind_near_gpu(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = synth.WORKLOADS[args.workload]
    thr, lthr = capi.score_threshold(wl.omega, wl.sigma, wl.k)
    filt = capi.FILTER_MIF0 if wl.filter == "mif0" else capi.FILTER_RANDOM
    b0, b1 = ipkdist.split_branches(wl.n_branches, world)[rank]

    # ---- inputs: this rank's ghost matrices, generated straight into pinned host memory, plus a resident device copy
    n_nodes = 2 * (b1 - b0)
    h_logp = torch.empty((n_nodes, wl.width, wl.sigma), dtype=torch.float32, pin_memory=True)
    _, node_branch = synth.posteriors(wl.n_taxa, wl.width, wl.sigma, seed=args.seed, branch_begin=b0, branch_end=b1, out=h_logp.numpy())
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    d_logp = h_logp.to(dev, non_blocking=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)          # > 126 MB L2
    ctx = capi.Context(local_rank, stream.cuda_stream)
    if os.environ.get("IPK_POOL_CAP") or os.environ.get("IPK_ENUM_MODE"):          # tuning knobs (defaults are the product)
        ctx.set_enumeration(int(os.environ.get("IPK_ENUM_MODE", "0")), int(os.environ.get("IPK_POOL_CAP", "0")))
    inp_dev, keep1 = capi.make_input(d_logp, node_branch, wl.width, wl.sigma, wl.bits, wl.k, lthr, thr, wl.n_tree_nodes, filt)
    inp_host, keep2 = capi.make_input(h_logp.numpy(), node_branch, wl.width, wl.sigma, wl.bits, wl.k, lthr, thr, wl.n_tree_nodes, filt)
    table = ipkdist.exchange_group_table(node_branch, dev) if world > 1 else None
    key_bits = wl.k * wl.bits

    arenas = {}          # N > 1: every rank's host result in a shared page-locked arena; rank 0 maps them all (set up after the warm-up)
    merged = {}
    exch_info = {}

    def step(host: bool):
        """one build; returns (n_entries of this rank's shard, n_kmers, stats dict, d2h bytes, launches)"""
        inp = inp_host if host else inp_dev
        if world == 1:
            res = ctx.build_raw(inp, result_on_device=not host, indexed=True, branch16=host)
            st = res.stats.as_dict()
            return int(res.n_entries), int(res.n_kmers), st, int(st["d2h_bytes"]), int(st["kernel_launches"]), 0
        res, info = ipkdist.sharded_build(ctx, inp, node_branch, key_bits, thr, wl.n_tree_nodes, filt, dev, result_on_device=not host, table=table,
                                          exchange=os.environ.get("IPK_EXCHANGE", "peer"), indexed=True, branch16=host)
        exch_info.update(exchange=info.get("exchange", "NCCL all-to-all"), bytes_per_record=info.get("bytes_per_record", 12),
                         n_received_rank0=int(info.get("n_received", 0)))
        if host and arenas:
            # the end-to-end build ends with ONE database in rank 0's memory: the shards (mapped in place) + their N-way merge order
            got = ipkdist.gather_database(res, arenas, dev, ctx=None if os.environ.get("IPK_HOST_MERGE") else ctx,
                                          key_ranges=str(info.get("exchange", "")).startswith("packed"), key_bits=key_bits)
            if got is not None:
                merged["shards"], merged["order"] = got
                merged["ms_merge_order"] = ipkdist.LAST_MERGE_ORDER_MS
        st = res.stats.as_dict()
        st1 = info["stage1"]
        for k_ in ("n_windows", "n_emitted", "ms_h2d", "ms_prefix", "ms_enumerate", "h2d_bytes"):
            st[k_] = st1[k_]
        st["ms_sort"] += st1["ms_sort"]; st["ms_merge"] += st1["ms_merge"]; st["sort_passes"] += st1["sort_passes"]
        st["ms_partition"] = info.get("ms_partition", 0.0); st["ms_exchange"] = info.get("ms_exchange", 0.0)
        launches = int(st["kernel_launches"]) + int(st1["kernel_launches"]) + 1
        return int(res.n_entries), int(res.n_kmers), st, int(st["d2h_bytes"]), launches, info["n_sent_remote"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(host: bool, steps: int, profile: bool):
        """K steps, each bracketed by CUDA events on the launching stream; L2 flushed between steps (untimed)"""
        tot_ms, last = 0.0, None
        if profile:
            ctx.kernel_profile(True)
        barrier()
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if world > 1:
                dist.barrier()
            e0.record(stream)
            last = step(host)
            e1.record(stream)
            e1.synchronize()
            tot_ms += e0.elapsed_time(e1)
        barrier()
        rows = ctx.kernel_profile_rows() if profile else []
        if profile:
            ctx.kernel_profile(False)
        return tot_ms, last, rows

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x: int) -> int:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return int(t.item())

    # ---- warm-up (both paths: sizes every grow-only buffer, pins the result staging, settles clocks)
    for _ in range(args.warmup):
        last_w = step(False)
    if world > 1:
        # shared arenas sized by the shard this rank just built (the workload is deterministic; 5 % head room)
        # (entries are staged by the RECEIVED record count, an upper bound of the unique entries known before the merge)
        cap_e, cap_k = int(max(last_w[0], exch_info.get("n_received_rank0", 0)) * 1.05) + 4096, int(last_w[1] * 1.05) + 4096
        caps = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(caps, torch.tensor([cap_e, cap_k], dtype=torch.int64, device=dev))
        caps = [[int(x) for x in c.cpu().tolist()] for c in caps]
        tag = os.environ.get("MASTER_PORT", "0")
        mine = ipkdist.SharedArena(f"ipkb200_{tag}_{rank}", cap_e, cap_k, create=True)
        mine.register()
        ctx.set_result_arena(mine.address, cap_e, cap_k)
        dist.barrier()
        for r in range(world):
            arenas[r] = mine if r == rank else (ipkdist.SharedArena(f"ipkb200_{tag}_{r}", caps[r][0], caps[r][1], create=False) if rank == 0 else None)
            if rank == 0 and r != rank:
                arenas[r].register_heads()          # rank 0 uploads every shard's filter values for the merge order
    step(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_dev, last_dev, rows = timed(False, args.steps, True)
    dig_dev = ctx.result_digest(lthr, 1e-5)                 # the database the last device-resident step left (this rank's shard)
    ms_e2e, last_e2e, _ = timed(True, args.steps, False)
    dig_e2e = ctx.result_digest(lthr, 1e-5)
    clocks = sampler.stop() if rank == 0 else None

    def reduce_digest(d: dict) -> dict:
        """digests of key shards add / xor up to the digest of the whole database (ipkgpu.h: ipkgpu_result_digest)"""
        if world == 1:
            return d
        names = list(d)
        t = torch.tensor([d[n] - (1 << 64) if d[n] >= (1 << 63) else d[n] for n in names], dtype=torch.int64, device=dev)
        parts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        out = {}
        for i, n in enumerate(names):
            vals = [int(p_[i].item()) & ((1 << 64) - 1) for p_ in parts]
            acc = 0
            for v in vals:
                acc = (acc ^ v) if n.endswith("_xor") else ((acc + v) & ((1 << 64) - 1))
            out[n] = acc
        return out

    dig_dev, dig_e2e = reduce_digest(dig_dev), reduce_digest(dig_e2e)
    if dig_dev != dig_e2e:
        raise SystemExit(f"bench.py: the host-bound build is not the database of the device-resident one: digest {dig_e2e} vs {dig_dev}")

    ms_dev, ms_e2e = reduce_max(ms_dev), reduce_max(ms_e2e)
    n_entries = reduce_sum(last_dev[0]); n_kmers = reduce_sum(last_dev[1])
    if (n_entries, n_kmers) != (reduce_sum(last_e2e[0]), reduce_sum(last_e2e[1])):
        raise SystemExit(f"bench.py: the host-bound build disagrees with the device-resident one: {reduce_sum(last_e2e[0])} entries / "
                         f"{reduce_sum(last_e2e[1])} k-mers vs {n_entries} / {n_kmers}")
    st = last_dev[2]
    n_emitted = reduce_sum(int(st["n_emitted"]))
    h2d = reduce_sum(int(last_e2e[2]["h2d_bytes"])); d2h = reduce_sum(last_e2e[3])
    launches = last_dev[4] * args.steps
    sent_remote = reduce_sum(last_dev[5])
    if rank != 0:
        if world > 1:
            _close_arenas(ctx, arenas, rank)
            dist.destroy_process_group()
        return 0

    # ---- the database itself: entry-for-entry identical at every N -- the digest of this run against the one
    # a single-GPU run of the same workload and seed committed under profiles/
    digest_path = os.path.join(ROOT, "profiles", "digests.json")
    digest_key = f"{wl.name}/seed{args.seed}"
    try:
        with open(digest_path) as f:
            known = json.load(f)
    except Exception:
        known = {}
    digest_check = "no committed digest for this workload"
    if digest_key in known and not args.write_digest:
        want = {k_: int(v) for k_, v in known[digest_key]["digest"].items()}
        if want != dig_dev:
            raise SystemExit(f"bench.py: the database built on {world} GPU(s) differs from the committed single-GPU one ({digest_key}): {dig_dev} vs {want}")
        digest_check = f"equal to profiles/digests.json[{digest_key}] (written by a {known[digest_key]['n_gpus']}-GPU run)"
    if args.write_digest:
        known[digest_key] = {"n_gpus": world, "digest": {k_: str(v) for k_, v in dig_dev.items()}}
        with open(digest_path, "w") as f:
            json.dump(known, f, indent=1, sort_keys=True)
        digest_check += "; written"

    # ---- roofline of the dominant kernel: algorithmic bytes (summed by the library at each launch site, DESIGN.md
    # kernel table) / device time measured live with CUDA events around every launch of the timed region
    peak, peak_kind = peaks()
    rows.sort(key=lambda r: -r["ms"])
    kernels = [{"kernel": r["kernel"], "launches": r["launches"], "ms": round(r["ms"], 4),
                "algo_gbs": round(r["bytes"] / max(r["ms"], 1e-9) * 1e-6, 1), "share": round(r["ms"] / max(ms_dev, 1e-9), 4)} for r in rows]
    roofline = None
    if rows:
        top = rows[0]
        achieved = top["bytes"] / max(top["ms"], 1e-9) * 1e-6              # GB/s
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                for t in json.load(f):
                    if t.get("kernel") == top["kernel"] and t.get("workload") == wl.name:
                        traffic = t.get("dram_bytes_per_launch")
        except Exception:
            pass
        roofline = {"bound": "hbm", "kernel": top["kernel"], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                    "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_kind,
                    "launches_per_step": top["launches"] // max(args.steps, 1),
                    "algo_bytes_per_launch": top["bytes"] // max(top["launches"], 1),
                    "ms_per_launch": round(top["ms"] / max(top["launches"], 1), 4),
                    "share_of_step": round(top["ms"] / max(ms_dev, 1e-9), 4)}
        if top["kernel"].startswith("k2_"):
            roofline["note"] = ("the enumeration is bound by instruction issue and shared-memory latency (ncu: 56 % issue slots, DRAM 7 %), "
                                "not by HBM: its bytes are the record write; `roofline_hbm_kernel` is the largest memory-bound kernel")
        hbm = [r for r in rows if r["kernel"].startswith(("ps_", "rs_onesweep", "mg_"))]
        if hbm:
            h0 = hbm[0]
            roofline_hbm = {"kernel": h0["kernel"], "achieved": round(h0["bytes"] / max(h0["ms"], 1e-9) * 1e-6, 1), "peak": peak, "unit": "GB/s",
                            "frac": round(h0["bytes"] / max(h0["ms"], 1e-9) * 1e-6 / peak, 4), "launches_per_step": h0["launches"] // max(args.steps, 1),
                            "ms_per_launch": round(h0["ms"] / max(h0["launches"], 1), 4), "share_of_step": round(h0["ms"] / max(ms_dev, 1e-9), 4)}
        else:
            roofline_hbm = None
    # whole-path roofline as north_star defines it: emitted-plus-sorted bytes over HBM vs all-to-all bytes over NVLink
    passes = int(st["sort_passes"])
    exch_bytes_per_record = float(exch_info.get("bytes_per_record", 12))     # 8: packed records; 12: (u64 composite key, f32 score)
    host_check = None
    if world > 1 and merged:
        # what arrived in rank 0's memory IS the database the devices built: digest of the mapped shards, and the merge order ascends
        from ipk_b200 import hostlib
        hd = hostlib.digest(merged["shards"])
        if any(hd[k_] != dig_e2e[k_] for k_ in hd):
            raise SystemExit(f"bench.py: the shards in rank 0's memory are not the database the devices built: {hd} vs {dig_e2e}")
        osh, oix = merged["order"]
        fv_m, key_m = np.zeros(len(osh), np.float32), np.zeros(len(osh), np.uint64)
        for r_, d_ in enumerate(merged["shards"]):
            m_ = osh == r_
            fv_m[m_] = np.asarray(d_.fv)[oix[m_]]; key_m[m_] = np.asarray(d_.keys)[oix[m_]]
        tie = fv_m[1:] == fv_m[:-1]
        if len(osh) != n_kmers or not np.all(fv_m[1:] >= fv_m[:-1]) or not np.all(key_m[1:][tie] > key_m[:-1][tie]):
            raise SystemExit("bench.py: the N-way merge order of the shards is not ascending in (filter value, key)")
        host_check = (f"{world} shards mapped into rank 0 (shared page-locked arenas), host digest == device digest, N-way merge order over "
                      f"{len(osh)} k-mers in {merged.get('ms_merge_order', 0.0):.1f} ms "
                      f"({'host cores' if os.environ.get('IPK_HOST_MERGE') else 'ipkgpu_merge_order on rank 0'}; inside the timed region)")
    path_bytes = sum(r["bytes"] for r in rows) / max(args.steps, 1)
    t_hbm = path_bytes / (peak * 1e9)                     # `rows` are rank 0's own kernels: its share of the path already
    t_nvl = (exch_bytes_per_record * sent_remote / world) / (NVLINK_GBS * 1e9) if world > 1 else 0.0
    t_bound = max(t_hbm, t_nvl)
    value = n_entries * args.steps / (ms_dev * 1e-3)
    e2e_value = n_entries * args.steps / (ms_e2e * 1e-3)
    line = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(wl, world, args.seed),
            "host": "rank processes bound to the CPUs next to their GPU (NVML affinity)" if numa_bound else "unbound",
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps, "build_time_s": ms_e2e / args.steps * 1e-3},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "roofline_hbm_kernel": roofline_hbm if rows else None,
            "path_roofline": {"algo_bytes_per_step": int(path_bytes), "t_hbm_ms": round(t_hbm * 1e3, 3), "t_nvlink_ms": round(t_nvl * 1e3, 3),
                              "frac": round(t_bound / (ms_dev / args.steps * 1e-3), 4), "nvlink_peak_gbs": NVLINK_GBS,
                              "bound": "nvlink" if t_nvl > t_hbm else "hbm"},
            "counts": {"windows": wl.n_windows, "emitted": n_emitted, "entries": n_entries, "kmers": n_kmers, "sort_passes": passes,
                       "tiles_rank0": int(st.get("n_tiles", 0)), "deferred_tiles_rank0": int(st.get("n_deferred_tiles", 0)),
                       "emitted_per_s": n_emitted * args.steps / (ms_dev * 1e-3)},
            "stages_e2e_ms_rank0": {k_: round(float(last_e2e[2][k_]), 3) for k_ in ("ms_h2d", "ms_prefix", "ms_enumerate", "ms_sort", "ms_merge", "ms_filter", "ms_order", "ms_d2h", "ms_total", "ms_partition", "ms_exchange") if k_ in last_e2e[2]},
            "stages_ms_rank0": {k_: round(float(st[k_]), 3) for k_ in ("ms_prefix", "ms_enumerate", "ms_sort", "ms_merge", "ms_filter", "ms_order", "ms_partition", "ms_exchange") if k_ in st},
            "kernels": kernels}
    line["digest"] = {**{k_: str(v) for k_, v in dig_dev.items()}, "check": digest_check,
                      "e2e_equal": True}        # (checked above: the run aborts otherwise)
    if world > 1:
        line["e2e"]["artefact"] = host_check or "shards only (no shared arenas)"
        line["exchange"] = exch_info
    else:
        line["e2e"]["artefact"] = "one indexed database in pinned host memory (ipkgpu_build, IPKGPU_RESULT_HOST_INDEXED | BRANCH16)"
    if world == 1 and not args.no_file:
        line["e2e_file"] = file_leg(ctx, inp_host, h_logp, node_branch, wl, n_entries, n_kmers)
    if not args.no_cpu_baseline and world == 1:
        def gpu_check(nb, cpu_db):
            """the same [0, nb) branch-group slice through ipkgpu_build, compared with the CPU arm's database entry for entry"""
            from parity import assert_db_equal
            got = ctx.build(h_logp.numpy()[: 2 * nb], node_branch[: 2 * nb], wl.width, wl.sigma, wl.bits, wl.k, lthr, thr, wl.n_tree_nodes, filt)
            assert got.stats["n_emitted"] == cpu_db.n_emitted, (got.stats["n_emitted"], cpu_db.n_emitted)
            assert_db_equal(got, cpu_db, exact_fv=(filt != capi.FILTER_MIF0))
            return (f"ipkgpu_build of branch groups [0,{nb}) == the CPU arm's database: {got.n_kmers} k-mers, {got.n_entries} entries, "
                    f"(k-mer, branch) set / order / scores bit-exact, filter values within 2e-5 relative")
        line["cpu_baseline"] = cpu_baseline(wl, args.seed, 15.0, args.cpu_sample_branches, gpu_check)
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)
    if world > 1:
        _close_arenas(ctx, arenas, rank)
        dist.destroy_process_group()
    return 0


def _close_arenas(ctx, arenas, rank):
    import torch.distributed as dist
    try:
        ctx.set_result_arena(None, 0, 0)
    except Exception:
        pass
    dist.barrier()
    for r, a in arenas.items():
        if a is not None and r != rank:
            a.close()
    dist.barrier()
    if arenas.get(rank) is not None:
        arenas[rank].close()


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
