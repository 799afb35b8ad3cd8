"""Multi-GPU database build: one process per GPU, partition by branch group, ONE all-to-all by k-mer hash.

Reference counterpart: none -- the reference is single-threaded (ipk/src/db_builder.cpp:602-605).  The shape is the one
its on-disk mode already has: k-mer space split into batches by key (branch_group.cpp:104-107, db_builder.cpp:654)
and a final N-way merge by filter value (merge_stage2, db_builder.cpp:392-458).

Per rank r of G:
  stage 1   enumerate the rank's own branch groups, records left as emitted      (ipkgpu_stage1_raw; no communication)
  partition bucket the records by owner = (mix64(kmer) >> 32) * G >> 32                 (ipkgpu_stage1_partition)
  exchange  counts all-gather, then one all-to-all-v of (u64 key, f32 score)     (NCCL over NVLink)
  stage 2   ONE sort on (kmer, group), per-branch max, CSR by k-mer, filter values, order   (ipkgpu_stage2)
Sorting after the exchange instead of before and after it costs 10 % more exchanged bytes (duplicates of a (k-mer,
branch) travel) and saves 3 of 9 radix passes and one merge pass per rank (r01e: N=2 ran 8 + 1 passes per rank).
`raw=False` keeps the earlier shape (local sort + max, exchange of unique records, stable regroup on the k-mer bits).
The final database is the G-way merge of the shards by (filter value, key): `merge_shards`.

All device work goes through a context object exposing stage1 / stage1_partition / stage2 (ipk_b200.capi.Context);
this module is only the torch.distributed plumbing around it and never computes on the records itself.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist


def bind_near_gpu(device_index: int) -> bool:
    """Pins the calling process to the CPUs next to its GPU (NVML's ideal affinity), so that the pinned host buffers it
    allocates afterwards -- posteriors, and the library's result staging -- land on that GPU's NUMA node.  With one process
    per GPU and all eight copying results back at once, unbound processes put every buffer wherever they happen to run and
    the copies share one socket's memory controllers.  Returns False (and changes nothing) when NVML cannot say."""
    try:
        import pynvml
        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device_index)
        bus = "%08x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        return True
    except Exception:
        return False


def split_branches(n_branches: int, world: int) -> list[tuple[int, int]]:
    """contiguous blocks of the branch-group processing order, sizes differing by at most one"""
    base, extra = divmod(n_branches, world)
    out, b = [], 0
    for r in range(world):
        e = b + base + (1 if r < extra else 0)
        out.append((b, e))
        b = e
    return out


def local_groups(node_branch: np.ndarray) -> np.ndarray:
    """branch ids in order of first appearance (group_ghost_ids, db_builder.cpp:524-553)"""
    nb = np.asarray(node_branch, dtype=np.uint32)
    _, first = np.unique(nb, return_index=True)
    return nb[np.sort(first)]


@dataclass
class GroupTable:
    group_base: int            # global index of this rank's first group
    group_bits: int            # bits of the global group field
    group_branch: np.ndarray   # [total groups] global group index -> branch id


def exchange_group_table(node_branch: np.ndarray, device, group=None) -> GroupTable:
    """every rank learns how many groups the ranks before it own and the full group -> branch table"""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    mine = local_groups(node_branch).astype(np.int64)
    counts = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([len(mine)], dtype=torch.int64, device=device), group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    pad = torch.zeros(cap, dtype=torch.int64, device=device)
    pad[: len(mine)] = torch.from_numpy(mine).to(device)
    parts = [torch.zeros(cap, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    table = np.concatenate([p[:c].cpu().numpy() for p, c in zip(parts, counts)]).astype(np.uint32)
    total = int(sum(counts))
    bits = 0
    while (1 << bits) < total:
        bits += 1
    return GroupTable(int(sum(counts[:rank])), bits, table)


def exchange_records(send_keys: torch.Tensor, send_scores: torch.Tensor, counts: np.ndarray, group=None):
    """all-to-all-v of the partitioned records.  counts[p] = records this rank sends to rank p (bucket order).
    Returns (recv_keys, recv_scores, recv_counts); chunks arrive in source-rank order = ascending group ranges."""
    world = dist.get_world_size(group)
    dev = send_keys.device
    mine = torch.tensor([int(c) for c in counts], dtype=torch.int64, device=dev)
    rows = [torch.zeros(world, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(rows, mine, group=group)
    matrix = torch.stack(rows).cpu().numpy()                 # matrix[src][dst]
    rank = dist.get_rank(group)
    recv_counts = matrix[:, rank].astype(np.int64)
    n_recv = int(recv_counts.sum())
    recv_keys = torch.empty(max(n_recv, 1), dtype=send_keys.dtype, device=dev)[:n_recv]
    recv_scores = torch.empty(max(n_recv, 1), dtype=send_scores.dtype, device=dev)[:n_recv]
    in_split, out_split = [int(c) for c in counts], [int(c) for c in recv_counts]
    n_send = int(sum(in_split))
    dist.all_to_all_single(recv_keys, send_keys[:n_send], out_split, in_split, group=group)
    dist.all_to_all_single(recv_scores, send_scores[:n_send], out_split, in_split, group=group)
    return recv_keys, recv_scores, recv_counts


def peer_exchange(ctx, device, group=None):
    """The fused partition + exchange (include/ipkgpu.h): owner counts all-gathered, every rank's receive buffers mapped
    into every other rank through CUDA IPC, then ONE scatter pass per rank that stores each bucket straight into its
    owner's buffers over NVLink, and a barrier.  No NCCL call touches the records.
    Returns (address of received keys, of received scores, n_received, counts this rank sent per destination)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    counts = ctx.stage1_owner_counts(world)
    mine = torch.tensor([int(c) for c in counts], dtype=torch.int64, device=device)
    rows = [torch.zeros(world, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(rows, mine, group=group)            # also orders this build after every rank's previous stage 2
    matrix = torch.stack(rows).cpu().numpy()            # matrix[src][dst]
    n_recv = int(matrix[:, rank].sum())
    rk, rs, hk, hs = ctx.recv_buffers(max(n_recv, 1))
    h_mine = torch.frombuffer(bytearray(hk + hs), dtype=torch.uint8).to(device)
    h_all = [torch.zeros(128, dtype=torch.uint8, device=device) for _ in range(world)]
    dist.all_gather(h_all, h_mine, group=group)
    peer_keys, peer_scores = [], []
    for r in range(world):
        if r == rank:
            peer_keys.append(rk); peer_scores.append(rs)
        else:
            hb = bytes(h_all[r].cpu().numpy().tobytes())
            peer_keys.append(ctx.open_peer(hb[:64])); peer_scores.append(ctx.open_peer(hb[64:]))
    base = [int(matrix[:rank, d].sum()) for d in range(world)]
    ctx.stage1_scatter(peer_keys, peer_scores, base)    # returns once this rank's stores have completed
    done = torch.ones(1, dtype=torch.int32, device=device)
    dist.all_reduce(done, group=group)
    assert int(done.item()) == world                    # host-side wait: every rank's stores have landed
    return rk, rs, n_recv, counts


def packed_eligible(key_bits: int, group_bits: int) -> bool:
    """the exchange on packed records (include/ipkgpu.h) needs the low k-mer bits and the group to fit 32 bits"""
    return key_bits >= 8 and key_bits + group_bits - 8 <= 32


def packed_exchange(ctx, device, group=None):
    """The exchange as the first pass of the merge sort (include/ipkgpu.h): per-rank counts of the 256 top-byte buckets
    all-gathered, owners = contiguous k-mer ranges of about equal size (ipkgpu_exchange_plan, host arithmetic in the
    library), every rank's MSD pass stores 8-byte packed records straight into the owners' receive buffers over NVLink.
    Returns (bucket_start[257] of this rank's buffer, n_received, group_ordered over all ranks, records sent to other ranks)."""
    from . import capi
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    counts = ctx.stage1_bucket_counts()
    mine = torch.from_numpy(counts.astype(np.int64)).to(device)
    rows = [torch.zeros(256, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(rows, mine, group=group)             # also orders this build after every rank's previous stage 2
    matrix = torch.stack(rows).cpu().numpy().astype(np.uint64)
    owner, base, start, n_recv = capi.exchange_plan(world, rank, matrix)
    rp, handle = ctx.recv_packed(max(n_recv, 1))
    h_all = [torch.zeros(64, dtype=torch.uint8, device=device) for _ in range(world)]
    dist.all_gather(h_all, torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(device), group=group)
    handles = torch.stack(h_all).cpu().numpy()           # one device -> host copy for all ranks' handles
    peers = [rp if r == rank else ctx.open_peer_slot(r, handles[r].tobytes()) for r in range(world)]
    ordered = ctx.stage1_scatter_packed(peers, owner, base)        # returns once this rank's stores have completed
    flag = torch.tensor([1 if ordered else 0], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)       # the barrier: every rank's stores have landed; and the AND of the flags
    sent_remote = int(counts[owner != rank].sum())
    return start, n_recv, bool(int(flag.item())), sent_remote


def sharded_build(ctx, inp, node_branch: np.ndarray, key_bits: int, threshold: float, total_num_groups: int, filter: int,
                  device, result_on_device: bool = False, group=None, table: GroupTable | None = None, raw: bool = True,
                  exchange: str = "peer", indexed: bool = False, branch16: bool = False):
    """The whole multi-GPU path for this rank.  ctx: ipk_b200.capi.Context (or anything with its stage methods);
    inp: the rank's ipkgpu_input (its own ghost nodes only).  Returns (stage-2 result, info dict)."""
    world = dist.get_world_size(group)
    if table is None:
        table = exchange_group_table(node_branch, device, group)
    cuda = torch.device(device).type == "cuda"
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if cuda else None
    n_unique, st1 = ctx.stage1(inp, table.group_base, table.group_bits, raw)
    if exchange in ("peer", "packed") and raw and cuda and hasattr(ctx, "stage1_scatter_packed") and packed_eligible(key_bits, table.group_bits) \
            and exchange != "peer-pairs":
        ev[0].record()
        start, n_recv, ordered, sent_remote = packed_exchange(ctx, device, group)
        ev[2].record()
        res = ctx.stage2_packed(n_recv, start, table.group_bits, ordered, table.group_branch, key_bits, threshold, total_num_groups, filter,
                                result_on_device, **({"indexed": True, "branch16": branch16} if indexed else {}))
        info = dict(stage1=st1, n_unique_local=int(n_unique), n_received=n_recv, keep=None, ms_partition=0.0, n_sent_remote=sent_remote,
                    ms_exchange=ev[0].elapsed_time(ev[2]), exchange="packed MSD pass into peer memory", group_ordered=ordered, bytes_per_record=8)
        return res, info
    if exchange in ("peer", "peer-pairs") and cuda and hasattr(ctx, "stage1_scatter"):
        ev[0].record()
        rk, rs, n_recv, counts = peer_exchange(ctx, device, group)
        ev[2].record()
        res = ctx.stage2(rk, rs, n_recv, table.group_bits, table.group_branch, key_bits, threshold, total_num_groups, filter,
                         not raw, result_on_device, **({"indexed": True, "branch16": branch16} if indexed else {}))
        info = dict(stage1=st1, n_unique_local=int(n_unique), n_received=n_recv, keep=None, ms_partition=0.0,
                    n_sent_remote=int(sum(int(c) for i, c in enumerate(counts) if i != dist.get_rank(group))),
                    ms_exchange=ev[0].elapsed_time(ev[2]), exchange="peer-memory scatter")
        return res, info
    send_keys = torch.empty(max(n_unique, 1), dtype=torch.int64, device=device)
    send_scores = torch.empty(max(n_unique, 1), dtype=torch.float32, device=device)
    if cuda:
        ev[0].record()
    counts = ctx.stage1_partition(world, send_keys, send_scores)
    if cuda:
        ev[1].record()
    recv_keys, recv_scores, recv_counts = exchange_records(send_keys, send_scores, counts, group)
    if cuda:
        ev[2].record()
    del send_keys, send_scores
    n_recv = int(recv_counts.sum())
    res = ctx.stage2(recv_keys, recv_scores, n_recv, table.group_bits, table.group_branch, key_bits, threshold,
                     total_num_groups, filter, not raw, result_on_device, **({"indexed": True, "branch16": branch16} if indexed else {}))
    info = dict(stage1=st1, n_unique_local=int(n_unique), n_sent_remote=int(sum(int(c) for i, c in enumerate(counts) if i != dist.get_rank(group))),
                n_received=n_recv, keep=(recv_keys, recv_scores))
    if cuda:           # stage 2 has synchronised the stream by now
        info["ms_partition"] = ev[0].elapsed_time(ev[1])
        info["ms_exchange"] = ev[1].elapsed_time(ev[2])
    return res, info


def merge_shards(shards):
    """G-way merge of per-rank databases (each in ascending (fv, key) order) into one, the shape of merge_stage2
    (db_builder.cpp:392-458): the native N-way merge of libipkhost (host/ipk_file.cpp: ipkhost_merge_order, worker threads
    over disjoint filter-value ranges) materialised as the five arrays of the merged database.  Shards may be contiguous
    (offsets[i+1] ends k-mer i) or indexed (a `counts` array)."""
    from . import hostlib
    return hostlib.merge_shards(shards)


# ---- one database in one address space: every rank's host result lands in a shared, page-locked arena that rank 0 maps ----
class SharedArena:
    """A POSIX shared-memory segment laid out by ipkgpu_result_arena_layout.  The owner rank creates it, page-locks it
    (cudaHostRegister) and hands it to its context (ipkgpu_set_result_arena): the D2H copies of its shard land there.  Any
    other process of the node attaches by name and reads the six arrays in place -- no second copy, no collective."""

    def __init__(self, name: str, cap_entries: int, cap_kmers: int, create: bool):
        from multiprocessing import shared_memory
        from . import capi
        self.cap_entries, self.cap_kmers = int(cap_entries), int(cap_kmers)
        self.nbytes, self.offsets = capi.result_arena_layout(self.cap_entries, self.cap_kmers)
        if create:
            try:
                self.shm = shared_memory.SharedMemory(name=name, create=True, size=self.nbytes)
            except FileExistsError:                    # left behind by a run that died: replace it
                stale = shared_memory.SharedMemory(name=name, create=False)
                stale.close(); stale.unlink()
                self.shm = shared_memory.SharedMemory(name=name, create=True, size=self.nbytes)
        else:
            self.shm = shared_memory.SharedMemory(name=name, create=False)
        self.buf = np.frombuffer(self.shm.buf, dtype=np.uint8, count=self.nbytes)
        self.owner, self.registered = create, False

    @property
    def address(self) -> int:
        return self.buf.ctypes.data

    def register(self):
        rc = torch.cuda.cudart().cudaHostRegister(self.address, self.nbytes, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister of the shared arena failed: {rc}")
        self.registered = True

    def register_heads(self):
        """page-lock only the k-mer heads (keys and filter values): what a process that attaches needs pinned for
        ipkgpu_merge_order to read them at the full PCIe rate"""
        lo = (self.address + self.offsets[2]) & ~4095
        hi = (self.address + self.offsets[4] + 4095) & ~4095
        rc = torch.cuda.cudart().cudaHostRegister(lo, hi - lo, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister of the shared arena's heads failed: {rc}")
        self.heads_registered = lo

    def shard(self, n_kmers: int, n_entries: int, branch16: bool):
        """the arrays of a finished indexed host result, viewed in place"""
        from types import SimpleNamespace
        o = self.offsets
        view = lambda off, dtype, n: np.frombuffer(self.shm.buf, dtype=dtype, count=n, offset=off)
        d = SimpleNamespace(keys=view(o[2], np.uint64, n_kmers), fv=view(o[3], np.float32, n_kmers), offsets=view(o[4], np.uint64, n_kmers + 1),
                            counts=view(o[5], np.uint32, n_kmers), score=view(o[1], np.float32, n_entries),
                            branch=None if branch16 else view(o[0], np.uint32, n_entries), branch16=view(o[0], np.uint16, n_entries) if branch16 else None)
        return d

    def close(self):
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.address)
            self.registered = False
        if getattr(self, "heads_registered", None):
            torch.cuda.cudart().cudaHostUnregister(self.heads_registered)
            self.heads_registered = None
        self.buf = None
        try:
            self.shm.close()
            if self.owner:
                self.shm.unlink()
        except Exception:
            pass


LAST_MERGE_ORDER_MS = 0.0


def gather_database(res, arenas, device, group=None, threads: int = 0, ctx=None, key_ranges: bool = False, key_bits: int = 64):
    """After sharded_build with host results in shared arenas: rank 0 learns every shard's sizes (one tiny all-gather) and
    merges the k-mer heads of all shards by (filter value, key) -- the N-way merge of merge_stage2 (db_builder.cpp:392-458),
    on rank 0's device when a context is given (ipkgpu_merge_order), else natively on the host cores (hostlib.merge_order).
    key_ranges: the shards are ascending k-mer ranges (the packed exchange).  Returns on rank 0 (shards, (order_shard, order_index)): the whole database in rank 0's
    address space, k-mer i of the build = k-mer order_index[i] of shards[order_shard[i]]; None elsewhere."""
    from . import hostlib
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    mine = torch.tensor([int(res.n_kmers), int(res.n_entries), 1 if res.branch16 else 0], dtype=torch.int64, device=device)
    rows = [torch.zeros(3, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(rows, mine, group=group)            # every rank's copies have completed (ipkgpu_stage2* is blocking) before it reports
    if rank != 0:
        return None
    sizes = [[int(x) for x in r.cpu().tolist()] for r in rows]
    shards = [arenas[r].shard(sizes[r][0], sizes[r][1], bool(sizes[r][2])) for r in range(world)]
    import time
    global LAST_MERGE_ORDER_MS
    t0 = time.perf_counter()
    if ctx is not None and hasattr(ctx, "merge_order") and torch.device(device).type == "cuda":
        # the same order from one stable device sort of the filter values (ipkgpu_merge_order): the heads go up from the
        # page-locked arenas, (shard, index) per merged k-mer comes back
        order = ctx.merge_order(shards, key_ranges, key_bits)
    else:
        order = hostlib.merge_order(shards, threads)
    LAST_MERGE_ORDER_MS = (time.perf_counter() - t0) * 1e3          # the order alone, without the wait for the other ranks' copies
    return shards, order
