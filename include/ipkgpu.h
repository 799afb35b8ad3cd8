/* ipkgpu.h -- C ABI of the B200 (sm_100a) phylo-k-mer database build.
 *
 * Drop-in boundary for ONE path of phylo42/IPK: what `ipk::build(...)` (ipk/include/db_builder.h:22-29, called
 * from ipk/src/main.cpp:182-198) computes between "posterior matrices are available" and "k-mers are handed to
 * save_phylo_kmer in filter-value order" (ipk/src/db_builder.cpp:182-338).  The reference has no FFI of its own;
 * these entry points are what a maintainer would bind from db_builder.cpp.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success or a negative
 * IPKGPU_E* code and never throws; the message for the last failure of a context is ipkgpu_last_error().
 * A context is bound to one CUDA device and one stream; calls on one context must not overlap in time.
 * There is NO CPU fallback: without a usable CUDA device ipkgpu_create() fails with IPKGPU_ENODEV.
 */
#ifndef IPKGPU_H
#define IPKGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IPKGPU_OK        0
#define IPKGPU_EINVAL   -1   /* bad argument (k, sigma, bits, null pointer, key does not fit 64 bits ...) */
#define IPKGPU_ENODEV   -2   /* no CUDA device / device is not sm_100 */
#define IPKGPU_ENOMEM   -3   /* device or pinned-host memory exhausted (emitted volume is data dependent) */
#define IPKGPU_ECUDA    -4   /* any other CUDA runtime failure */

#define IPKGPU_FILTER_MIF0   0   /* ipk::filter_type::mif0   (ipk/include/filter.h:12-18) */
#define IPKGPU_FILTER_RANDOM 1   /* ipk::filter_type::random */

typedef struct ipkgpu_ctx ipkgpu_ctx;

/* Replaces the arguments of ipk::build that matter to the computation (db_builder.h:22-29):
 * `proba_matrix& matrix` + `ghost_mapping` + `ar::mapping` + ghost_strategy collapse into (logp, node_branch):
 * the caller lists the ghost nodes it wants explored (is_ghost, db_builder.cpp:495-507; root skipped, :535-539)
 * in processing order; equal branch ids form one branch group (group_ghost_ids, :524-553). */
typedef struct ipkgpu_input
{
    const float*    logp;          /* [n_nodes][width][sigma] fp32 log10 posteriors, site-major like
                                      ipk::matrix (window.h:23,67); host memory, or device memory if logp_on_device */
    const uint32_t* node_branch;   /* [n_nodes] post-order id of the ORIGINAL node (db_builder.cpp:613-618); host */
    uint32_t n_nodes;
    uint32_t width;                /* alignment columns W; windows per node = W-k+1 (window.cpp:159-197) */
    uint32_t sigma;                /* i2l::seq_traits::alphabet_size: 4 (DNA) or 20 (amino acids); 2..32 accepted */
    uint32_t bits;                 /* i2l::bit_length<seq_type>(): 2 or 5 (pk_compute.cpp:99,103) */
    uint32_t k;                    /* k-mer size; k*bits + ceil(log2(#groups)) must fit 64 bits */
    float    log_threshold;        /* log10(score_threshold(omega,k)), db_builder.cpp:640 -- computed by the caller */
    float    threshold;            /* score_threshold(omega,k) as score_type, db_builder.cpp:260 (MIF0 only) */
    uint32_t total_num_groups;     /* original_tree.get_node_count(), db_builder.cpp:261 (MIF0 only) */
    int32_t  filter;               /* IPKGPU_FILTER_* */
    int32_t  logp_on_device;       /* 0: logp is a host pointer (copied H2D inside the call); 1: device pointer */
} ipkgpu_input;

typedef struct ipkgpu_stats
{
    uint64_t n_windows;            /* n_nodes * (W-k+1) */
    uint64_t n_emitted;            /* candidates DCLA would emit over all windows (pk_compute.cpp:105) */
    uint64_t n_entries;            /* unique (k-mer, branch) after the per-branch max (branch_group.cpp:88-101) */
    uint64_t n_kmers;              /* distinct k-mers */
    uint32_t n_groups;             /* branch groups */
    uint32_t sort_passes;          /* radix passes of the merge sort */
    uint32_t enum_launches;        /* enumeration launches (2 when the first attempt overflowed its buffer) */
    uint32_t kernel_launches;      /* all kernels of this library launched by the call */
    float ms_h2d, ms_prefix, ms_enumerate, ms_sort, ms_merge, ms_filter, ms_order, ms_d2h, ms_total;  /* CUDA events */
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t n_tiles;              /* (node, 128-window) tiles enumerated */
    uint64_t n_deferred_tiles;     /* ... of which the depth-first kernel handled (list pool overflow, k == 1, mode 1) */
} ipkgpu_stats;

/* The database in the order db_builder.cpp:323-332 serializes it: k-mers by ascending filter value (ties: ascending
 * key), each with its (branch, score) entries in branch-group processing order (db_builder.cpp:685-693).
 * Arrays are owned by the context and stay valid until the next ipkgpu_build / ipkgpu_free_result / ipkgpu_destroy. */
typedef struct ipkgpu_result
{
    uint64_t n_kmers, n_entries;
    const uint64_t* keys;          /* [n_kmers]   i2l::phylo_kmer::key_type */
    const float*    fv;            /* [n_kmers]   kmer_fv::filter_value */
    const uint64_t* offsets;       /* [n_kmers+1] CSR offsets into branch/score */
    const uint32_t* branch;        /* [n_entries] pkdb_value::branch (post-order id) */
    const float*    score;         /* [n_entries] pkdb_value::score (log10) */
    int32_t on_device;             /* 1: the arrays are device pointers; 0: pinned host memory */
    const uint32_t* counts;        /* NULL: contiguous CSR, k-mer q owns [offsets[q], offsets[q+1]).  Indexed results
                                    * (IPKGPU_RESULT_*_INDEXED): [n_kmers], k-mer q owns [offsets[q], offsets[q] + counts[q])
                                    * and branch/score hold the k-mers in ascending KEY order -- the db.at(kmer) walk of
                                    * db_builder.cpp:323-332 as a look-up instead of a device-side gather */
    const uint16_t* branch16;      /* IPKGPU_RESULT_BRANCH16 honoured: the branch ids as 16-bit values and branch == NULL */
    ipkgpu_stats stats;
} ipkgpu_result;

/* i2l::score_threshold(omega, k) = (omega / sigma)^k evaluated in score_type (float) arithmetic, and the
 * log10 of it that explore_group compares against (db_builder.cpp:260,640).  Host-only helper so that every caller
 * (C++ host, Python harness, oracle drivers) feeds the kernels the same two constants.  UNVERIFIED(i2l): the
 * submodule source is absent; float powf / log10f is what the call sites' types imply. */
void ipkgpu_score_threshold(float omega, uint32_t sigma, uint32_t k, float* threshold, float* log_threshold);

/* device: CUDA ordinal.  stream: a cudaStream_t to run on (e.g. the caller's timing stream), or NULL for a private one. */
int  ipkgpu_create(ipkgpu_ctx** out, int device, void* stream);
void ipkgpu_destroy(ipkgpu_ctx* ctx);
const char* ipkgpu_last_error(const ipkgpu_ctx* ctx);      /* ctx may be NULL: message of the last failed create */
void* ipkgpu_stream(const ipkgpu_ctx* ctx);                /* the cudaStream_t all work of ctx is ordered on */
int  ipkgpu_set_workspace_limit(ipkgpu_ctx* ctx, uint64_t bytes);   /* cap on device scratch; 0 = 85% of free memory */

/* Page-locked host memory for the posteriors a caller is about to load (matrix rows of window.h:67 in node order): only
 * pinned input lets ipkgpu_build upload in node chunks under the enumeration kernels; pageable input works, unpipelined.
 * No context needed (the loader runs before the device path is set up).  Release with ipkgpu_host_free. */
int  ipkgpu_host_alloc(int device, size_t bytes, void** out);
void ipkgpu_host_free(void* p);

/* Enumeration strategy (no reference counterpart; the output is identical either way).  mode 0: list-join kernel, the
 * depth-first kernel takes the tiles whose lists overflow the shared-memory pool; mode 1: depth-first kernel only.
 * pool_cap: list-pool capacity in 8-byte elements (0 = default for the alphabet, at most 65000). */
int  ipkgpu_set_enumeration(ipkgpu_ctx* ctx, int mode, uint32_t pool_cap);

/* Per-kernel device timing (measurement aid; no reference counterpart).  While enabled, every kernel launch of ctx is
 * bracketed by CUDA events on ctx's stream; rows accumulate per kernel name until the next ipkgpu_kernel_profile().
 * ipkgpu_kernel_profile_get(index) -> name, launches, summed device ms, summed ALGORITHMIC bytes (DESIGN.md kernel
 * table); returns IPKGPU_EINVAL once index is past the last row. */
int  ipkgpu_kernel_profile(ipkgpu_ctx* ctx, int enable);
int  ipkgpu_kernel_profile_get(const ipkgpu_ctx* ctx, uint32_t index, char* name, size_t name_cap, uint32_t* launches,
                               double* ms, uint64_t* bytes);

/* values of result_on_device (ipkgpu_build, ipkgpu_stage2) */
#define IPKGPU_RESULT_HOST           0   /* pinned host arrays, contiguous CSR in filter-value order */
#define IPKGPU_RESULT_DEVICE         1   /* the same, left in device memory */
#define IPKGPU_RESULT_HOST_INDEXED   2   /* pinned host arrays, indexed (see ipkgpu_result.counts): no gather kernel, and the
                                          * entries cross PCIe under the filter / ordering kernels */
#define IPKGPU_RESULT_DEVICE_INDEXED 3
#define IPKGPU_RESULT_BRANCH16       4   /* OR-ed to IPKGPU_RESULT_HOST_INDEXED: if every branch id is below 65,536 the branch
                                          * ids come back as uint16_t (ipkgpu_result.branch16, branch == NULL) -- the result
                                          * crossing PCIe is the end-to-end bottleneck, and this takes 2 of its 8 bytes per entry */

/* The whole path: H2D (if needed) -> best_prefix -> enumerate -> sort + per-branch max -> CSR by k-mer -> filter
 * values -> order by filter value -> gather (-> D2H unless result_on_device).  Blocking. */
int  ipkgpu_build(ipkgpu_ctx* ctx, const ipkgpu_input* in, int result_on_device, ipkgpu_result* out);
void ipkgpu_free_result(ipkgpu_ctx* ctx, ipkgpu_result* res);

/* KEEP_POSITIONS builds (the reference's `ipk-aa-pos` binary: ipk::put of ipk/src/branch_group.cpp:72-86, entries
 * {branch, score, position}, db_builder.cpp:655-690).  While switched on, ipkgpu_build also records, per (k-mer, branch) entry,
 * the alignment position (window start) of the FIRST window in processing order -- the group's ghost nodes one after the
 * other, windows ascending -- that reached the entry's score (the strict `<` of put keeps the first maximum).
 * ipkgpu_result_positions hands back the array that runs parallel to ipkgpu_result.branch / score of the last build (same
 * memory space, same order; NULL when the last build kept none).  The path is the reference's rarely used variant and is
 * built for clarity, not speed (positions.cuh); 16-bit branches, streamed parts and the multi-GPU stages do not apply. */
int  ipkgpu_set_keep_positions(ipkgpu_ctx* ctx, int on);
int  ipkgpu_result_positions(ipkgpu_ctx* ctx, const uint32_t** positions, int32_t* on_device);

/* Order-independent digest of the database the last ipkgpu_build / ipkgpu_stage2 of ctx produced (it is still resident
 * on the device whatever result mode was asked for).  No reference counterpart -- the reference compares databases entry by
 * entry (tools/src/diff.cpp:129-295); the digest does the same for builds that never share a memory: sums (mod 2^64) and
 * xors of mix64 over every (k-mer, branch, score bits), every (k-mer, entry count) and every (k-mer, filter-value bits).
 * The digests of the key shards of a multi-GPU build add / xor up to the digest of the single-GPU database.
 * n_boundary_above: entries with log_threshold < score < log_threshold + band -- the emitted side of the boundary cases
 * BASELINE.json names (|score - threshold| < 1e-5); the full two-sided list is ipkgpu_boundary_list. */
typedef struct ipkgpu_digest
{
    uint64_t entries_sum, entries_xor;   /* over (key, branch, score bits) */
    uint64_t kmers_sum, kmers_xor;       /* over (key, number of entries) */
    uint64_t fv_sum, fv_xor;             /* over (key, filter-value bits) */
    uint64_t n_boundary_above;
    uint64_t n_kmers, n_entries;
} ipkgpu_digest;
int  ipkgpu_result_digest(ipkgpu_ctx* ctx, float log_threshold, float band, ipkgpu_digest* out);

/* Boundary cases (BASELINE.json: "k-mers whose reference score lies within 1e-5 of the threshold are listed separately"; the
 * reference's own differ does it by hand, diff-plain-text.py:36-45).  log_threshold comes out of i2l::score_threshold
 * (db_builder.cpp:640), whose rounding cannot be pinned here (source absent): an entry whose presence or value flips when the
 * threshold moves by less than `band` cannot be held against either build.  The list is exact by construction: the input is
 * enumerated and merged twice, at log_threshold - band and + band, and every (k-mer, branch) whose entry differs between
 * the two -- missing from the tight build (score_tight = NaN), or there with another per-branch maximum -- is reported.
 * Up to cap entries are written (any order), *n_out is their true number.  Costs two enumerations: a diagnostic, not a
 * stage of the build. */
int  ipkgpu_boundary_list(ipkgpu_ctx* ctx, const ipkgpu_input* in, float band, uint64_t cap, uint64_t* kmer_out, uint32_t* branch_out,
                          float* score_loose_out, float* score_tight_out, uint64_t* n_out);

/* ---- stage-level entry points (parity tests; multi-GPU build, where the caller runs the exchange) ---------------- */

/* matrix::preprocess (window.cpp:16-27) for every node: B_out[n_nodes][W+1] (host). */
int  ipkgpu_best_prefix(ipkgpu_ctx* ctx, const ipkgpu_input* in, float* B_out);

/* `for window in to_windows(m,k): DCLA(window,k).run(eps)` (db_builder.cpp:645-649) for every node, unmerged.
 * Writes up to cap records (any order) to the HOST arrays and the true total to *n_out. */
int  ipkgpu_enumerate(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint64_t cap,
                      uint32_t* node_out, uint32_t* window_out, uint64_t* key_out, float* score_out, uint64_t* n_out);

/* Stage 1 of one rank: enumerate + per-branch max over this rank's nodes (the rank's branch groups are numbered
 * group_base, group_base+1, ... in order of first appearance).  Leaves the unique records on the device as
 * composite keys  (kmer << group_bits) | global group index  with their scores; chunks are sorted by (kmer, group).
 * *n_out = their number.  stats may be NULL. */
int  ipkgpu_stage1(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint32_t group_base, uint32_t group_bits, uint64_t* n_out,
                   ipkgpu_stats* stats);
/* Same, without the local sort and per-branch max: the records are left as emitted (duplicates of a (k-mer, group)
 * included, any order), so that they are sorted ONCE, by their owner rank after the exchange -- stage 2 with
 * chunks_group_ordered = 0 takes the maximum over equal composite keys.  If the emitted volume does not fit the
 * workspace the call falls back to batches of whole groups merged locally (records then unique, as ipkgpu_stage1). */
int  ipkgpu_stage1_raw(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint32_t group_base, uint32_t group_bits, uint64_t* n_out,
                       ipkgpu_stats* stats);
/* Buckets the stage-1 records by owner = (mix64(kmer) >> 32) * n_parts >> 32, stably, into caller-provided DEVICE buffers of n_out
 * records (the send buffers of the all-to-all); counts[n_parts] (host) receives the bucket sizes. */
int  ipkgpu_stage1_partition(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* d_keys, float* d_scores, uint64_t* counts);
/* ---- fused partition + exchange over NVLink peer memory (one process per GPU; no NCCL on the data path) ------------
 * 1. every rank:  ipkgpu_stage1_owner_counts -> counts[dst];   the caller all-gathers the G x G matrix
 * 2. every rank:  ipkgpu_recv_buffers(sum over src of matrix[src][me]) -> its receive buffers + their CUDA IPC handles;
 *                 the caller all-gathers the handles (only when they changed) and maps them with ipkgpu_open_peer
 * 3. every rank:  ipkgpu_stage1_scatter: ONE pass over its stage-1 records; bucket d is stored straight into rank d's
 *                 receive buffers at element offset base_offsets[d] = sum over src < me of matrix[src][d]  (chunks land
 *                 in source-rank order = ascending group ranges).  Returns when the stores have completed.
 * 4. a barrier among the ranks, then ipkgpu_stage2 on the receive buffers. */
#define IPKGPU_IPC_HANDLE_BYTES 64
int  ipkgpu_stage1_owner_counts(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* counts);
int  ipkgpu_recv_buffers(ipkgpu_ctx* ctx, uint64_t n_records, uint64_t** d_keys, float** d_scores,
                         unsigned char* handle_keys /*[64] or NULL*/, unsigned char* handle_scores /*[64] or NULL*/);
int  ipkgpu_open_peer(ipkgpu_ctx* ctx, const unsigned char* handle /*[64]*/, void** d_ptr);
int  ipkgpu_stage1_scatter(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* const* peer_keys, float* const* peer_scores,
                           const uint64_t* base_offsets);

/* ---- the same exchange on PACKED records: the first pass of the merge sort IS the exchange -------------------------
 * The merge sort of the single-GPU build starts with a most-significant-digit pass on the top 8 bits of the k-mer that
 * turns (u64 key, f32 score) pairs into 8-byte packed records (packed_sort.cuh).  With G ranks the 256 buckets of that
 * pass are dealt to the ranks as contiguous k-mer ranges of about equal size and every rank's MSD pass stores each bucket
 * straight into its owner's receive buffer over NVLink: no separate partition pass, 8 instead of 12 bytes per record on
 * the wire, and the owner sorts 8 bits less.  A rank reads its emission in branch-group order (the runs the enumeration
 * recorded) and the sources land in rank order inside every bucket, so the owner's sort skips the group bits altogether.
 * Needs k*bits >= 8, k*bits + group_bits - 8 <= 32 and < 2^32 records per rank; otherwise use the pair exchange above.
 * 1. every rank:  ipkgpu_stage1_raw, then ipkgpu_stage1_bucket_counts -> counts[256]; the caller all-gathers G x 256
 * 2. every rank:  ipkgpu_exchange_plan (host arithmetic only) -> owner[256], base[256] (where THIS rank's records of each
 *                 bucket start inside the owner's buffer), bucket_start[257] and n_recv (this rank's own buffer)
 * 3. every rank:  ipkgpu_recv_packed(n_recv) -> buffer + CUDA IPC handle; handles all-gathered, mapped with
 *                 ipkgpu_open_peer_slot(rank of the peer, ...) (a changed handle replaces the stale mapping)
 * 4. every rank:  ipkgpu_stage1_scatter_packed; *group_ordered tells whether this rank's records went out in group order
 *                 (false after a depth-first or batched enumeration): the caller ANDs it over the ranks
 * 5. a barrier, then ipkgpu_stage2_packed on the rank's own receive buffer. */
/* ranks that are contexts of ONE process (ipk-b200 --devices): direct access from ctx's device to memory of peer_device,
 * so that the receive buffer pointers of ipkgpu_recv_packed can be handed over as they are (no IPC handle) */
int  ipkgpu_enable_peer(ipkgpu_ctx* ctx, int peer_device);
int  ipkgpu_stage1_bucket_counts(ipkgpu_ctx* ctx, uint64_t* counts /*[256]*/);
int  ipkgpu_exchange_plan(uint32_t n_parts, uint32_t rank, const uint64_t* counts /*[n_parts][256]*/, uint8_t* owner /*[256]*/,
                          uint64_t* base /*[256]*/, uint64_t* bucket_start /*[257]*/, uint64_t* n_recv);
int  ipkgpu_recv_packed(ipkgpu_ctx* ctx, uint64_t n_records, uint64_t** d_recs, unsigned char* handle /*[64] or NULL*/);
int  ipkgpu_open_peer_slot(ipkgpu_ctx* ctx, uint32_t slot, const unsigned char* handle /*[64]*/, void** d_ptr);
int  ipkgpu_stage1_scatter_packed(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* const* peer_recs, const uint8_t* owner, const uint64_t* base,
                                  int32_t* group_ordered);
int  ipkgpu_stage2_packed(ipkgpu_ctx* ctx, uint64_t n, const uint64_t* bucket_start /*[257]*/, uint32_t group_bits, int32_t group_ordered,
                          const uint32_t* group_branch, uint32_t n_groups_total, uint32_t key_bits, float threshold,
                          uint32_t total_num_groups, int32_t filter, int result_mode, ipkgpu_result* out);

/* Host results of ctx (every array of ipkgpu_result) go to ONE caller-provided PAGE-LOCKED arena instead of the context's
 * own staging -- e.g. a shared-memory segment that another process maps, so that the shards of a multi-GPU build meet in
 * one address space without a second copy.  ipkgpu_result_arena_layout gives the arena size for the capacities and the byte
 * offsets of its six arrays: [0] branch (u32, or u16 with IPKGPU_RESULT_BRANCH16) [1] score f32 [2] keys u64 [3] fv f32
 * [4] offsets u64 [5] counts u32 -- any process can address a mapped arena with it.  The arena is never grown or freed by
 * the library (a result that does not fit is IPKGPU_ENOMEM); base = NULL returns to the context's own staging. */
size_t ipkgpu_result_arena_layout(uint64_t cap_entries, uint64_t cap_kmers, size_t* offsets /*[6] or NULL*/);
int  ipkgpu_set_result_arena(ipkgpu_ctx* ctx, void* base, uint64_t cap_entries, uint64_t cap_kmers);

/* The N-way merge of merge_stage2 (db_builder.cpp:392-458) over the k-mer heads, on the device: shard s holds n_kmers[s]
 * k-mers in ascending (filter value, key) -- HOST arrays fv[s] / keys[s] (page-locked ones copy fastest) -- and the k-mer at
 * merged position i is k-mer out_index[i] of shard out_shard[i] (host arrays of sum(n_kmers) elements; entries stay where
 * they are).  The order-preserving 32-bit images of all filter values are sorted stably with their positions as payload.
 * key_ranges != 0 promises that shard s holds smaller keys than shard s + 1 (what the packed exchange's k-mer-range owners
 * produce): equal filter values then fall in key order by stability alone and `keys` may be NULL; otherwise the positions are
 * first put into key order by LSD passes on the low key_bits bits of the keys.  Same order as ipkhost_merge_order
 * (ipk_b200/host/ipkhost.h), which stays the host-only form; <= 256 shards, < 2^32 k-mers in all. */
int  ipkgpu_merge_order(ipkgpu_ctx* ctx, uint32_t n_shards, const float* const* fv, const uint64_t* const* keys, const uint64_t* n_kmers,
                        int32_t key_ranges, uint32_t key_bits, uint8_t* out_shard, uint32_t* out_index);

/* Stage 2 of one rank: DEVICE records (composite keys as above, unique) -> regroup sort -> CSR by k-mer -> filter
 * values -> order (equal composite keys are merged by maximum score).  d_keys/d_scores are overwritten (one side of the sort's ping-pong).
 * chunks_group_ordered != 0 promises that the records are a concatenation of chunks in ascending group ranges, each
 * sorted by (kmer, group) -- what the all-to-all of stably partitioned stage-1 records delivers -- so that a stable
 * sort on the k-mer bits alone suffices; 0 sorts on all key_bits + group_bits.
 * group_branch[n_groups_total] (host) maps global group index -> branch id. */
int  ipkgpu_stage2(ipkgpu_ctx* ctx, uint64_t* d_keys, float* d_scores, uint64_t n, uint32_t group_bits,
                   const uint32_t* group_branch, uint32_t n_groups_total, uint32_t key_bits,
                   float threshold, uint32_t total_num_groups, int32_t filter, int32_t chunks_group_ordered,
                   int result_on_device, ipkgpu_result* out);

/* ---- allocation-free entry point ---------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * ipkgpu_best_prefix_noalloc = the two kernels of ipkgpu_best_prefix (column maxima, then the sequential fp32 prefix
 * sums of matrix::preprocess) on DEVICE arrays of the caller: d_logp[n_nodes][width][sigma] in, d_B[n_nodes][width+1]
 * out, both on `device`, work ordered on `stream` (a cudaStream_t of that device, NULL = its default stream).  The call
 * returns once the kernels are queued: synchronize the stream before reading d_B.  Same bits as ipkgpu_best_prefix.
 * It needs no context and calls no cudaMalloc / cudaFree / allocating routine: the column maxima between the two
 * kernels live in a file-scope __device__ array of IPKGPU_NOALLOC_MAX_COLUMNS floats (16 MiB) that every GPU using
 * this library holds until the process exits.  n_nodes * width > IPKGPU_NOALLOC_MAX_COLUMNS is IPKGPU_EINVAL.
 * Every call on a device shares that array (each call overwrites the part it reads, nothing needs clearing): calls that
 * overlap on one GPU -- two streams, two threads -- race; order them on one stream.  With sigma == 4, d_logp must be
 * 16-byte aligned.  The message of the calling thread's last failure: ipkgpu_noalloc_last_error(). */
#define IPKGPU_NOALLOC_MAX_COLUMNS (1u << 22)
int  ipkgpu_best_prefix_noalloc(int device, void* stream, const float* d_logp, uint32_t n_nodes, uint32_t width, uint32_t sigma,
                                float* d_B);
const char* ipkgpu_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* IPKGPU_H */
