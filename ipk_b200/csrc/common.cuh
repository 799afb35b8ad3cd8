// common.cuh -- context, error plumbing and device scratch management of libipkgpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <cstring>
#include <string>
#include <vector>
#include <utility>
#include <cstdlib>
#include <algorithm>
#include <chrono>
#include "../../include/ipkgpu.h"

namespace ipkgpu
{
    struct error
    {
        int code;
        std::string what;
    };

    [[noreturn]] inline void fail(int code, const char* fmt, ...)
    {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        throw error{ code, buf };
    }

#define IPK_CUDA(expr)                                                                                       \
    do {                                                                                                     \
        cudaError_t e_ = (expr);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            ::ipkgpu::fail(e_ == cudaErrorMemoryAllocation ? IPKGPU_ENOMEM : IPKGPU_ECUDA, "%s failed: %s (%s:%d)", \
                           #expr, cudaGetErrorString(e_), __FILE__, __LINE__);                               \
    } while (0)

    /// Grow-only device buffer, reused across builds so that steady-state calls do no cudaMalloc.
    struct dbuf
    {
        void* p = nullptr;
        size_t cap = 0;
        template<class T> T* as() const { return static_cast<T*>(p); }
    };

    /// Grow-only pinned host buffer (results handed back to the caller live here).
    struct hbuf
    {
        void* p = nullptr;
        size_t cap = 0;
        bool external = false;             // the caller's page-locked memory (ipkgpu_set_entry_staging): never grown or freed here
        template<class T> T* as() const { return static_cast<T*>(p); }
    };

    constexpr int NUM_SMS_B200 = 148;

    /// Page-locked host memory.  cudaMallocHost pins 4 KB pages one by one (measured on the B200 hosts: 4.4 s for 8 GiB, most of
    /// a cold build); large blocks are instead anonymous mappings advised to transparent huge pages, first touched by several
    /// threads and then registered (0.12 s + 0.25 s for the same 8 GiB, tools/probe/pin_probe.cu).  Small blocks, and hosts where
    /// the mapping or the registration fails, take cudaMallocHost.  pinned_free takes either kind.
    cudaError_t pinned_alloc(void** out, size_t bytes);
    void pinned_free(void* p);
}

struct ipkgpu_ctx
{
    int device = 0;
    int num_sms = ipkgpu::NUM_SMS_B200;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint64_t workspace_limit = 0;      // 0 = 85 % of free memory, evaluated per build
    size_t dev_bytes = 0;              // bytes currently held by dbufs
    std::string last_error;
    uint32_t launches = 0;             // kernels launched by the current call

    // device scratch
    ipkgpu::dbuf logp, colmax, bestpref, node_group, node_order, group_branch, counters, node_counts;
    ipkgpu::dbuf keys_a, keys_b, vals_a, vals_b, hist, scan_tmp, block_counts, block_offs;
    ipkgpu::dbuf dbg_node, dbg_win, deferred, deferred2;
    ipkgpu::dbuf tile_status;          // ordered emission: look-back status word per tile of the join kernel's first level
    ipkgpu::dbuf side_keys_a, side_keys_b, side_vals_a, side_vals_b;   // ... and the side stream of the tiles it deferred
    ipkgpu::dbuf mg_split;             // merge: B-elements before every tile boundary of the merged (main, side) sequence
    int ordered = 0;                   // 1: ordered emission + k-mer-only sort (env IPKGPU_ORDERED=1); measured slower, see DESIGN.md
    ipkgpu::dbuf mif_tab;              // MIF0: bucket table of mif_log2 (mif_math.cuh)
    ipkgpu::dbuf mif_chunks;           // MIF0: first k-mer of every chunk of entries
    ipkgpu::dbuf rs_status;            // radix sort: look-back status words [tile][digit]
    ipkgpu::dbuf ps_hist;              // packed sort: per-bucket digit bases of the LSD passes [256][passes][256] + tickets
    ipkgpu::dbuf mg_status;            // merge: look-back status words [tile][2]
    uint32_t mg_epoch = 0;
    int rs_shape = 2;                  // tile shape of the sweep kernel (index into RS_SHAPES; env IPKGPU_RS_SHAPE)
    uint32_t rs_epoch = 0;             // ... and the pass counter that tags them
    int enum_mode = 0;                 // 0: list-join kernel + depth-first kernel on deferred tiles; 1: depth-first only
    std::vector<std::pair<uint32_t, uint32_t>> levels;   // join levels (pool, windows per item) from IPKGPU_LEVELS; empty = defaults
    uint32_t group_min = 20;           // join kernel: shortest list a lane group builds (env IPKGPU_GROUP_MIN)
    uint32_t small_list = 64;          // join kernel: longest list its owner thread fills and sorts alone (env IPKGPU_SMALL_LIST)
    uint32_t pool_cap = 0;             // list-pool capacity of the join kernel in elements; 0 = default per alphabet
    uint64_t last_tiles = 0, last_deferred_tiles = 0;   // of the current call
    ipkgpu::dbuf join_hist;            // per-SM partial top-byte histograms of the join kernel's emission (enum_args::top_hist)
    bool top_hist_on = false;          // ... counted by the current enumeration (valid as long as runs_valid is)
    ipkgpu::dbuf runs, runs_scratch;   // commits of the join kernel (enumerate.cuh: emit_run) and their group-ordered index (packed_sort.cuh)
    uint64_t runs_n = 0;               // runs recorded by the current enumeration
    bool runs_valid = false;           // ... and whether they cover every emitted record (join kernel only, nothing deferred to the depth-first kernel)
    uint64_t enum_join_emitted = 0;    // records the join kernel of the last enumeration launch emitted
    ipkgpu::dbuf u_keys, u_scores;                 // unique (kmer,group) records, composite keys
    ipkgpu::dbuf kmer_keys, kmer_offs, ent_branch, ent_score, fv, ord_keys_a, ord_keys_b, ord_vals_a, ord_vals_b;
    ipkgpu::dbuf out_keys, out_fv, out_offs, out_branch, out_score, out_cnt;
    ipkgpu::dbuf mo_buf;                           // ipkgpu_merge_order: filter values, sort ping-pong and the split order of all shards
    // stage-1 state kept for ipkgpu_stage1_partition
    uint64_t s1_n = 0;
    uint32_t s1_group_bits = 0;
    ipkgpu::dbuf recv_keys, recv_scores, peer_base;   // receive side of the fused partition + exchange (exported through CUDA IPC)
    struct ipc_mapping { unsigned char handle[64]; void* ptr = nullptr; };
    std::vector<ipc_mapping> ipc_open;  // peers' receive buffers mapped into this process
    std::vector<ipc_mapping> ipc_slots; // ... one per peer rank for the packed exchange: a changed handle replaces the mapping (ipkgpu_open_peer_slot)
    uint32_t s1_key_bits = 0;          // k * bits of the stage-1 build
    ipkgpu::dbuf recv_old2;            // (the score buffer of the pair exchange, same rule)
    ipkgpu::dbuf recv_old;             // a receive buffer that was outgrown: freed one exchange later, once no peer maps it any more
    uint64_t* s1_keys = nullptr;       // the records ipkgpu_stage1_partition buckets: u_keys (merged) or keys_a (raw)
    uint32_t* s1_scores = nullptr;
    // KEEP_POSITIONS builds (positions.cuh; ipkgpu_set_keep_positions)
    bool keep_positions = false;
    ipkgpu::dbuf ent_pos, out_pos, pos_flags, node_rank;
    ipkgpu::hbuf h_pos;
    const uint32_t* res_perm = nullptr;             // k-mer order of the last result (fv order -> index in key order), device
    const uint32_t* res_pos = nullptr;              // positions aligned with the last result's entries (host or device like the result), or null
    int res_pos_on_device = 0;
    uint64_t res_n_kmers = 0, res_n_entries = 0;      // the CSR the last build / stage 2 left in kmer_keys / kmer_offs / ent_* / fv
    bool res_valid = false;
    uint64_t hint_side = 0;            // ... and of its side stream
    uint64_t hint_emitted = 0;         // emitted volume of the previous build: sizes the first enumeration attempt
    uint64_t hint_kmers = 0, hint_kmers_records = 0;   // k-mers of the previous result and the records it was merged from (parts_eligible)
    // pinned host staging
    ipkgpu::hbuf h_small, h_keys, h_fv, h_offs, h_branch, h_score, h_cnt;
    uint32_t max_branch = 0;           // largest branch id of the current build (16-bit branch results)
    cudaEvent_t ev_merged = nullptr, ev_copied = nullptr;
    // host-bound indexed results: sort + merge in n_parts k-mer ranges once there are part_min records (env IPKGPU_PARTS,
    // IPKGPU_PART_MIN), each range's entries queued on the copy stream as soon as it is merged
    // host posteriors of at least h2d_chunk_min bytes go up in h2d_chunks node ranges (env IPKGPU_H2D_CHUNKS, IPKGPU_H2D_CHUNK_MIN)
    uint32_t h2d_chunks = 4;
    uint64_t h2d_chunk_min = 256ull << 20;
    std::vector<cudaEvent_t> ev_h2d;
    uint32_t n_parts = 8;
    uint32_t first_part_div = 1;       // first part this much smaller than the others (env IPKGPU_FIRST_PART_DIV); 4 and 8 measured 4 ms SLOWER on cfg3 (r02e)
    uint64_t part_min = 1ull << 27;
    bool entries_streamed = false, streamed_branch16 = false;
    cudaStream_t copy_stream = nullptr;   // indexed host results: the entries leave under the filter / ordering kernels

    std::vector<cudaEvent_t> events;

    // optional per-kernel timing (ipkgpu_kernel_profile): one event after every launch of the current call
    struct kprof_mark { const char* name; uint64_t bytes; };
    struct kprof_row { std::string name; uint32_t launches = 0; double ms = 0; uint64_t bytes = 0; };
    bool kprof_on = false;
    std::vector<cudaEvent_t> kprof_events;
    std::vector<kprof_mark> kprof_marks;       // kprof_marks[i] ended at kprof_events[i]; [0] = start of the call
    std::vector<kprof_row> kprof_rows;         // accumulated since the last reset

    void kprof_push(const char* name, uint64_t bytes)
    {
        if (!kprof_on) return;
        const size_t i = kprof_marks.size();
        while (kprof_events.size() <= i)
        {
            cudaEvent_t e;
            IPK_CUDA(cudaEventCreate(&e));
            kprof_events.push_back(e);
        }
        IPK_CUDA(cudaEventRecord(kprof_events[i], stream));
        kprof_marks.push_back({ name, bytes });
    }
    void kprof_add_bytes(const char* name, uint64_t bytes)
    {
        for (size_t i = kprof_marks.size(); i-- > 0;)
            if (kprof_marks[i].name && !strcmp(kprof_marks[i].name, name)) { kprof_marks[i].bytes += bytes; return; }
    }
    /// fold the marks of the finished call into kprof_rows (stream must be idle)
    void kprof_collect()
    {
        if (!kprof_on || kprof_marks.empty()) { kprof_marks.clear(); return; }
        IPK_CUDA(cudaEventSynchronize(kprof_events[kprof_marks.size() - 1]));
        for (size_t i = 1; i < kprof_marks.size(); ++i)
        {
            if (!kprof_marks[i].name) continue;            // a separator (host work / copies before it are not a kernel's)
            float ms = 0;
            IPK_CUDA(cudaEventElapsedTime(&ms, kprof_events[i - 1], kprof_events[i]));
            kprof_row* row = nullptr;
            for (auto& r : kprof_rows) if (r.name == kprof_marks[i].name) { row = &r; break; }
            if (!row) { kprof_rows.push_back({}); row = &kprof_rows.back(); row->name = kprof_marks[i].name; }
            row->launches++; row->ms += ms; row->bytes += kprof_marks[i].bytes;
        }
        kprof_marks.clear();
    }

    void reserve(ipkgpu::dbuf& b, size_t bytes)
    {
        if (bytes <= b.cap) return;
        if (b.p) { cudaFree(b.p); dev_bytes -= b.cap; b.p = nullptr; b.cap = 0; }
        // round up a little so that slowly growing inputs do not realloc on every call
        size_t want = bytes + bytes / 16 + 256;
        static const bool trace = getenv("IPKGPU_TRACE") != nullptr;
        const auto t_0 = std::chrono::steady_clock::now();
        cudaError_t e = cudaMalloc(&b.p, want);
        if (e != cudaSuccess) { want = bytes; e = cudaMalloc(&b.p, want); }
        if (trace && want >= (size_t(64) << 20))
            fprintf(stderr, "[ipkgpu] cudaMalloc %.2f GB: %.1f ms\n", double(want) * 1e-9, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_0).count());
        if (e != cudaSuccess)
        {
            cudaGetLastError();
            b.p = nullptr;
            ipkgpu::fail(IPKGPU_ENOMEM, "device allocation of %zu bytes failed (%s); %zu bytes already held",
                         bytes, cudaGetErrorString(e), dev_bytes);
        }
        b.cap = want;
        dev_bytes += want;
    }
    /// grow while preserving the first keep_bytes (accumulators that are appended to batch after batch)
    void reserve_keep(ipkgpu::dbuf& b, size_t bytes, size_t keep_bytes)
    {
        if (bytes <= b.cap) return;
        if (!b.p || keep_bytes == 0) { reserve(b, bytes); return; }
        ipkgpu::dbuf nb;
        reserve(nb, std::max(bytes, b.cap * 2));
        cudaError_t e = cudaMemcpyAsync(nb.p, b.p, keep_bytes, cudaMemcpyDeviceToDevice, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) { release(nb); ipkgpu::fail(IPKGPU_ECUDA, "device copy while growing a buffer failed: %s", cudaGetErrorString(e)); }
        release(b);
        b = nb;
    }
    void release(ipkgpu::dbuf& b)
    {
        if (b.p) { cudaFree(b.p); dev_bytes -= b.cap; }
        b.p = nullptr; b.cap = 0;
    }
    void reserve_host(ipkgpu::hbuf& b, size_t bytes)
    {
        if (bytes <= b.cap) return;
        if (b.external) ipkgpu::fail(IPKGPU_ENOMEM, "the caller's staging buffer holds %zu bytes, the result needs %zu (ipkgpu_set_entry_staging)", b.cap, bytes);
        if (b.p) ipkgpu::pinned_free(b.p);
        b.p = nullptr; b.cap = 0;
        size_t want = bytes + bytes / 16 + 256;
        cudaError_t e = ipkgpu::pinned_alloc(&b.p, want);
        if (e != cudaSuccess)
        {
            cudaGetLastError();
            b.p = nullptr;
            ipkgpu::fail(IPKGPU_ENOMEM, "pinned host allocation of %zu bytes failed (%s)", want, cudaGetErrorString(e));
        }
        b.cap = want;
    }
    cudaEvent_t event(size_t i)
    {
        while (events.size() <= i)
        {
            cudaEvent_t e;
            IPK_CUDA(cudaEventCreate(&e));
            events.push_back(e);
        }
        return events[i];
    }
};

namespace ipkgpu
{
    /// call right after a kernel launch; algo_bytes = algorithmic DRAM bytes of that launch (kernel table)
    inline void check_launch(ipkgpu_ctx* ctx, const char* name, uint64_t algo_bytes = 0)
    {
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) fail(IPKGPU_ECUDA, "launch of %s failed: %s", name, cudaGetErrorString(e));
        ctx->launches++;
        ctx->kprof_push(name, algo_bytes);
    }
    /// call right BEFORE a kernel launch that follows copies / memsets / host waits, so that those are not billed to it
    inline void kprof_fence(ipkgpu_ctx* ctx) { ctx->kprof_push(nullptr, 0); }
/// fence + launch + check: IPK_LAUNCH(ctx, "name", algorithmic_bytes, kernel<<<grid, block, smem, stream>>>(args...));
#define IPK_LAUNCH(ctx_, name_, bytes_, ...)                 \
    do {                                                     \
        ::ipkgpu::kprof_fence(ctx_);                         \
        __VA_ARGS__;                                         \
        ::ipkgpu::check_launch(ctx_, name_, bytes_);         \
    } while (0)

    inline uint32_t ceil_log2(uint64_t x)
    {
        uint32_t b = 0;
        while ((1ull << b) < x) ++b;
        return b;
    }
    template<class T> inline T div_up(T a, T b) { return (a + b - 1) / b; }
}
