// ipkgpu.cu -- C-ABI implementation (include/ipkgpu.h) of the B200 phylo-k-mer database build.  sm_100a only.
//
// Pipeline of one build (reference: db_builder::run, ipk/src/db_builder.cpp:182-338):
//   H2D posteriors -> K1 best_prefix -> [per batch of branch groups: K2 enumerate -> radix sort on (kmer|group) ->
//   segmented max (K3)] -> (regroup sort, only when there were several batches / an exchange) -> CSR by k-mer (K3')
//   -> K4 filter values -> K5 order by filter value -> K6 gather -> D2H.
// There is no CPU fallback anywhere in this file: every stage is a kernel from the .cuh files next to it.
#include "common.cuh"
#include "enumerate.cuh"
#include "join.cuh"
#include "radix_sort.cuh"
#include "merge.cuh"
#include "packed_sort.cuh"
#include "filter.cuh"
#include "digest.cuh"
#include "positions.cuh"

#include <chrono>
#include <cmath>
#include <numeric>
#include <unordered_map>

#include <mutex>
#include <thread>
#include <sys/mman.h>

using namespace ipkgpu;

namespace ipkgpu
{
    static std::mutex g_pin_mutex;
    static std::unordered_map<void*, std::pair<void*, size_t>> g_pin_maps;        // registered pointer -> (mapping base, mapping length)

    cudaError_t pinned_alloc(void** out, size_t bytes)
    {
        *out = nullptr;
        static const bool thp_off = getenv("IPKGPU_PIN_THP") && atoi(getenv("IPKGPU_PIN_THP")) == 0;
        const size_t HUGE = size_t(2) << 20;
        if (bytes >= (size_t(64) << 20) && !thp_off)
        {
            const size_t len = (bytes + HUGE - 1) & ~(HUGE - 1);
            void* base = mmap(nullptr, len + HUGE, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
            if (base != MAP_FAILED)
            {
                char* q = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(base) + HUGE - 1) & ~uintptr_t(HUGE - 1));
                madvise(q, len, MADV_HUGEPAGE);
                static const bool trace = getenv("IPKGPU_TRACE") != nullptr;          // where a cold build's host time goes (stderr)
                const auto t_0 = std::chrono::steady_clock::now();
                const unsigned T = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
                std::vector<std::thread> th;
                for (unsigned t = 0; t < T; ++t)
                    th.emplace_back([=] {
                        const size_t lo = len / T * t, hi = t + 1 == T ? len : len / T * (t + 1);
                        for (size_t o = lo; o < hi; o += 4096) q[o] = 0;
                    });
                for (auto& x : th) x.join();
                const auto t_1 = std::chrono::steady_clock::now();
                const cudaError_t reg = cudaHostRegister(q, len, cudaHostRegisterPortable);
                if (trace)
                {
                    // how much of the mapping the kernel backed with huge pages (smaps of the VMA that holds q)
                    double huge_kb = -1.0, size_kb = 0.0;
                    if (FILE* f = fopen("/proc/self/smaps", "r"))
                    {
                        char line[256];
                        bool mine = false;
                        while (fgets(line, sizeof line, f))
                        {
                            unsigned long lo = 0, hi = 0;
                            if (sscanf(line, "%lx-%lx ", &lo, &hi) == 2 && strchr(line, '-') && !strstr(line, "kB")) mine = uintptr_t(q) >= lo && uintptr_t(q) < hi;
                            else if (mine && !strncmp(line, "Size:", 5)) size_kb = atof(line + 5);
                            else if (mine && !strncmp(line, "AnonHugePages:", 14)) { huge_kb = atof(line + 14); break; }
                        }
                        fclose(f);
                    }
                    fprintf(stderr, "[ipkgpu] pinned_alloc: %.0f of %.0f MB of the mapping are huge pages\n", huge_kb / 1024.0, size_kb / 1024.0);
                }
                if (trace)
                    fprintf(stderr, "[ipkgpu] pinned_alloc %.2f GB: touch %.1f ms, cudaHostRegister %.1f ms\n", double(len) * 1e-9,
                            std::chrono::duration<double, std::milli>(t_1 - t_0).count(),
                            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_1).count());
                if (reg == cudaSuccess)
                {
                    std::lock_guard<std::mutex> l(g_pin_mutex);
                    g_pin_maps[q] = { base, len + HUGE };
                    *out = q;
                    return cudaSuccess;
                }
                cudaGetLastError();
                munmap(base, len + HUGE);
            }
        }
        return cudaMallocHost(out, bytes ? bytes : 1);
    }

    void pinned_free(void* p)
    {
        if (!p) return;
        std::pair<void*, size_t> m{ nullptr, 0 };
        {
            std::lock_guard<std::mutex> l(g_pin_mutex);
            const auto it = g_pin_maps.find(p);
            if (it != g_pin_maps.end()) { m = it->second; g_pin_maps.erase(it); }
        }
        if (m.first)
        {
            if (cudaHostUnregister(p) != cudaSuccess) cudaGetLastError();
            munmap(m.first, m.second);
        }
        else if (cudaFreeHost(p) != cudaSuccess) cudaGetLastError();
    }
}

namespace
{
    thread_local std::string g_create_error;

    /// group_ghost_ids (db_builder.cpp:524-553): groups in order of first appearance of the branch id; plus the
    /// node order the enumeration launches in (nodes sorted by group, stable).
    struct group_tables
    {
        std::vector<uint32_t> node_group, group_branch, node_order, group_first_slot;   // group_first_slot[g..n_groups]
    };

    group_tables make_groups(const uint32_t* node_branch, uint32_t n_nodes, uint32_t group_base)
    {
        group_tables t;
        t.node_group.resize(n_nodes);
        std::unordered_map<uint32_t, uint32_t> index;
        index.reserve(n_nodes * 2 + 1);
        for (uint32_t n = 0; n < n_nodes; ++n)
        {
            auto it = index.find(node_branch[n]);
            if (it == index.end())
            {
                it = index.emplace(node_branch[n], uint32_t(t.group_branch.size())).first;
                t.group_branch.push_back(node_branch[n]);
            }
            t.node_group[n] = it->second;
        }
        const uint32_t n_groups = uint32_t(t.group_branch.size());
        t.group_first_slot.assign(n_groups + 1, 0);
        for (uint32_t n = 0; n < n_nodes; ++n) t.group_first_slot[t.node_group[n] + 1]++;
        for (uint32_t g = 0; g < n_groups; ++g) t.group_first_slot[g + 1] += t.group_first_slot[g];
        t.node_order.resize(n_nodes);
        std::vector<uint32_t> cursor(t.group_first_slot.begin(), t.group_first_slot.end() - 1);
        for (uint32_t n = 0; n < n_nodes; ++n) t.node_order[cursor[t.node_group[n]]++] = n;
        for (auto& g : t.node_group) g += group_base;
        return t;
    }

    void validate(const ipkgpu_input* in)
    {
        if (!in) fail(IPKGPU_EINVAL, "input is null");
        if (in->n_nodes && (!in->logp || !in->node_branch)) fail(IPKGPU_EINVAL, "logp / node_branch is null");
        if (in->sigma < 2 || in->sigma > 32) fail(IPKGPU_EINVAL, "sigma = %u outside [2, 32]", in->sigma);
        if (in->bits < 1 || in->bits > 8 || (1u << in->bits) < in->sigma)
            fail(IPKGPU_EINVAL, "bits = %u cannot encode %u states", in->bits, in->sigma);
        if (in->k < 1 || in->k > uint32_t(MAX_K)) fail(IPKGPU_EINVAL, "k = %u outside [1, %d]", in->k, MAX_K);
        if (in->k * in->bits > 63) fail(IPKGPU_EINVAL, "k * bits = %u does not fit the 64-bit k-mer key", in->k * in->bits);
        if (in->filter != IPKGPU_FILTER_MIF0 && in->filter != IPKGPU_FILTER_RANDOM) fail(IPKGPU_EINVAL, "unknown filter %d", in->filter);
        if (uint64_t(in->n_nodes) * (uint64_t(in->width) + 1) > 0xffffffffull) fail(IPKGPU_EINVAL, "n_nodes * width too large");
        // k1a_colmax reads a DNA column as one float4
        if (in->sigma == 4 && in->n_nodes && (reinterpret_cast<uintptr_t>(in->logp) & 15) != 0) fail(IPKGPU_EINVAL, "logp must be 16-byte aligned for 4 states");
    }

    struct stage_timer
    {
        ipkgpu_ctx* ctx;
        size_t next = 0;
        explicit stage_timer(ipkgpu_ctx* c) : ctx(c) {}
        size_t mark()
        {
            cudaEvent_t e = ctx->event(next);
            IPK_CUDA(cudaEventRecord(e, ctx->stream));
            return next++;
        }
        float ms(size_t a, size_t b)
        {
            float t = 0;
            IPK_CUDA(cudaEventElapsedTime(&t, ctx->event(a), ctx->event(b)));
            return t;
        }
    };

    size_t free_device_bytes()
    {
        size_t f = 0, t = 0;
        IPK_CUDA(cudaMemGetInfo(&f, &t));
        return f;
    }

    /// device pointer to the posteriors (copying them in when they are on the host) + group tables on the device
    __global__ void k0_nop() {}

    /// node_bounds (optional): filled with c + 1 node indices when the host posteriors are uploaded in c > 1 chunks on the
    /// copy stream (ctx->ev_h2d[i] fires when chunk i has landed), so that K1 / K2 of a chunk run under the next upload.
    /// That needs launch slots == node indices (nodes arrive grouped by branch, the reference's own order).
    const float* stage_inputs(ipkgpu_ctx* ctx, const ipkgpu_input* in, const group_tables& t, ipkgpu_stats* st,
                              std::vector<uint32_t>* node_bounds = nullptr)
    {
        const size_t node_bytes = size_t(in->width) * in->sigma * sizeof(float);
        const size_t logp_bytes = size_t(in->n_nodes) * node_bytes;
        const float* d_logp = in->logp;
        // the small tables first: same-direction copies share one DMA queue, and these must not wait behind the posteriors
        const size_t nb = size_t(in->n_nodes) * 4;
        ctx->reserve(ctx->node_group, nb);
        ctx->reserve(ctx->node_order, nb);
        ctx->reserve(ctx->group_branch, t.group_branch.size() * 4 + 4);
        IPK_CUDA(cudaMemcpyAsync(ctx->node_group.p, t.node_group.data(), nb, cudaMemcpyHostToDevice, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(ctx->node_order.p, t.node_order.data(), nb, cudaMemcpyHostToDevice, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(ctx->group_branch.p, t.group_branch.data(), t.group_branch.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        ctx->max_branch = 0;
        for (uint32_t b : t.group_branch) ctx->max_branch = std::max(ctx->max_branch, b);
        if (st) st->h2d_bytes += 2 * nb + t.group_branch.size() * 4;
        if (!in->logp_on_device)
        {
            ctx->reserve(ctx->logp, logp_bytes);
            d_logp = ctx->logp.as<float>();
            bool in_order = node_bounds != nullptr && ctx->h2d_chunks > 1 && logp_bytes >= ctx->h2d_chunk_min && in->n_nodes >= ctx->h2d_chunks;
            for (uint32_t i = 0; in_order && i < in->n_nodes; ++i) in_order = t.node_order[i] == i;
            if (in_order)
            {
                if (!ctx->copy_stream) IPK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
                while (ctx->ev_h2d.size() < ctx->h2d_chunks)
                {
                    cudaEvent_t e;
                    IPK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    ctx->ev_h2d.push_back(e);
                }
                // (whatever read ctx->logp before has finished: every entry point ends with a stream synchronize)
                // The uploads are held back until the small tables above are through: the copy engine serves one channel's
                // queue at a time in no promised order, and the first kernels wait for those tables.
                while (ctx->ev_h2d.size() < ctx->h2d_chunks + 1)
                {
                    cudaEvent_t e;
                    IPK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    ctx->ev_h2d.push_back(e);
                }
                // (an empty kernel, so that the stream's next event records are not the copy engine's to process either)
                k0_nop<<<1, 32, 0, ctx->stream>>>();
                IPK_CUDA(cudaEventRecord(ctx->ev_h2d[ctx->h2d_chunks], ctx->stream));
                IPK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_h2d[ctx->h2d_chunks], 0));
                node_bounds->assign(1, 0u);
                for (uint32_t c = 0; c < ctx->h2d_chunks; ++c)
                {
                    const uint32_t n0 = node_bounds->back(), n1 = uint32_t(uint64_t(in->n_nodes) * (c + 1) / ctx->h2d_chunks);
                    IPK_CUDA(cudaMemcpyAsync(ctx->logp.as<char>() + size_t(n0) * node_bytes, reinterpret_cast<const char*>(in->logp) + size_t(n0) * node_bytes,
                                             size_t(n1 - n0) * node_bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
                    IPK_CUDA(cudaEventRecord(ctx->ev_h2d[c], ctx->copy_stream));
                    node_bounds->push_back(n1);
                }
            }
            else IPK_CUDA(cudaMemcpyAsync(ctx->logp.p, in->logp, logp_bytes, cudaMemcpyHostToDevice, ctx->stream));
            if (st) st->h2d_bytes += logp_bytes;
        }
        // the host vectors above are pageable: the copies are staged synchronously by the runtime, safe to drop after
        return d_logp;
    }

    /// K1: matrix::preprocess for every node (window.cpp:16-27)
    /// nodes [node0, node1) (default: all); a range keeps its transposed column maxima in its own slice of ctx->colmax
    void run_best_prefix(ipkgpu_ctx* ctx, const float* d_logp, const ipkgpu_input* in, uint32_t node0 = 0, uint32_t node1 = ~0u)
    {
        node1 = std::min(node1, in->n_nodes);
        ctx->reserve(ctx->colmax, uint64_t(in->n_nodes) * in->width * 4 + 4);
        ctx->reserve(ctx->bestpref, uint64_t(in->n_nodes) * (in->width + 1) * 4 + 4);
        if (node1 <= node0) return;
        const uint32_t nn = node1 - node0;
        const uint64_t cols = uint64_t(nn) * in->width;
        float* colmax = ctx->colmax.as<float>() + uint64_t(node0) * in->width;
        if (cols)
        {
            const uint32_t grid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(cols, 256), uint64_t(ctx->num_sms) * 16));
            IPK_LAUNCH(ctx, "k1a_colmax", cols * (in->sigma + 1) * 4,
                       k1a_colmax<<<grid, 256, 0, ctx->stream>>>(d_logp + uint64_t(node0) * in->width * in->sigma, colmax, nn, in->width, in->sigma));
        }
        IPK_LAUNCH(ctx, "k1b_prefix", cols * 8 + uint64_t(nn) * 4,
                   k1b_prefix<<<div_up<uint32_t>(nn, 64), 64, 0, ctx->stream>>>(colmax, ctx->bestpref.as<float>() + uint64_t(node0) * (in->width + 1), nn, in->width));
    }
}

namespace
{
    struct enum_launch
    {
        uint32_t slot_begin, slot_end;     // launch slots (positions in node_order)
        uint32_t group_sub, group_bits;    // batch-local group field
        uint64_t cap;
        uint64_t* out_keys; float* out_scores;
        uint32_t* out_node = nullptr; uint32_t* out_win = nullptr;     // debug
        bool debug = false;
        // ordered emission: level 0 of the join kernel writes (out_keys, out_scores) in tile order; the tiles it defers are
        // enumerated into this side stream (any order).  side_keys == nullptr: one stream, any order.
        uint64_t* side_keys = nullptr; float* side_scores = nullptr; uint64_t side_cap = 0;
        uint64_t emitted_before = 0;       // records already in (out_keys, out_scores): a later node chunk of the same enumeration
        bool want_runs = false;            // record every commit of the join kernel in ctx->runs (counters[5] = their number, zeroed by the caller)
    };

    uint64_t read_u64(ipkgpu_ctx* ctx, const void* d);

    /// largest list pool one CTA of the join kernel can have (the whole shared memory of an SM)
    uint32_t max_pool_cap(const ipkgpu_input* in, int n_fam)
    {
        const size_t fixed = make_join_layout(int(in->k), int(in->sigma), n_fam, 0).total;
        const size_t avail = size_t(225) * 1024 > fixed ? size_t(225) * 1024 - fixed : 0;      // (227 KB per CTA less the kernel's 1.5 KB of static shared memory)
        return uint32_t(std::min<size_t>(avail / 8, 65000));
    }

    struct enum_level { uint32_t pool, width; };       // pool 0 = the largest one CTA can hold

    /// levels of the join kernel: IPKGPU_LEVELS="pool:width,pool:width,..." (tuning knob), an explicit pool from
    /// ipkgpu_set_enumeration (a single level), or the defaults per alphabet
    std::vector<enum_level> enumeration_levels(const ipkgpu_ctx* ctx, const ipkgpu_input* in)
    {
        if (ctx->pool_cap) return { { ctx->pool_cap, uint32_t(ENUM_THREADS) } };
        if (!ctx->levels.empty())
        {
            std::vector<enum_level> v;
            for (auto& p : ctx->levels) v.push_back({ p.first, p.second });
            return v;
        }
        // (the first retry takes the 1.3 % of cfg3's tiles whose lists overflow 3072 elements: with a 12288-element pool they ran at
        // 2 CTAs/SM for 2.6 ms, with 4608 at 4 CTAs/SM for 1.5 ms -- profiles/r03w_sweep_levels.txt)
        if (in->sigma <= 4) return { { 3072, 128 }, { 4608, 128 }, { 12288, 128 }, { 0, 128 }, { 0, 16 }, { 0, 1 } };
        return { { 20000, 128 }, { 0, 128 }, { 0, 16 }, { 0, 1 } };
    }

    /// K2 over a range of launch slots: the list-join kernel, then the depth-first kernel on the tiles it deferred
    /// (every tile when k == 1 or ctx->enum_mode == 1).  counters[0] must have been zeroed; node_counts zeroed by the
    /// caller if used.  Blocks until the join kernel's deferred count is known.
    void launch_enumerate(ipkgpu_ctx* ctx, const float* d_logp, const ipkgpu_input* in, const enum_program& P, const enum_launch& L,
                          bool with_node_counts)
    {
        if (in->width < in->k || L.slot_end <= L.slot_begin) return;
        enum_args A;
        memset(&A, 0, sizeof A);
        A.logp = d_logp;
        A.B = ctx->bestpref.as<float>();
        A.node_order = ctx->node_order.as<uint32_t>();
        A.node_group = ctx->node_group.as<uint32_t>();
        A.out_keys = L.out_keys;
        A.out_scores = L.out_scores;
        A.out_node = L.out_node;
        A.out_win = L.out_win;
        A.counter = ctx->counters.as<unsigned long long>();
        A.node_counts = with_node_counts ? ctx->node_counts.as<unsigned long long>() : nullptr;
        A.cap = L.cap;
        A.slot_base = L.slot_begin;
        A.group_sub = L.group_sub;
        A.n_nodes = in->n_nodes; A.W = in->width; A.sigma = in->sigma; A.bits = in->bits;
        A.group_bits = L.group_bits;
        A.nwin = in->width - in->k + 1;
        A.tiles_per_node = div_up<uint32_t>(A.nwin, ENUM_THREADS);
        A.log_threshold = in->log_threshold;
        // TMA bulk copies need 16-byte aligned source, destination and size
        A.use_tma = ((in->sigma * 4) % 16 == 0) && ((reinterpret_cast<uintptr_t>(d_logp) & 15) == 0);
        const uint64_t grid = uint64_t(L.slot_end - L.slot_begin) * A.tiles_per_node;
        if (grid > 0x7fffffffull) fail(IPKGPU_EINVAL, "enumeration grid of %llu tiles is too large", (unsigned long long)grid);
        // algorithmic bytes: the posterior columns and prefix sums of the launched nodes; the emitted records are added
        // once their number is known (kprof_add_bytes)
        const uint64_t in_bytes = uint64_t(L.slot_end - L.slot_begin) * (uint64_t(in->width) * in->sigma + in->width + 1) * 4;

        uint64_t n_deferred = grid;
        ctx->enum_join_emitted = 0;
        const bool use_join = in->k >= 2 && ctx->enum_mode != 1;
        A.item_windows = ENUM_THREADS; A.items_per_tile = 1; A.parent_ratio = 1;
        A.group_min = std::min<uint32_t>(std::max<uint32_t>(ctx->group_min, 1), 31);
        A.small_list = std::min<uint32_t>(ctx->small_list, 8 * JOIN_GROUP);      // build_list_group holds 8 elements per lane
        static const bool phase_dbg = getenv("IPKGPU_PHASE_CLOCKS") != nullptr;
        if (phase_dbg && use_join)
        {
            ctx->reserve(ctx->peer_base, 1024);
            IPK_CUDA(cudaMemsetAsync(ctx->peer_base.p, 0, 256, ctx->stream));
            A.phase_clocks = ctx->peer_base.as<unsigned long long>();
        }
        const bool ordered = use_join && !L.debug && L.side_keys != nullptr;
        unsigned long long* const counters = ctx->counters.as<unsigned long long>();
        const bool record_runs = L.want_runs && use_join && !ordered && !L.debug;
        if (L.want_runs && !record_runs) ctx->runs_valid = false;
        if (ordered)
        {
            ctx->reserve(ctx->tile_status, grid * 8);
            IPK_CUDA(cudaMemsetAsync(ctx->tile_status.p, 0, grid * 8, ctx->stream));
            IPK_CUDA(cudaMemsetAsync(counters + 8, 0, 16, ctx->stream));
            A.tile_status = ctx->tile_status.as<unsigned long long>();
        }
        if (use_join)
        {
            // Levels: every tile with a small pool (several CTAs per SM); what did not fit is retried with larger pools, then
            // in narrower runs of windows (shorter lists); what is left after the last level goes to the depth-first kernel.
            const join_program J = make_join_program(int(in->k));
            const uint32_t min_pool = join_min_pool(int(in->k), int(in->sigma));
            const uint32_t pool_max = std::max(max_pool_cap(in, J.n_fam), min_pool);
            std::vector<enum_level> levels = enumeration_levels(ctx, in);
            ctx->reserve(ctx->deferred, grid * 4);
            uint32_t* lists[2] = { nullptr, nullptr };
            uint64_t n_items = grid, emitted_before = L.emitted_before;            // items of the level about to run
            uint32_t prev_width = ENUM_THREADS;
            for (size_t level = 0; level < levels.size() && n_items > 0; ++level)
            {
                const uint32_t width = std::min<uint32_t>(levels[level].width, prev_width);
                const uint32_t ratio = prev_width / width;
                const uint64_t lvl_grid = n_items * ratio;
                if (lvl_grid > 0x7fffffffull) break;
                A.pool_cap = std::min(pool_max, std::max(levels[level].pool ? levels[level].pool : pool_max, min_pool));
                A.item_windows = width; A.items_per_tile = ENUM_THREADS / width; A.parent_ratio = ratio;
                A.tile_list = level == 0 ? nullptr : lists[(level - 1) & 1];
                dbuf& out_list = (level & 1) ? ctx->deferred2 : ctx->deferred;
                ctx->reserve(out_list, lvl_grid * 4);
                lists[level & 1] = out_list.as<uint32_t>();
                A.deferred_tiles = lists[level & 1];
                if (record_runs)
                {
                    // every item of this level commits at most once
                    ctx->reserve_keep(ctx->runs, (ctx->runs_n + lvl_grid) * sizeof(emit_run), ctx->runs_n * sizeof(emit_run));
                    A.runs = ctx->runs.as<emit_run>();
                    A.run_counter = counters + 5;
                    A.run_cap = uint32_t(std::min<uint64_t>(ctx->runs.cap / sizeof(emit_run), 0xffffffffull));
                    if (ctx->top_hist_on)
                    {
                        A.top_hist = ctx->join_hist.as<uint32_t>();
                        A.top_shift = uint32_t(J.root_hl) * in->bits - 8;
                    }
                }
                IPK_CUDA(cudaMemsetAsync(A.counter + 1, 0, 8, ctx->stream));
                const size_t smem = make_join_layout(int(in->k), int(in->sigma), J.n_fam, A.pool_cap).total;
                if (smem > 227 * 1024) fail(IPKGPU_EINVAL, "join kernel needs %zu bytes of shared memory", smem);
                const uint64_t lvl_in_bytes = level == 0 ? in_bytes : lvl_grid * uint64_t(width + in->k) * (in->sigma + 1) * 4;
                // (the retries of deferred items are their own launch class in the kernel profile: a few thousand CTAs against half a million)
                static const char* const level_names[] = { "k2_join", "k2_join_retry1", "k2_join_retry2", "k2_join_retry3", "k2_join_retry4" };
                const char* const kname = level_names[std::min<size_t>(level, 4)];
                if (L.debug)
                {
                    IPK_CUDA(cudaFuncSetAttribute(k2_join<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
                    IPK_LAUNCH(ctx, kname, lvl_in_bytes, k2_join<true><<<uint32_t(lvl_grid), JOIN_THREADS, smem, ctx->stream>>>(A, P, J));
                }
                else
                {
                    IPK_CUDA(cudaFuncSetAttribute(k2_join<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
                    IPK_LAUNCH(ctx, kname, lvl_in_bytes, k2_join<false><<<uint32_t(lvl_grid), JOIN_THREADS, smem, ctx->stream>>>(A, P, J));
                }
                uint64_t h[6] = { 0, 0, 0, 0, 0, 0 };
                IPK_CUDA(cudaMemcpyAsync(h, A.counter, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
                IPK_CUDA(cudaStreamSynchronize(ctx->stream));
                if (record_runs) ctx->runs_n = h[5];
                ctx->kprof_add_bytes(kname, 12 * (std::min<uint64_t>(h[0], A.cap) - std::min<uint64_t>(emitted_before, A.cap)));
                emitted_before = h[0];
                n_items = h[1];
                prev_width = width;
                A.tile_list = lists[level & 1];
                if (ordered && level == 0)
                {
                    // everything after the ordered level goes to the side stream, with its own counters
                    A.tile_status = nullptr;
                    A.out_keys = L.side_keys; A.out_scores = L.side_scores; A.cap = L.side_cap;
                    A.counter = counters + 8;
                    emitted_before = 0;
                }
            }
            if (A.phase_clocks)
            {
                unsigned long long pc[32];
                IPK_CUDA(cudaMemcpyAsync(pc, A.phase_clocks, sizeof pc, cudaMemcpyDeviceToHost, ctx->stream));
                IPK_CUDA(cudaStreamSynchronize(ctx->stream));
                fprintf(stderr, "[ipkgpu] join phases, mean cycles of thread 0 per tile (%llu tiles):", pc[31]);
                for (int i = 0; i < 16; ++i) fprintf(stderr, " %d:%.0f", i, pc[31] ? double(pc[i]) / double(pc[31]) : 0.0);
                fprintf(stderr, "\n");
            }
            // the depth-first kernel takes the remaining items as they are
            A.parent_ratio = 1;
            n_deferred = n_items;
            ctx->enum_join_emitted = ordered ? 0 : emitted_before;      // (only used for the byte accounting of the depth-first kernel)
            ctx->last_deferred_tiles += n_deferred;
            ctx->last_tiles += grid;
        }
        if (!use_join) { ctx->last_deferred_tiles += grid; ctx->last_tiles += grid; }
        if (n_deferred == 0) return;
        ctx->runs_valid = false;           // the depth-first kernel commits per warp, unrecorded
        const size_t smem = enum_smem_bytes(int(in->k), int(in->sigma), L.debug);
        const uint64_t dfs_in_bytes = use_join ? n_deferred * uint64_t(ENUM_THREADS + in->k) * (in->sigma + 1) * 4 : in_bytes;
        if (L.debug)
        {
            IPK_CUDA(cudaFuncSetAttribute(k2_enumerate<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            IPK_LAUNCH(ctx, "k2_enumerate", dfs_in_bytes, k2_enumerate<true><<<uint32_t(n_deferred), ENUM_THREADS, smem, ctx->stream>>>(A, P));
        }
        else
        {
            IPK_CUDA(cudaFuncSetAttribute(k2_enumerate<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            IPK_LAUNCH(ctx, "k2_enumerate", dfs_in_bytes, k2_enumerate<false><<<uint32_t(n_deferred), ENUM_THREADS, smem, ctx->stream>>>(A, P));
        }
    }

    uint64_t read_u64(ipkgpu_ctx* ctx, const void* d)
    {
        uint64_t v = 0;
        IPK_CUDA(cudaMemcpyAsync(&v, d, 8, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        return v;
    }

    struct merge_out
    {
        uint64_t n_records = 0, n_kmers = 0;
    };

    /// K3 on n sorted records (keys, scores): segmented max over equal composite keys, compaction, and -- when
    /// `csr` -- the CSR by k-mer (K3').  Unique records go to (u_keys, u_scores) + u_off when `unique`.
    merge_out run_merge(ipkgpu_ctx* ctx, const uint64_t* keys, const float* scores, uint64_t n_main, uint32_t group_bits,
                        uint32_t out_group_bits, uint32_t group_add, bool unique, uint64_t u_off, bool csr, uint32_t key_bits,
                        const uint64_t* side_keys = nullptr, const float* side_scores = nullptr, uint64_t n_side = 0,
                        uint64_t ent_base = 0, uint64_t kmer_base = 0, uint64_t n_total = 0)
    {
        // (ent_base, kmer_base, n_total): this call appends one k-mer range to a CSR of at most n_total entries that earlier
        // calls began -- the outputs start at those bases and the buffers are sized once, for the whole
        merge_out r;
        const uint64_t n = n_main + n_side;          // the kernel walks the merge of the two sorted streams
        if (n == 0) return r;
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu records exceed the 2^32-1 limit of one merge on one device: split the build over devices (ipk-b200 --devices, key shards stay below the limit)", (unsigned long long)n);
        const uint32_t tiles = uint32_t(div_up<uint64_t>(n, MG_TILE));
        // single pass: the outputs are sized by their upper bounds (records <= n, k-mers <= min(n, 2^key_bits))
        const uint64_t n_cap = std::max(n_total, ent_base + n);
        const uint64_t kmer_bound = key_bits < 40 ? std::min<uint64_t>(n_cap, 1ull << key_bits) : n_cap;
        if (unique)
        {
            ctx->reserve_keep(ctx->u_keys, (u_off + n) * 8, u_off * 8);
            ctx->reserve_keep(ctx->u_scores, (u_off + n) * 4, u_off * 4);
        }
        if (csr)
        {
            ctx->reserve(ctx->kmer_keys, kmer_bound * 8 + 8);
            ctx->reserve(ctx->kmer_offs, (kmer_bound + 1) * 8);
            ctx->reserve(ctx->ent_branch, n_cap * 4 + 4);
            ctx->reserve(ctx->ent_score, n_cap * 4 + 4);
        }
        ctx->reserve(ctx->counters, 128);
        const size_t status_bytes = size_t(tiles) * 16;
        if (status_bytes > ctx->mg_status.cap)
        {
            ctx->reserve(ctx->mg_status, status_bytes);
            IPK_CUDA(cudaMemsetAsync(ctx->mg_status.p, 0, ctx->mg_status.cap, ctx->stream));
            ctx->mg_epoch = 0;
        }
        if (++ctx->mg_epoch >= (1u << 30))
        {
            IPK_CUDA(cudaMemsetAsync(ctx->mg_status.p, 0, ctx->mg_status.cap, ctx->stream));
            ctx->mg_epoch = 1;
        }
        unsigned long long* totals = ctx->counters.as<unsigned long long>() + 2;      // [2], [3]: totals; [4]: ticket
        IPK_CUDA(cudaMemsetAsync(totals, 0, 24, ctx->stream));
        merge_args M;
        M.ckeys = keys; M.scores = scores; M.n = n;
        M.skeys = side_keys; M.sscores = side_scores; M.ns = n_side; M.split = nullptr;
        if (n_side)
        {
            ctx->reserve(ctx->mg_split, (size_t(tiles) + 1) * 4);
            M.split = ctx->mg_split.as<uint32_t>();
            IPK_LAUNCH(ctx, "mg_split", (uint64_t(tiles) + 1) * 20,
                       mg_split<<<div_up<uint32_t>(tiles + 1, 256), 256, 0, ctx->stream>>>(keys, n_main, side_keys, n_side, tiles, ctx->mg_split.as<uint32_t>()));
        }
        M.group_bits = group_bits; M.out_group_bits = out_group_bits; M.group_add = group_add;
        M.u_keys = unique ? ctx->u_keys.as<uint64_t>() + u_off : nullptr;
        M.u_scores = unique ? ctx->u_scores.as<float>() + u_off : nullptr;
        M.group_branch = ctx->group_branch.as<uint32_t>();
        M.kmer_keys = csr ? ctx->kmer_keys.as<uint64_t>() + kmer_base : nullptr;
        M.kmer_offs = ctx->kmer_offs.as<uint64_t>() + kmer_base;
        M.ent_branch = ctx->ent_branch.as<uint32_t>() + ent_base;
        M.ent_score = ctx->ent_score.as<float>() + ent_base;
        M.offs_add = ent_base;
        M.status = ctx->mg_status.as<uint64_t>();
        M.ticket = reinterpret_cast<uint32_t*>(totals + 2);
        M.totals = totals;
        M.epoch = ctx->mg_epoch;
        IPK_LAUNCH(ctx, "mg_merge", n * 12, mg_merge<<<tiles, MG_THREADS, 0, ctx->stream>>>(M));
        uint64_t h[2];
        IPK_CUDA(cudaMemcpyAsync(h, totals, 16, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        r.n_records = h[0];
        r.n_kmers = h[1];
        ctx->kprof_add_bytes("mg_merge", (unique ? r.n_records * 12 : 0) + (csr ? r.n_records * 8 + r.n_kmers * 16 : 0));
        if (csr)
        {
            const uint64_t end = ent_base + r.n_records;
            IPK_CUDA(cudaMemcpyAsync(ctx->kmer_offs.as<uint64_t>() + kmer_base + r.n_kmers, &end, 8, cudaMemcpyHostToDevice, ctx->stream));
        }
        return r;
    }
}

namespace
{
    struct parts_out
    {
        merge_out m;
        uint32_t passes = 0;
        float ms_sort = 0, ms_merge = 0;
    };

    /// Host-bound indexed results of at least ctx->part_min records are sorted and merged in k-mer ranges -- when the ENTRIES
    /// are what crosses PCIe.  The k-mer heads (24 bytes each) only exist after the ordering, at the very end; a database
    /// of few entries per k-mer (cfg5: 2.5) gains nothing from streaming its entries and pays for the shorter launches
    /// (110 -> 119 ms), so ranges need >= 8 records per expected k-mer: the previous build's count on this context if it was
    /// of similar size, else the bound the caller knows (sigma^k or 2^key_bits).  r02i / r02k A/B runs under profiles/.
    bool parts_eligible(const ipkgpu_ctx* ctx, uint64_t n, uint32_t key_bits, uint32_t n_bits, int result_mode, uint64_t kmer_bound)
    {
        if ((result_mode & 3) != IPKGPU_RESULT_HOST_INDEXED || ctx->n_parts <= 1 || n < ctx->part_min || key_bits < 8 || n_bits <= 8 || n > 0xffffffffull)
            return false;
        uint64_t kmers = std::min(n, kmer_bound);
        if (ctx->hint_kmers_records && n / 2 <= ctx->hint_kmers_records && ctx->hint_kmers_records <= 2 * n)
            kmers = std::min<uint64_t>(kmers, uint64_t(double(ctx->hint_kmers) * double(n) / double(ctx->hint_kmers_records)) + 1);
        return n / 8 >= kmers || ctx->part_min <= 1;
    }

    /// Sort + max-merge + CSR of n records in k-mer RANGES, so that the entries start crossing PCIe while later ranges are
    /// still being sorted: one most-significant-digit pass splits (ka, va) into 256 ranges of the composite key in (kb, vb);
    /// the ranges are then sorted on the bits below (in aligned units, see inside), and runs of about n / n_parts records
    /// ("parts") are max-merged onto the end of the CSR the earlier parts began and handed to the copy stream.  The pass
    /// count equals the plain sort's (1 + ceil((n_bits - 8) / 8)); the extra cost is a second read of the keys for the
    /// per-unit digit histograms and the shorter launches.  Needs key_bits >= 8, so that no k-mer spans two ranges.
    /// On return every entry copy is queued behind ctx->ev_copied and ctx->entries_streamed tells finish_from_csr so.
    parts_out sort_merge_parts(ipkgpu_ctx* ctx, uint64_t* ka, uint32_t* va, uint64_t* kb, uint32_t* vb, uint64_t n, uint32_t key_bits,
                               uint32_t gbits, uint32_t out_group_bits, uint32_t group_add, int result_mode, stage_timer& tm)
    {
        parts_out R;
        const uint32_t low_bits = key_bits + gbits - 8;
        const bool b16 = (result_mode & 4) != 0 && ctx->max_branch < 65536u;
        // everything the copies touch is sized now, by the record count (an upper bound of the entry count)
        if (!ctx->copy_stream) IPK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        if (!ctx->ev_merged) { IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_merged, cudaEventDisableTiming)); IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming)); }
        ctx->reserve_host(ctx->h_branch, n * (b16 ? 2 : 4) + 8);
        ctx->reserve_host(ctx->h_score, n * 4 + 4);
        if (b16) ctx->reserve(ctx->out_branch, n * 2 + 8);

        uint32_t start[257];
        const size_t e0 = tm.mark();
        radix_partition_top(ctx, ka, va, kb, vb, n, low_bits, ctx->hist, start);
        start[256] = uint32_t(n);
        const size_t e1 = tm.mark();
        R.passes = 1;
        struct mark3 { size_t a, b, c; };
        std::vector<mark3> marks;
        // The passes below the top digit cover 8 * lsd_passes >= low_bits bits; the `slack` bits they have to spare are spent
        // on the lowest bits of the top digit, so a sort UNIT is an aligned run of 2^slack digit ranges, sorted on exactly
        // 8 * lsd_passes bits.  A part = a run of whole units of about n / n_parts records: one merge, one copy.
        const uint32_t lsd_passes = div_up<uint32_t>(low_bits, 8), unit_bits = 8 * lsd_passes;
        const uint32_t unit = 1u << (unit_bits - low_bits);                 // digit ranges per unit (1 .. 128)
        const uint64_t target = div_up<uint64_t>(n, ctx->n_parts);
        uint64_t ent_base = 0, kmer_base = 0;
        for (uint32_t d0 = 0; d0 < 256;)
        {
            uint32_t d1 = d0 + unit;
            // (a smaller first part would start the copy chain earlier; on cfg3 the copy engine then idles until the second part is merged and the build ends later -- r02e)
            while (d1 < 256 && uint64_t(start[d1]) - start[d0] < (ent_base == 0 && d0 == 0 ? target / ctx->first_part_div : target)) d1 += unit;
            const uint64_t o = start[d0], c = uint64_t(start[d1]) - start[d0];
            const uint32_t first = d0;
            d0 = d1;
            if (c == 0) continue;
            const size_t ea = tm.mark();
            for (uint32_t u = first; u < d1; u += unit)
            {
                const uint64_t uo = start[u], uc = uint64_t(start[u + unit]) - start[u];
                radix_sort_pairs(ctx, kb + uo, vb + uo, ka + uo, va + uo, uc, unit_bits, ctx->hist, ctx->scan_tmp);
            }
            const bool in_a = (lsd_passes & 1) != 0;                           // every unit ping-pongs the same number of times
            const size_t eb = tm.mark();
            const uint64_t* sk = in_a ? ka + o : kb + o;
            const float* sv = reinterpret_cast<const float*>(in_a ? va + o : vb + o);
            const merge_out m = run_merge(ctx, sk, sv, c, gbits, out_group_bits, group_add, false, 0, true, key_bits, nullptr, nullptr, 0,
                                          ent_base, kmer_base, n);
            marks.push_back({ ea, eb, tm.mark() });
            if (m.n_records)
            {
                IPK_CUDA(cudaEventRecord(ctx->ev_merged, ctx->stream));
                IPK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_merged, 0));
                if (b16)
                {
                    k6_narrow_branch<<<uint32_t(std::min<uint64_t>(div_up<uint64_t>(m.n_records, 1024), 1u << 20)), 256, 0, ctx->copy_stream>>>(
                        ctx->ent_branch.as<uint32_t>(), ent_base, ent_base + m.n_records, ctx->out_branch.as<uint16_t>());
                    check_launch(ctx, "k6_narrow_branch", 0);
                    IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.as<uint16_t>() + ent_base, ctx->out_branch.as<uint16_t>() + ent_base, m.n_records * 2,
                                             cudaMemcpyDeviceToHost, ctx->copy_stream));
                }
                else
                    IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.as<uint32_t>() + ent_base, ctx->ent_branch.as<uint32_t>() + ent_base, m.n_records * 4,
                                             cudaMemcpyDeviceToHost, ctx->copy_stream));
                IPK_CUDA(cudaMemcpyAsync(ctx->h_score.as<float>() + ent_base, ctx->ent_score.as<float>() + ent_base, m.n_records * 4,
                                         cudaMemcpyDeviceToHost, ctx->copy_stream));
            }
            ent_base += m.n_records;
            kmer_base += m.n_kmers;
        }
        IPK_CUDA(cudaEventRecord(ctx->ev_copied, ctx->copy_stream));
        ctx->entries_streamed = true;
        ctx->streamed_branch16 = b16;
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        R.passes += lsd_passes;
        R.m.n_records = ent_base;
        R.m.n_kmers = kmer_base;
        R.ms_sort = tm.ms(e0, e1);
        for (const mark3& k : marks) { R.ms_sort += tm.ms(k.a, k.b); R.ms_merge += tm.ms(k.b, k.c); }
        return R;
    }
}

namespace
{
    /// queues the D2H copy of the entries [ent_base, ent_base + n) of the CSR on the copy stream, behind whatever the main
    /// stream has done so far (host-bound indexed results: the entries leave while later k-mer ranges are still sorted)
    void queue_entry_copy(ipkgpu_ctx* ctx, uint64_t ent_base, uint64_t n, bool b16)
    {
        if (n == 0) return;
        IPK_CUDA(cudaEventRecord(ctx->ev_merged, ctx->stream));
        IPK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_merged, 0));
        if (b16)
        {
            k6_narrow_branch<<<uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, 1024), 1u << 20)), 256, 0, ctx->copy_stream>>>(
                ctx->ent_branch.as<uint32_t>(), ent_base, ent_base + n, ctx->out_branch.as<uint16_t>());
            check_launch(ctx, "k6_narrow_branch", 0);
            IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.as<uint16_t>() + ent_base, ctx->out_branch.as<uint16_t>() + ent_base, n * 2,
                                     cudaMemcpyDeviceToHost, ctx->copy_stream));
        }
        else
            IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.as<uint32_t>() + ent_base, ctx->ent_branch.as<uint32_t>() + ent_base, n * 4,
                                     cudaMemcpyDeviceToHost, ctx->copy_stream));
        IPK_CUDA(cudaMemcpyAsync(ctx->h_score.as<float>() + ent_base, ctx->ent_score.as<float>() + ent_base, n * 4,
                                 cudaMemcpyDeviceToHost, ctx->copy_stream));
    }

    /// K3 + K3' over the sorted packed records of buckets [d0, d1) (packed_sort.cuh), appended to the CSR at (ent_base, kmer_base)
    merge_out run_merge_packed(ipkgpu_ctx* ctx, const uint64_t* recs, ps_table& T, uint32_t d0, uint32_t d1, uint32_t key_bits, uint32_t gbits,
                               uint32_t group_add, uint64_t ent_base, uint64_t kmer_base, uint64_t n_total)
    {
        merge_out r;
        const uint64_t n = uint64_t(T.start[d1]) - T.start[d0];
        if (n == 0) return r;
        ps_fill_tiles(T, d0, d1, MP_TILE);
        const uint32_t tiles = T.tile_first[d1];
        const uint64_t n_cap = std::max(n_total, ent_base + n);
        const uint64_t kmer_bound = key_bits < 40 ? std::min<uint64_t>(n_cap, 1ull << key_bits) : n_cap;
        ctx->reserve(ctx->kmer_keys, kmer_bound * 8 + 8);
        ctx->reserve(ctx->kmer_offs, (kmer_bound + 1) * 8);
        ctx->reserve(ctx->ent_branch, n_cap * 4 + 4);
        ctx->reserve(ctx->ent_score, n_cap * 4 + 4);
        ctx->reserve(ctx->counters, 128);
        const size_t status_bytes = size_t(tiles) * 16;
        if (status_bytes > ctx->mg_status.cap)
        {
            ctx->reserve(ctx->mg_status, status_bytes);
            IPK_CUDA(cudaMemsetAsync(ctx->mg_status.p, 0, ctx->mg_status.cap, ctx->stream));
            ctx->mg_epoch = 0;
        }
        if (++ctx->mg_epoch >= (1u << 30))
        {
            IPK_CUDA(cudaMemsetAsync(ctx->mg_status.p, 0, ctx->mg_status.cap, ctx->stream));
            ctx->mg_epoch = 1;
        }
        unsigned long long* totals = ctx->counters.as<unsigned long long>() + 2;      // [2], [3]: totals; [4]: ticket
        IPK_CUDA(cudaMemsetAsync(totals, 0, 24, ctx->stream));
        mergep_args M;
        M.recs = recs; M.group_bits = gbits; M.kmer_low_bits = key_bits - 8; M.group_add = group_add;
        M.group_branch = ctx->group_branch.as<uint32_t>();
        M.kmer_keys = ctx->kmer_keys.as<uint64_t>() + kmer_base;
        M.kmer_offs = ctx->kmer_offs.as<uint64_t>() + kmer_base;
        M.ent_branch = ctx->ent_branch.as<uint32_t>() + ent_base;
        M.ent_score = ctx->ent_score.as<float>() + ent_base;
        M.status = ctx->mg_status.as<uint64_t>();
        M.ticket = reinterpret_cast<uint32_t*>(totals + 2);
        M.totals = totals;
        M.offs_add = ent_base;
        M.epoch = ctx->mg_epoch;
        M.n_tiles = tiles;
        IPK_LAUNCH(ctx, "mg_merge_packed", n * 8, mg_merge_packed<<<tiles, MP_THREADS, 0, ctx->stream>>>(M, T));
        uint64_t h[2];
        IPK_CUDA(cudaMemcpyAsync(h, totals, 16, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        r.n_records = h[0];
        r.n_kmers = h[1];
        ctx->kprof_add_bytes("mg_merge_packed", r.n_records * 8 + r.n_kmers * 16);
        const uint64_t end = ent_base + r.n_records;
        IPK_CUDA(cudaMemcpyAsync(ctx->kmer_offs.as<uint64_t>() + kmer_base + r.n_kmers, &end, 8, cudaMemcpyHostToDevice, ctx->stream));
        return r;
    }

    parts_out sort_merge_buckets(ipkgpu_ctx* ctx, uint64_t* p0, uint64_t* p1, ps_table& T, uint64_t n, uint32_t key_bits, uint32_t gbits,
                                 uint32_t skip_bits, uint32_t group_add, int result_mode, bool stream_parts, bool b16, stage_timer& tm,
                                 size_t e_before0, size_t e_before1, uint32_t passes_before);

    /// copy stream, events and pinned staging for entries that leave part by part
    void prepare_entry_stream(ipkgpu_ctx* ctx, uint64_t n, bool b16)
    {
        if (!ctx->copy_stream) IPK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        if (!ctx->ev_merged) { IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_merged, cudaEventDisableTiming)); IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming)); }
        ctx->reserve_host(ctx->h_branch, n * (b16 ? 2 : 4) + 8);
        ctx->reserve_host(ctx->h_score, n * 4 + 4);
        if (b16) ctx->reserve(ctx->out_branch, n * 2 + 8);
    }

    /// The merge sort on packed records (packed_sort.cuh): top-byte histogram -> MSD pass (ka, va) -> 8-byte records in p0 ->
    /// segmented LSD passes p0 <-> p1 (p1 = ka's memory, free once the MSD pass has read it) -> max-merge + CSR.
    /// stream_parts: host-bound indexed result -- buckets are sorted and merged in runs of about n / n_parts records and each
    /// run's entries are queued on the copy stream at once (the role sort_merge_parts has for the pair sort).
    parts_out sort_merge_packed(ipkgpu_ctx* ctx, uint64_t* ka, uint32_t* va, uint64_t* p0, uint64_t n, uint32_t key_bits, uint32_t gbits,
                                uint32_t group_add, int result_mode, bool stream_parts, stage_timer& tm, bool use_runs = false)
    {
        // use_runs: ctx->runs lists the commits of the enumeration (one branch group each, together all n records).  The MSD
        // pass reads them in group order, so its output is group-ordered inside every bucket and the LSD passes only have the
        // k-mer bits left: 1 + ceil((key_bits - 8) / 8) passes instead of 1 + ceil((key_bits + gbits - 8) / 8).
        const uint32_t low_bits = key_bits + gbits - 8;
        const uint32_t skip_bits = use_runs ? gbits : 0;
        uint64_t* p1 = ka;
        const bool b16 = stream_parts && (result_mode & 4) != 0 && ctx->max_branch < 65536u;
        if (stream_parts) prepare_entry_stream(ctx, n, b16);
        ps_table T;
        memset(&T, 0, sizeof T);
        const size_t e0 = tm.mark();
        ps_runs Rn;
        memset(&Rn, 0, sizeof Rn);
        if (use_runs) Rn = ps_order_runs(ctx, n, gbits);          // (uses ctx->hist as sort scratch: before the top histogram)
        if (use_runs && ctx->top_hist_on) ps_top_histogram_from_join(ctx);
        else ps_top_histogram(ctx, ka, n, low_bits);
        IPK_LAUNCH(ctx, "rs_scan_hist", 2048, rs_scan_hist<<<1, RS_RADIX, 0, ctx->stream>>>(ctx->hist.as<uint32_t>()));
        IPK_CUDA(cudaMemcpyAsync(T.start, ctx->hist.p, RS_RADIX * 4, cudaMemcpyDeviceToHost, ctx->stream));
        ps_dest D;
        memset(&D, 0, sizeof D);
        D.out[0] = p0;
        ps_msd_pack(ctx, ka, va, n, low_bits, ctx->hist.as<uint32_t>(), D, false, Rn);
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));           // T.start is on the host now (and ka is free: it becomes p1)
        T.start[PS_BUCKETS] = uint32_t(n);
        const size_t e1 = tm.mark();
        return sort_merge_buckets(ctx, p0, p1, T, n, key_bits, gbits, skip_bits, group_add, result_mode, stream_parts, b16, tm, e0, e1, 1);
    }

    /// Second half of the packed sort: packed records bucketed by their top k-mer byte (T.start) -> segmented LSD passes on
    /// the bits above skip_bits -> max-merge + CSR, in runs of buckets of about n / n_parts records when stream_parts.
    /// Also what an owner rank runs on the records its peers stored into its receive buffer (ipkgpu_stage2_packed).
    parts_out sort_merge_buckets(ipkgpu_ctx* ctx, uint64_t* p0, uint64_t* p1, ps_table& T, uint64_t n, uint32_t key_bits, uint32_t gbits,
                                 uint32_t skip_bits, uint32_t group_add, int result_mode, bool stream_parts, bool b16, stage_timer& tm,
                                 size_t e_before0, size_t e_before1, uint32_t passes_before)
    {
        parts_out R;
        const uint32_t low_bits = key_bits + gbits - 8;
        R.passes = passes_before + div_up<uint32_t>(low_bits - skip_bits, 8);
        struct mark3 { size_t a, b, c; };
        std::vector<mark3> marks;
        const uint64_t target = stream_parts ? div_up<uint64_t>(n, ctx->n_parts) : n;
        uint64_t ent_base = 0, kmer_base = 0;
        for (uint32_t d0 = 0; d0 < PS_BUCKETS;)
        {
            uint32_t d1 = d0 + 1;
            while (d1 < PS_BUCKETS && uint64_t(T.start[d1]) - T.start[d0] < target) ++d1;
            if (!stream_parts) d1 = PS_BUCKETS;
            const uint32_t first = d0;
            d0 = d1;
            if (T.start[d1] == T.start[first]) continue;
            const size_t ea = tm.mark();
            const bool in1 = ps_sort_buckets(ctx, p0, p1, T, first, d1, low_bits, skip_bits);
            const size_t eb = tm.mark();
            const merge_out m = run_merge_packed(ctx, in1 ? p1 : p0, T, first, d1, key_bits, gbits, group_add, ent_base, kmer_base, n);
            marks.push_back({ ea, eb, tm.mark() });
            if (stream_parts) queue_entry_copy(ctx, ent_base, m.n_records, b16);
            ent_base += m.n_records;
            kmer_base += m.n_kmers;
        }
        if (stream_parts)
        {
            IPK_CUDA(cudaEventRecord(ctx->ev_copied, ctx->copy_stream));
            ctx->entries_streamed = true;
            ctx->streamed_branch16 = b16;
        }
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        R.m.n_records = ent_base;
        R.m.n_kmers = kmer_base;
        R.ms_sort = tm.ms(e_before0, e_before1);
        for (const mark3& k : marks) { R.ms_sort += tm.ms(k.a, k.b); R.ms_merge += tm.ms(k.b, k.c); }
        return R;
    }
}

namespace
{
    struct stage1_out
    {
        uint64_t n_emitted = 0, n_unique = 0, n_kmers = 0;
        uint32_t n_groups = 0, sort_passes = 0, enum_launches = 0, n_batches = 0;
        bool csr_done = false;             // single batch: the CSR by k-mer was written together with the max-merge
        float ms_h2d = 0, ms_prefix = 0, ms_enumerate = 0, ms_sort = 0, ms_merge = 0;
    };

    uint64_t device_budget(ipkgpu_ctx* ctx)
    {
        if (ctx->workspace_limit) return ctx->workspace_limit;
        const size_t reusable = ctx->keys_a.cap + ctx->keys_b.cap + ctx->vals_a.cap + ctx->vals_b.cap + ctx->u_keys.cap + ctx->u_scores.cap;
        return uint64_t(double(free_device_bytes() + reusable) * 0.85);
    }

    /// Stage 1 of one rank: K1, then per batch of branch groups K2 -> sort -> K3.
    /// With want_csr and a single batch the CSR by k-mer comes out of the same K3 pass; otherwise the unique
    /// records are left in ctx->u_keys/u_scores as (kmer << group_bits) | global group, batch after batch.
    stage1_out run_stage1(ipkgpu_ctx* ctx, const ipkgpu_input* in, const group_tables& T, uint32_t group_base, uint32_t group_bits,
                          bool want_csr, ipkgpu_stats* st, stage_timer& tm, bool raw = false, int result_mode = 0)
    {
        stage1_out o;
        ctx->res_valid = false; ctx->res_pos = nullptr;
        o.n_groups = uint32_t(T.group_branch.size());
        const uint32_t key_bits = in->k * in->bits;
        if (key_bits + group_bits > 64) fail(IPKGPU_EINVAL, "k*bits + group bits = %u exceeds 64", key_bits + group_bits);
        ctx->reserve(ctx->counters, 128);
        ctx->reserve(ctx->node_counts, size_t(in->n_nodes) * 8 + 8);
        IPK_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 128, ctx->stream));
        IPK_CUDA(cudaMemsetAsync(ctx->node_counts.p, 0, size_t(in->n_nodes) * 8 + 8, ctx->stream));
        const size_t eh = tm.mark();
        // host posteriors of some size go up in node chunks on the copy stream, K1 + K2 of a chunk under the next upload
        std::vector<uint32_t> chunk_bounds;
        const float* d_logp = stage_inputs(ctx, in, T, st, ctx->ordered || in->width < in->k ? nullptr : &chunk_bounds);
        const bool chunked = chunk_bounds.size() > 2;
        const size_t e0 = tm.mark();
        if (!chunked) run_best_prefix(ctx, d_logp, in);
        const size_t e1 = tm.mark();
        ctx->s1_n = 0;
        if (in->n_nodes == 0 || in->width < in->k)
        {
            IPK_CUDA(cudaStreamSynchronize(ctx->stream));
            o.ms_h2d = tm.ms(eh, e0);
            o.ms_prefix = tm.ms(e0, e1);
            return o;
        }

        const enum_program P = make_enum_program(int(in->k));
        const uint64_t n_windows = uint64_t(in->n_nodes) * (in->width - in->k + 1);
        const uint64_t budget = device_budget(ctx);
        const uint64_t batch_cap = std::max<uint64_t>(1 << 16, std::min<uint64_t>(budget / 2 / 24, 0xfff00000ull));
        uint64_t cap0 = ctx->hint_emitted ? ctx->hint_emitted + ctx->hint_emitted / 16 + 4096 : std::max<uint64_t>(1 << 20, 64 * n_windows);
        cap0 = std::min(cap0, batch_cap);

        struct batch { uint32_t g0, g1; uint64_t n; };
        std::vector<batch> batches;
        const uint32_t all_bits = ceil_log2(o.n_groups);

        ctx->reserve(ctx->keys_a, cap0 * 8);
        ctx->reserve(ctx->vals_a, cap0 * 4);
        // raw: the records leave this rank unsorted and unmerged, so they carry the GLOBAL group field from the start
        enum_launch L{ 0, in->n_nodes, raw ? 0u : group_base, raw ? group_bits : all_bits, cap0, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<float>() };
        // Ordered emission (single batch, local sort): the join kernel's first level writes its records in tile order = group
        // order, so a STABLE sort on the k-mer bits alone brings equal (k-mer, group) together -- key_bits / 8 radix passes
        // instead of (key_bits + group_bits) / 8.  The few tiles that level defers go to a side stream that is sorted on
        // all bits and merged in by the max-merge kernel.
        const bool ordered = ctx->ordered && !raw && in->k >= 2 && ctx->enum_mode != 1;
        // commits of the join kernel recorded as runs of one branch group each: read in group order by the first pass of the
        // packed sort they spare it every pass on the group bits (packed_sort.cuh)
        static const bool runs_off = getenv("IPKGPU_RUNS") && atoi(getenv("IPKGPU_RUNS")) == 0;
        L.want_runs = !runs_off && !ordered && (raw || want_csr);
        ctx->runs_n = 0;
        ctx->runs_valid = L.want_runs;
        // the join kernel also counts its records by the top byte of their k-mer (the histogram of the merge sort's first pass)
        // when that byte lies inside the root's left half and the counters fit the memory of the list thresholds
        {
            static const bool off = getenv("IPKGPU_JOIN_HIST") && atoi(getenv("IPKGPU_JOIN_HIST")) == 0;
            const int NCP_ = (ENUM_THREADS + int(in->k) - 1 + 3) & ~3;
            ctx->top_hist_on = !off && L.want_runs && in->k >= 2 && (in->k / 2) * in->bits >= 8 && in->k * in->bits + L.group_bits <= 40 &&
                               size_t(make_join_program(int(in->k)).n_fam) * NCP_ * 4 >= 1024;
            if (ctx->top_hist_on)
            {
                ctx->reserve(ctx->join_hist, size_t(JOIN_HIST_SLOTS) * 256 * 4);
                IPK_CUDA(cudaMemsetAsync(ctx->join_hist.p, 0, size_t(JOIN_HIST_SLOTS) * 256 * 4, ctx->stream));
            }
        }
        uint64_t side_cap = 0;
        if (ordered)
        {
            side_cap = std::max<uint64_t>({ uint64_t(1) << 20, cap0 / 16, ctx->hint_side + ctx->hint_side / 2 });
            ctx->reserve(ctx->side_keys_a, side_cap * 8);
            ctx->reserve(ctx->side_vals_a, side_cap * 4);
            L.side_keys = ctx->side_keys_a.as<uint64_t>(); L.side_scores = ctx->side_vals_a.as<float>(); L.side_cap = side_cap;
        }
        if (chunked)
        {
            for (size_t c = 0; c + 1 < chunk_bounds.size(); ++c)
            {
                IPK_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d[c], 0));
                run_best_prefix(ctx, d_logp, in, chunk_bounds[c], chunk_bounds[c + 1]);
                enum_launch Lc = L;
                Lc.slot_begin = chunk_bounds[c]; Lc.slot_end = chunk_bounds[c + 1];
                Lc.emitted_before = c ? read_u64(ctx, ctx->counters.p) : 0;
                launch_enumerate(ctx, d_logp, in, P, Lc, true);
            }
        }
        else launch_enumerate(ctx, d_logp, in, P, L, true);
        o.enum_launches = 1;
        uint64_t n_main = read_u64(ctx, ctx->counters.p), n_side = 0;
        if (ordered) n_side = read_u64(ctx, ctx->counters.as<unsigned long long>() + 8);
        o.n_emitted = n_main + n_side;
        ctx->kprof_add_bytes("k2_enumerate", 12 * (ordered ? std::min(n_side, side_cap) : std::min(o.n_emitted, cap0) - std::min(ctx->enum_join_emitted, cap0)));
        ctx->hint_emitted = o.n_emitted;
        ctx->hint_side = n_side;
        // a side stream that overflowed its buffer sends the build down the batch path, which enumerates unordered
        bool first_valid = n_main <= cap0 && n_side <= side_cap;
        const bool use_side = ordered && first_valid;
        if (first_valid) batches.push_back({ 0, o.n_groups, n_main });
        else
        {
            ctx->runs_valid = false;
            // the first attempt overflowed its buffer: it still counted every node exactly -> plan batches of whole groups
            std::vector<unsigned long long> counts(in->n_nodes);
            IPK_CUDA(cudaMemcpyAsync(counts.data(), ctx->node_counts.p, size_t(in->n_nodes) * 8, cudaMemcpyDeviceToHost, ctx->stream));
            IPK_CUDA(cudaStreamSynchronize(ctx->stream));
            uint32_t g0 = 0;
            uint64_t acc = 0;
            for (uint32_t g = 0; g < o.n_groups; ++g)
            {
                uint64_t c = 0;
                for (uint32_t s = T.group_first_slot[g]; s < T.group_first_slot[g + 1]; ++s) c += counts[s];
                if (c > batch_cap)
                    fail(IPKGPU_ENOMEM, "branch group %u alone emits %llu phylo-k-mers, above the per-batch capacity of %llu records "
                         "(device budget %llu bytes)", g, (unsigned long long)c, (unsigned long long)batch_cap, (unsigned long long)budget);
                if (acc + c > batch_cap) { batches.push_back({ g0, g, acc }); g0 = g; acc = 0; }
                acc += c;
            }
            batches.push_back({ g0, o.n_groups, acc });
        }
        size_t e_prev = tm.mark();
        ctx->s1_keys = nullptr; ctx->s1_scores = nullptr;
        if (raw && first_valid)
        {
            IPK_CUDA(cudaEventSynchronize(ctx->event(e_prev)));
            o.ms_h2d = tm.ms(eh, e0);
            o.ms_prefix = tm.ms(e0, e1);
            o.ms_enumerate = tm.ms(e1, e_prev);
            o.n_unique = o.n_emitted;
            o.n_batches = 1;
            ctx->s1_n = o.n_emitted;
            ctx->s1_key_bits = key_bits;
            ctx->s1_group_bits = group_bits;
            ctx->s1_keys = ctx->keys_a.as<uint64_t>();
            ctx->s1_scores = ctx->vals_a.as<uint32_t>();
            return o;
        }
        o.ms_h2d = tm.ms(eh, e0);
        o.ms_prefix = tm.ms(e0, e1);
        float ms_enum = 0, ms_sort = 0, ms_merge = 0;
        if (!first_valid) { IPK_CUDA(cudaEventSynchronize(ctx->event(e_prev))); ms_enum += tm.ms(e1, e_prev); }
        bool timed_first = false;
        o.n_batches = uint32_t(batches.size());
        const bool single = batches.size() == 1;
        uint64_t u_off = 0;
        for (const batch& b : batches)
        {
            uint32_t bbits = all_bits;
            size_t ea = e_prev;
            if (!first_valid)
            {
                bbits = ceil_log2(b.g1 - b.g0);
                ctx->reserve(ctx->keys_a, b.n * 8 + 8);
                ctx->reserve(ctx->vals_a, b.n * 4 + 4);
                IPK_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 8, ctx->stream));
                ea = tm.mark();
                enum_launch Lb{ T.group_first_slot[b.g0], T.group_first_slot[b.g1], group_base + b.g0, bbits, b.n,
                                ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<float>() };
                launch_enumerate(ctx, d_logp, in, P, Lb, false);
                o.enum_launches++;
            }
            else if (!timed_first) { ea = e1; timed_first = true; }
            const size_t eb = tm.mark();
            ctx->reserve(ctx->keys_b, b.n * 8 + 8);
            if (single && want_csr && !use_side && ps_eligible(key_bits, bbits, b.n))
            {
                // 8-byte packed records: one MSD pass, then segmented LSD passes and the merge (packed_sort.cuh)
                const bool parts = parts_eligible(ctx, b.n, key_bits, key_bits + bbits, result_mode, uint64_t(std::min(1e18, std::pow(double(in->sigma), double(in->k)))));
                const parts_out po = sort_merge_packed(ctx, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(), ctx->keys_b.as<uint64_t>(), b.n, key_bits, bbits,
                                                       group_base + b.g0, result_mode, parts, tm, ctx->runs_valid && ctx->runs_n > 0);
                o.sort_passes += po.passes;
                o.csr_done = true; o.n_kmers = po.m.n_kmers;
                u_off += po.m.n_records;
                ms_enum += tm.ms(ea, eb); ms_sort += po.ms_sort; ms_merge += po.ms_merge;
                e_prev = tm.mark();
                continue;
            }
            ctx->reserve(ctx->vals_b, b.n * 4 + 4);
            if (single && want_csr && !use_side && parts_eligible(ctx, b.n, key_bits, key_bits + bbits, result_mode, uint64_t(std::min(1e18, std::pow(double(in->sigma), double(in->k))))))
            {
                const parts_out po = sort_merge_parts(ctx, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(), ctx->keys_b.as<uint64_t>(),
                                                      ctx->vals_b.as<uint32_t>(), b.n, key_bits, bbits, group_bits, group_base + b.g0, result_mode, tm);
                o.sort_passes += po.passes;
                o.csr_done = true; o.n_kmers = po.m.n_kmers;
                u_off += po.m.n_records;
                ms_enum += tm.ms(ea, eb); ms_sort += po.ms_sort; ms_merge += po.ms_merge;
                e_prev = tm.mark();
                continue;
            }
            uint32_t passes = 0;
            bool in_b;
            if (single && !use_side && first_valid && !raw && ctx->runs_valid && ctx->runs_n > 0 && b.n <= 0xffffffffull && bbits > 0)
            {
                // records that do not pack: ONE copy of the emission into group order through the run index (packed_sort.cuh), then a
                // stable sort on the k-mer bits alone
                ps_gather_in_group_order(ctx, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(), ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(), b.n, bbits);
                in_b = !radix_sort_pairs(ctx, ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(), ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(),
                                         b.n, key_bits, ctx->hist, ctx->scan_tmp, &passes, bbits);
            }
            else
            // ordered main stream: stable sort on the k-mer bits only (bit_begin = width of the group field)
            in_b = radix_sort_pairs(ctx, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(), ctx->keys_b.as<uint64_t>(),
                                               ctx->vals_b.as<uint32_t>(), b.n, use_side ? key_bits : key_bits + bbits, ctx->hist, ctx->scan_tmp,
                                               &passes, use_side ? bbits : 0);
            o.sort_passes += passes;
            const uint64_t* side_k = nullptr;
            const float* side_v = nullptr;
            if (use_side && n_side)
            {
                ctx->reserve(ctx->side_keys_b, n_side * 8 + 8);
                ctx->reserve(ctx->side_vals_b, n_side * 4 + 4);
                const bool side_in_b = radix_sort_pairs(ctx, ctx->side_keys_a.as<uint64_t>(), ctx->side_vals_a.as<uint32_t>(), ctx->side_keys_b.as<uint64_t>(),
                                                        ctx->side_vals_b.as<uint32_t>(), n_side, key_bits + bbits, ctx->hist, ctx->scan_tmp);
                side_k = side_in_b ? ctx->side_keys_b.as<uint64_t>() : ctx->side_keys_a.as<uint64_t>();
                side_v = side_in_b ? ctx->side_vals_b.as<float>() : ctx->side_vals_a.as<float>();
            }
            const size_t ec = tm.mark();
            const uint64_t* sk = in_b ? ctx->keys_b.as<uint64_t>() : ctx->keys_a.as<uint64_t>();
            const float* sv = in_b ? ctx->vals_b.as<float>() : ctx->vals_a.as<float>();
            const bool csr = single && want_csr;
            const merge_out m = run_merge(ctx, sk, sv, b.n, bbits, group_bits, group_base + b.g0, !csr, u_off, csr, key_bits,
                                          side_k, side_v, side_k ? n_side : 0);
            if (csr) { o.csr_done = true; o.n_kmers = m.n_kmers; }
            u_off += m.n_records;
            const size_t ed = tm.mark();
            IPK_CUDA(cudaEventSynchronize(ctx->event(ed)));
            ms_enum += tm.ms(ea, eb); ms_sort += tm.ms(eb, ec); ms_merge += tm.ms(ec, ed);
            e_prev = ed;
        }
        o.n_unique = u_off;
        o.ms_enumerate = ms_enum; o.ms_sort = ms_sort; o.ms_merge = ms_merge;
        ctx->s1_n = o.csr_done ? 0 : u_off;
        ctx->s1_key_bits = key_bits;
        ctx->s1_group_bits = group_bits;
        if (!o.csr_done) ctx->runs_valid = false;          // merged records: no longer the emission the runs describe
        ctx->s1_keys = ctx->u_keys.as<uint64_t>();
        ctx->s1_scores = ctx->u_scores.as<uint32_t>();
        return o;
    }
}

namespace
{
    struct stage2_params
    {
        uint32_t group_bits, key_bits, total_num_groups;
        float threshold;
        int32_t filter;
    };

    /// lanes that share one k-mer in K4 / K6: short entry lists keep one thread per k-mer, long ones get a lane group
    int lanes_per_kmer(uint64_t n_kmers, uint64_t n_entries)
    {
        const double avg = n_kmers ? double(n_entries) / double(n_kmers) : 0.0;
        return avg < 2.0 ? 1 : avg < 8.0 ? 4 : avg < 512.0 ? 8 : 32;
    }

    /// K4 -> K5 -> K6 on the CSR in ctx (kmer_keys / kmer_offs / ent_branch / ent_score) and hand-over of the result.
    void finish_from_csr(ipkgpu_ctx* ctx, uint64_t n_kmers, uint64_t n_entries, const stage2_params& sp, int result_on_device,
                         ipkgpu_result* out, stage_timer& tm)
    {
        if (n_entries > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu entries exceed the 2^32-1 limit of one device shard: split the build over more devices (ipk-b200 --devices)", (unsigned long long)n_entries);
        ctx->hint_kmers = n_kmers;
        ipkgpu_stats& st = out->stats;
        const bool on_device = (result_on_device & 1) != 0, indexed = (result_on_device & 2) != 0;
        // 16-bit branch ids for host results whose every branch id fits: 6 instead of 8 bytes per entry over PCIe
        // entries that sort_merge_parts already queued on the copy stream, k-mer range by k-mer range
        const bool streamed = ctx->entries_streamed && indexed && !on_device;
        ctx->entries_streamed = false;
        const bool branch16 = streamed ? ctx->streamed_branch16 : (result_on_device & 4) != 0 && indexed && !on_device && ctx->max_branch < 65536u;
        // Indexed results leave the entries where the max-merge wrote them (k-mers in ascending key order) and hand back,
        // per k-mer in filter-value order, where its entries start and how many there are -- the db.at(kmer) walk of
        // db_builder.cpp:323-332 without a gather.  For host results the 8 bytes per entry start crossing PCIe right now, on
        // a second stream, under the filter and ordering kernels.
        cudaEvent_t copy_done = nullptr;
        if (streamed)
        {
            copy_done = ctx->ev_copied;
            st.d2h_bytes += n_entries * (branch16 ? 6 : 8);
        }
        else if (indexed && !on_device && n_entries)
        {
            if (!ctx->copy_stream) IPK_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
            ctx->reserve_host(ctx->h_branch, n_entries * 4 + 4); ctx->reserve_host(ctx->h_score, n_entries * 4 + 4);
            if (!ctx->ev_merged) { IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_merged, cudaEventDisableTiming)); IPK_CUDA(cudaEventCreateWithFlags(&ctx->ev_copied, cudaEventDisableTiming)); }
            cudaEvent_t merged = ctx->ev_merged, done = ctx->ev_copied;
            IPK_CUDA(cudaEventRecord(merged, ctx->stream));
            IPK_CUDA(cudaStreamWaitEvent(ctx->copy_stream, merged, 0));
            if (branch16)
            {
                // narrowed on the copy stream (it only reads ent_branch), into the buffer the contiguous mode gathers into
                ctx->reserve(ctx->out_branch, n_entries * 2 + 4);
                k6_narrow_branch<<<uint32_t(std::min<uint64_t>(div_up<uint64_t>(n_entries, 1024), 1u << 20)), 256, 0, ctx->copy_stream>>>(
                    ctx->ent_branch.as<uint32_t>(), 0, n_entries, ctx->out_branch.as<uint16_t>());
                check_launch(ctx, "k6_narrow_branch", 0);
                IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.p, ctx->out_branch.p, n_entries * 2, cudaMemcpyDeviceToHost, ctx->copy_stream));
            }
            else IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.p, ctx->ent_branch.p, n_entries * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
            IPK_CUDA(cudaMemcpyAsync(ctx->h_score.p, ctx->ent_score.p, n_entries * 4, cudaMemcpyDeviceToHost, ctx->copy_stream));
            IPK_CUDA(cudaEventRecord(done, ctx->copy_stream));
            copy_done = done;
            st.d2h_bytes += n_entries * (branch16 ? 6 : 8);
        }
        const size_t e0 = tm.mark();
        ctx->reserve(ctx->fv, n_kmers * 4 + 4);
        ctx->reserve(ctx->out_keys, n_kmers * 8 + 8);
        ctx->reserve(ctx->out_fv, n_kmers * 4 + 4);
        ctx->reserve(ctx->out_offs, (n_kmers + 1) * 8);
        ctx->reserve(ctx->out_cnt, n_kmers * 4 + 4);
        if (!indexed)
        {
            ctx->reserve(ctx->out_branch, n_entries * 4 + 4);
            ctx->reserve(ctx->out_score, n_entries * 4 + 4);
        }
        size_t e1 = e0, e2 = e0;
        if (n_kmers)
        {
            const uint32_t grid = uint32_t(div_up<uint64_t>(n_kmers, 256));
            if (sp.filter == IPKGPU_FILTER_MIF0)
            {
                // CTA b owns the k-mers whose entries start in [b, b+1) * MIF_CHUNK
                const uint64_t n_chunks = div_up<uint64_t>(std::max<uint64_t>(n_entries, 1), MIF_CHUNK);
                ctx->reserve(ctx->mif_chunks, (n_chunks + 1) * 4);
                IPK_LAUNCH(ctx, "k4_chunk_first", (n_chunks + 1) * 4,
                           k4_chunk_first<<<uint32_t(div_up<uint64_t>(n_chunks + 1, 256)), 256, 0, ctx->stream>>>(ctx->kmer_offs.as<uint64_t>(), n_kmers, n_chunks,
                                                                                                                  ctx->mif_chunks.as<uint32_t>()));
                if (!ctx->mif_tab.p)
                {
                    mif_log2_entry tab[MIF_LOG2_BUCKETS];
                    mif_log2_table(tab);
                    ctx->reserve(ctx->mif_tab, sizeof tab);
                    IPK_CUDA(cudaMemcpyAsync(ctx->mif_tab.p, tab, sizeof tab, cudaMemcpyHostToDevice, ctx->stream));
                    IPK_CUDA(cudaStreamSynchronize(ctx->stream));
                }
                IPK_CUDA(cudaFuncSetAttribute(k4_mif0, cudaFuncAttributeMaxDynamicSharedMemorySize, int(MIF_SMEM)));
                IPK_LAUNCH(ctx, "k4_mif0", n_kmers * 12 + n_entries * 4,
                           k4_mif0<<<uint32_t(n_chunks), MIF_THREADS, MIF_SMEM, ctx->stream>>>(ctx->kmer_offs.as<uint64_t>(), ctx->ent_score.as<float>(),
                                                                                        ctx->mif_chunks.as<uint32_t>(), mif0_make_consts(sp.threshold, sp.total_num_groups), ctx->fv.as<float>(),
                                                                                        ctx->mif_tab.as<mif_log2_entry>()));
            }
            else
            {
                IPK_LAUNCH(ctx, "k4_random", n_kmers * 12, k4_random<<<grid, 256, 0, ctx->stream>>>(n_kmers, ctx->kmer_keys.as<uint64_t>(), ctx->fv.as<float>()));
            }
            e1 = tm.mark();
            // K5: stable sort on the 32-bit order-preserving image of fv; k-mers are in ascending key order, so ties keep it
            ctx->reserve(ctx->ord_keys_a, n_kmers * 8); ctx->reserve(ctx->ord_keys_b, n_kmers * 8);
            ctx->reserve(ctx->ord_vals_a, n_kmers * 4); ctx->reserve(ctx->ord_vals_b, n_kmers * 4);
            IPK_LAUNCH(ctx, "k5_order_keys", n_kmers * 16,
                       k5_order_keys<<<grid, 256, 0, ctx->stream>>>(ctx->fv.as<float>(), n_kmers, ctx->ord_keys_a.as<uint64_t>(), ctx->ord_vals_a.as<uint32_t>()));
            const bool in_b = radix_sort_pairs(ctx, ctx->ord_keys_a.as<uint64_t>(), ctx->ord_vals_a.as<uint32_t>(), ctx->ord_keys_b.as<uint64_t>(),
                                               ctx->ord_vals_b.as<uint32_t>(), n_kmers, 32, ctx->hist, ctx->scan_tmp);
            const uint32_t* perm = in_b ? ctx->ord_vals_b.as<uint32_t>() : ctx->ord_vals_a.as<uint32_t>();
            ctx->res_perm = perm;
            // K6
            IPK_LAUNCH(ctx, "k6_gather_heads", n_kmers * 44,
                       k6_gather_heads<<<grid, 256, 0, ctx->stream>>>(perm, ctx->kmer_keys.as<uint64_t>(), ctx->kmer_offs.as<uint64_t>(), ctx->fv.as<float>(),
                                                                      n_kmers, ctx->out_keys.as<uint64_t>(), ctx->out_fv.as<float>(), ctx->out_cnt.as<uint32_t>(),
                                                                      indexed ? ctx->out_offs.as<uint64_t>() : nullptr, n_entries));
            if (!indexed)
            {
                exclusive_scan_u32(ctx, ctx->out_cnt.as<uint32_t>(), ctx->out_cnt.as<uint32_t>(), n_kmers, ctx->scan_tmp);
                const int G = lanes_per_kmer(n_kmers, n_entries);
                const uint32_t ggrid = uint32_t(div_up<uint64_t>(n_kmers * uint64_t(G), 256));
#define IPK_GATHER(G_) IPK_LAUNCH(ctx, "k6_gather_entries", n_kmers * 32 + n_entries * 16,                                                  \
                                  k6_gather_entries<G_><<<ggrid, 256, 0, ctx->stream>>>(perm, ctx->kmer_offs.as<uint64_t>(), ctx->out_cnt.as<uint32_t>(), \
                                      n_kmers, n_entries, ctx->ent_branch.as<uint32_t>(), ctx->ent_score.as<float>(),                     \
                                      ctx->out_offs.as<uint64_t>(), ctx->out_branch.as<uint32_t>(), ctx->out_score.as<float>()))
                if (G == 1) IPK_GATHER(1); else if (G == 4) IPK_GATHER(4); else if (G == 8) IPK_GATHER(8); else IPK_GATHER(32);
#undef IPK_GATHER
            }
            e2 = tm.mark();
        }
        else
        {
            IPK_CUDA(cudaMemsetAsync(ctx->out_offs.p, 0, 8, ctx->stream));
            e1 = e2 = tm.mark();
        }
        out->n_kmers = n_kmers;
        out->n_entries = n_entries;
        out->on_device = on_device ? 1 : 0;
        out->counts = nullptr;
        out->branch16 = nullptr;
        size_t e3 = e2;
        if (on_device)
        {
            out->keys = ctx->out_keys.as<uint64_t>(); out->fv = ctx->out_fv.as<float>(); out->offsets = ctx->out_offs.as<uint64_t>();
            out->branch = indexed ? ctx->ent_branch.as<uint32_t>() : ctx->out_branch.as<uint32_t>();
            out->score = indexed ? ctx->ent_score.as<float>() : ctx->out_score.as<float>();
            if (indexed) out->counts = ctx->out_cnt.as<uint32_t>();
        }
        else
        {
            ctx->reserve_host(ctx->h_keys, n_kmers * 8 + 8); ctx->reserve_host(ctx->h_fv, n_kmers * 4 + 4);
            ctx->reserve_host(ctx->h_offs, (n_kmers + 1) * 8);
            if (!streamed) { ctx->reserve_host(ctx->h_branch, n_entries * 4 + 4); ctx->reserve_host(ctx->h_score, n_entries * 4 + 4); }
            IPK_CUDA(cudaMemcpyAsync(ctx->h_keys.p, ctx->out_keys.p, n_kmers * 8, cudaMemcpyDeviceToHost, ctx->stream));
            IPK_CUDA(cudaMemcpyAsync(ctx->h_fv.p, ctx->out_fv.p, n_kmers * 4, cudaMemcpyDeviceToHost, ctx->stream));
            IPK_CUDA(cudaMemcpyAsync(ctx->h_offs.p, ctx->out_offs.p, (n_kmers + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
            st.d2h_bytes += n_kmers * 12 + (n_kmers + 1) * 8;
            if (indexed)
            {
                ctx->reserve_host(ctx->h_cnt, n_kmers * 4 + 4);
                IPK_CUDA(cudaMemcpyAsync(ctx->h_cnt.p, ctx->out_cnt.p, n_kmers * 4, cudaMemcpyDeviceToHost, ctx->stream));
                st.d2h_bytes += n_kmers * 4;
                out->counts = ctx->h_cnt.as<uint32_t>();
                if (copy_done) IPK_CUDA(cudaStreamWaitEvent(ctx->stream, copy_done, 0));      // the entries: under way since the merge
            }
            else
            {
                IPK_CUDA(cudaMemcpyAsync(ctx->h_branch.p, ctx->out_branch.p, n_entries * 4, cudaMemcpyDeviceToHost, ctx->stream));
                IPK_CUDA(cudaMemcpyAsync(ctx->h_score.p, ctx->out_score.p, n_entries * 4, cudaMemcpyDeviceToHost, ctx->stream));
                st.d2h_bytes += n_entries * 8;
            }
            e3 = tm.mark();
            out->keys = ctx->h_keys.as<uint64_t>(); out->fv = ctx->h_fv.as<float>(); out->offsets = ctx->h_offs.as<uint64_t>();
            out->branch = ctx->h_branch.as<uint32_t>(); out->score = ctx->h_score.as<float>();
            if (branch16 && n_entries) { out->branch16 = ctx->h_branch.as<uint16_t>(); out->branch = nullptr; }
        }
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        st.ms_filter = tm.ms(e0, e1);
        st.ms_order = tm.ms(e1, e2);
        st.ms_d2h = tm.ms(e2, e3);
        st.n_kmers = n_kmers;
        st.n_entries = n_entries;
        ctx->res_n_kmers = n_kmers; ctx->res_n_entries = n_entries; ctx->res_valid = true;
    }

    /// Stage 2 on unique records (composite keys with the GLOBAL group field): regroup by k-mer -> CSR -> finish.
    /// (keys, scores) are overwritten (they are one side of the sort's ping-pong).
    void run_stage2(ipkgpu_ctx* ctx, uint64_t* keys, float* scores, uint64_t n, bool group_ordered, const stage2_params& sp,
                    int result_on_device, ipkgpu_result* out, stage_timer& tm)
    {
        ctx->res_valid = false; ctx->res_pos = nullptr;
        if (n > 0xffffffffull)
            fail(IPKGPU_EINVAL, "%llu unique (k-mer, branch) records exceed the 2^32-1 limit of one device's regroup sort (about 51 GB of records, near the "
                 "memory bound of the device anyway): split the build over devices (ipk-b200 --devices; every key shard has its own limit)", (unsigned long long)n);
        const size_t e0 = tm.mark();
        const uint64_t* sk = keys;
        const float* sv = scores;
        if (n)
        {
            ctx->reserve(ctx->keys_b, n * 8 + 8);
            if (ps_eligible(sp.key_bits, sp.group_bits, n))
            {
                const bool parts = parts_eligible(ctx, n, sp.key_bits, sp.key_bits + sp.group_bits, result_on_device, sp.key_bits < 63 ? 1ull << sp.key_bits : ~0ull);
                const parts_out po = sort_merge_packed(ctx, keys, reinterpret_cast<uint32_t*>(scores), ctx->keys_b.as<uint64_t>(), n, sp.key_bits, sp.group_bits, 0,
                                                       result_on_device, parts, tm);
                out->stats.sort_passes += po.passes;
                ctx->hint_kmers_records = n;
                finish_from_csr(ctx, po.m.n_kmers, po.m.n_records, sp, result_on_device, out, tm);
                out->stats.ms_sort += po.ms_sort;
                out->stats.ms_merge += po.ms_merge;
                return;
            }
            ctx->reserve(ctx->vals_b, n * 4 + 4);
            if (!group_ordered && parts_eligible(ctx, n, sp.key_bits, sp.key_bits + sp.group_bits, result_on_device, sp.key_bits < 63 ? 1ull << sp.key_bits : ~0ull))
            {
                const parts_out po = sort_merge_parts(ctx, keys, reinterpret_cast<uint32_t*>(scores), ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(),
                                                      n, sp.key_bits, sp.group_bits, sp.group_bits, 0, result_on_device, tm);
                out->stats.sort_passes += po.passes;
                ctx->hint_kmers_records = n;
                finish_from_csr(ctx, po.m.n_kmers, po.m.n_records, sp, result_on_device, out, tm);
                out->stats.ms_sort += po.ms_sort;
                out->stats.ms_merge += po.ms_merge;
                return;
            }
            uint32_t passes = 0;
            // chunks that arrive in group order, each sorted by (kmer, group): a STABLE sort on the k-mer bits alone regroups them
            const bool in_b = group_ordered
                ? radix_sort_pairs(ctx, keys, reinterpret_cast<uint32_t*>(scores), ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(), n,
                                   sp.key_bits, ctx->hist, ctx->scan_tmp, &passes, sp.group_bits)
                : radix_sort_pairs(ctx, keys, reinterpret_cast<uint32_t*>(scores), ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(), n,
                                   sp.key_bits + sp.group_bits, ctx->hist, ctx->scan_tmp, &passes, 0);
            out->stats.sort_passes += passes;
            if (in_b) { sk = ctx->keys_b.as<uint64_t>(); sv = ctx->vals_b.as<float>(); }
        }
        const size_t e1 = tm.mark();
        const merge_out m = run_merge(ctx, sk, sv, n, sp.group_bits, sp.group_bits, 0, false, 0, true, sp.key_bits);
        if (n == 0)
        {
            ctx->reserve(ctx->kmer_offs, 8);
            IPK_CUDA(cudaMemsetAsync(ctx->kmer_offs.p, 0, 8, ctx->stream));
        }
        const size_t e2 = tm.mark();
        ctx->hint_kmers_records = n;
        finish_from_csr(ctx, m.n_kmers, m.n_records, sp, result_on_device, out, tm);
        out->stats.ms_sort += tm.ms(e0, e1);
        out->stats.ms_merge += tm.ms(e1, e2);
    }

    /// The KEEP_POSITIONS build (positions.cuh): enumeration with per-record (node, window), sequence numbers below the group
    /// field, pair sort on (k-mer, group), one thread per run for the first maximum, then the common filter / order / hand-over.
    void run_build_positions(ipkgpu_ctx* ctx, const ipkgpu_input* in, const group_tables& T, int result_mode, ipkgpu_result* out, stage_timer& tm)
    {
        ipkgpu_stats& st = out->stats;
        const uint32_t key_bits = in->k * in->bits, n_groups = uint32_t(T.group_branch.size()), group_bits = ceil_log2(n_groups);
        const uint32_t nwin = in->width >= in->k ? in->width - in->k + 1 : 0;
        uint32_t max_rank = 0;
        std::vector<uint32_t> node_rank(in->n_nodes, 0);
        for (uint32_t g = 0; g < n_groups; ++g)
            for (uint32_t s = T.group_first_slot[g]; s < T.group_first_slot[g + 1]; ++s)
            {
                node_rank[T.node_order[s]] = s - T.group_first_slot[g];           // processing order inside the group (db_builder.cpp:642)
                max_rank = std::max(max_rank, s - T.group_first_slot[g]);
            }
        const uint32_t seq_bits = ceil_log2(uint64_t(max_rank + 1) * std::max<uint32_t>(nwin, 1));
        if (key_bits + group_bits + seq_bits > 64) fail(IPKGPU_EINVAL, "k*bits %u + group bits %u + position bits %u exceed 64", key_bits, group_bits, seq_bits);
        const stage2_params sp{ group_bits, key_bits, in->total_num_groups, in->threshold, in->filter };
        ctx->res_valid = false; ctx->res_pos = nullptr;
        ctx->reserve(ctx->counters, 128);
        const size_t eh = tm.mark();
        const float* d_logp = stage_inputs(ctx, in, T, &st);
        ctx->reserve(ctx->node_rank, size_t(in->n_nodes) * 4 + 4);
        IPK_CUDA(cudaMemcpyAsync(ctx->node_rank.p, node_rank.data(), size_t(in->n_nodes) * 4, cudaMemcpyHostToDevice, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));                       // node_rank is a local vector
        const size_t e0 = tm.mark();
        run_best_prefix(ctx, d_logp, in);
        const size_t e1 = tm.mark();
        uint64_t n = 0;
        if (in->n_nodes && nwin)
        {
            const enum_program P = make_enum_program(int(in->k));
            uint64_t cap = std::max<uint64_t>(1 << 20, ctx->hint_emitted ? ctx->hint_emitted + ctx->hint_emitted / 16 : 32ull * in->n_nodes * nwin);
            for (int attempt = 0; attempt < 2; ++attempt)
            {
                ctx->reserve(ctx->keys_a, cap * 8 + 8); ctx->reserve(ctx->vals_a, cap * 4 + 4);
                ctx->reserve(ctx->dbg_node, cap * 4 + 4); ctx->reserve(ctx->dbg_win, cap * 4 + 4);
                IPK_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 128, ctx->stream));
                enum_launch L{ 0, in->n_nodes, 0, 0, cap, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<float>(),
                               ctx->dbg_node.as<uint32_t>(), ctx->dbg_win.as<uint32_t>(), true };
                launch_enumerate(ctx, d_logp, in, P, L, false);
                n = read_u64(ctx, ctx->counters.p);
                st.enum_launches++;
                if (n <= cap) break;
                if (attempt == 1) fail(IPKGPU_ECUDA, "the enumeration emitted %llu records into a buffer sized for %llu by its own count", (unsigned long long)n, (unsigned long long)cap);
                cap = n;
            }
            ctx->hint_emitted = n;
        }
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu records exceed the 2^32-1 limit of a KEEP_POSITIONS build", (unsigned long long)n);
        st.n_emitted = n;
        const size_t e2 = tm.mark();
        uint64_t n_entries = 0, n_kmers = 0;
        size_t e3 = e2;
        if (n)
        {
            const uint32_t grid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, 256), uint64_t(ctx->num_sms) * 32));
            IPK_LAUNCH(ctx, "kp_compose", n * 24, kp_compose<<<grid, 256, 0, ctx->stream>>>(ctx->keys_a.as<uint64_t>(), ctx->dbg_node.as<uint32_t>(), ctx->dbg_win.as<uint32_t>(),
                                                                                         ctx->node_group.as<uint32_t>(), ctx->node_rank.as<uint32_t>(), n, group_bits, seq_bits, nwin));
            ctx->reserve(ctx->keys_b, n * 8 + 8); ctx->reserve(ctx->vals_b, n * 4 + 4);
            uint32_t passes = 0;
            const bool in_b = radix_sort_pairs(ctx, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<uint32_t>(), ctx->keys_b.as<uint64_t>(), ctx->vals_b.as<uint32_t>(), n,
                                               key_bits + group_bits, ctx->hist, ctx->scan_tmp, &passes, seq_bits);
            st.sort_passes = passes;
            e3 = tm.mark();
            const uint64_t* sk = in_b ? ctx->keys_b.as<uint64_t>() : ctx->keys_a.as<uint64_t>();
            const float* sv = in_b ? ctx->vals_b.as<float>() : ctx->vals_a.as<float>();
            // flags and their scans live in one scratch: head_e, idx_e, head_k, idx_k
            ctx->reserve(ctx->pos_flags, n * 16 + 64);
            uint32_t* head_e = ctx->pos_flags.as<uint32_t>(); uint32_t* idx_e = head_e + n; uint32_t* head_k = idx_e + n; uint32_t* idx_k = head_k + n;
            IPK_LAUNCH(ctx, "kp_flags", n * 16, kp_flags<<<grid, 256, 0, ctx->stream>>>(sk, n, group_bits, seq_bits, head_e, head_k));
            unsigned long long* totals = ctx->counters.as<unsigned long long>() + 2;
            exclusive_scan_u32(ctx, head_e, idx_e, n, ctx->scan_tmp, totals);
            exclusive_scan_u32(ctx, head_k, idx_k, n, ctx->scan_tmp, totals + 1);
            uint64_t h[2];
            IPK_CUDA(cudaMemcpyAsync(h, totals, 16, cudaMemcpyDeviceToHost, ctx->stream));
            IPK_CUDA(cudaStreamSynchronize(ctx->stream));
            n_entries = h[0]; n_kmers = h[1];
            ctx->reserve(ctx->kmer_keys, n_kmers * 8 + 8); ctx->reserve(ctx->kmer_offs, (n_kmers + 1) * 8);
            ctx->reserve(ctx->ent_branch, n_entries * 4 + 4); ctx->reserve(ctx->ent_score, n_entries * 4 + 4); ctx->reserve(ctx->ent_pos, n_entries * 4 + 4);
            IPK_LAUNCH(ctx, "kp_write", n * 12 + n_entries * 12 + n_kmers * 16,
                       kp_write<<<grid, 256, 0, ctx->stream>>>(sk, sv, n, group_bits, seq_bits, nwin, head_e, idx_e, head_k, idx_k, ctx->group_branch.as<uint32_t>(),
                                                               ctx->kmer_keys.as<uint64_t>(), ctx->kmer_offs.as<uint64_t>(), ctx->ent_branch.as<uint32_t>(),
                                                               ctx->ent_score.as<float>(), ctx->ent_pos.as<uint32_t>()));
            IPK_CUDA(cudaMemcpyAsync(ctx->kmer_offs.as<uint64_t>() + n_kmers, &n_entries, 8, cudaMemcpyHostToDevice, ctx->stream));
            IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        }
        else
        {
            ctx->reserve(ctx->kmer_offs, 8);
            IPK_CUDA(cudaMemsetAsync(ctx->kmer_offs.p, 0, 8, ctx->stream));
        }
        const size_t e4 = tm.mark();
        const int mode = result_mode & 3;                                    // 16-bit branches / streaming do not apply here
        finish_from_csr(ctx, n_kmers, n_entries, sp, mode, out, tm);
        // the positions follow their entries: in place for indexed results, gathered by the same k-mer order otherwise
        const bool on_device = (mode & 1) != 0, indexed = (mode & 2) != 0;
        const uint32_t* d_pos = ctx->ent_pos.as<uint32_t>();
        if (!indexed && n_kmers)
        {
            ctx->reserve(ctx->out_pos, n_entries * 4 + 4);
            const uint32_t ggrid = uint32_t(div_up<uint64_t>(n_kmers * 8, 256));
            IPK_LAUNCH(ctx, "kp_gather_pos", n_entries * 8, kp_gather_pos<8><<<ggrid, 256, 0, ctx->stream>>>(ctx->res_perm, ctx->kmer_offs.as<uint64_t>(), ctx->out_offs.as<uint64_t>(),
                                                                                                           n_kmers, ctx->ent_pos.as<uint32_t>(), ctx->out_pos.as<uint32_t>()));
            d_pos = ctx->out_pos.as<uint32_t>();
        }
        if (on_device) ctx->res_pos = n_entries ? d_pos : nullptr;
        else
        {
            ctx->reserve_host(ctx->h_pos, n_entries * 4 + 4);
            if (n_entries) IPK_CUDA(cudaMemcpyAsync(ctx->h_pos.p, d_pos, n_entries * 4, cudaMemcpyDeviceToHost, ctx->stream));
            st.d2h_bytes += n_entries * 4;
            ctx->res_pos = n_entries ? ctx->h_pos.as<uint32_t>() : nullptr;
        }
        ctx->res_pos_on_device = on_device ? 1 : 0;
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        st.ms_h2d = tm.ms(eh, e0); st.ms_prefix = tm.ms(e0, e1); st.ms_enumerate = tm.ms(e1, e2); st.ms_sort = tm.ms(e2, e3); st.ms_merge = tm.ms(e3, e4);
    }

    template<class F> int guarded(ipkgpu_ctx* ctx, F&& f)
    {
        try
        {
            if (!ctx) return IPKGPU_EINVAL;
            IPK_CUDA(cudaSetDevice(ctx->device));
            ctx->launches = 0;
            ctx->entries_streamed = false;
            ctx->last_tiles = ctx->last_deferred_tiles = 0;
            ctx->kprof_marks.clear();
            f();
            ctx->kprof_collect();
            return IPKGPU_OK;
        }
        catch (const error& e) { if (ctx) ctx->last_error = e.what; cudaGetLastError(); return e.code; }
        catch (const std::bad_alloc&) { if (ctx) ctx->last_error = "host allocation failed"; return IPKGPU_ENOMEM; }
        catch (const std::exception& e) { if (ctx) ctx->last_error = e.what(); return IPKGPU_ECUDA; }
        catch (...) { if (ctx) ctx->last_error = "unknown failure"; return IPKGPU_ECUDA; }
    }
}

// =====================================================================================================================
// C ABI
// =====================================================================================================================
extern "C"
{

void ipkgpu_score_threshold(float omega, uint32_t sigma, uint32_t k, float* threshold, float* log_threshold)
{
    const float t = powf(omega / float(sigma), float(k));
    if (threshold) *threshold = t;
    if (log_threshold) *log_threshold = log10f(t);
}

int ipkgpu_create(ipkgpu_ctx** out, int device, void* stream)
{
    if (!out) return IPKGPU_EINVAL;
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
    {
        cudaGetLastError();
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (libipkgpu has no CPU fallback)";
        return IPKGPU_ENODEV;
    }
    if (device < 0 || device >= n_dev) { g_create_error = "device ordinal out of range"; return IPKGPU_EINVAL; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10)
    {
        cudaGetLastError();
        g_create_error = "device is not sm_100 (this library is built for sm_100a only)";
        return IPKGPU_ENODEV;
    }
    ipkgpu_ctx* ctx = new (std::nothrow) ipkgpu_ctx();
    if (!ctx) { g_create_error = "host allocation failed"; return IPKGPU_ENOMEM; }
    ctx->device = device;
    ctx->num_sms = prop.multiProcessorCount;
    if (const char* e = getenv("IPKGPU_RS_SHAPE")) ctx->rs_shape = atoi(e);
    if (const char* e = getenv("IPKGPU_H2D_CHUNKS")) ctx->h2d_chunks = uint32_t(std::min(64, std::max(1, atoi(e))));
    if (const char* e = getenv("IPKGPU_H2D_CHUNK_MIN")) ctx->h2d_chunk_min = uint64_t(std::max(1ll, atoll(e)));
    if (const char* e = getenv("IPKGPU_FIRST_PART_DIV")) ctx->first_part_div = uint32_t(std::max(1, atoi(e)));
    if (const char* e = getenv("IPKGPU_PARTS")) ctx->n_parts = uint32_t(std::max(1, atoi(e)));
    if (const char* e = getenv("IPKGPU_PART_MIN")) ctx->part_min = uint64_t(std::max(1ll, atoll(e)));
    if (const char* e = getenv("IPKGPU_ORDERED")) ctx->ordered = atoi(e);
    if (const char* e = getenv("IPKGPU_GROUP_MIN")) ctx->group_min = uint32_t(std::max(1, atoi(e)));
    if (const char* e = getenv("IPKGPU_SMALL_LIST")) ctx->small_list = uint32_t(std::max(1, atoi(e)));
    if (const char* e = getenv("IPKGPU_LEVELS"))
    {
        // "pool:width,pool:width,..."; width a power of two <= 128
        const char* p = e;
        while (*p)
        {
            char* q = nullptr;
            const unsigned long pool = strtoul(p, &q, 10);
            unsigned long width = ENUM_THREADS;
            if (*q == ':') width = strtoul(q + 1, &q, 10);
            if (width >= 1 && width <= uint32_t(ENUM_THREADS) && (width & (width - 1)) == 0) ctx->levels.push_back({ uint32_t(pool), uint32_t(width) });
            p = *q == ',' ? q + 1 : q;
            if (q == p && *p) break;
        }
    }      // tuning knob; the default is the product
    cudaSetDevice(device);
    if (stream) ctx->stream = static_cast<cudaStream_t>(stream);
    else
    {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess)
        {
            g_create_error = "cudaStreamCreate failed";
            delete ctx;
            return IPKGPU_ECUDA;
        }
        ctx->own_stream = true;
    }
    *out = ctx;
    return IPKGPU_OK;
}

void ipkgpu_destroy(ipkgpu_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ipkgpu::dbuf* all[] = { &ctx->logp, &ctx->colmax, &ctx->bestpref, &ctx->node_group, &ctx->node_order, &ctx->group_branch, &ctx->counters,
                            &ctx->node_counts, &ctx->keys_a, &ctx->keys_b, &ctx->vals_a, &ctx->vals_b, &ctx->hist, &ctx->scan_tmp,
                            &ctx->block_counts, &ctx->block_offs, &ctx->rs_status, &ctx->mg_status, &ctx->mif_chunks, &ctx->dbg_node, &ctx->dbg_win, &ctx->deferred, &ctx->deferred2, &ctx->u_keys, &ctx->u_scores,
                            &ctx->kmer_keys, &ctx->kmer_offs, &ctx->ent_branch, &ctx->ent_score, &ctx->fv, &ctx->ord_keys_a, &ctx->ord_keys_b,
                            &ctx->ord_vals_a, &ctx->ord_vals_b, &ctx->out_keys, &ctx->out_fv, &ctx->out_offs, &ctx->out_branch,
                            &ctx->out_score, &ctx->out_cnt, &ctx->mo_buf, &ctx->recv_keys, &ctx->recv_scores, &ctx->peer_base, &ctx->mif_tab, &ctx->tile_status,
                            &ctx->side_keys_a, &ctx->side_keys_b, &ctx->side_vals_a, &ctx->side_vals_b, &ctx->mg_split, &ctx->ps_hist,
                            &ctx->runs, &ctx->runs_scratch, &ctx->join_hist, &ctx->recv_old, &ctx->recv_old2, &ctx->ent_pos, &ctx->out_pos, &ctx->pos_flags, &ctx->node_rank };
    for (auto& m : ctx->ipc_open) cudaIpcCloseMemHandle(m.ptr);
    ctx->ipc_open.clear();
    for (auto& m : ctx->ipc_slots) if (m.ptr) cudaIpcCloseMemHandle(m.ptr);
    ctx->ipc_slots.clear();
    for (auto* b : all) ctx->release(*b);
    ipkgpu::hbuf* hall[] = { &ctx->h_small, &ctx->h_keys, &ctx->h_fv, &ctx->h_offs, &ctx->h_branch, &ctx->h_score, &ctx->h_cnt, &ctx->h_pos };
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->ev_merged) { cudaEventDestroy(ctx->ev_merged); cudaEventDestroy(ctx->ev_copied); }
    for (auto* b : hall) if (b->p && !b->external) ipkgpu::pinned_free(b->p);
    for (auto e : ctx->events) cudaEventDestroy(e);
    for (auto e : ctx->ev_h2d) cudaEventDestroy(e);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int ipkgpu_merge_order(ipkgpu_ctx* ctx, uint32_t n_shards, const float* const* fv, const uint64_t* const* keys, const uint64_t* n_kmers,
                       int32_t key_ranges, uint32_t key_bits, uint8_t* out_shard, uint32_t* out_index)
{
    return guarded(ctx, [&] {
        if (n_shards < 1 || n_shards > 256) fail(IPKGPU_EINVAL, "n_shards = %u outside [1, 256]", n_shards);
        if (!fv || !n_kmers || (!key_ranges && !keys)) fail(IPKGPU_EINVAL, "fv / keys / n_kmers is null");
        if (!key_ranges && (key_bits < 1 || key_bits > 64)) fail(IPKGPU_EINVAL, "key_bits = %u outside [1, 64]", key_bits);
        mo_table T;
        memset(&T, 0, sizeof T);
        T.n_shards = n_shards;
        for (uint32_t s = 0; s < n_shards; ++s)
        {
            if (n_kmers[s] && (!fv[s] || (!key_ranges && !keys[s]))) fail(IPKGPU_EINVAL, "shard %u: array is null", s);
            T.prefix[s + 1] = T.prefix[s] + n_kmers[s];
        }
        for (uint32_t s = n_shards + 1; s <= 256; ++s) T.prefix[s] = T.prefix[n_shards];
        const uint64_t n = T.prefix[n_shards];
        if (n == 0) return;
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu k-mers over all shards: the merge order indexes them with 32 bits", (unsigned long long)n);
        if (!out_shard || !out_index) fail(IPKGPU_EINVAL, "output pointer is null");
        // one allocation: fv | keys a | keys b | vals a | vals b | shard | index
        const uint64_t n8 = (n + 31) & ~31ull;
        ctx->reserve(ctx->mo_buf, n8 * (4 + 8 + 8 + 4 + 4 + 4 + 1) + 256);
        unsigned char* base = ctx->mo_buf.as<unsigned char>();
        uint64_t* ka = reinterpret_cast<uint64_t*>(base);
        uint64_t* kb = ka + n8;
        float* d_fv = reinterpret_cast<float*>(kb + n8);
        uint32_t* va = reinterpret_cast<uint32_t*>(d_fv + n8);
        uint32_t* vb = va + n8;
        uint32_t* d_index = vb + n8;
        uint8_t* d_shard = reinterpret_cast<uint8_t*>(d_index + n8);
        for (uint32_t s = 0; s < n_shards; ++s)
            if (n_kmers[s]) IPK_CUDA(cudaMemcpyAsync(d_fv + T.prefix[s], fv[s], n_kmers[s] * 4, cudaMemcpyHostToDevice, ctx->stream));
        const uint32_t grid = uint32_t(div_up<uint64_t>(n, 256));
        const uint32_t* perm = nullptr;
        uint64_t *k0 = ka, *k1 = kb;                  // where the image keys go / the other side of the ping-pong
        uint32_t *v0 = va, *v1 = vb;
        if (!key_ranges)
        {
            // shards with interleaved key sets: ties between equal filter values must fall in key order, so the positions are
            // put into key order first (stable LSD passes on the key bits), then sorted stably on the filter values
            for (uint32_t s = 0; s < n_shards; ++s)
                if (n_kmers[s]) IPK_CUDA(cudaMemcpyAsync(ka + T.prefix[s], keys[s], n_kmers[s] * 8, cudaMemcpyHostToDevice, ctx->stream));
            IPK_LAUNCH(ctx, "k9_iota", n * 4, k9_iota<<<grid, 256, 0, ctx->stream>>>(va, n));
            const bool in_b = radix_sort_pairs(ctx, ka, va, kb, vb, n, key_bits, ctx->hist, ctx->scan_tmp);
            perm = in_b ? vb : va;
            if (!in_b) { k0 = kb; k1 = ka; v0 = vb; v1 = va; }
        }
        IPK_LAUNCH(ctx, "k9_image_keys", n * 20, k9_image_keys<<<grid, 256, 0, ctx->stream>>>(d_fv, perm, n, k0, v0));
        // key-range shards (the packed exchange): shard s holds smaller keys than shard s + 1 and is itself in (fv, key) order,
        // so the stable sort on the filter values alone leaves equal values in key order
        const bool in_1 = radix_sort_pairs(ctx, k0, v0, k1, v1, n, 32, ctx->hist, ctx->scan_tmp);
        IPK_LAUNCH(ctx, "k9_split_positions", n * 9, k9_split_positions<<<grid, 256, 0, ctx->stream>>>(in_1 ? v1 : v0, n, T, d_shard, d_index));
        IPK_CUDA(cudaMemcpyAsync(out_index, d_index, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(out_shard, d_shard, n, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

const char* ipkgpu_last_error(const ipkgpu_ctx* ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }
void* ipkgpu_stream(const ipkgpu_ctx* ctx) { return ctx ? static_cast<void*>(ctx->stream) : nullptr; }

int ipkgpu_set_enumeration(ipkgpu_ctx* ctx, int mode, uint32_t pool_cap)
{
    if (!ctx || mode < 0 || mode > 1 || pool_cap > 65000u) return IPKGPU_EINVAL;
    ctx->enum_mode = mode;
    ctx->pool_cap = pool_cap;
    return IPKGPU_OK;
}

int ipkgpu_kernel_profile(ipkgpu_ctx* ctx, int enable)
{
    if (!ctx) return IPKGPU_EINVAL;
    ctx->kprof_on = enable != 0;
    ctx->kprof_rows.clear();
    ctx->kprof_marks.clear();
    return IPKGPU_OK;
}

int ipkgpu_kernel_profile_get(const ipkgpu_ctx* ctx, uint32_t index, char* name, size_t name_cap, uint32_t* launches, double* ms, uint64_t* bytes)
{
    if (!ctx || index >= ctx->kprof_rows.size()) return IPKGPU_EINVAL;
    const auto& r = ctx->kprof_rows[index];
    if (name && name_cap) { strncpy(name, r.name.c_str(), name_cap - 1); name[name_cap - 1] = 0; }
    if (launches) *launches = r.launches;
    if (ms) *ms = r.ms;
    if (bytes) *bytes = r.bytes;
    return IPKGPU_OK;
}

int ipkgpu_set_workspace_limit(ipkgpu_ctx* ctx, uint64_t bytes)
{
    if (!ctx) return IPKGPU_EINVAL;
    ctx->workspace_limit = bytes;
    return IPKGPU_OK;
}

int ipkgpu_host_alloc(int device, size_t bytes, void** out)
{
    if (!out) return IPKGPU_EINVAL;
    *out = nullptr;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return IPKGPU_ENODEV; }
    if (ipkgpu::pinned_alloc(out, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); *out = nullptr; return IPKGPU_ENOMEM; }
    return IPKGPU_OK;
}

void ipkgpu_host_free(void* p)
{
    ipkgpu::pinned_free(p);
}

int ipkgpu_build(ipkgpu_ctx* ctx, const ipkgpu_input* in, int result_on_device, ipkgpu_result* out)
{
    return guarded(ctx, [&] {
        validate(in);
        if (!out) fail(IPKGPU_EINVAL, "result is null");
        memset(out, 0, sizeof *out);
        stage_timer tm(ctx);
        const size_t t0 = tm.mark();
        const group_tables T = make_groups(in->node_branch, in->n_nodes, 0);
        const uint32_t n_groups = uint32_t(T.group_branch.size());
        const uint32_t group_bits = ceil_log2(n_groups);
        const stage2_params sp{ group_bits, in->k * in->bits, in->total_num_groups, in->threshold, in->filter };
        ipkgpu_stats& st = out->stats;
        st.n_windows = in->width >= in->k ? uint64_t(in->n_nodes) * (in->width - in->k + 1) : 0;
        st.n_groups = n_groups;
        if (ctx->keep_positions)
        {
            run_build_positions(ctx, in, T, result_on_device, out, tm);
            const size_t t1p = tm.mark();
            IPK_CUDA(cudaEventSynchronize(ctx->event(t1p)));
            st.ms_total = tm.ms(t0, t1p);
            st.kernel_launches = ctx->launches;
            st.n_tiles = ctx->last_tiles; st.n_deferred_tiles = ctx->last_deferred_tiles;
            return;
        }
        ctx->res_pos = nullptr;
        const stage1_out s1 = run_stage1(ctx, in, T, 0, group_bits, true, &st, tm, false, result_on_device);
        st.n_emitted = s1.n_emitted;
        st.sort_passes = s1.sort_passes;
        st.enum_launches = s1.enum_launches;
        st.ms_h2d = s1.ms_h2d; st.ms_prefix = s1.ms_prefix; st.ms_enumerate = s1.ms_enumerate; st.ms_sort = s1.ms_sort; st.ms_merge = s1.ms_merge;
        if (s1.csr_done) ctx->hint_kmers_records = s1.n_emitted;
        if (s1.csr_done) finish_from_csr(ctx, s1.n_kmers, s1.n_unique, sp, result_on_device, out, tm);
        else run_stage2(ctx, ctx->u_keys.as<uint64_t>(), ctx->u_scores.as<float>(), s1.n_unique, true, sp, result_on_device, out, tm);
        const size_t t1 = tm.mark();
        IPK_CUDA(cudaEventSynchronize(ctx->event(t1)));
        st.ms_total = tm.ms(t0, t1);
        st.kernel_launches = ctx->launches;
        st.n_tiles = ctx->last_tiles; st.n_deferred_tiles = ctx->last_deferred_tiles;
    });
}

int ipkgpu_set_keep_positions(ipkgpu_ctx* ctx, int on)
{
    if (!ctx) return IPKGPU_EINVAL;
    ctx->keep_positions = on != 0;
    return IPKGPU_OK;
}

int ipkgpu_result_positions(ipkgpu_ctx* ctx, const uint32_t** positions, int32_t* on_device)
{
    if (!ctx || !positions) return IPKGPU_EINVAL;
    *positions = ctx->res_valid ? ctx->res_pos : nullptr;
    if (on_device) *on_device = ctx->res_pos_on_device;
    return IPKGPU_OK;
}

int ipkgpu_result_digest(ipkgpu_ctx* ctx, float log_threshold, float band, ipkgpu_digest* out)
{
    return guarded(ctx, [&] {
        if (!out) fail(IPKGPU_EINVAL, "digest is null");
        if (!ctx->res_valid) fail(IPKGPU_EINVAL, "no database on this context: call ipkgpu_build / ipkgpu_stage2 first");
        memset(out, 0, sizeof *out);
        out->n_kmers = ctx->res_n_kmers; out->n_entries = ctx->res_n_entries;
        if (ctx->res_n_kmers == 0) return;
        ctx->reserve(ctx->counters, 128);
        unsigned long long* acc = ctx->counters.as<unsigned long long>() + 8;
        IPK_CUDA(cudaMemsetAsync(acc, 0, 56, ctx->stream));
        const uint64_t groups = div_up<uint64_t>(ctx->res_n_kmers * DG_LANES, DG_THREADS);
        const uint32_t grid = uint32_t(std::min<uint64_t>(groups, uint64_t(ctx->num_sms) * 32));
        IPK_LAUNCH(ctx, "k7_digest", ctx->res_n_kmers * 20 + ctx->res_n_entries * 8,
                   k7_digest<<<grid, DG_THREADS, 0, ctx->stream>>>(ctx->kmer_keys.as<uint64_t>(), ctx->kmer_offs.as<uint64_t>(), ctx->ent_branch.as<uint32_t>(),
                                                                   ctx->ent_score.as<float>(), ctx->fv.as<float>(), ctx->res_n_kmers, log_threshold, band, acc));
        uint64_t h[7];
        IPK_CUDA(cudaMemcpyAsync(h, acc, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        out->entries_sum = h[0]; out->entries_xor = h[1]; out->kmers_sum = h[2]; out->kmers_xor = h[3];
        out->fv_sum = h[4]; out->fv_xor = h[5]; out->n_boundary_above = h[6];
    });
}

int ipkgpu_boundary_list(ipkgpu_ctx* ctx, const ipkgpu_input* in, float band, uint64_t cap, uint64_t* kmer_out, uint32_t* branch_out,
                         float* score_loose_out, float* score_tight_out, uint64_t* n_out)
{
    return guarded(ctx, [&] {
        validate(in);
        if (!n_out || (cap && (!kmer_out || !branch_out || !score_loose_out || !score_tight_out))) fail(IPKGPU_EINVAL, "output pointer is null");
        if (!(band > 0.0f)) fail(IPKGPU_EINVAL, "band must be positive");
        *n_out = 0;
        const group_tables T = make_groups(in->node_branch, in->n_nodes, 0);
        const uint32_t group_bits = ceil_log2(uint32_t(T.group_branch.size()));
        stage_timer tm(ctx);
        ipkgpu_input run = *in;
        // loose: everything a threshold up to `band` lower would let in
        run.log_threshold = in->log_threshold - band;
        ctx->hint_emitted = 0;
        const stage1_out lo = run_stage1(ctx, &run, T, 0, group_bits, false, nullptr, tm);
        if (lo.n_batches > 1) fail(IPKGPU_ENOMEM, "the boundary analysis needs the emitted records of the input in one batch; pass fewer ghost nodes at a time");
        const uint64_t n_loose = lo.n_unique;
        ctx->reserve(ctx->side_keys_a, n_loose * 8 + 8);
        ctx->reserve(ctx->side_vals_a, n_loose * 4 + 4);
        IPK_CUDA(cudaMemcpyAsync(ctx->side_keys_a.p, ctx->u_keys.p, n_loose * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(ctx->side_vals_a.p, ctx->u_scores.p, n_loose * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        // tight: what survives a threshold up to `band` higher
        run.log_threshold = in->log_threshold + band;
        const stage1_out hi = run_stage1(ctx, &run, T, 0, group_bits, false, nullptr, tm);
        if (hi.n_batches > 1) fail(IPKGPU_ENOMEM, "the boundary analysis needs the emitted records of the input in one batch");
        ctx->hint_emitted = 0;
        if (n_loose == 0) return;
        ctx->reserve(ctx->out_keys, cap * 8 + 8); ctx->reserve(ctx->out_branch, cap * 4 + 4);
        ctx->reserve(ctx->out_score, cap * 4 + 4); ctx->reserve(ctx->out_fv, cap * 4 + 4);
        unsigned long long* counter = ctx->counters.as<unsigned long long>() + 8;
        IPK_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
        const uint32_t grid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n_loose, 256), uint64_t(ctx->num_sms) * 16));
        IPK_LAUNCH(ctx, "k8_boundary_diff", n_loose * 12,
                   k8_boundary_diff<<<grid, 256, 0, ctx->stream>>>(ctx->side_keys_a.as<uint64_t>(), ctx->side_vals_a.as<float>(), n_loose, ctx->u_keys.as<uint64_t>(),
                                                                   ctx->u_scores.as<float>(), hi.n_unique, group_bits, ctx->group_branch.as<uint32_t>(), cap,
                                                                   ctx->out_keys.as<uint64_t>(), ctx->out_branch.as<uint32_t>(), ctx->out_score.as<float>(),
                                                                   ctx->out_fv.as<float>(), counter));
        const uint64_t n = read_u64(ctx, counter);
        *n_out = n;
        const uint64_t m = std::min(n, cap);
        IPK_CUDA(cudaMemcpyAsync(kmer_out, ctx->out_keys.p, m * 8, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(branch_out, ctx->out_branch.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(score_loose_out, ctx->out_score.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(score_tight_out, ctx->out_fv.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

void ipkgpu_free_result(ipkgpu_ctx* ctx, ipkgpu_result* res)
{
    (void)ctx;
    if (res) { res->keys = nullptr; res->fv = nullptr; res->offsets = nullptr; res->branch = nullptr; res->score = nullptr; res->n_kmers = res->n_entries = 0; }
}

}   // extern "C"

// ---- stage-level entry points ---------------------------------------------------------------------------------------
extern "C"
{

int ipkgpu_best_prefix(ipkgpu_ctx* ctx, const ipkgpu_input* in, float* B_out)
{
    return guarded(ctx, [&] {
        validate(in);
        if (!B_out) fail(IPKGPU_EINVAL, "B_out is null");
        const group_tables T = make_groups(in->node_branch, in->n_nodes, 0);
        const float* d_logp = stage_inputs(ctx, in, T, nullptr);
        run_best_prefix(ctx, d_logp, in);
        IPK_CUDA(cudaMemcpyAsync(B_out, ctx->bestpref.p, size_t(in->n_nodes) * (in->width + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

int ipkgpu_enumerate(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint64_t cap,
                     uint32_t* node_out, uint32_t* window_out, uint64_t* key_out, float* score_out, uint64_t* n_out)
{
    return guarded(ctx, [&] {
        validate(in);
        if (!n_out || (cap && (!node_out || !window_out || !key_out || !score_out))) fail(IPKGPU_EINVAL, "output pointer is null");
        const group_tables T = make_groups(in->node_branch, in->n_nodes, 0);
        const float* d_logp = stage_inputs(ctx, in, T, nullptr);
        run_best_prefix(ctx, d_logp, in);
        ctx->reserve(ctx->counters, 128);
        IPK_CUDA(cudaMemsetAsync(ctx->counters.p, 0, 128, ctx->stream));
        ctx->reserve(ctx->keys_a, cap * 8 + 8); ctx->reserve(ctx->vals_a, cap * 4 + 4);
        ctx->reserve(ctx->dbg_node, cap * 4 + 4); ctx->reserve(ctx->dbg_win, cap * 4 + 4);
        *n_out = 0;
        if (in->n_nodes == 0 || in->width < in->k) return;
        const enum_program P = make_enum_program(int(in->k));
        enum_launch L{ 0, in->n_nodes, 0, 0, cap, ctx->keys_a.as<uint64_t>(), ctx->vals_a.as<float>(),
                       ctx->dbg_node.as<uint32_t>(), ctx->dbg_win.as<uint32_t>(), true };
        launch_enumerate(ctx, d_logp, in, P, L, false);
        const uint64_t n = read_u64(ctx, ctx->counters.p);
        *n_out = n;
        const uint64_t m = std::min(n, cap);
        IPK_CUDA(cudaMemcpyAsync(key_out, ctx->keys_a.p, m * 8, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(score_out, ctx->vals_a.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(node_out, ctx->dbg_node.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaMemcpyAsync(window_out, ctx->dbg_win.p, m * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
    });
}

static int stage1_entry(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint32_t group_base, uint32_t group_bits, uint64_t* n_out, ipkgpu_stats* stats, bool raw)
{
    return guarded(ctx, [&] {
        validate(in);
        if (!n_out) fail(IPKGPU_EINVAL, "n_out is null");
        ipkgpu_stats st;
        memset(&st, 0, sizeof st);
        stage_timer tm(ctx);
        const size_t t0 = tm.mark();
        const group_tables T = make_groups(in->node_branch, in->n_nodes, group_base);
        if (group_base + T.group_branch.size() > (1ull << group_bits) && !(group_bits >= 32))
            fail(IPKGPU_EINVAL, "group_base %u + %zu groups do not fit %u group bits", group_base, T.group_branch.size(), group_bits);
        const stage1_out s1 = run_stage1(ctx, in, T, group_base, group_bits, false, &st, tm, raw);
        const size_t t1 = tm.mark();
        IPK_CUDA(cudaEventSynchronize(ctx->event(t1)));
        *n_out = s1.n_unique;
        st.n_windows = in->width >= in->k ? uint64_t(in->n_nodes) * (in->width - in->k + 1) : 0;
        st.n_emitted = s1.n_emitted; st.n_entries = s1.n_unique; st.n_groups = s1.n_groups;
        st.sort_passes = s1.sort_passes; st.enum_launches = s1.enum_launches;
        st.ms_h2d = s1.ms_h2d; st.ms_prefix = s1.ms_prefix; st.ms_enumerate = s1.ms_enumerate; st.ms_sort = s1.ms_sort; st.ms_merge = s1.ms_merge;
        st.ms_total = tm.ms(t0, t1);
        st.kernel_launches = ctx->launches;
        st.n_tiles = ctx->last_tiles; st.n_deferred_tiles = ctx->last_deferred_tiles;
        if (stats) *stats = st;
    });
}

int ipkgpu_stage1(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint32_t group_base, uint32_t group_bits, uint64_t* n_out, ipkgpu_stats* stats)
{
    return stage1_entry(ctx, in, group_base, group_bits, n_out, stats, false);
}

int ipkgpu_stage1_raw(ipkgpu_ctx* ctx, const ipkgpu_input* in, uint32_t group_base, uint32_t group_bits, uint64_t* n_out, ipkgpu_stats* stats)
{
    return stage1_entry(ctx, in, group_base, group_bits, n_out, stats, true);
}

int ipkgpu_stage1_partition(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* d_keys, float* d_scores, uint64_t* counts)
{
    return guarded(ctx, [&] {
        if (n_parts < 1 || n_parts > RS_RADIX) fail(IPKGPU_EINVAL, "n_parts = %u outside [1, %d]", n_parts, RS_RADIX);
        if (!counts) fail(IPKGPU_EINVAL, "counts is null");
        const uint64_t n = ctx->s1_n;
        for (uint32_t p = 0; p < n_parts; ++p) counts[p] = 0;
        if (n == 0) return;
        if (!d_keys || !d_scores) fail(IPKGPU_EINVAL, "send buffer is null");
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu records exceed one shard's 2^32-1 limit", (unsigned long long)n);
        const digit_owner own{ ctx->s1_group_bits, n_parts };
        radix_pass(ctx, ctx->s1_keys, ctx->s1_scores, d_keys, reinterpret_cast<uint32_t*>(d_scores), n,
                   own, ctx->hist, ctx->scan_tmp);
        // bucket sizes: radix_pass leaves the exclusive digit starts in the first 256 words of the table
        std::vector<uint32_t> starts(n_parts);
        IPK_CUDA(cudaMemcpyAsync(starts.data(), ctx->hist.p, size_t(n_parts) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        for (uint32_t p = 0; p < n_parts; ++p) counts[p] = (p + 1 < n_parts ? starts[p + 1] : n) - starts[p];
    });
}

int ipkgpu_stage1_owner_counts(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* counts)
{
    return guarded(ctx, [&] {
        if (n_parts < 1 || n_parts > uint32_t(RS_MAX_PEERS)) fail(IPKGPU_EINVAL, "n_parts = %u outside [1, %d]", n_parts, RS_MAX_PEERS);
        if (!counts) fail(IPKGPU_EINVAL, "counts is null");
        for (uint32_t p = 0; p < n_parts; ++p) counts[p] = 0;
        const uint64_t n = ctx->s1_n;
        if (n == 0) return;
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu records exceed one shard's 2^32-1 limit", (unsigned long long)n);
        owner_histogram(ctx, ctx->s1_keys, n, digit_owner{ ctx->s1_group_bits, n_parts }, ctx->hist);
        std::vector<uint32_t> h(n_parts);
        IPK_CUDA(cudaMemcpyAsync(h.data(), ctx->hist.p, size_t(n_parts) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        for (uint32_t p = 0; p < n_parts; ++p) counts[p] = h[p];
    });
}

int ipkgpu_recv_buffers(ipkgpu_ctx* ctx, uint64_t n_records, uint64_t** d_keys, float** d_scores, unsigned char* handle_keys,
                        unsigned char* handle_scores)
{
    return guarded(ctx, [&] {
        if (!d_keys || !d_scores) fail(IPKGPU_EINVAL, "output pointer is null");
        static_assert(sizeof(cudaIpcMemHandle_t) == IPKGPU_IPC_HANDLE_BYTES, "IPC handle size");
        // peers hold mappings of these allocations: grow rarely (25 % head room) so that handles stay valid across builds
        // (a buffer that is outgrown stays alive for one more exchange: peers that still map it re-open on the changed handle first)
        if (n_records * 8 + 8 > ctx->recv_keys.cap || n_records * 4 + 4 > ctx->recv_scores.cap)
        {
            ctx->release(ctx->recv_old); ctx->release(ctx->recv_old2);
            ctx->recv_old = ctx->recv_keys; ctx->recv_old2 = ctx->recv_scores;
            ctx->recv_keys = dbuf(); ctx->recv_scores = dbuf();
            ctx->reserve(ctx->recv_keys, (n_records + n_records / 4) * 8 + 8);
            ctx->reserve(ctx->recv_scores, (n_records + n_records / 4) * 4 + 4);
        }
        *d_keys = ctx->recv_keys.as<uint64_t>();
        *d_scores = ctx->recv_scores.as<float>();
        if (handle_keys)
        {
            cudaIpcMemHandle_t h;
            IPK_CUDA(cudaIpcGetMemHandle(&h, ctx->recv_keys.p));
            memcpy(handle_keys, &h, sizeof h);
        }
        if (handle_scores)
        {
            cudaIpcMemHandle_t h;
            IPK_CUDA(cudaIpcGetMemHandle(&h, ctx->recv_scores.p));
            memcpy(handle_scores, &h, sizeof h);
        }
    });
}

int ipkgpu_open_peer(ipkgpu_ctx* ctx, const unsigned char* handle, void** d_ptr)
{
    return guarded(ctx, [&] {
        if (!handle || !d_ptr) fail(IPKGPU_EINVAL, "null argument");
        for (auto& m : ctx->ipc_open)
            if (!memcmp(m.handle, handle, sizeof m.handle)) { *d_ptr = m.ptr; return; }
        cudaIpcMemHandle_t h;
        memcpy(&h, handle, sizeof h);
        void* p = nullptr;
        IPK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        ipkgpu_ctx::ipc_mapping m;
        memcpy(m.handle, handle, sizeof m.handle);
        m.ptr = p;
        ctx->ipc_open.push_back(m);
        *d_ptr = p;
    });
}

int ipkgpu_stage1_scatter(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* const* peer_keys, float* const* peer_scores, const uint64_t* base_offsets)
{
    return guarded(ctx, [&] {
        if (n_parts < 1 || n_parts > uint32_t(RS_MAX_PEERS)) fail(IPKGPU_EINVAL, "n_parts = %u outside [1, %d]", n_parts, RS_MAX_PEERS);
        if (!peer_keys || !peer_scores || !base_offsets) fail(IPKGPU_EINVAL, "null argument");
        const uint64_t n = ctx->s1_n;
        if (n == 0) return;
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "%llu records exceed one shard's 2^32-1 limit", (unsigned long long)n);
        rs_peer_table PT;
        memset(&PT, 0, sizeof PT);
        uint32_t base[RS_RADIX];
        memset(base, 0, sizeof base);
        for (uint32_t p = 0; p < n_parts; ++p)
        {
            if (!peer_keys[p] || !peer_scores[p]) fail(IPKGPU_EINVAL, "receive buffer of rank %u is null", p);
            if (base_offsets[p] > 0xffffffffull) fail(IPKGPU_EINVAL, "offset into rank %u's receive buffer exceeds 2^32-1", p);
            PT.keys[p] = peer_keys[p];
            PT.vals[p] = reinterpret_cast<uint32_t*>(peer_scores[p]);
            base[p] = uint32_t(base_offsets[p]);
        }
        ctx->reserve(ctx->peer_base, sizeof base);
        IPK_CUDA(cudaMemcpyAsync(ctx->peer_base.p, base, sizeof base, cudaMemcpyHostToDevice, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));          // `base` is a stack array
        scatter_to_peers(ctx, ctx->s1_keys, ctx->s1_scores, n, digit_owner{ ctx->s1_group_bits, n_parts }, ctx->peer_base.as<uint32_t>(), PT, ctx->hist);
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));          // the caller's barrier may only follow completed stores
    });
}

int ipkgpu_stage2(ipkgpu_ctx* ctx, uint64_t* d_keys, float* d_scores, uint64_t n, uint32_t group_bits,
                  const uint32_t* group_branch, uint32_t n_groups_total, uint32_t key_bits,
                  float threshold, uint32_t total_num_groups, int32_t filter, int32_t chunks_group_ordered,
                  int result_on_device, ipkgpu_result* out)
{
    return guarded(ctx, [&] {
        if (!out) fail(IPKGPU_EINVAL, "result is null");
        if (n && (!d_keys || !d_scores)) fail(IPKGPU_EINVAL, "record buffer is null");
        if (!group_branch && n_groups_total) fail(IPKGPU_EINVAL, "group_branch is null");
        if (key_bits + group_bits > 64 || key_bits == 0) fail(IPKGPU_EINVAL, "key_bits %u + group_bits %u outside (0, 64]", key_bits, group_bits);
        if (filter != IPKGPU_FILTER_MIF0 && filter != IPKGPU_FILTER_RANDOM) fail(IPKGPU_EINVAL, "unknown filter %d", filter);
        memset(out, 0, sizeof *out);
        stage_timer tm(ctx);
        const size_t t0 = tm.mark();
        ctx->reserve(ctx->group_branch, size_t(n_groups_total) * 4 + 4);
        IPK_CUDA(cudaMemcpyAsync(ctx->group_branch.p, group_branch, size_t(n_groups_total) * 4, cudaMemcpyHostToDevice, ctx->stream));
        ctx->max_branch = 0;
        for (uint32_t g = 0; g < n_groups_total; ++g) ctx->max_branch = std::max(ctx->max_branch, group_branch[g]);
        const stage2_params sp{ group_bits, key_bits, total_num_groups, threshold, filter };
        run_stage2(ctx, d_keys, d_scores, n, chunks_group_ordered != 0, sp, result_on_device, out, tm);
        const size_t t1 = tm.mark();
        IPK_CUDA(cudaEventSynchronize(ctx->event(t1)));
        out->stats.ms_total = tm.ms(t0, t1);
        out->stats.n_groups = n_groups_total;
        out->stats.kernel_launches = ctx->launches;
    });
}


// ---- multi-GPU build on packed records: the MSD pass of the merge sort is the exchange (include/ipkgpu.h) ------------

size_t ipkgpu_result_arena_layout(uint64_t cap_entries, uint64_t cap_kmers, size_t* offsets)
{
    const size_t sizes[6] = { size_t(cap_entries) * 4 + 16, size_t(cap_entries) * 4 + 16, size_t(cap_kmers) * 8 + 16, size_t(cap_kmers) * 4 + 16,
                              (size_t(cap_kmers) + 1) * 8 + 16, size_t(cap_kmers) * 4 + 16 };
    size_t at = 0;
    for (int i = 0; i < 6; ++i)
    {
        if (offsets) offsets[i] = at;
        at += (sizes[i] + 255) & ~size_t(255);
    }
    return at;
}

int ipkgpu_set_result_arena(ipkgpu_ctx* ctx, void* base, uint64_t cap_entries, uint64_t cap_kmers)
{
    return guarded(ctx, [&] {
        ipkgpu::hbuf* bufs[6] = { &ctx->h_branch, &ctx->h_score, &ctx->h_keys, &ctx->h_fv, &ctx->h_offs, &ctx->h_cnt };
        for (ipkgpu::hbuf* b : bufs)
        {
            if (b->p && !b->external) ipkgpu::pinned_free(b->p);
            b->p = nullptr; b->cap = 0; b->external = false;
        }
        if (!base) return;
        size_t off[6];
        const size_t total = ipkgpu_result_arena_layout(cap_entries, cap_kmers, off);
        for (int i = 0; i < 6; ++i)
        {
            bufs[i]->p = static_cast<char*>(base) + off[i];
            bufs[i]->cap = (i + 1 < 6 ? off[i + 1] : total) - off[i];
            bufs[i]->external = true;
        }
    });
}

int ipkgpu_enable_peer(ipkgpu_ctx* ctx, int peer_device)
{
    return guarded(ctx, [&] {
        if (peer_device == ctx->device) return;
        int can = 0;
        IPK_CUDA(cudaDeviceCanAccessPeer(&can, ctx->device, peer_device));
        if (!can) fail(IPKGPU_ENODEV, "device %d cannot access device %d directly (no NVLink / PCIe peer path)", ctx->device, peer_device);
        const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return; }
        IPK_CUDA(e);
    });
}

int ipkgpu_stage1_bucket_counts(ipkgpu_ctx* ctx, uint64_t* counts)
{
    return guarded(ctx, [&] {
        if (!counts) fail(IPKGPU_EINVAL, "counts is null");
        for (int d = 0; d < PS_BUCKETS; ++d) counts[d] = 0;
        const uint64_t n = ctx->s1_n;
        if (n == 0) return;
        if (!ps_eligible(ctx->s1_key_bits, ctx->s1_group_bits, n))
            fail(IPKGPU_EINVAL, "packed exchange needs k*bits >= 8, k*bits + group bits - 8 <= 32 and fewer than 2^32 records per rank (have %u + %u bits, %llu records)",
                 ctx->s1_key_bits, ctx->s1_group_bits, (unsigned long long)n);
        if (ctx->runs_valid && ctx->top_hist_on && ctx->runs_n > 0) ps_top_histogram_from_join(ctx);
        else ps_top_histogram(ctx, ctx->s1_keys, n, ctx->s1_key_bits + ctx->s1_group_bits - 8);
        uint32_t h[PS_BUCKETS];
        IPK_CUDA(cudaMemcpyAsync(h, ctx->hist.p, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int d = 0; d < PS_BUCKETS; ++d) counts[d] = h[d];
    });
}

int ipkgpu_exchange_plan(uint32_t n_parts, uint32_t rank, const uint64_t* counts, uint8_t* owner, uint64_t* base, uint64_t* bucket_start, uint64_t* n_recv)
{
    if (!counts || !owner || !base || !bucket_start || !n_recv || n_parts < 1 || n_parts > uint32_t(RS_MAX_PEERS) || rank >= n_parts) return IPKGPU_EINVAL;
    uint64_t total[PS_BUCKETS], grand = 0;
    for (int d = 0; d < PS_BUCKETS; ++d)
    {
        total[d] = 0;
        for (uint32_t r = 0; r < n_parts; ++r) total[d] += counts[size_t(r) * PS_BUCKETS + d];
        grand += total[d];
    }
    // contiguous k-mer ranges of about grand / n_parts records: a bucket goes to the rank its midpoint falls to
    uint64_t cum = 0;
    for (int d = 0; d < PS_BUCKETS; ++d)
    {
        const long double mid = (long double)cum + (long double)total[d] / 2;
        uint32_t o = grand ? uint32_t(mid * n_parts / (long double)grand) : 0;
        owner[d] = uint8_t(std::min(o, n_parts - 1));
        if (d && owner[d] < owner[d - 1]) owner[d] = owner[d - 1];
        cum += total[d];
    }
    // an owner's buffer: its buckets in ascending order, inside a bucket the source ranks in ascending order (= ascending groups)
    uint64_t fill[RS_MAX_PEERS];
    for (uint32_t r = 0; r < n_parts; ++r) fill[r] = 0;
    for (int d = 0; d < PS_BUCKETS; ++d)
    {
        const uint32_t o = owner[d];
        uint64_t before = 0;
        for (uint32_t r = 0; r < rank; ++r) before += counts[size_t(r) * PS_BUCKETS + d];
        base[d] = fill[o] + before;
        bucket_start[d] = fill[rank];            // where bucket d begins in THIS rank's buffer (empty range when another rank owns it)
        fill[o] += total[d];
    }
    bucket_start[PS_BUCKETS] = fill[rank];
    *n_recv = fill[rank];
    for (uint32_t r = 0; r < n_parts; ++r) if (fill[r] > 0xffffffffull) return IPKGPU_EINVAL;
    return IPKGPU_OK;
}

int ipkgpu_recv_packed(ipkgpu_ctx* ctx, uint64_t n_records, uint64_t** d_recs, unsigned char* handle)
{
    return guarded(ctx, [&] {
        if (!d_recs) fail(IPKGPU_EINVAL, "output pointer is null");
        if (n_records * 8 + 8 > ctx->recv_keys.cap)
        {
            // peers may still map the buffer being outgrown: it is kept for one more exchange (they re-open on a changed
            // handle before they store, ipkgpu_open_peer_slot) and only then freed
            ctx->release(ctx->recv_old);
            ctx->recv_old = ctx->recv_keys;
            ctx->recv_keys = dbuf();
            ctx->reserve(ctx->recv_keys, (n_records + n_records / 4) * 8 + 8);
        }
        *d_recs = ctx->recv_keys.as<uint64_t>();
        if (handle)
        {
            cudaIpcMemHandle_t h;
            IPK_CUDA(cudaIpcGetMemHandle(&h, ctx->recv_keys.p));
            memcpy(handle, &h, sizeof h);
        }
    });
}

int ipkgpu_open_peer_slot(ipkgpu_ctx* ctx, uint32_t slot, const unsigned char* handle, void** d_ptr)
{
    return guarded(ctx, [&] {
        if (!handle || !d_ptr || slot >= uint32_t(RS_MAX_PEERS)) fail(IPKGPU_EINVAL, "bad argument");
        if (ctx->ipc_slots.size() <= slot) ctx->ipc_slots.resize(slot + 1);
        auto& m = ctx->ipc_slots[slot];
        if (m.ptr && !memcmp(m.handle, handle, sizeof m.handle)) { *d_ptr = m.ptr; return; }
        if (m.ptr) { cudaIpcCloseMemHandle(m.ptr); m.ptr = nullptr; }      // the peer reallocated: drop the stale mapping first
        cudaIpcMemHandle_t h;
        memcpy(&h, handle, sizeof h);
        void* p = nullptr;
        IPK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        memcpy(m.handle, handle, sizeof m.handle);
        m.ptr = p;
        *d_ptr = p;
    });
}

int ipkgpu_stage1_scatter_packed(ipkgpu_ctx* ctx, uint32_t n_parts, uint64_t* const* peer_recs, const uint8_t* owner, const uint64_t* base,
                                 int32_t* group_ordered)
{
    return guarded(ctx, [&] {
        if (n_parts < 1 || n_parts > uint32_t(RS_MAX_PEERS)) fail(IPKGPU_EINVAL, "n_parts = %u outside [1, %d]", n_parts, RS_MAX_PEERS);
        if (!peer_recs || !owner || !base || !group_ordered) fail(IPKGPU_EINVAL, "null argument");
        const uint64_t n = ctx->s1_n;
        const bool ordered = n == 0 || (ctx->runs_valid && ctx->runs_n > 0);
        *group_ordered = ordered ? 1 : 0;
        if (n == 0) return;
        if (!ps_eligible(ctx->s1_key_bits, ctx->s1_group_bits, n)) fail(IPKGPU_EINVAL, "records do not fit the packed form");
        ps_dest D;
        memset(&D, 0, sizeof D);
        uint32_t b32[PS_BUCKETS];
        for (int d = 0; d < PS_BUCKETS; ++d)
        {
            if (owner[d] >= n_parts) fail(IPKGPU_EINVAL, "owner of bucket %d is rank %u of %u", d, owner[d], n_parts);
            if (base[d] > 0xffffffffull) fail(IPKGPU_EINVAL, "offset into rank %u's receive buffer exceeds 2^32-1", owner[d]);
            D.owner[d] = owner[d];
            b32[d] = uint32_t(base[d]);
        }
        for (uint32_t p = 0; p < n_parts; ++p)
        {
            if (!peer_recs[p]) fail(IPKGPU_EINVAL, "receive buffer of rank %u is null", p);
            D.out[p] = peer_recs[p];
        }
        ps_runs Rn;
        memset(&Rn, 0, sizeof Rn);
        if (ordered) Rn = ps_order_runs(ctx, n, ctx->s1_group_bits);
        ctx->reserve(ctx->peer_base, sizeof b32);
        IPK_CUDA(cudaMemcpyAsync(ctx->peer_base.p, b32, sizeof b32, cudaMemcpyHostToDevice, ctx->stream));
        ps_msd_pack(ctx, ctx->s1_keys, ctx->s1_scores, n, ctx->s1_key_bits + ctx->s1_group_bits - 8, ctx->peer_base.as<uint32_t>(), D, true, Rn);
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));          // the caller's barrier may only follow completed stores (and b32 is a stack array)
    });
}

int ipkgpu_stage2_packed(ipkgpu_ctx* ctx, uint64_t n, const uint64_t* bucket_start, uint32_t group_bits, int32_t group_ordered,
                         const uint32_t* group_branch, uint32_t n_groups_total, uint32_t key_bits, float threshold, uint32_t total_num_groups,
                         int32_t filter, int result_mode, ipkgpu_result* out)
{
    return guarded(ctx, [&] {
        if (!out || !bucket_start) fail(IPKGPU_EINVAL, "null argument");
        if (!group_branch && n_groups_total) fail(IPKGPU_EINVAL, "group_branch is null");
        if (filter != IPKGPU_FILTER_MIF0 && filter != IPKGPU_FILTER_RANDOM) fail(IPKGPU_EINVAL, "unknown filter %d", filter);
        if (n && !ps_eligible(key_bits, group_bits, n)) fail(IPKGPU_EINVAL, "records do not fit the packed form");
        if (n * 8 > ctx->recv_keys.cap) fail(IPKGPU_EINVAL, "%llu records exceed the receive buffer (ipkgpu_recv_packed)", (unsigned long long)n);
        memset(out, 0, sizeof *out);
        ctx->res_valid = false; ctx->res_pos = nullptr;
        stage_timer tm(ctx);
        const size_t t0 = tm.mark();
        ctx->reserve(ctx->group_branch, size_t(n_groups_total) * 4 + 4);
        IPK_CUDA(cudaMemcpyAsync(ctx->group_branch.p, group_branch, size_t(n_groups_total) * 4, cudaMemcpyHostToDevice, ctx->stream));
        ctx->max_branch = 0;
        for (uint32_t g = 0; g < n_groups_total; ++g) ctx->max_branch = std::max(ctx->max_branch, group_branch[g]);
        const stage2_params sp{ group_bits, key_bits, total_num_groups, threshold, filter };
        if (n == 0)
        {
            ctx->reserve(ctx->kmer_offs, 8);
            IPK_CUDA(cudaMemsetAsync(ctx->kmer_offs.p, 0, 8, ctx->stream));
            finish_from_csr(ctx, 0, 0, sp, result_mode, out, tm);
        }
        else
        {
            ps_table T;
            memset(&T, 0, sizeof T);
            for (int d = 0; d <= PS_BUCKETS; ++d)
            {
                if (bucket_start[d] > n || (d && bucket_start[d] < bucket_start[d - 1])) fail(IPKGPU_EINVAL, "bucket_start is not an ascending table ending at n");
                T.start[d] = uint32_t(bucket_start[d]);
            }
            if (bucket_start[PS_BUCKETS] != n) fail(IPKGPU_EINVAL, "bucket_start[256] must equal n");
            ctx->reserve(ctx->keys_a, n * 8 + 8);
            const bool parts = parts_eligible(ctx, n, key_bits, key_bits + group_bits, result_mode, key_bits < 63 ? 1ull << key_bits : ~0ull);
            const bool b16 = parts && (result_mode & 4) != 0 && ctx->max_branch < 65536u;
            if (parts) prepare_entry_stream(ctx, n, b16);
            const size_t e0 = tm.mark();
            const parts_out po = sort_merge_buckets(ctx, ctx->recv_keys.as<uint64_t>(), ctx->keys_a.as<uint64_t>(), T, n, key_bits, group_bits,
                                                    group_ordered ? group_bits : 0, 0, result_mode, parts, b16, tm, e0, e0, 0);
            out->stats.sort_passes = po.passes;
            ctx->hint_kmers_records = n;
            finish_from_csr(ctx, po.m.n_kmers, po.m.n_records, sp, result_mode, out, tm);
            out->stats.ms_sort = po.ms_sort;
            out->stats.ms_merge = po.ms_merge;
        }
        const size_t t1 = tm.mark();
        IPK_CUDA(cudaEventSynchronize(ctx->event(t1)));
        out->stats.ms_total = tm.ms(t0, t1);
        out->stats.n_groups = n_groups_total;
        out->stats.kernel_launches = ctx->launches;
    });
}

}   // extern "C"

// ---- allocation-free entry point (include/ipkgpu.h: no reference counterpart) ---------
//
// The column maxima between k1a_colmax and k1b_prefix live in a file-scope device array instead of the context's workspace.
// In host code `g_noalloc_colmax` is only the host-side stand-in of the symbol: the device address comes from
// cudaGetSymbolAddress, taken once per device with that device current, and only that pointer is given to the kernels.
static __device__ __align__(256) float g_noalloc_colmax[IPKGPU_NOALLOC_MAX_COLUMNS];

namespace
{
    constexpr int NOALLOC_MAX_DEVICES = 64;
    struct noalloc_device { float* colmax = nullptr; int num_sms = 0; bool ready = false; };
    noalloc_device g_noalloc_dev[NOALLOC_MAX_DEVICES];
    std::mutex g_noalloc_mutex;
    thread_local char g_noalloc_error[256] = "";

    int noalloc_fail(int code, const char* fmt, ...)
    {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(g_noalloc_error, sizeof g_noalloc_error, fmt, ap);
        va_end(ap);
        return code;
    }

    /// current device for the scope of one call; the caller's device is put back on the way out
    struct noalloc_device_scope
    {
        int prev = -1;
        bool switched = false;
        cudaError_t enter(int device)
        {
            cudaError_t e = cudaGetDevice(&prev);
            if (e != cudaSuccess) return e;
            if (prev != device) { e = cudaSetDevice(device); switched = e == cudaSuccess; }
            return e;
        }
        ~noalloc_device_scope() { if (switched) cudaSetDevice(prev); }
    };

    /// set-up of one device (its device current): address of the scratch symbol there, SM count for the grid
    int noalloc_setup(int device, noalloc_device** out)
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        noalloc_device& d = g_noalloc_dev[device];
        if (!d.ready)
        {
            cudaDeviceProp prop;
            cudaError_t e = cudaGetDeviceProperties(&prop, device);
            if (e != cudaSuccess) { cudaGetLastError(); return noalloc_fail(IPKGPU_ECUDA, "cudaGetDeviceProperties(%d): %s", device, cudaGetErrorString(e)); }
            if (prop.major != 10) return noalloc_fail(IPKGPU_ENODEV, "device %d is sm_%d%d, not sm_100 (this library is built for sm_100a only)", device, prop.major, prop.minor);
            float* p = nullptr;
            e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_colmax);
            if (e != cudaSuccess || !p) { cudaGetLastError(); return noalloc_fail(IPKGPU_ECUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(e)); }
            d.colmax = p;
            d.num_sms = prop.multiProcessorCount;
            d.ready = true;
        }
        *out = &d;
        return IPKGPU_OK;
    }

    bool noalloc_on_device(const void* p, int device)
    {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
        return (a.type == cudaMemoryTypeDevice && a.device == device) || a.type == cudaMemoryTypeManaged;
    }
}

extern "C" {

const char* ipkgpu_noalloc_last_error(void) { return g_noalloc_error; }

int ipkgpu_best_prefix_noalloc(int device, void* stream, const float* d_logp, uint32_t n_nodes, uint32_t width, uint32_t sigma, float* d_B)
{
    using namespace ipkgpu;
    g_noalloc_error[0] = 0;
    if (sigma < 2 || sigma > 32) return noalloc_fail(IPKGPU_EINVAL, "sigma = %u outside [2, 32]", sigma);
    const uint64_t cols = uint64_t(n_nodes) * width;
    if (cols > IPKGPU_NOALLOC_MAX_COLUMNS)
        return noalloc_fail(IPKGPU_EINVAL, "n_nodes * width = %llu exceeds the static scratch of %u columns (IPKGPU_NOALLOC_MAX_COLUMNS); use ipkgpu_best_prefix",
                            (unsigned long long)cols, IPKGPU_NOALLOC_MAX_COLUMNS);
    if (n_nodes == 0) return IPKGPU_OK;
    if (!d_B || (cols && !d_logp)) return noalloc_fail(IPKGPU_EINVAL, "d_logp / d_B is null");
    if (sigma == 4 && (reinterpret_cast<uintptr_t>(d_logp) & 15) != 0) return noalloc_fail(IPKGPU_EINVAL, "d_logp must be 16-byte aligned for 4 states");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) { cudaGetLastError(); return noalloc_fail(IPKGPU_ENODEV, "no CUDA device: %s", e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"); }
    if (device < 0 || device >= n_dev || device >= NOALLOC_MAX_DEVICES) return noalloc_fail(IPKGPU_EINVAL, "device ordinal %d out of range", device);

    noalloc_device_scope scope;
    e = scope.enter(device);
    if (e != cudaSuccess) { cudaGetLastError(); return noalloc_fail(IPKGPU_ECUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e)); }
    noalloc_device* dev = nullptr;
    if (const int rc = noalloc_setup(device, &dev)) return rc;
    if ((cols && !noalloc_on_device(d_logp, device)) || !noalloc_on_device(d_B, device))
        return noalloc_fail(IPKGPU_EINVAL, "d_logp / d_B must be device memory of device %d", device);

    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (cols)
    {
        const uint32_t grid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(cols, 256), uint64_t(dev->num_sms) * 16));
        k1a_colmax<<<grid, 256, 0, s>>>(d_logp, dev->colmax, n_nodes, width, sigma);
        e = cudaGetLastError();
        if (e != cudaSuccess) return noalloc_fail(IPKGPU_ECUDA, "k1a_colmax launch: %s", cudaGetErrorString(e));
    }
    k1b_prefix<<<div_up<uint32_t>(n_nodes, 64), 64, 0, s>>>(dev->colmax, d_B, n_nodes, width);
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail(IPKGPU_ECUDA, "k1b_prefix launch: %s", cudaGetErrorString(e));
    return IPKGPU_OK;
}

}   // extern "C"
