// enumerate.cuh -- K1 (best-score prefix sums) and K2 (phylo-k-mer enumeration) for sm_100a.
//
// Replaces, with a different algorithm but bit-identical output:
//   matrix::preprocess / range_max_sum      ipk/src/window.cpp:16-27,69-72
//   to_windows                               ipk/src/window.cpp:159-197,289-301
//   as_column, DCLA::DC, DCLA::run           ipk/src/pk_compute.cpp:14-114
//
// What DCLA emits for a window is order-free: with T the floor(h/2) | h-floor(h/2) split tree of
// [0,k), eps(root) = log_threshold, eps(left) = eps(v) - rms(right interval), eps(right) = eps(v) - rms(left interval)
// (each one fp32 subtraction of one fp32 difference of the whole-alignment prefix sums B), S(leaf j) = M[s+j][c_j],
// S(v) = S(left) + S(right) in fp32:   k-mer emitted  <=>  for all v in T: S(v) > eps(v);   score = S(root).
//
// K2 evaluates exactly that predicate, but walks the k-mer space as a depth-first branch-and-bound over positions
// 0..k-1 instead of building and joining per-node lists: one thread per window, states of every column visited in
// descending score order, tree nodes checked the moment their last leaf is placed.  All checks are monotone in the
// score of the state being placed, so the first failing state ends the scan of a column.  One extra bound that the
// reference does not have -- prefix sum + best completion < log_threshold - MARGIN -- prunes branches that cannot
// pass the ROOT check; MARGIN is three orders of magnitude above the fp32 error of those short local sums, so it
// never removes a k-mer the predicate accepts.
#pragma once
#include "common.cuh"

namespace ipkgpu
{
#ifndef IPK_ENUM_THREADS
#define IPK_ENUM_THREADS 128
#endif
    constexpr int ENUM_THREADS = IPK_ENUM_THREADS;        // windows per CTA tile
    constexpr int ENUM_WARPS = ENUM_THREADS / 32;
    constexpr int MAX_K = 32;
    constexpr int STAGE_CAP = 96;            // per-warp staging records; flushed when fewer than 32 slots remain
    constexpr float PRUNE_MARGIN = 1e-3f;

    /// Static description of the split tree of [0,k), built once per call on the host.
    struct enum_program
    {
        int k, n_eps_ops, n_nodes;
        struct { uint8_t dst, src, hi, lo; } eps_ops[2 * MAX_K];      // eps[dst] = eps[src] - (B[s+hi] - B[s+lo])
        uint8_t leaf_eps[MAX_K];                                      // tree-node id of leaf p (indexes eps[] and val[])
        uint8_t merge_begin[MAX_K + 1];                               // merges completed by placing position p
        struct { uint8_t id, left; } merges[MAX_K];                   // S = val[left] + val[left+1]; check S > eps[id]; val[id] = S
    };

    namespace detail
    {
        // Every tree node owns one partial-sum cell val[id]: a depth-first search revisits earlier positions, so a
        // cell may not be recycled between subtrees (the left operand of a merge must survive the backtracking).
        inline void build_node(enum_program& P, int j, int h, int id, int& next_id, int& n_merges)
        {
            if (h == 1)
            {
                P.leaf_eps[j] = uint8_t(id);
                P.merge_begin[j] = uint8_t(n_merges);
                return;
            }
            const int hl = h / 2;                 // pk_compute.cpp:57-58: DC(j, h/2), DC(j + h/2, h - h/2)
            const int l = next_id++, r = next_id++;
            auto& a = P.eps_ops[P.n_eps_ops++];   // :54  eps_l = eps - rms(j + h/2, h - h/2)
            a.dst = uint8_t(l); a.src = uint8_t(id); a.hi = uint8_t(j + h); a.lo = uint8_t(j + hl);
            auto& b = P.eps_ops[P.n_eps_ops++];   // :55  eps_r = eps - rms(j, h/2)
            b.dst = uint8_t(r); b.src = uint8_t(id); b.hi = uint8_t(j + hl); b.lo = uint8_t(j);
            build_node(P, j, hl, l, next_id, n_merges);
            build_node(P, j + hl, h - hl, r, next_id, n_merges);
            auto& m = P.merges[n_merges++];
            m.id = uint8_t(id); m.left = uint8_t(l);                  // right child = l + 1
        }
    }

    inline enum_program make_enum_program(int k)
    {
        if (k < 1 || k > MAX_K) fail(IPKGPU_EINVAL, "k = %d outside [1, %d]", k, MAX_K);
        enum_program P;
        memset(&P, 0, sizeof P);
        P.k = k;
        int next_id = 1, n_merges = 0;
        detail::build_node(P, 0, k, 0, next_id, n_merges);
        P.merge_begin[k] = uint8_t(n_merges);
        P.n_nodes = next_id;
        return P;
    }

    // ----------------------------------------------------------------------------------------------------------
    // K1: best-score prefix sums.  The fp32 accumulation order is part of the reference's result:
    // column maxima are taken fully in parallel, the running sum is ONE sequential fp32 chain per node.
    // ----------------------------------------------------------------------------------------------------------

    /// colmaxT[j][n] = max_i logp[n][j][i]   (transposed so that K1b reads and K2 never needs it)
    __global__ void k1a_colmax(const float* __restrict__ logp, float* __restrict__ colmaxT,
                               uint32_t n_nodes, uint32_t W, uint32_t sigma)
    {
        const uint64_t total = uint64_t(n_nodes) * W;
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < total; i += uint64_t(gridDim.x) * blockDim.x)
        {
            const float* col = logp + i * sigma;
            float best;
            if (sigma == 4)
            {
                const float4 v = *reinterpret_cast<const float4*>(col);
                best = v.x;                                   // std::max_element: first maximum; value is all we need
                if (best < v.y) best = v.y;
                if (best < v.z) best = v.z;
                if (best < v.w) best = v.w;
            }
            else
            {
                best = col[0];
                for (uint32_t s = 1; s < sigma; ++s) { const float x = col[s]; if (best < x) best = x; }
            }
            const uint32_t n = uint32_t(i / W), j = uint32_t(i % W);
            colmaxT[uint64_t(j) * n_nodes + n] = best;
        }
    }

    /// B[n][0] = 0; B[n][j+1] = B[n][j] + colmax[n][j]  -- one thread per node, strictly sequential fp32 (window.cpp:19-26)
    __global__ void k1b_prefix(const float* __restrict__ colmaxT, float* __restrict__ B, uint32_t n_nodes, uint32_t W)
    {
        const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
        if (n >= n_nodes) return;
        float* row = B + uint64_t(n) * (W + 1);
        float product = 0.0f;
        row[0] = 0.0f;
        for (uint32_t j = 0; j < W; ++j)
        {
            product = __fadd_rn(product, colmaxT[uint64_t(j) * n_nodes + n]);
            row[j + 1] = product;
        }
    }

    // ----------------------------------------------------------------------------------------------------------
    // K2: enumeration
    // ----------------------------------------------------------------------------------------------------------

    /// One commit of the join kernel: `count` records at `offset` of the emission stream, all of branch group `group`
    /// (a tile lies inside one node).  Read in group order, the runs ARE the stream sorted by group: the first pass of the
    /// merge sort walks them in that order (packed_sort.cuh), so no radix pass is ever spent on the group bits.
    struct emit_run
    {
        unsigned long long offset;
        uint32_t count, group;
    };

    struct enum_args
    {
        const float* logp;            // [n_nodes][W][sigma]
        const float* B;               // [n_nodes][W+1]
        const uint32_t* node_order;   // [n_nodes] node ids sorted by group (stable): launch slot -> node
        const uint32_t* node_group;   // [n_nodes] group index of a node (global, i.e. already offset by group_base)
        uint64_t* out_keys;           // composite (kmer << group_bits) | group, or plain kmer when DEBUG
        float* out_scores;
        uint32_t* out_node;           // DEBUG only
        uint32_t* out_win;            // DEBUG only
        unsigned long long* counter;  // [0] = total emitted
        unsigned long long* node_counts;   // [n_nodes] emitted per launch slot (may be null)
        uint64_t cap;
        uint32_t slot_base;           // first launch slot of this batch
        uint32_t group_sub;           // subtracted from the group index: batch-local group field of group_bits bits
        uint32_t n_nodes, W, sigma, bits, group_bits, tiles_per_node, nwin;
        float log_threshold;
        int use_tma;
        // Work items.  An item is a run of item_windows consecutive windows (a power of two <= ENUM_THREADS) inside one
        // (node, ENUM_THREADS-window) tile: item id = tile * items_per_tile + sub.  blockIdx.x is the item id itself, or -- when
        // tile_list is set -- child blockIdx.x % parent_ratio of the coarser item tile_list[blockIdx.x / parent_ratio].
        const uint32_t* tile_list;
        uint32_t item_windows, items_per_tile, parent_ratio;
        uint32_t* deferred_tiles;     // join kernel: items it could not finish (lists too long for the pool ...); their number is counter[1]
        uint32_t pool_cap;            // join kernel: list-pool capacity in elements (<= 65535)
        // ordered emission (level 0 of the join kernel): status word per tile, flag << 62 | records; a tile's output offset is the
        // sum over all tiles BEFORE it (decoupled look-back), so the record stream is in tile order = branch-group order
        unsigned long long* tile_status;
        unsigned long long* phase_clocks;   // debug (IPKGPU_PHASE_CLOCKS): cycles of thread 0 per phase of the join kernel, summed over tiles
        emit_run* runs;               // join kernel: one entry per committing item (null: not recorded) ...
        unsigned long long* run_counter;   // ... their number so far
        uint32_t run_cap;
        uint32_t* top_hist;           // join kernel: [JOIN_HIST_SLOTS][256] counts of the emitted records by the top byte of their k-mer (null: not
                                      // counted) -- the histogram the first pass of the merge sort needs, taken where the records are made
        uint32_t top_shift;           // ... that byte = (key of the root's left element) >> top_shift
        uint32_t group_min;           // join kernel: lists of at least this many elements are built by a lane group
        uint32_t small_list;          // join kernel: lists up to this length are filled and sorted by their owner thread
    };

    struct tile_geom
    {
        uint32_t item, lslot, node, s0;
        int nwin;
    };

    /// false: the item lies past the node's last window (nothing to do)
    __device__ __forceinline__ bool tile_geometry(const enum_args& A, tile_geom& g)
    {
        uint32_t item = blockIdx.x;
        if (A.tile_list) item = A.tile_list[blockIdx.x / A.parent_ratio] * A.parent_ratio + blockIdx.x % A.parent_ratio;
        const uint32_t tile = item / A.items_per_tile, sub = item % A.items_per_tile;
        g.item = item;
        g.lslot = A.slot_base + tile / A.tiles_per_node;
        g.node = A.node_order[g.lslot];
        g.s0 = (tile % A.tiles_per_node) * ENUM_THREADS + sub * A.item_windows;
        if (g.s0 >= A.nwin) { g.nwin = 0; return false; }
        g.nwin = int(min(A.item_windows, A.nwin - g.s0));
        return true;
    }

    __device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

    /// Dynamic shared memory layout (floats unless noted), NC = ENUM_THREADS + k - 1 columns:
    ///   sval [sigma][NCP]   column states' scores, descending          NCP = NC rounded up to a multiple of 4
    ///   bs   [NCP+4]        B[n][s0 .. s0+NC]
    ///   pt   [6k][ENUM_THREADS]  per-thread arrays (eps 2k, lim k, ps k, val 2k), index-major => bank = lane, never conflicts
    ///        (aliased, before the DFS starts, by the raw TMA-staged tile [NC][sigma])
    ///   stage keys u64 [ENUM_WARPS][STAGE_CAP], scores f32 [...], (DEBUG: win u32 [...])
    ///   sidx u8 [sigma][NCP]
    inline size_t enum_smem_bytes(int k, int sigma, bool debug)
    {
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;
        size_t pt = size_t(6 * k) * ENUM_THREADS * 4;
        size_t raw = size_t(NC) * sigma * 4;
        size_t b = 0;
        b += size_t(sigma) * NCP * 4;                 // sval
        b += size_t(NCP + 4) * 4;                     // bs
        b += std::max(pt, raw);                       // pt / raw
        b = (b + 15) & ~size_t(15);
        b += size_t(ENUM_WARPS) * STAGE_CAP * 8;      // stage keys
        b += size_t(ENUM_WARPS) * STAGE_CAP * 4;      // stage scores
        if (debug) b += size_t(ENUM_WARPS) * STAGE_CAP * 4;
        b += size_t(sigma) * NCP;                     // sidx
        return b + 16;
    }

    /// Common prologue of both enumeration kernels: the tile's ncols*sigma contiguous floats of logp[node][s0..] go to
    /// `raw` (one 1-D TMA bulk copy when alignment allows), the prefix sums B[n][s0 .. s0+ncols] to `bs`, and every
    /// column's states are ordered by descending score into sval / sidx (ties by state index).
    /// Ends WITHOUT a barrier: callers __syncthreads() before reusing `raw` or reading sval.
    __device__ __forceinline__ void stage_tile(const enum_args& A, uint32_t node, uint32_t s0, int ncols, int NCP, float* raw, float* bs,
                                               float* sval, uint8_t* sidx, unsigned long long* mbar_p, const int nthreads = ENUM_THREADS)
    {
        const int tid = threadIdx.x;
        const int sigma = int(A.sigma);
        unsigned long long& mbar = *mbar_p;
        // ---- stage the tile: ncols*sigma contiguous floats of logp[node][s0..] ------------------------------------
        const float* gsrc = A.logp + (uint64_t(node) * A.W + s0) * sigma;
        const uint32_t tile_bytes = uint32_t(ncols) * sigma * 4;
        if (A.use_tma)
        {
            // TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
            if (tid == 0)
            {
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            }
            __syncthreads();
            if (tid == 0)
            {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&mbar)), "r"(tile_bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(raw)), "l"(gsrc), "r"(tile_bytes), "r"(smem_u32(&mbar)) : "memory");
            }
        }
        else
        {
            for (int i = tid; i < ncols * sigma; i += nthreads) raw[i] = gsrc[i];
        }
        {
            const float* brow = A.B + uint64_t(node) * (A.W + 1) + s0;
            for (int i = tid; i <= ncols; i += nthreads) bs[i] = brow[i];
        }
        if (A.use_tma)
        {
            uint32_t done = 0;
            while (!done)
            {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done) : "r"(smem_u32(&mbar)) : "memory");
            }
        }
        __syncthreads();

        // ---- per column: order the states by descending score (ties by state index; tied states behave identically)
        for (int c = tid; c < ncols; c += nthreads)
        {
            const float* col = raw + c * sigma;
            if (sigma == 4)
            {
                float v0 = col[0], v1 = col[1], v2 = col[2], v3 = col[3];
                int i0 = 0, i1 = 1, i2 = 2, i3 = 3;
#define IPK_CSWAP(a, b, ia, ib) if (a < b) { float t = a; a = b; b = t; int ti = ia; ia = ib; ib = ti; }
                IPK_CSWAP(v0, v1, i0, i1) IPK_CSWAP(v2, v3, i2, i3) IPK_CSWAP(v0, v2, i0, i2) IPK_CSWAP(v1, v3, i1, i3) IPK_CSWAP(v1, v2, i1, i2)
#undef IPK_CSWAP
                sval[0 * NCP + c] = v0; sval[1 * NCP + c] = v1; sval[2 * NCP + c] = v2; sval[3 * NCP + c] = v3;
                sidx[0 * NCP + c] = uint8_t(i0); sidx[1 * NCP + c] = uint8_t(i1); sidx[2 * NCP + c] = uint8_t(i2); sidx[3 * NCP + c] = uint8_t(i3);
            }
            else
            {
                // rank of each state by counting: O(sigma^2) compares, no local arrays
                for (int i = 0; i < sigma; ++i)
                {
                    const float v = col[i];
                    int r = 0;
                    for (int j = 0; j < sigma; ++j)
                    {
                        const float w = col[j];
                        r += (w > v) || (w == v && j < i) || (v != v && (w == w || j < i));   // NaN sorts last
                    }
                    sval[r * NCP + c] = v;
                    sidx[r * NCP + c] = uint8_t(i);
                }
            }
        }
    }

    template<bool DEBUG>
    __global__ void __launch_bounds__(ENUM_THREADS)
    k2_enumerate(const __grid_constant__ enum_args A, const __grid_constant__ enum_program P)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        __shared__ __align__(8) unsigned long long mbar;
        __shared__ uint32_t s_pos[MAX_K];      // per position: leaf node id | merge begin << 16 | merge end << 24
        __shared__ uint32_t s_merge[MAX_K];    // per merge: node id | left child id << 8

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int k = P.k, sigma = int(A.sigma);
        tile_geom G;
        if (!tile_geometry(A, G)) return;
        const uint32_t lslot = G.lslot, node = G.node, s0 = G.s0;
        const int nwin_tile = G.nwin;
        const int ncols = nwin_tile + k - 1;
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;

        float* sval = reinterpret_cast<float*>(smem_raw);
        float* bs = sval + sigma * NCP;
        float* pt = bs + NCP + 4;
        const size_t pt_bytes = size_t(6 * k) * ENUM_THREADS * 4, raw_bytes = size_t(NC) * sigma * 4;
        size_t off = (size_t(sigma) * NCP * 4 + size_t(NCP + 4) * 4 + (pt_bytes > raw_bytes ? pt_bytes : raw_bytes) + 15) & ~size_t(15);
        unsigned long long* st_keys = reinterpret_cast<unsigned long long*>(smem_raw + off) + warp * STAGE_CAP;
        off += size_t(ENUM_WARPS) * STAGE_CAP * 8;
        float* st_scores = reinterpret_cast<float*>(smem_raw + off) + warp * STAGE_CAP;
        off += size_t(ENUM_WARPS) * STAGE_CAP * 4;
        uint32_t* st_win = nullptr;
        if (DEBUG) { st_win = reinterpret_cast<uint32_t*>(smem_raw + off) + warp * STAGE_CAP; off += size_t(ENUM_WARPS) * STAGE_CAP * 4; }
        uint8_t* sidx = smem_raw + off;
        float* raw = pt;                                   // aliased: consumed before the per-thread arrays are written

        if (tid < MAX_K)
        {
            // lanes sit at different depths of their searches: keep the tree tables where divergent reads are cheap
            s_pos[tid] = tid < k ? (uint32_t(P.leaf_eps[tid]) | uint32_t(P.merge_begin[tid]) << 16 | uint32_t(P.merge_begin[tid + 1]) << 24) : 0u;
            s_merge[tid] = uint32_t(P.merges[tid].id) | uint32_t(P.merges[tid].left) << 8;
        }
        stage_tile(A, node, s0, ncols, NCP, raw, bs, sval, sidx, &mbar);
        __syncthreads();      // raw fully consumed; pt may now be written

        // ---- per-thread tables (index-major in shared memory) ------------------------------------------------------
        float* EPS = pt + tid;                                   // [n_nodes <= 2k-1]   EPS[i*ENUM_THREADS]
        float* LIM = pt + (2 * k) * ENUM_THREADS + tid;          // [k]
        float* PS = pt + (3 * k) * ENUM_THREADS + tid;           // [k]
        float* VAL = pt + (4 * k) * ENUM_THREADS + tid;          // [n_nodes <= 2k-1]   partial sum of tree node i
        bool active = tid < nwin_tile;
        if (active)
        {
            EPS[0] = A.log_threshold;
            for (int i = 0; i < P.n_eps_ops; ++i)
            {
                const auto op = P.eps_ops[i];
                const float r = __fsub_rn(bs[tid + op.hi], bs[tid + op.lo]);          // range_max_sum, window.cpp:69-72
                EPS[op.dst * ENUM_THREADS] = __fsub_rn(EPS[op.src * ENUM_THREADS], r);   // pk_compute.cpp:54-55
            }
            float suf = 0.0f;
            for (int p = k - 1; p >= 0; --p)
            {
                LIM[p * ENUM_THREADS] = (A.log_threshold - PRUNE_MARGIN) - suf;      // bound for the prefix ending at p
                suf += sval[tid + p];                                                 // best state of column s+p
            }
        }

        // ---- depth-first branch and bound ---------------------------------------------------------------------------
        const unsigned full = 0xffffffffu;
        const uint32_t bits = A.bits;
        const uint64_t state_mask = (1ull << bits) - 1;
        const uint64_t group = DEBUG ? 0 : uint64_t(A.node_group[node] - A.group_sub);
        int p = 0, r = 0;
        uint64_t rk = 0, key = 0;
        int wcount = 0;                        // records staged by this warp (warp-uniform)
        unsigned long long emitted = 0;        // by this warp (warp-uniform)

        while (__any_sync(full, active))
        {
            bool emit = false;
            uint64_t ekey = 0;
            float escore = 0.0f;
            if (active)
            {
                bool ok = r < sigma;
                float v = 0.0f, ps = 0.0f, S = 0.0f;
                int c = 0;
                const uint32_t pi = s_pos[p];
                if (ok)
                {
                    const int col = tid + p;
                    v = sval[r * NCP + col];
                    c = sidx[r * NCP + col];
                    ps = p ? __fadd_rn(PS[(p - 1) * ENUM_THREADS], v) : v;
                    ok = (v > EPS[(pi & 0xff) * ENUM_THREADS]) && (ps >= LIM[p * ENUM_THREADS]);   // as_column: strict >
                }
                if (ok)
                {
                    VAL[(pi & 0xff) * ENUM_THREADS] = v;
                    S = v;
                    for (uint32_t m = (pi >> 16) & 0xff; m < (pi >> 24); ++m)
                    {
                        const uint32_t mg = s_merge[m];
                        const uint32_t left = mg >> 8;
                        S = __fadd_rn(VAL[left * ENUM_THREADS], VAL[(left + 1) * ENUM_THREADS]);         // pk_compute.cpp:90
                        if (!(S > EPS[(mg & 0xff) * ENUM_THREADS])) { ok = false; break; }                // :91 `score <= eps` => drop
                        VAL[(mg & 0xff) * ENUM_THREADS] = S;
                    }
                }
                if (ok)
                {
                    if (p == k - 1)
                    {
                        emit = true;
                        ekey = (key << bits) | uint64_t(c);          // position 0 = most significant digits (:96-104)
                        escore = S;
                        ++r;
                    }
                    else
                    {
                        PS[p * ENUM_THREADS] = ps;
                        rk = (rk << bits) | uint64_t(r);
                        key = (key << bits) | uint64_t(c);
                        ++p;
                        r = 0;
                    }
                }
                else
                {
                    // states are visited in descending score and every check is monotone: nothing further at p can pass
                    if (p == 0) active = false;
                    else
                    {
                        --p;
                        r = int(rk & state_mask) + 1;
                        rk >>= bits;
                        key >>= bits;
                    }
                }
            }
            // ---- warp-aggregated staging in shared memory, flushed with ONE global atomic per <= STAGE_CAP records --
            const unsigned em = __ballot_sync(full, emit);
            if (em)
            {
                const int n = __popc(em);
                if (wcount + n > STAGE_CAP)
                {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(A.counter, (unsigned long long)wcount);
                    base = __shfl_sync(full, base, 0);
                    __syncwarp();
                    for (int i = lane; i < wcount; i += 32)
                    {
                        const unsigned long long o = base + i;
                        if (o < A.cap)
                        {
                            A.out_keys[o] = st_keys[i];
                            A.out_scores[o] = st_scores[i];
                            if (DEBUG) { A.out_node[o] = node; A.out_win[o] = st_win[i]; }
                        }
                    }
                    __syncwarp();
                    wcount = 0;
                }
                if (emit)
                {
                    const int slot = wcount + __popc(em & ((1u << lane) - 1));
                    st_keys[slot] = DEBUG ? ekey : ((ekey << A.group_bits) | group);
                    st_scores[slot] = escore;
                    if (DEBUG) st_win[slot] = s0 + tid;
                }
                wcount += n;
                emitted += n;
            }
        }
        __syncwarp();
        if (wcount)
        {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(A.counter, (unsigned long long)wcount);
            base = __shfl_sync(full, base, 0);
            for (int i = lane; i < wcount; i += 32)
            {
                const unsigned long long o = base + i;
                if (o < A.cap)
                {
                    A.out_keys[o] = st_keys[i];
                    A.out_scores[o] = st_scores[i];
                    if (DEBUG) { A.out_node[o] = node; A.out_win[o] = st_win[i]; }
                }
            }
        }
        if (A.node_counts && lane == 0 && emitted) atomicAdd(A.node_counts + lslot, emitted);
    }
}
