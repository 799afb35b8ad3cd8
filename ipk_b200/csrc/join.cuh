// join.cuh -- K2 fast path: phylo-k-mer enumeration of a (node, 128-window) tile by shared list joins, sm_100a.
//
// Replaces, with bit-identical output:  to_windows + DCLA::run per window
//   ipk/src/window.cpp:159-197, ipk/src/pk_compute.cpp:14-114 (driven by ipk/src/db_builder.cpp:645-649)
//
// The reference builds, per window, the list of every tree node v of the floor(h/2) | h-floor(h/2) split tree:
//   L_v(s) = { x : every node u under v has S_u(x) > eps_u(s) }   and emits L_root(s).
// S_u(x) depends only on the ABSOLUTE columns u covers and on the split shape, not on the window; only eps_u(s) is
// per window.  So one list per (length h, absolute column j) serves every window that has a length-h node there:
//   G_h(j) = { (a,b) : a in G_hl(j), b in G_hr(j+hl), fl(S(a)+S(b)) > g_h(j) },  g_h(j) = min over the tile's windows s
//            and nodes v (length h, column j) of eps_v(s)            -- a superset of every L_v(s), same fp32 sums.
// Per window, L_left(s) / L_right(s) are then cut out of the two top lists (sorted by descending score):
//   S(a) > eps_left(s) is tested exactly; the deeper checks of `a` are implied once S(a) clears eps_left(s) by more than
//   DELTA (the rounding slack of the reference's fp32 threshold chain, bounded from |B|), and are evaluated one by one
//   (verify_subtree: the predicate itself) for the rare elements inside that band.  The root join uses the reference's own
//   formula fl(S(a)+S(b)) > eps (pk_compute.cpp:90-94), so scores and the emitted set are bit-identical.
// A tile whose lists do not fit the shared-memory pool is not enumerated here: its id goes to a deferred list and the
// depth-first kernel (enumerate.cuh) handles it.
#pragma once
#include "common.cuh"
#include "enumerate.cuh"

namespace ipkgpu
{
    constexpr int JOIN_MAX_FAM = 12;
    constexpr int JOIN_MAX_USES = 32;

    /// Families = distinct lengths h of the non-root tree nodes, ascending.  Built once per call on the host.
    struct join_program
    {
        int n_fam;
        int root_left, root_right;                 // family index of the root's children
        int root_hl;                               // length of the root's left child
        struct
        {
            uint8_t h, hl;                         // node length, left-child length (h/2); h == 1: leaf
            uint8_t left, right;                   // child families (h > 1)
            uint8_t n_uses;
            uint8_t use_node[JOIN_MAX_USES];       // tree-node ids (index into eps[]) of the nodes of this length
            uint8_t use_off[JOIN_MAX_USES];        // ... and their offsets in the window
        } fam[JOIN_MAX_FAM];
    };

    namespace detail
    {
        inline int fam_index(join_program& J, int h)
        {
            for (int f = 0; f < J.n_fam; ++f) if (J.fam[f].h == h) return f;
            if (J.n_fam >= JOIN_MAX_FAM) fail(IPKGPU_EINVAL, "split tree has more than %d distinct node lengths", JOIN_MAX_FAM);
            auto& F = J.fam[J.n_fam];
            memset(&F, 0, sizeof F);
            F.h = uint8_t(h);
            F.hl = uint8_t(h / 2);
            return J.n_fam++;
        }
        // walks the tree in the same order as build_node (enumerate.cuh), so node ids agree with enum_program
        inline void join_walk(join_program& J, int j, int h, int id, int& next_id, bool is_root)
        {
            if (!is_root)
            {
                const int f = fam_index(J, h);
                auto& F = J.fam[f];
                if (F.n_uses >= JOIN_MAX_USES) fail(IPKGPU_EINVAL, "more than %d tree nodes of length %d", JOIN_MAX_USES, h);
                F.use_node[F.n_uses] = uint8_t(id);
                F.use_off[F.n_uses] = uint8_t(j);
                F.n_uses++;
            }
            if (h == 1) return;
            const int hl = h / 2;
            const int l = next_id++, r = next_id++;
            join_walk(J, j, hl, l, next_id, false);
            join_walk(J, j + hl, h - hl, r, next_id, false);
        }
    }

    /// k >= 2 (k == 1 has no join: the depth-first kernel handles it)
    inline join_program make_join_program(int k)
    {
        join_program J;
        memset(&J, 0, sizeof J);
        int next_id = 1;
        detail::join_walk(J, 0, k, 0, next_id, true);
        // ascending h, so that children are built before parents
        for (int a = 0; a < J.n_fam; ++a)
            for (int b = a + 1; b < J.n_fam; ++b)
                if (J.fam[b].h < J.fam[a].h) std::swap(J.fam[a], J.fam[b]);
        auto find = [&](int h) { for (int f = 0; f < J.n_fam; ++f) if (J.fam[f].h == h) return f; return -1; };
        for (int f = 0; f < J.n_fam; ++f)
        {
            auto& F = J.fam[f];
            if (F.h > 1) { F.left = uint8_t(find(F.hl)); F.right = uint8_t(find(F.h - F.hl)); }
        }
        J.root_hl = k / 2;
        J.root_left = find(k / 2);
        J.root_right = find(k - k / 2);
        return J;
    }

    // ----------------------------------------------------------------------------------------------------------
    // shared-memory layout of k2_join (NC = ENUM_THREADS + k - 1 local columns, NCP = NC rounded up to 4):
    //   sval  f32 [sigma][NCP]          column states' scores, descending
    //   bs    f32 [NCP+4]               B[n][s0 .. s0+NC]
    //   eps   f32 [2k][ENUM_THREADS]    eps of every tree node for every window of the tile (index-major)
    //   fcnt  u16 [n_fam][NCP]          length of G_h(c)
    //   foff  u16 [n_fam][NCP]          its start in the pool
    //   gthr  f32 [NCP]                 g_h(c) of the family being built
    //   pool_score f32 [pool_cap], pool_key u32 [pool_cap]     (the raw TMA-staged tile aliases the pool's start)
    //   stage keys u64 [ENUM_WARPS][STAGE_CAP], scores f32 [...], (DEBUG: win u32 [...])
    //   sidx  u8 [sigma][NCP]
    // ----------------------------------------------------------------------------------------------------------
    struct join_layout
    {
        size_t sval, bs, eps, fcnt, foff, gthr, pool_score, pool_key, st_keys, st_scores, st_win, sidx, total;
    };

    __host__ __device__ inline join_layout make_join_layout(int k, int sigma, int n_fam, uint32_t pool_cap, bool debug)
    {
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;
        join_layout L;
        size_t b = 0;
        L.sval = b; b += size_t(sigma) * NCP * 4;
        L.bs = b; b += size_t(NCP + 4) * 4;
        L.eps = b; b += size_t(2 * k) * ENUM_THREADS * 4;
        L.fcnt = b; b += size_t(n_fam) * NCP * 2;
        L.foff = b; b += size_t(n_fam) * NCP * 2;
        b = (b + 3) & ~size_t(3);
        L.gthr = b; b += size_t(NCP) * 4;
        b = (b + 15) & ~size_t(15);
        L.pool_score = b; b += size_t(pool_cap) * 4;
        L.pool_key = b; b += size_t(pool_cap) * 4;
        b = (b + 15) & ~size_t(15);
        L.st_keys = b; b += size_t(ENUM_WARPS) * STAGE_CAP * 8;
        L.st_scores = b; b += size_t(ENUM_WARPS) * STAGE_CAP * 4;
        L.st_win = b; if (debug) b += size_t(ENUM_WARPS) * STAGE_CAP * 4;
        L.sidx = b; b += size_t(sigma) * NCP;
        L.total = b + 16;
        return L;
    }

    /// smallest pool the raw tile fits in (it is staged over the pool's start before any list exists)
    inline uint32_t join_min_pool(int k, int sigma) { return uint32_t((size_t(ENUM_THREADS + k - 1) * sigma * 4 + 7) / 8 + 16); }

    /// insertion sort of one list by descending score (lists are short; the owner thread sorts its own list)
    __device__ __forceinline__ void sort_list_desc(float* ps, uint32_t* pk, int n)
    {
        for (int a = 1; a < n; ++a)
        {
            const float s = ps[a];
            const uint32_t kk = pk[a];
            int b = a;
            while (b > 0 && ps[b - 1] < s) { ps[b] = ps[b - 1]; pk[b] = pk[b - 1]; --b; }
            ps[b] = s; pk[b] = kk;
        }
    }

    /// The predicate itself for the subtree of h leaves that starts at window offset o, for window t of the
    /// tile: every leaf and every internal node of the subtree is checked against its own eps.  Used only for list
    /// elements whose score lies within DELTA of their node's threshold.
    __device__ __noinline__ bool verify_subtree(uint32_t key, int o, int h, int t, int bits, int sigma, int NCP,
                                                const float* sval, const uint8_t* sidx, const float* eps,
                                                const uint32_t* s_pos, const uint32_t* s_merge)
    {
        float val[2 * MAX_K];
        int merges_left = h - 1;                       // internal nodes of the subtree; they complete before any ancestor's merge
        const uint32_t mask = (1u << bits) - 1;
        for (int q = 0; q < h; ++q)
        {
            const int p = o + q, col = t + p;
            const uint32_t state = (key >> (bits * (h - 1 - q))) & mask;
            float v = 0.0f;
            bool found = false;
            for (int r = 0; r < sigma; ++r)
                if (sidx[r * NCP + col] == state) { v = sval[r * NCP + col]; found = true; break; }
            if (!found) return false;
            const uint32_t pi = s_pos[p];
            if (!(v > eps[(pi & 0xff) * ENUM_THREADS + t])) return false;
            val[pi & 0xff] = v;
            for (uint32_t m = (pi >> 16) & 0xff; m < (pi >> 24) && merges_left > 0; ++m, --merges_left)
            {
                const uint32_t mg = s_merge[m];
                const uint32_t left = mg >> 8, id = mg & 0xff;
                const float S = __fadd_rn(val[left], val[left + 1]);
                if (!(S > eps[id * ENUM_THREADS + t])) return false;
                val[id] = S;
            }
        }
        return true;
    }

    template<bool DEBUG>
    __global__ void __launch_bounds__(ENUM_THREADS)
    k2_join(const __grid_constant__ enum_args A, const __grid_constant__ enum_program P, const __grid_constant__ join_program J)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        __shared__ __align__(8) unsigned long long mbar;
        __shared__ uint32_t s_pos[MAX_K];
        __shared__ uint32_t s_merge[MAX_K];
        __shared__ uint32_t s_pool_used, s_defer;

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int k = P.k, sigma = int(A.sigma), bits = int(A.bits);
        const uint32_t tile = blockIdx.x;
        const uint32_t lslot = A.slot_base + tile / A.tiles_per_node;
        const uint32_t node = A.node_order[lslot];
        const uint32_t s0 = (tile % A.tiles_per_node) * ENUM_THREADS;
        const int nwin_tile = int(min(uint32_t(ENUM_THREADS), A.nwin - s0));
        const int ncols = nwin_tile + k - 1;
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;
        const join_layout L = make_join_layout(k, sigma, J.n_fam, A.pool_cap, DEBUG);

        float* sval = reinterpret_cast<float*>(smem_raw + L.sval);
        float* bs = reinterpret_cast<float*>(smem_raw + L.bs);
        float* eps = reinterpret_cast<float*>(smem_raw + L.eps);
        uint16_t* fcnt = reinterpret_cast<uint16_t*>(smem_raw + L.fcnt);
        uint16_t* foff = reinterpret_cast<uint16_t*>(smem_raw + L.foff);
        float* gthr = reinterpret_cast<float*>(smem_raw + L.gthr);
        float* pool_s = reinterpret_cast<float*>(smem_raw + L.pool_score);
        uint32_t* pool_k = reinterpret_cast<uint32_t*>(smem_raw + L.pool_key);
        unsigned long long* st_keys = reinterpret_cast<unsigned long long*>(smem_raw + L.st_keys) + warp * STAGE_CAP;
        float* st_scores = reinterpret_cast<float*>(smem_raw + L.st_scores) + warp * STAGE_CAP;
        uint32_t* st_win = DEBUG ? reinterpret_cast<uint32_t*>(smem_raw + L.st_win) + warp * STAGE_CAP : nullptr;
        uint8_t* sidx = smem_raw + L.sidx;
        float* raw = pool_s;                               // the staged tile is consumed before the first list is written

        if (tid < MAX_K)
        {
            s_pos[tid] = tid < k ? (uint32_t(P.leaf_eps[tid]) | uint32_t(P.merge_begin[tid]) << 16 | uint32_t(P.merge_begin[tid + 1]) << 24) : 0u;
            s_merge[tid] = uint32_t(P.merges[tid].id) | uint32_t(P.merges[tid].left) << 8;
        }
        if (tid == 0) { s_pool_used = 0; s_defer = 0; }
        stage_tile(A, node, s0, ncols, NCP, raw, bs, sval, sidx, &mbar);

        // ---- eps of every tree node for this thread's window (pk_compute.cpp:54-55; window.cpp:69-72) ---------------
        const bool has_window = tid < nwin_tile;
        float delta = 0.0f;
        if (has_window)
        {
            eps[tid] = A.log_threshold;
            for (int i = 0; i < P.n_eps_ops; ++i)
            {
                const auto op = P.eps_ops[i];
                const float r = __fsub_rn(bs[tid + op.hi], bs[tid + op.lo]);
                eps[op.dst * ENUM_THREADS + tid] = __fsub_rn(eps[op.src * ENUM_THREADS + tid], r);
            }
            // slack of the threshold chain: every level rounds once at the magnitude of B and once locally
            float bmax = 0.0f;
            for (int i = 0; i <= k; ++i) bmax = fmaxf(bmax, fabsf(bs[tid + i]));
            delta = 2.0f * float(k) * 1.2e-7f * bmax + 1e-4f;
        }
        __syncthreads();                                   // raw consumed, sval / eps complete


        // ---- G_h(c) for every family, shortest first -----------------------------------------------------------------
        const unsigned full = 0xffffffffu;
        for (int f = 0; f < J.n_fam; ++f)
        {
            const int h = J.fam[f].h, hl = J.fam[f].hl, hr = h - hl;
            const int fa = J.fam[f].left, fb = J.fam[f].right;
            const int nlist = ncols - h + 1;
            uint16_t* cnt_f = fcnt + f * NCP;
            uint16_t* off_f = foff + f * NCP;
            // pass 1: g_h(c) and the length of G_h(c)
            for (int c = tid; c < nlist; c += ENUM_THREADS)
            {
                float g = __int_as_float(0x7f800000);      // +inf: no window of this tile has a length-h node at column c
                for (int u = 0; u < J.fam[f].n_uses; ++u)
                {
                    const int t = c - int(J.fam[f].use_off[u]);
                    if (t >= 0 && t < nwin_tile) g = fminf(g, eps[int(J.fam[f].use_node[u]) * ENUM_THREADS + t]);
                }
                gthr[c] = g;
                uint32_t n = 0;
                if (h == 1)
                {
                    while (n < uint32_t(sigma) && sval[n * NCP + c] > g) ++n;            // as_column (pk_compute.cpp:14-26): strict >
                }
                else
                {
                    const int na = cnt_f[(fa - f) * NCP + c], nb = cnt_f[(fb - f) * NCP + c + hl];
                    const float* as = pool_s + off_f[(fa - f) * NCP + c];
                    const float* bsc = pool_s + off_f[(fb - f) * NCP + c + hl];
                    for (int i = 0; i < na && n <= A.pool_cap; ++i)
                    {
                        const float sa = as[i];
                        int l = 0;
                        while (l < nb && __fadd_rn(sa, bsc[l]) > g) ++l;                 // both lists descend: first failure ends the row
                        if (l == 0) break;                                               // ... and the first empty row ends the join
                        n += l;
                    }
                }
                cnt_f[c] = uint16_t(min(n, 0xffffu));
            }
            __syncthreads();
            // exclusive scan of the lengths -> pool offsets (warp 0; nlist <= NCP <= 160)
            if (warp == 0)
            {
                const int per = (nlist + 31) / 32;
                uint32_t sum = 0;
                for (int i = 0; i < per; ++i) { const int c = lane * per + i; if (c < nlist) sum += cnt_f[c]; }
                uint32_t inc = sum;
                for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(full, inc, o); if (lane >= o) inc += v; }
                uint32_t run = s_pool_used + inc - sum;
                const uint32_t total = __shfl_sync(full, inc, 31);
                const bool fits = s_pool_used + total <= A.pool_cap;
                if (fits)
                    for (int i = 0; i < per; ++i) { const int c = lane * per + i; if (c < nlist) { off_f[c] = uint16_t(run); run += cnt_f[c]; } }
                __syncwarp();
                if (lane == 0) { if (fits) s_pool_used += total; else s_defer = 1; }
            }
            __syncthreads();
            if (s_defer) break;
            // pass 2: fill and sort
            for (int c = tid; c < nlist; c += ENUM_THREADS)
            {
                const int n = cnt_f[c];
                if (n == 0) continue;
                float* os = pool_s + off_f[c];
                uint32_t* ok = pool_k + off_f[c];
                if (h == 1)
                {
                    for (int r = 0; r < n; ++r) { os[r] = sval[r * NCP + c]; ok[r] = sidx[r * NCP + c]; }
                }
                else
                {
                    const float g = gthr[c];
                    const int na = cnt_f[(fa - f) * NCP + c], nb = cnt_f[(fb - f) * NCP + c + hl];
                    const uint32_t oa = off_f[(fa - f) * NCP + c], ob = off_f[(fb - f) * NCP + c + hl];
                    int w = 0;
                    for (int i = 0; i < na; ++i)
                    {
                        const float sa = pool_s[oa + i];
                        const uint32_t ka = pool_k[oa + i] << (hr * bits);               // prefix = most significant digits (pk_compute.cpp:96-104)
                        int l = 0;
                        for (; l < nb; ++l)
                        {
                            const float S = __fadd_rn(sa, pool_s[ob + l]);
                            if (!(S > g)) break;
                            os[w] = S; ok[w] = ka | pool_k[ob + l]; ++w;
                        }
                        if (l == 0) break;
                    }
                    sort_list_desc(os, ok, n);
                }
            }
            __syncthreads();
        }
        if (s_defer)
        {
            // the lists of this tile do not fit: hand the whole tile to the depth-first kernel (nothing was emitted yet)
            if (tid == 0) A.deferred_tiles[atomicAdd(A.counter + 1, 1ull)] = tile;
            return;
        }


        // ---- per window: cut L_left(s), L_right(s) out of the two top lists -------------------------------------------
        const int hl0 = J.root_hl, hr0 = k - hl0;
        int nl = 0, nr = 0;
        uint32_t oa = 0, ob = 0;
        bool irregular = false;
        if (has_window)
        {
            const int fa = J.root_left, fb = J.root_right;
            const int na = fcnt[fa * NCP + tid], nb = fcnt[fb * NCP + tid + hl0];
            oa = foff[fa * NCP + tid]; ob = foff[fb * NCP + tid + hl0];
            const float eps_l = eps[1 * ENUM_THREADS + tid], eps_r = eps[2 * ENUM_THREADS + tid];   // nodes 1, 2 = the root's children
            while (nl < na)
            {
                const float sa = pool_s[oa + nl];
                if (!(sa > eps_l)) break;
                if (hl0 > 1 && !(sa > eps_l + delta) &&
                    !verify_subtree(pool_k[oa + nl], 0, hl0, tid, bits, sigma, NCP, sval, sidx, eps, s_pos, s_merge)) { irregular = true; break; }
                ++nl;
            }
            while (nr < nb && !irregular)
            {
                const float sb = pool_s[ob + nr];
                if (!(sb > eps_r)) break;
                if (hr0 > 1 && !(sb > eps_r + delta) &&
                    !verify_subtree(pool_k[ob + nr], hl0, hr0, tid, bits, sigma, NCP, sval, sidx, eps, s_pos, s_merge)) { irregular = true; break; }
                ++nr;
            }
        }
        // an element inside the DELTA band failed a deeper check: L(s) is not a prefix of the sorted list.  Rare enough to
        // give the whole tile to the depth-first kernel (nothing was emitted yet).
        if (__syncthreads_or(irregular ? 1 : 0))
        {
            if (tid == 0) A.deferred_tiles[atomicAdd(A.counter + 1, 1ull)] = tile;
            return;
        }

        // ---- root join (pk_compute.cpp:75-105) + warp-aggregated emission --------------------------------------------
        const float eps0 = A.log_threshold;
        const uint64_t group = DEBUG ? 0 : uint64_t(A.node_group[node] - A.group_sub);
        const int shift0 = hr0 * bits;
        bool active = has_window && nl > 0 && nr > 0;
        int i = 0, l = 0;
        float sa = active ? pool_s[oa] : 0.0f;
        uint64_t ka = active ? uint64_t(pool_k[oa]) << shift0 : 0;
        int wcount = 0;
        unsigned long long emitted = 0;
        while (__any_sync(full, active))
        {
            bool emit = false;
            uint64_t ekey = 0;
            float escore = 0.0f;
            if (active)
            {
                bool ok = l < nr;
                if (ok)
                {
                    escore = __fadd_rn(sa, pool_s[ob + l]);                               // pk_compute.cpp:90
                    ok = escore > eps0;                                                   // :91 `score <= eps` => break
                }
                if (ok)
                {
                    emit = true;
                    ekey = ka | uint64_t(pool_k[ob + l]);
                    ++l;
                }
                else if (l == 0) active = false;                                          // rows only get shorter (both lists descend)
                else
                {
                    ++i; l = 0;
                    if (i < nl) { sa = pool_s[oa + i]; ka = uint64_t(pool_k[oa + i]) << shift0; }
                    else active = false;
                }
            }
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
                    for (int q = lane; q < wcount; q += 32)
                    {
                        const unsigned long long o = base + q;
                        if (o < A.cap)
                        {
                            A.out_keys[o] = st_keys[q];
                            A.out_scores[o] = st_scores[q];
                            if (DEBUG) { A.out_node[o] = node; A.out_win[o] = st_win[q]; }
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
            for (int q = lane; q < wcount; q += 32)
            {
                const unsigned long long o = base + q;
                if (o < A.cap)
                {
                    A.out_keys[o] = st_keys[q];
                    A.out_scores[o] = st_scores[q];
                    if (DEBUG) { A.out_node[o] = node; A.out_win[o] = st_win[q]; }
                }
            }
        }
        if (A.node_counts && lane == 0 && emitted) atomicAdd(A.node_counts + lslot, emitted);
    }
}
