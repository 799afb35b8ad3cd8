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
            uint8_t top;                           // list storage: 1 = taken from the top end of the pool, 0 = from the bottom
            uint8_t reset;                         // 1 = every list at the bottom end is dead when this family is built: start over at 0
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
        // Two-ended pool.  Only the root's children (the one or two longest families) are needed by the root join, and
        // while THEY are built only their own children must still exist.  So: children of the root's children live at the
        // top end; every other family is built at the bottom end first and is dead by the time the first root child is
        // built, which therefore starts the bottom end over.  (A root child that is also a child of the other one -- k = 3,
        // 5 -- stays at the top.)
        auto is_root_child = [&](int f) { return f == J.root_left || f == J.root_right; };
        for (int f = 0; f < J.n_fam; ++f)
            if (is_root_child(f) && J.fam[f].h > 1) { J.fam[J.fam[f].left].top = 1; J.fam[J.fam[f].right].top = 1; }
        for (int f = 0; f < J.n_fam; ++f)
            if (is_root_child(f) && !J.fam[f].top) { J.fam[f].reset = 1; break; }
        return J;
    }


    // ----------------------------------------------------------------------------------------------------------
    // shared-memory layout of k2_join (NC = ENUM_THREADS + k - 1 local columns, NCP = NC rounded up to 4):
    //   pool  u64 [pool_cap]            list elements: fp32 score bits << 32 | key   (the raw TMA-staged tile aliases its start)
    //   sval  f32 [sigma][NCP]          column states' scores, descending
    //   bs    f32 [NCP+4]               B[n][s0 .. s0+NC]
    //   (eps  f32 [2k][ENUM_THREADS]    eps of every tree node for every window of the tile, index-major: lives in the POOL,
    //                                   behind the raw tile, only until gall / epsl / epsr have been derived from it)
    //   gall  f32 [n_fam][NCP]          g_h(c) of every family
    //   epsl  f32 [ENUM_THREADS]        eps of the root's left child per window;  epsr: of its right child
    //   wbase u32 [ENUM_THREADS+1]      exclusive scan of the windows' output counts
    //   rbase u32 [ENUM_THREADS+1]      exclusive scan of the windows' row counts (rows = left elements with a non-empty row)
    //   fcnt  u16 [n_fam][NCP]          length of G_h(c)
    //   foff  u16 [n_fam][NCP]          its start in the pool
    //   wnr   u16 [ENUM_THREADS]        |L_right(s)|
    //   queue u16 [NCP]                 columns whose list is built by a whole warp (long lists)
    //   perm  u16 [NCP]                 columns with a short list, by descending length (phase A2)
    //   sidx  u8  [sigma][NCP]
    // ----------------------------------------------------------------------------------------------------------
    struct join_layout
    {
        size_t pool, sval, bs, eps, gall, epsl, epsr, wbase, rbase, fcnt, foff, wnr, queue, perm, sidx, total;
    };

    __host__ __device__ inline join_layout make_join_layout(int k, int sigma, int n_fam, uint32_t pool_cap)
    {
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;
        join_layout L;
        size_t b = 0;
        L.pool = b; b += size_t(pool_cap) * 8;
        L.sval = b; b += size_t(sigma) * NCP * 4;
        L.bs = b; b += size_t(NCP + 4) * 4;
        L.eps = L.pool + ((size_t(NC) * sigma * 4 + 15) & ~size_t(15));      // inside the pool, behind the raw tile
        L.gall = b; b += size_t(n_fam) * NCP * 4;
        L.epsl = b; b += size_t(ENUM_THREADS) * 4;
        L.epsr = b; b += size_t(ENUM_THREADS) * 4;
        L.wbase = b; b += size_t(ENUM_THREADS + 1) * 4;
        L.rbase = b; b += size_t(ENUM_THREADS + 1) * 4;
        L.fcnt = b; b += size_t(n_fam) * NCP * 2;
        L.foff = b; b += size_t(n_fam) * NCP * 2;
        L.wnr = b; b += size_t(ENUM_THREADS) * 2;
        L.queue = b; b += size_t(NCP) * 2;
        L.perm = b; b += size_t(NCP) * 2;
        L.sidx = b; b += size_t(sigma) * NCP;
        L.total = ((b + 15) & ~size_t(15)) + 16;
        return L;
    }

    /// smallest pool: the raw tile and the transient eps table are staged in it before any list exists
    inline uint32_t join_min_pool(int k, int sigma)
    {
        return uint32_t((((size_t(ENUM_THREADS + k - 1) * sigma * 4 + 15) & ~size_t(15)) + size_t(2 * k) * ENUM_THREADS * 4 + 7) / 8 + 16);
    }

    // threads of a k2_join CTA: one per window of the tile (ENUM_THREADS) plus one warp, so that the tile's
    // ENUM_THREADS + k - 1 columns (k <= 33) are covered by ONE pass of every per-column loop (phase clocks r01m: the second
    // pass over the last k - 1 columns cost as much latency as the first)
    constexpr int JOIN_HIST_SLOTS = 256;     // partial top-byte histograms of the emission, one per SM id (mod 256)
    constexpr int JOIN_THREADS = ENUM_THREADS + 32;
    constexpr int JOIN_WARPS = JOIN_THREADS / 32;
#ifndef IPK_JOIN_GROUP
#define IPK_JOIN_GROUP 8
#endif
    constexpr int JOIN_GROUP = IPK_JOIN_GROUP;   // lanes that share one of the longer lists in phase A2
    constexpr int JOIN_GROUP_MIN = 20;       // ... lists of at least this many elements (and at most 8 * JOIN_GROUP, see small_list)
    // r01x/r01y sweeps on cfg3 (ms of enumeration): 4 lanes, lists 14..32: 48.7; 8 lanes, lists 20..64: 44.8; 8 lanes, 24..64: 45.6;
    // 8 lanes, 31..64: 57.9 (a single thread insertion-sorting 30 elements); cfg4 (amino acids) 157 -> 125
    constexpr int JOIN_SMALL_LIST = 24;      // lists up to this length are filled and insertion-sorted by their owner thread

    __device__ __forceinline__ unsigned long long el_pack(float s, uint32_t key) { return (unsigned long long)__float_as_uint(s) << 32 | key; }
    __device__ __forceinline__ float el_score(unsigned long long e) { return __uint_as_float(uint32_t(e >> 32)); }
    __device__ __forceinline__ uint32_t el_key(unsigned long long e) { return uint32_t(e); }

    /// insertion sort of one list by descending score (short lists; the owner thread sorts its own list)
    __device__ __forceinline__ void sort_list_desc(unsigned long long* p, int n)
    {
        for (int a = 1; a < n; ++a)
        {
            const unsigned long long e = p[a];
            const float s = el_score(e);
            int b = a;
            while (b > 0 && el_score(p[b - 1]) < s) { p[b] = p[b - 1]; --b; }
            p[b] = e;
        }
    }

    /// One warp sorts p[0..n) by descending score in shared memory: bitonic network in its mirrored form (every
    /// comparison moves the larger element to the lower index), so the virtual -inf padding above n never moves.
    __device__ __forceinline__ void warp_sort_desc(unsigned long long* p, int n, int lane)
    {
        int lg = 0;
        while ((1 << lg) < n) ++lg;                       // pw = 1 << lg >= n
        const int half = (1 << lg) >> 1;
        for (int ls = 1; ls <= lg; ++ls)                   // merge blocks of size 1 << ls
        {
            const int size = 1 << ls, lh = ls - 1;
            for (int t = lane; t < half; t += 32)
            {
                const int blk = t >> lh, j = t & ((1 << lh) - 1);
                const int lo = (blk << ls) + j, hi = (blk << ls) + size - 1 - j;
                if (hi < n)
                {
                    const unsigned long long a = p[lo], b = p[hi];
                    if (el_score(b) > el_score(a)) { p[lo] = b; p[hi] = a; }
                }
            }
            __syncwarp();
            for (int ld = lh - 1; ld >= 0; --ld)           // half-cleaners at distance 1 << ld
            {
                for (int t = lane; t < half; t += 32)
                {
                    const int lo = ((t >> ld) << (ld + 1)) + (t & ((1 << ld) - 1)), hi = lo + (1 << ld);
                    if (hi < n)
                    {
                        const unsigned long long a = p[lo], b = p[hi];
                        if (el_score(b) > el_score(a)) { p[lo] = b; p[hi] = a; }
                    }
                }
                __syncwarp();
            }
        }
    }

    /// number of leading elements b of the descending list q[0..n) with fl(sa + S(b)) > thr  (fp32 addition is monotone,
    /// so they form a prefix)
    __device__ __forceinline__ int row_length(const unsigned long long* q, int n, float sa, float thr)
    {
        int lo = 0, hi = n;                     // invariant: elements < lo pass, elements >= hi fail
        while (lo < hi)
        {
            const int mid = (lo + hi) >> 1;
            if (__fadd_rn(sa, el_score(q[mid])) > thr) lo = mid + 1; else hi = mid;
        }
        return lo;
    }

    __device__ __forceinline__ uint32_t warp_incl_scan_u32(uint32_t v, int lane)
    {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        return v;
    }

    /// lane that owns output o of a batch: the last lane whose exclusive offset is <= o (incl = inclusive scan of the
    /// per-lane row lengths; rows of length 0 are skipped automatically)
    __device__ __forceinline__ int owner_lane(uint32_t incl, uint32_t o)
    {
        int lo = 0;                              // first lane with incl > o
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1)
        {
            const uint32_t v = __shfl_sync(0xffffffffu, incl, lo + step - 1);
            if (v <= o) lo += step;
        }
        return lo;
    }

    /// The predicate itself for the subtree of h leaves that starts at window offset o, for window t of the
    /// tile: every leaf and every internal node of the subtree is checked against its own eps.  Used only for list
    /// elements whose score lies within DELTA of their node's threshold.
    __device__ __noinline__ bool verify_subtree(uint32_t key, int o, int h, int t, int bits, int sigma, int NCP,
                                                const float* sval, const uint8_t* sidx, const float* bs, float log_threshold,
                                                const enum_program& P, const uint32_t* s_pos, const uint32_t* s_merge)
    {
        float val[2 * MAX_K], eps[2 * MAX_K];
        eps[0] = log_threshold;                        // the window's threshold chain again (pk_compute.cpp:54-55); this path is rare
        for (int i = 0; i < P.n_eps_ops; ++i)
        {
            const auto op = P.eps_ops[i];
            eps[op.dst] = __fsub_rn(eps[op.src], __fsub_rn(bs[t + op.hi], bs[t + op.lo]));
        }
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
            if (!(v > eps[pi & 0xff])) return false;
            val[pi & 0xff] = v;
            for (uint32_t m = (pi >> 16) & 0xff; m < (pi >> 24) && merges_left > 0; ++m, --merges_left)
            {
                const uint32_t mg = s_merge[m];
                const uint32_t left = mg >> 8, id = mg & 0xff;
                const float S = __fadd_rn(val[left], val[left + 1]);
                if (!(S > eps[id])) return false;
                val[id] = S;
            }
        }
        return true;
    }

    /// n elements from the bottom (grows up) or the top (grows down) end of the pool; false when the two ends would meet.
    /// All allocations of one family go to the same end, so the other cursor is stable while they race.
    __device__ __forceinline__ bool pool_alloc(int* lo, int* hi, bool top, int n, uint32_t& off)
    {
        if (top)
        {
            const int o = atomicSub(hi, n) - n;
            off = uint32_t(o);
            return o >= *reinterpret_cast<volatile int*>(lo);
        }
        const int o = atomicAdd(lo, n);
        off = uint32_t(o);
        return o + n <= *reinterpret_cast<volatile int*>(hi);
    }

    /// first index of the descending list p[0..n) whose score is NOT > thr  (= number of elements above thr)
    __device__ __forceinline__ int cut_above(const unsigned long long* p, int n, float thr)
    {
        int lo = 0, hi = n;
        while (lo < hi)
        {
            const int mid = (lo + hi) >> 1;
            if (el_score(p[mid]) > thr) lo = mid + 1; else hi = mid;
        }
        return lo;
    }

    /// G lanes (a power of two, aligned in the warp) build and sort ONE list of n <= 8*G elements: rows by binary search,
    /// row offsets by a scan inside the group, then a rank sort -- every lane counts, for each of its elements, how many
    /// elements precede it (descending score, ties by fill position), all from independent loads, and the elements move to
    /// their ranks in place.  Replaces the owner thread's serial fill + insertion sort for the longer lists: the critical
    /// path of a 24-element list drops from ~150 dependent moves to ~6 rank loops of 24 independent compares.
    template<int G>
    __device__ __forceinline__ void build_list_group(const unsigned long long* la, int na, const unsigned long long* lb, int nb, float g,
                                                     unsigned long long* os, int n, int shift, int gl, unsigned gmask)
    {
        int carry = 0;
        for (int i0 = 0; i0 < na; i0 += G)
        {
            const int i = i0 + gl;
            float sa = 0.0f;
            uint32_t ka = 0;
            int len = 0;
            if (i < na)
            {
                const unsigned long long ea = la[i];
                sa = el_score(ea);
                ka = el_key(ea) << shift;
                len = row_length(lb, nb, sa, g);
            }
            int inc = len;
#pragma unroll
            for (int o = 1; o < G; o <<= 1)
            {
                const int t = __shfl_up_sync(gmask, inc, o, G);
                if (gl >= o) inc += t;
            }
            const int tot = __shfl_sync(gmask, inc, G - 1, G);
            unsigned long long* w = os + carry + inc - len;
            for (int j = 0; j < len; ++j)
            {
                const unsigned long long eb = lb[j];
                w[j] = el_pack(__fadd_rn(sa, el_score(eb)), ka | el_key(eb));
            }
            carry += tot;
            if (tot == 0) break;                           // rows only get shorter
        }
        __syncwarp(gmask);
        // rank sort: every element of the list is loaded once and compared against the lane's (up to 8) own elements
        unsigned long long ev[8];
        float v[8];
        int rk[8];
#pragma unroll
        for (int q = 0; q < 8; ++q)
        {
            const int e = gl + q * G;
            ev[q] = e < n ? os[e] : 0ull;
            v[q] = e < n ? el_score(ev[q]) : __int_as_float(0x7f800000);
            rk[q] = 0;
        }
#pragma unroll 4
        for (int j = 0; j < n; ++j)
        {
            const float sj = el_score(os[j]);
#pragma unroll
            for (int q = 0; q < 8; ++q) rk[q] += int((sj > v[q]) | ((sj == v[q]) & (j < gl + q * G)));
        }
        __syncwarp(gmask);
#pragma unroll
        for (int q = 0; q < 8; ++q)
            if (gl + q * G < n) os[rk[q]] = ev[q];
    }

    constexpr unsigned long long JOIN_ST_LOCAL = 1ull << 62, JOIN_ST_INCL = 2ull << 62, JOIN_ST_MASK = (1ull << 62) - 1;

    /// records emitted by all tiles before `tile` (one warp; 32 predecessors per step; every lane gets the sum).  Tiles are
    /// block indices of one launch, dispatched in order, so every predecessor is running or done.
    __device__ __forceinline__ unsigned long long join_look_back(const volatile unsigned long long* st, uint32_t tile, int lane)
    {
        unsigned long long sum = 0;
        long long base = (long long)tile - 1;
        while (base >= 0)
        {
            const long long t = base - lane;
            unsigned long long w = JOIN_ST_INCL;                       // lanes that fall off the front see an inclusive zero
            if (t >= 0) do { w = st[t]; } while ((w >> 62) == 0);
            const unsigned incl = __ballot_sync(0xffffffffu, (w >> 62) == 2);
            const int first = incl ? __ffs(incl) - 1 : 32;
            unsigned long long v = lane <= first ? (w & JOIN_ST_MASK) : 0ull;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            sum += v;
            if (incl) break;
            base -= 32;
        }
        return sum;
    }

    template<bool DEBUG>
    __global__ void __launch_bounds__(JOIN_THREADS)
    k2_join(const __grid_constant__ enum_args A, const __grid_constant__ enum_program P, const __grid_constant__ join_program J)
    {
        extern __shared__ __align__(128) unsigned char smem_raw[];
        __shared__ __align__(8) unsigned long long mbar;
        __shared__ unsigned long long s_base;
        __shared__ uint32_t s_pos[MAX_K];
        __shared__ uint32_t s_merge[MAX_K];
        __shared__ uint32_t s_defer, s_qn[2], s_bh[32], s_nwork, s_nheavy;
        __shared__ int s_lo, s_hi;                           // pool cursors: bottom end grows up, top end grows down
        __shared__ uint32_t s_wsum[2][JOIN_WARPS];
        __shared__ uint32_t s_hist[256];                     // the tile's records by the top byte of their k-mer (enum_args::top_hist)

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int k = P.k, sigma = int(A.sigma), bits = int(A.bits);
        long long t_phase = A.phase_clocks ? clock64() : 0;
#define IPK_PHASE(i_) do { if (A.phase_clocks && tid == 0) { const long long t_ = clock64(); atomicAdd(A.phase_clocks + (i_), (unsigned long long)(t_ - t_phase)); \
                                                            if ((i_) == 0) atomicAdd(A.phase_clocks + 31, 1ull); t_phase = t_; } } while (0)
        tile_geom G;
        if (!tile_geometry(A, G)) return;
        const uint32_t lslot = G.lslot, node = G.node, s0 = G.s0, tile = G.item;
        const int nwin_tile = G.nwin;
        const int ncols = nwin_tile + k - 1;
        const int NC = ENUM_THREADS + k - 1, NCP = (NC + 3) & ~3;
        const join_layout L = make_join_layout(k, sigma, J.n_fam, A.pool_cap);

        unsigned long long* pool = reinterpret_cast<unsigned long long*>(smem_raw + L.pool);
        float* sval = reinterpret_cast<float*>(smem_raw + L.sval);
        float* bs = reinterpret_cast<float*>(smem_raw + L.bs);
        float* eps = reinterpret_cast<float*>(smem_raw + L.eps);
        float* gall = reinterpret_cast<float*>(smem_raw + L.gall);
        float* epsl = reinterpret_cast<float*>(smem_raw + L.epsl);
        float* epsr = reinterpret_cast<float*>(smem_raw + L.epsr);
        uint32_t* wbase = reinterpret_cast<uint32_t*>(smem_raw + L.wbase);
        uint32_t* rbase = reinterpret_cast<uint32_t*>(smem_raw + L.rbase);
        uint16_t* fcnt = reinterpret_cast<uint16_t*>(smem_raw + L.fcnt);
        uint16_t* foff = reinterpret_cast<uint16_t*>(smem_raw + L.foff);
        uint16_t* wnr = reinterpret_cast<uint16_t*>(smem_raw + L.wnr);
        uint16_t* queue = reinterpret_cast<uint16_t*>(smem_raw + L.queue);
        uint16_t* perm = reinterpret_cast<uint16_t*>(smem_raw + L.perm);
        uint8_t* sidx = smem_raw + L.sidx;
        float* raw = reinterpret_cast<float*>(pool);         // the staged tile is consumed before the first list is written

        if (tid < MAX_K)
        {
            s_pos[tid] = tid < k ? (uint32_t(P.leaf_eps[tid]) | uint32_t(P.merge_begin[tid]) << 16 | uint32_t(P.merge_begin[tid + 1]) << 24) : 0u;
            s_merge[tid] = uint32_t(P.merges[tid].id) | uint32_t(P.merges[tid].left) << 8;
        }
        if (tid == 0) { s_lo = 0; s_hi = int(A.pool_cap); s_defer = 0; s_qn[0] = 0; s_qn[1] = 0; }
        if (!DEBUG && A.top_hist)
            for (int i = tid; i < 256; i += JOIN_THREADS) s_hist[i] = 0;
        stage_tile(A, node, s0, ncols, NCP, raw, bs, sval, sidx, &mbar, JOIN_THREADS);

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
        // everything the rest of the kernel needs from the eps table: g_h(c) of every family and the root children's eps
        for (int c = tid; c < ncols; c += JOIN_THREADS)
            for (int f = 0; f < J.n_fam; ++f)
            {
                float g = __int_as_float(0x7f800000);      // +inf: no window of this tile has a length-h node at column c
                for (int u = 0; u < J.fam[f].n_uses; ++u)
                {
                    const int t = c - int(J.fam[f].use_off[u]);
                    if (t >= 0 && t < nwin_tile) g = fminf(g, eps[int(J.fam[f].use_node[u]) * ENUM_THREADS + t]);
                }
                gall[f * NCP + c] = g;
            }
        if (has_window) { epsl[tid] = eps[1 * ENUM_THREADS + tid]; epsr[tid] = eps[2 * ENUM_THREADS + tid]; }   // nodes 1, 2 = the root's children
        __syncthreads();                                   // the pool may now be overwritten
        IPK_PHASE(0);

        // ---- G_h(c) for every family, shortest first -----------------------------------------------------------------
        const unsigned full = 0xffffffffu;
        for (int f = 0; f < J.n_fam; ++f)
        {
            const int h = J.fam[f].h, hl = J.fam[f].hl, hr = h - hl;
            const int fa = J.fam[f].left, fb = J.fam[f].right;
            const int nlist = ncols - h + 1;
            uint16_t* cnt_f = fcnt + f * NCP;
            uint16_t* off_f = foff + f * NCP;
            const bool top = J.fam[f].top != 0;
            const float* gthr = gall + f * NCP;
            if (J.fam[f].reset)
            {
                if (tid == 0) s_lo = 0;                        // every list at the bottom end is dead by now (make_join_program)
                __syncthreads();
            }
            if (tid == 0) s_qn[(f + 1) & 1] = 0;              // the other family's queue counter: last read before the previous barrier
            // phase A1, one thread per column: threshold and exact length (two-pointer walk over the staircase).  Leaf lists
            // are filled on the spot; long lists are queued for phase B; the others are ranked by length for phase A2.
            if (tid < 32) s_bh[tid] = 0;
            __syncthreads();
            for (int c = tid; c < nlist; c += JOIN_THREADS)
            {
                const float g = gthr[c];
                uint32_t n = 0, off = 0;
                if (h == 1)
                {
                    while (n < uint32_t(sigma) && sval[n * NCP + c] > g) ++n;            // as_column (pk_compute.cpp:14-26): strict >
                    if (n)
                    {
                        if (!pool_alloc(&s_lo, &s_hi, top, int(n), off)) s_defer = 1;
                        else for (uint32_t r = 0; r < n; ++r) pool[off + r] = el_pack(sval[r * NCP + c], sidx[r * NCP + c]);
                    }
                }
                else
                {
                    const int na = cnt_f[(fa - f) * NCP + c], nb = cnt_f[(fb - f) * NCP + c + hl];
                    const unsigned long long* la = pool + off_f[(fa - f) * NCP + c];
                    const unsigned long long* lb = pool + off_f[(fb - f) * NCP + c + hl];
                    int l = nb;
                    for (int i = 0; i < na && l > 0; ++i)                                // both lists descend: rows only get shorter
                    {
                        const float sa = el_score(la[i]);
                        while (l > 0 && !(__fadd_rn(sa, el_score(lb[l - 1])) > g)) --l;
                        n += uint32_t(l);
                    }
                    if (n > A.pool_cap) { s_defer = 1; n = 0; }
                    else if (n > A.small_list) queue[atomicAdd(&s_qn[f & 1], 1u)] = uint16_t(c);
                    else if (n) atomicAdd(&s_bh[31 - min(n, 31u)], 1u);                   // longest lists first
                }
                cnt_f[c] = uint16_t(n);
                off_f[c] = uint16_t(off);
            }
            __syncthreads();
            IPK_PHASE(1 + 3 * min(f, 3));
            if (h > 1)
            {
                // phase A2: the short lists are filled and sorted by one thread each, IN ORDER OF DESCENDING LENGTH, so that the
                // lanes of a warp run loops of similar trip counts (r01e ncu: 3.5-10 of 32 lanes active in these loops when
                // thread c simply took column c) and the light warps retire early.
                if (warp == 0)
                {
                    const uint32_t v = s_bh[lane];
                    const uint32_t inc = warp_incl_scan_u32(v, lane);
                    s_bh[lane] = inc - v;
                    if (lane == 31) s_nwork = inc;
                    if (lane == 31 - (int(A.group_min) - 1)) s_nheavy = inc - v;        // lists of >= JOIN_GROUP_MIN elements come first
                }
                __syncthreads();
                for (int c = tid; c < nlist; c += JOIN_THREADS)
                {
                    const uint32_t n = cnt_f[c];
                    if (n && n <= A.small_list) perm[atomicAdd(&s_bh[31 - min(n, 31u)], 1u)] = uint16_t(c);
                }
                __syncthreads();
                const int nwork = int(s_nwork), nheavy = int(s_nheavy);
                // the longer lists: JOIN_GROUP lanes each
                for (int w0 = warp * (32 / JOIN_GROUP); w0 < nheavy; w0 += JOIN_THREADS / JOIN_GROUP)
                {
                    const int w = w0 + lane / JOIN_GROUP, gl = lane & (JOIN_GROUP - 1);
                    const unsigned gmask = ((1u << JOIN_GROUP) - 1) << (lane & ~(JOIN_GROUP - 1));
                    if (w >= nheavy) continue;                                    // uniform within a group
                    const int c = perm[w];
                    const int n = cnt_f[c];
                    uint32_t off = 0;
                    int ok = 0;
                    if (gl == 0) ok = pool_alloc(&s_lo, &s_hi, top, n, off) ? 1 : 0;
                    off = __shfl_sync(gmask, off, 0, JOIN_GROUP);
                    ok = __shfl_sync(gmask, ok, 0, JOIN_GROUP);
                    if (!ok) { if (gl == 0) s_defer = 1; continue; }
                    build_list_group<JOIN_GROUP>(pool + off_f[(fa - f) * NCP + c], cnt_f[(fa - f) * NCP + c],
                                                 pool + off_f[(fb - f) * NCP + c + hl], cnt_f[(fb - f) * NCP + c + hl], gthr[c],
                                                 pool + off, n, hr * bits, gl, gmask);
                    if (gl == 0) off_f[c] = uint16_t(off);
                }
                // the short ones: one thread each
                for (int w = nheavy + tid; w < nwork; w += JOIN_THREADS)
                {
                    const int c = perm[w];
                    const uint32_t n = cnt_f[c];
                    const float g = gthr[c];
                    const int na = cnt_f[(fa - f) * NCP + c], nb = cnt_f[(fb - f) * NCP + c + hl];
                    const unsigned long long* la = pool + off_f[(fa - f) * NCP + c];
                    const unsigned long long* lb = pool + off_f[(fb - f) * NCP + c + hl];
                    uint32_t off = 0;
                    if (!pool_alloc(&s_lo, &s_hi, top, int(n), off)) { s_defer = 1; continue; }
                    unsigned long long* os = pool + off;
                    int wr = 0, l = nb;
                    for (int i = 0; i < na && l > 0; ++i)
                    {
                        const unsigned long long ea = la[i];
                        const float sa = el_score(ea);
                        const uint32_t ka = el_key(ea) << (hr * bits);           // prefix = most significant digits (pk_compute.cpp:96-104)
                        while (l > 0 && !(__fadd_rn(sa, el_score(lb[l - 1])) > g)) --l;
                        for (int j = 0; j < l; ++j)
                        {
                            const unsigned long long eb = lb[j];
                            os[wr++] = el_pack(__fadd_rn(sa, el_score(eb)), ka | el_key(eb));
                        }
                    }
                    sort_list_desc(os, int(n));
                    off_f[c] = uint16_t(off);
                }
                __syncthreads();
            }
            IPK_PHASE(2 + 3 * min(f, 3));
            // phase B, one warp per long list: row lengths by binary search, staircase expanded 32 elements at a time,
            // bitonic sort in place
            const int qn = int(s_qn[f & 1]);
            for (int q = warp; q < qn; q += JOIN_WARPS)
            {
                if (__shfl_sync(full, *reinterpret_cast<volatile uint32_t*>(&s_defer), 0)) break;
                const int c = queue[q];
                const int n = cnt_f[c];
                const int na = cnt_f[(fa - f) * NCP + c], nb = cnt_f[(fb - f) * NCP + c + hl];
                const unsigned long long* la = pool + off_f[(fa - f) * NCP + c];
                const unsigned long long* lb = pool + off_f[(fb - f) * NCP + c + hl];
                const float g = gthr[c];
                uint32_t off = 0;
                int ok = 0;
                if (lane == 0) ok = pool_alloc(&s_lo, &s_hi, top, n, off) ? 1 : 0;
                off = __shfl_sync(full, off, 0);
                ok = __shfl_sync(full, ok, 0);
                if (!ok) { if (lane == 0) s_defer = 1; break; }
                unsigned long long* os = pool + off;
                uint32_t carry = 0;
                for (int i0 = 0; i0 < na; i0 += 32)
                {
                    const int i = i0 + lane;
                    float sa = 0.0f;
                    uint32_t ka = 0, len = 0;
                    if (i < na)
                    {
                        const unsigned long long ea = la[i];
                        sa = el_score(ea);
                        ka = el_key(ea) << (hr * bits);
                        len = uint32_t(row_length(lb, nb, sa, g));
                    }
                    const uint32_t incl = warp_incl_scan_u32(len, lane);
                    const uint32_t total = __shfl_sync(full, incl, 31);
                    if (total == 0) break;
                    const uint32_t excl = incl - len;
                    for (uint32_t o0 = 0; o0 < total; o0 += 32)
                    {
                        const uint32_t o = o0 + lane;
                        const bool act = o < total;
                        const int r = owner_lane(incl, act ? o : 0u);
                        const float sa_r = __shfl_sync(full, sa, r);
                        const uint32_t ka_r = __shfl_sync(full, ka, r);
                        const uint32_t ex_r = __shfl_sync(full, excl, r);
                        if (act)
                        {
                            const unsigned long long eb = lb[o - ex_r];
                            os[carry + o] = el_pack(__fadd_rn(sa_r, el_score(eb)), ka_r | el_key(eb));
                        }
                    }
                    carry += total;
                }
                __syncwarp();
                warp_sort_desc(os, n, lane);
                if (lane == 0) off_f[c] = uint16_t(off);
            }
            __syncthreads();
            IPK_PHASE(3 + 3 * min(f, 3));
            if (s_defer) break;
        }
        if (s_defer)
        {
            // the lists of this tile do not fit: hand the whole tile to the next level (nothing was emitted yet)
            if (tid == 0)
            {
                A.deferred_tiles[atomicAdd(A.counter + 1, 1ull)] = tile;
                if (A.tile_status) *reinterpret_cast<volatile unsigned long long*>(A.tile_status + tile) = JOIN_ST_LOCAL;     // no records from here
            }
            return;
        }

        // ---- per window: cut L_left(s), L_right(s) out of the two top lists, count the root join ----------------------
        const int hl0 = J.root_hl, hr0 = k - hl0;
        const int fa0 = J.root_left, fb0 = J.root_right;
        const float eps0 = A.log_threshold;
        uint32_t my_out = 0, my_rows = 0, la_off = 0, lb_off = 0;
        int l0 = 0;
        bool irregular = false;
        if (has_window)
        {
            const int na = fcnt[fa0 * NCP + tid], nb = fcnt[fb0 * NCP + tid + hl0];
            const unsigned long long* la = pool + foff[fa0 * NCP + tid];
            const unsigned long long* lb = pool + foff[fb0 * NCP + tid + hl0];
            const float eps_l = epsl[tid], eps_r = epsr[tid];
            // S(a) > eps_left(s) is exact; the deeper checks of `a` are implied unless S(a) lies within delta of the
            // threshold -- those few elements (the tail of the cut) are put through the predicate itself
            const int nl = cut_above(la, na, eps_l);
            if (hl0 > 1)
                for (int q = nl - 1; q >= 0 && !(el_score(la[q]) > eps_l + delta); --q)
                    if (!verify_subtree(el_key(la[q]), 0, hl0, tid, bits, sigma, NCP, sval, sidx, bs, A.log_threshold, P, s_pos, s_merge)) { irregular = true; break; }
            int nr = cut_above(lb, nb, eps_r);
            if (hr0 > 1 && !irregular)
                for (int q = nr - 1; q >= 0 && !(el_score(lb[q]) > eps_r + delta); --q)
                    if (!verify_subtree(el_key(lb[q]), hl0, hr0, tid, bits, sigma, NCP, sval, sidx, bs, A.log_threshold, P, s_pos, s_merge)) { irregular = true; break; }
            // root join (pk_compute.cpp:75-105): row i = the b with fl(S(a_i) + S(b)) > eps, a prefix that only shrinks with i.
            // Here only the number of non-empty rows (the a_i whose sum with the best b passes) and the first row's length;
            // the row lengths are summed below, flat over the warp's rows.
            if (nl > 0 && nr > 0)
            {
                const float sb0 = el_score(lb[0]);
                int lo = 0, hi = nl;                       // rows [0, lo) are non-empty
                while (lo < hi)
                {
                    const int mid = (lo + hi) >> 1;
                    if (__fadd_rn(el_score(la[mid]), sb0) > eps0) lo = mid + 1; else hi = mid;
                }
                my_rows = uint32_t(lo);
                if (my_rows) nr = row_length(lb, nr, el_score(la[0]), eps0);      // no row is longer than the first
            }
            wnr[tid] = uint16_t(nr);
            la_off = uint32_t(la - pool);
            lb_off = uint32_t(lb - pool);
            l0 = my_rows ? nr : 0;
        }
        // sum of the row lengths per window: the warp's rows are spread over its lanes (a window with many rows no longer
        // holds its warp up: phase clocks r01m, 16k of 142k cycles per tile)
        if (warp < ENUM_WARPS)
        {
            wbase[tid] = 0;
            const uint32_t incl = warp_incl_scan_u32(my_rows, lane);
            const uint32_t total_rows = __shfl_sync(full, incl, 31);
            const uint32_t excl = incl - my_rows;
            __syncwarp();
            for (uint32_t r0 = 0; r0 < total_rows; r0 += 32)
            {
                const uint32_t rho = r0 + lane;
                const bool act = rho < total_rows;
                const int s = owner_lane(incl, act ? rho : 0u);
                const uint32_t la_s = __shfl_sync(full, la_off, s), lb_s = __shfl_sync(full, lb_off, s);
                const int l0_s = __shfl_sync(full, l0, s);
                const uint32_t ex_s = __shfl_sync(full, excl, s);
                if (act)
                {
                    const unsigned long long ea = pool[la_s + (rho - ex_s)];
                    const uint32_t len = uint32_t(row_length(pool + lb_s, l0_s, el_score(ea), eps0));
                    atomicAdd(&wbase[(tid & ~31) + s], len);
                    // the records of this row share the top byte of their k-mer: it lies in the left element's key
                    if (!DEBUG && A.top_hist && len) atomicAdd(&s_hist[el_key(ea) >> A.top_shift], len);
                }
            }
            __syncwarp();
            my_out = wbase[tid];
            if (my_out > (1u << 24)) irregular = true;     // keeps the tile total inside 32 bits; the depth-first kernel has no such limit
        }
        // an element inside the DELTA band failed a deeper check: L(s) is not a prefix of the sorted list.  Rare enough to
        // give the whole tile to the depth-first kernel (nothing was emitted yet).
        if (__syncthreads_or(irregular ? 1 : 0))
        {
            if (tid == 0)
            {
                A.deferred_tiles[atomicAdd(A.counter + 1, 1ull)] = tile;
                if (A.tile_status) *reinterpret_cast<volatile unsigned long long*>(A.tile_status + tile) = JOIN_ST_LOCAL;
            }
            return;
        }

        if (!DEBUG && A.top_hist)
        {
            // the tile will emit what it counted: one RED per non-empty bin into the partial histogram of this SM (the slots keep
            // the contention per address low); the first pass of the merge sort then needs no pass over the keys for it
            uint32_t smid;
            asm("mov.u32 %0, %%smid;" : "=r"(smid));
            uint32_t* const g = A.top_hist + size_t(smid & (JOIN_HIST_SLOTS - 1)) * 256;
            for (int i = tid; i < 256; i += JOIN_THREADS)
                if (const uint32_t v = s_hist[i]) atomicAdd(g + i, v);
        }
        IPK_PHASE(13);
        // ---- exclusive scans over the tile's windows: output offsets and row offsets ----------------------------------
        {
            const uint32_t io = warp_incl_scan_u32(my_out, lane), ir = warp_incl_scan_u32(my_rows, lane);
            if (lane == 31) { s_wsum[0][warp] = io; s_wsum[1][warp] = ir; }
            __syncthreads();
            uint32_t bo = 0, br = 0;
            for (int w = 0; w < warp; ++w) { bo += s_wsum[0][w]; br += s_wsum[1][w]; }
            if (tid < ENUM_THREADS)
            {
                wbase[tid] = bo + io - my_out;
                rbase[tid] = br + ir - my_rows;
            }
            if (tid == ENUM_THREADS - 1)
            {
                const uint32_t total = bo + io;
                wbase[ENUM_THREADS] = total;
                rbase[ENUM_THREADS] = br + ir;
                // ONE global atomic per tile reserves the tile's exact output range ...
                if (!A.tile_status)
                {
                    s_base = total ? atomicAdd(A.counter, (unsigned long long)total) : 0ull;
                    if (A.runs && total)
                    {
                        const unsigned long long r = atomicAdd(A.run_counter, 1ull);
                        if (r < A.run_cap) A.runs[r] = emit_run{ s_base, total, A.node_group[node] - A.group_sub };
                    }
                }
                else
                {
                    // ... or, in ordered mode, only counts; the range follows from the look-back below
                    *reinterpret_cast<volatile unsigned long long*>(A.tile_status + tile) = (tile == 0 ? JOIN_ST_INCL : JOIN_ST_LOCAL) | total;
                    if (total) atomicAdd(A.counter, (unsigned long long)total);
                }
                if (A.node_counts && total) atomicAdd(A.node_counts + lslot, (unsigned long long)total);
            }
            __syncthreads();
            if (A.tile_status)
            {
                if (warp == 0)
                {
                    unsigned long long before = 0;
                    if (tile > 0)
                    {
                        before = join_look_back(A.tile_status, tile, lane);
                        if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(A.tile_status + tile) = JOIN_ST_INCL | (before + wbase[ENUM_THREADS]);
                    }
                    if (lane == 0) s_base = before;
                }
                __syncthreads();
            }
            IPK_PHASE(14);
        }

        // ---- emission: every warp expands the rows of its 32 windows, 32 records per step, straight to global memory ---
        {
            const int ws = warp * 32, we = min(ws + 32, nwin_tile);
            if (ws >= we) return;
            const uint32_t rbeg = rbase[ws], rend = rbase[we];
            unsigned long long pos = s_base + wbase[ws];
            const uint64_t group = DEBUG ? 0 : uint64_t(A.node_group[node] - A.group_sub);
            const int shift0 = hr0 * bits;
            for (uint32_t r0 = rbeg; r0 < rend; r0 += 32)
            {
                const uint32_t rho = r0 + lane;
                float sa = 0.0f;
                uint32_t ka = 0, len = 0, lbo = 0, win = 0;
                if (rho < rend)
                {
                    int s = ws;                            // last window of [ws, we) whose first row is <= rho
#pragma unroll
                    for (int step = 16; step >= 1; step >>= 1)
                        if (s + step < we && rbase[s + step] <= rho) s += step;
                    const unsigned long long ea = pool[uint32_t(foff[fa0 * NCP + s]) + (rho - rbase[s])];
                    sa = el_score(ea);
                    ka = el_key(ea);
                    lbo = foff[fb0 * NCP + s + hl0];
                    len = uint32_t(row_length(pool + lbo, wnr[s], sa, eps0));
                    win = uint32_t(s);
                }
                const uint32_t incl = warp_incl_scan_u32(len, lane);
                const uint32_t total = __shfl_sync(full, incl, 31);
                const uint32_t excl = incl - len;
#pragma unroll 2
                for (uint32_t o0 = 0; o0 < total; o0 += 32)
                {
                    const uint32_t o = o0 + lane;
                    const bool act = o < total;
                    // every row of the batch is non-empty (rbase counts rows with a non-empty join only), so the row of output o is
                    // (rows that start at or before o0) - 1 + (row starts in (o0, o]): one ballot and one warp OR-reduction
                    const int first = __popc(__ballot_sync(full, len > 0 && excl <= o0)) - 1;
                    const uint32_t starts = __reduce_or_sync(full, (len > 0 && excl > o0 && excl < o0 + 32) ? 1u << (excl - o0) : 0u);
                    const int r = first + __popc(starts & (0xffffffffu >> (31 - lane)));
                    const float sa_r = __shfl_sync(full, sa, r);
                    const uint32_t ka_r = __shfl_sync(full, ka, r);
                    const uint32_t ex_r = __shfl_sync(full, excl, r);
                    const uint32_t lbo_r = __shfl_sync(full, lbo, r);
                    const uint32_t win_r = DEBUG ? __shfl_sync(full, win, r) : 0u;
                    const unsigned long long gpos = pos + o;
                    if (act && gpos < A.cap)
                    {
                        const unsigned long long eb = pool[lbo_r + (o - ex_r)];
                        const float score = __fadd_rn(sa_r, el_score(eb));                          // pk_compute.cpp:90
                        const uint64_t key = (uint64_t(ka_r) << shift0) | uint64_t(el_key(eb));     // :96-104
                        A.out_keys[gpos] = DEBUG ? key : ((key << A.group_bits) | group);
                        A.out_scores[gpos] = score;
                        if (DEBUG) { A.out_node[gpos] = node; A.out_win[gpos] = s0 + win_r; }
                    }
                }
                pos += total;
            }
            IPK_PHASE(15);
        }
#undef IPK_PHASE
    }
}
