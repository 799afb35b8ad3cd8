// merge.cuh -- per-branch max-merge and regrouping by k-mer over SORTED composite keys, sm_100a.
//
// Input: records sorted by  ckey = (kmer << group_bits) | group  (group = branch group in processing order).
//   * equal ckey  = the same k-mer emitted again for the same branch (another window, or the branch's other ghost
//     node): keep the maximum score                      -- ipk::put, ipk/src/branch_group.cpp:88-101
//   * equal kmer  = one database row: its (branch, score) entries in group processing order
//                                                        -- unsafe_insert loop, ipk/src/db_builder.cpp:685-693
//
// One pass: a tile finds its record heads and k-mer heads, gets the number of heads in all preceding tiles by a
// decoupled look-back (one warp per counter, 32 predecessors per step) and writes its compacted output.
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace ipkgpu
{
    constexpr int MG_THREADS = 256;
    constexpr int MG_ITEMS = 12;
    constexpr int MG_TILE = MG_THREADS * MG_ITEMS;
    constexpr int MG_WARPS = MG_THREADS / 32;

    /// Look-back of one warp over status words  epoch << 34 | flag << 32 | value  at status[tile * stride]: returns the
    /// sum of the values of all tiles before `tile` (every lane gets it).  Tiles take their ids from an atomic ticket, so
    /// every predecessor is running or done.
    __device__ __forceinline__ uint32_t warp_look_back(const uint64_t* status, uint32_t tile, uint32_t stride, uint32_t epoch, int lane)
    {
        uint32_t sum = 0;
        int64_t base = int64_t(tile) - 1;
        while (base >= 0)
        {
            const int64_t t = base - lane;
            uint64_t w = rs_status(epoch, RS_FLAG_INCL, 0);            // lanes that fall off the front see an inclusive zero
            if (t >= 0)
            {
                const volatile uint64_t* p = status + uint64_t(t) * stride;
                do { w = *p; } while (uint32_t(w >> 34) != epoch || ((w >> 32) & 3) == 0);
            }
            const unsigned incl = __ballot_sync(0xffffffffu, ((w >> 32) & 3) == RS_FLAG_INCL);
            const int first = incl ? __ffs(incl) - 1 : 32;             // nearest predecessor with an inclusive value
            uint32_t v = lane <= first ? uint32_t(w) : 0u;
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            sum += v;
            if (incl) break;
            base -= 32;
        }
        return sum;
    }

    struct merge_args
    {
        const uint64_t* ckeys;             // sorted composite keys (main stream)
        const float* scores;
        uint64_t n;                        // records of BOTH streams
        // side stream (the tiles the ordered enumeration level deferred), sorted too; the kernel walks the MERGE of the two
        // (equal keys: main first).  split[t] = side records among the first t * MG_TILE merged records (mg_split).
        const uint64_t* skeys; const float* sscores; uint64_t ns; const uint32_t* split;
        uint32_t group_bits;               // width of the (batch-local) group field of ckeys
        uint32_t out_group_bits;           // width of the group field of the compacted records
        uint32_t group_add;                // batch-local group + group_add = global group
        uint64_t* u_keys; float* u_scores; // compacted records (kmer << out_group_bits | global group), or null
        const uint32_t* group_branch;      // global group -> branch id
        uint64_t* kmer_keys; uint64_t* kmer_offs; uint32_t* ent_branch; float* ent_score;     // CSR by k-mer, or kmer_keys == null
        uint64_t* status;                  // [tiles][2]
        uint32_t* ticket;
        unsigned long long* totals;        // [0] records, [1] k-mers
        uint64_t offs_add = 0;             // added to the entry offsets written to kmer_offs (a k-mer range of a larger CSR)
        uint32_t epoch;
    };

    /// split[t] = number of side records among the first min(t * MG_TILE, n) records of the merge (equal keys: main first),
    /// t = 0 .. tiles: the smallest j with  main[m - j - 1] <= side[j]  (merge path).
    __global__ void mg_split(const uint64_t* __restrict__ mkeys, uint64_t nm, const uint64_t* __restrict__ skeys, uint64_t ns,
                             uint32_t tiles, uint32_t* __restrict__ split)
    {
        const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
        if (t > tiles) return;
        const uint64_t m = min(uint64_t(t) * MG_TILE, nm + ns);
        uint64_t lo = m > nm ? m - nm : 0, hi = min(m, ns);
        while (lo < hi)
        {
            const uint64_t j = (lo + hi) >> 1;
            // m - j main records precede: fine iff the last of them is not greater than the next side record
            if (m - j > 0 && mkeys[m - j - 1] > skeys[j]) lo = j + 1; else hi = j;
        }
        split[t] = uint32_t(lo);
    }

    __global__ void __launch_bounds__(MG_THREADS) mg_merge(const __grid_constant__ merge_args A)
    {
        __shared__ uint64_t sk[MG_TILE + 1];            // sk[0] = the key before the tile
        __shared__ float ss[MG_TILE];
        __shared__ uint32_t s_cnt[2][MG_ITEMS][MG_WARPS];
        __shared__ uint32_t s_tile, s_before[2], s_first;

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const unsigned full = 0xffffffffu;
        if (tid == 0) s_tile = atomicAdd(A.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        const uint64_t base = uint64_t(tile) * MG_TILE;
        const uint32_t tile_n = uint32_t(min(uint64_t(MG_TILE), A.n - base));
        const uint32_t j0 = A.ns ? A.split[tile] : 0u, j1 = A.ns ? A.split[tile + 1] : 0u;
        const uint64_t a0 = base - j0, na = A.n - A.ns;          // first main record of the tile; main records in all
        const uint32_t nst = j1 - j0, nat = tile_n - nst;        // side / main records of this tile
        if (nst == 0)
        {
            // all loads first (12 keys + 12 scores in flight per thread), then the shared-memory stores
            const uint64_t* __restrict__ gk = A.ckeys + a0;
            const float* __restrict__ gs = A.scores + a0;
            uint64_t rk[MG_ITEMS];
            float rs[MG_ITEMS];
#pragma unroll
            for (int j = 0; j < MG_ITEMS; ++j)
            {
                const uint32_t q = uint32_t(j) * MG_THREADS + tid;
                rk[j] = q < tile_n ? __ldg(gk + q) : 0ull;
            }
#pragma unroll
            for (int j = 0; j < MG_ITEMS; ++j)
            {
                const uint32_t q = uint32_t(j) * MG_THREADS + tid;
                rs[j] = q < tile_n ? __ldg(gs + q) : 0.0f;
            }
#pragma unroll
            for (int j = 0; j < MG_ITEMS; ++j)
            {
                const uint32_t q = uint32_t(j) * MG_THREADS + tid;
                if (q < tile_n) { sk[q + 1] = rk[j]; ss[q] = rs[j]; }
            }
        }
        else
        {
            // a tile with records of both streams (rare): every record finds its merged position by a binary search in the
            // other stream's part of the tile
            for (uint32_t q = tid; q < nat; q += MG_THREADS)
            {
                const uint64_t key = A.ckeys[a0 + q];
                uint32_t lo = 0, hi = nst;                       // side records of the tile with a key < this one
                while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (A.skeys[j0 + mid] < key) lo = mid + 1; else hi = mid; }
                sk[q + lo + 1] = key; ss[q + lo] = A.scores[a0 + q];
            }
            for (uint32_t q = tid; q < nst; q += MG_THREADS)
            {
                const uint64_t key = A.skeys[j0 + q];
                uint32_t lo = 0, hi = nat;                       // main records of the tile with a key <= this one
                while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (A.ckeys[a0 + mid] <= key) lo = mid + 1; else hi = mid; }
                sk[q + lo + 1] = key; ss[q + lo] = A.sscores[j0 + q];
            }
        }
        if (tid == 0)
        {
            // the merged record before the tile = the larger of the two streams' predecessors
            uint64_t prev = 0;
            bool any = false;
            if (a0 > 0) { prev = A.ckeys[a0 - 1]; any = true; }
            if (j0 > 0) { const uint64_t p = A.skeys[j0 - 1]; if (!any || p > prev) prev = p; any = true; }
            sk[0] = prev;
            s_first = any ? 0u : 1u;
        }
        __syncthreads();
        // heads: a record whose composite key differs from its predecessor's; a k-mer head: whose k-mer differs
        unsigned hb[MG_ITEMS], kb[MG_ITEMS];
#pragma unroll
        for (int j = 0; j < MG_ITEMS; ++j)
        {
            const uint32_t q = uint32_t(j) * MG_THREADS + tid;
            bool head = false, khead = false;
            if (q < tile_n)
            {
                const uint64_t cur = sk[q + 1], prev = sk[q];
                const bool very_first = q == 0 && s_first;
                head = cur != prev || very_first;
                khead = (cur >> A.group_bits) != (prev >> A.group_bits) || very_first;
            }
            hb[j] = __ballot_sync(full, head);
            kb[j] = __ballot_sync(full, khead);
            if (lane == 0) { s_cnt[0][j][warp] = __popc(hb[j]); s_cnt[1][j][warp] = __popc(kb[j]); }
        }
        __syncthreads();
        // exclusive scan of the MG_ITEMS * MG_WARPS partial counts (tile order = row, warp), one warp per counter; the
        // tile's totals are published at once so that the successors' look-backs find them early
        uint32_t total = 0;
        uint64_t* mine = A.status + uint64_t(tile) * 2 + (warp < 2 ? warp : 0);
        if (warp < 2)
        {
            uint32_t* c = &s_cnt[warp][0][0];
            constexpr int N = MG_ITEMS * MG_WARPS, PER = (N + 31) / 32;
            uint32_t v[PER], sum = 0;
#pragma unroll
            for (int i = 0; i < PER; ++i) { const int e = lane * PER + i; v[i] = e < N ? c[e] : 0; sum += v[i]; }
            const uint32_t inc = warp_incl_scan(sum, lane);
            total = __shfl_sync(full, inc, 31);
            uint32_t run = inc - sum;
#pragma unroll
            for (int i = 0; i < PER; ++i) { const int e = lane * PER + i; if (e < N) c[e] = run; run += v[i]; }
            if (lane == 0) *reinterpret_cast<volatile uint64_t*>(mine) = rs_status(A.epoch, tile == 0 ? RS_FLAG_INCL : RS_FLAG_LOCAL, total);
        }
        // maximum score of every run that starts in this tile (runs may leave the tile): needs no global offsets
        float best[MG_ITEMS];
#pragma unroll
        for (int j = 0; j < MG_ITEMS; ++j)
        {
            best[j] = 0.0f;
            if (!((hb[j] >> lane) & 1)) continue;
            const uint32_t q = uint32_t(j) * MG_THREADS + tid;
            const uint64_t cur = sk[q + 1];
            float m = ss[q];
            uint32_t r = q + 1;
            for (; r < tile_n && sk[r + 1] == cur; ++r) { const float v = ss[r]; if (m < v) m = v; }   // branch_group.cpp:92 `it->second < kmer.score`
            if (r == tile_n)
            {
                for (uint64_t g = a0 + nat; g < na && A.ckeys[g] == cur; ++g) { const float v = A.scores[g]; if (m < v) m = v; }
                for (uint64_t g = j1; g < A.ns && A.skeys[g] == cur; ++g) { const float v = A.sscores[g]; if (m < v) m = v; }
            }
            best[j] = m;
        }
        if (warp < 2)
        {
            uint32_t before = 0;
            if (tile > 0)
            {
                before = warp_look_back(A.status + warp, tile, 2, A.epoch, lane);
                if (lane == 0) *reinterpret_cast<volatile uint64_t*>(mine) = rs_status(A.epoch, RS_FLAG_INCL, before + total);
            }
            if (lane == 0)
            {
                s_before[warp] = before;
                if (base + tile_n == A.n) A.totals[warp] = (unsigned long long)(before) + total;      // the last tile knows the grand totals
            }
        }
        __syncthreads();
        const uint64_t group_mask = (1ull << A.group_bits) - 1;
        const uint32_t lt = (1u << lane) - 1;
#pragma unroll
        for (int j = 0; j < MG_ITEMS; ++j)
        {
            if (!((hb[j] >> lane) & 1)) continue;
            const uint32_t q = uint32_t(j) * MG_THREADS + tid;
            const uint64_t cur = sk[q + 1];
            const float m = best[j];
            const uint64_t r1 = uint64_t(s_before[0]) + s_cnt[0][j][warp] + __popc(hb[j] & lt);
            const uint32_t group = uint32_t(cur & group_mask) + A.group_add;
            if (A.u_keys) { A.u_keys[r1] = ((cur >> A.group_bits) << A.out_group_bits) | group; A.u_scores[r1] = m; }
            if (A.kmer_keys)
            {
                A.ent_branch[r1] = A.group_branch[group];
                A.ent_score[r1] = m;
                if ((kb[j] >> lane) & 1)
                {
                    const uint64_t r2 = uint64_t(s_before[1]) + s_cnt[1][j][warp] + __popc(kb[j] & lt);
                    A.kmer_keys[r2] = cur >> A.group_bits;
                    A.kmer_offs[r2] = r1 + A.offs_add;
                }
            }
        }
    }
}
