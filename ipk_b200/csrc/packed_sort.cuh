// packed_sort.cuh -- the merge sort of the emitted records on 8-byte PACKED records, sm_100a.
//
// Same job as radix_sort.cuh + merge.cuh (ipk::put, branch_group.cpp:88-101, and the unsafe_insert regroup,
// db_builder.cpp:685-693, as one sort on (k-mer, group) + segmented max), with less memory traffic:
//
//   1. ONE most-significant-digit pass on the top 8 bits of the composite key  ckey = (kmer << group_bits) | group
//      splits the (u64 key, f32 score) pairs into 256 buckets.  Inside a bucket the top byte is implied by the position,
//      so what is stored is  rec = (ckey & low_mask) << 32 | score bits : 8 bytes instead of 12 (needs
//      key_bits + group_bits - 8 <= 32; every DNA configuration of BASELINE.json qualifies, the amino-acid one keeps the
//      pair sort).
//   2. The remaining ceil(low_bits / 8) LSD passes run on ONE u64 array, segmented by bucket: tiles are cut at bucket
//      boundaries, every bucket has its own digit histogram and its own look-back chain, one launch per pass covers a
//      run of buckets.  16 bytes of traffic per record and pass instead of 24, and one staging round through shared
//      memory instead of two.
//   3. The max-merge reads the packed records (8 bytes) and writes the CSR by k-mer as merge.cuh does.
//
// With several GPUs pass 1 IS the exchange: the buckets are dealt to owner ranks as contiguous k-mer ranges and the MSD
// pass stores each bucket straight into its owner's receive buffer (peer memory over NVLink), so a rank pays no extra
// pass for the all-to-all and sorts 8 bits less afterwards (ipkgpu_stage1_scatter_packed).
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"
#include "merge.cuh"
#include "enumerate.cuh"

namespace ipkgpu
{
    constexpr int PS_BUCKETS = 256;
    // Tile shapes of the two sweeps and of the merge, from an A/B of builds on one box (profiles/r03n_ab_shapes.txt; ms per step on
    // cfg3): MSD pass 384x16 @2 CTAs/SM 16.7, 384x12 @3 15.4, 512x12 @2 15.8; LSD passes (two) 384x16 @2 29.9, 384x8 @4 29.0, 512x8 @2
    // 33.5; merge 256x12 @4 12.05, 256x8 @6 11.4, 256x16 @3 12.9.  All within 8 %: the kernels are bound by per-record work, not by
    // occupancy.  With the look-back after the shared-memory scatter (r03q_ab_lookback.txt): LSD 384x8 @4 26.7, 384x12 @3 26.0,
    // 384x16 @2 27.6; polling 2 / 4 / 8 predecessors at a time 27.3 / 26.7 / 26.7; sleeping 0 / 64 / 200 ns in the spin: no difference.
    constexpr int PS_THREADS = 384;
    constexpr int PS_MSD_ITEMS = 12, PS_MSD_MINB = 3, PS_MSD_TILE = PS_THREADS * PS_MSD_ITEMS;
#ifndef IPK_LSD_ITEMS
#define IPK_LSD_ITEMS 12
#define IPK_LSD_MINB 3
#endif
#ifndef IPK_PS_RANKB
#define IPK_PS_RANKB 12         // items whose matches are issued back to back (r04f: 0 / 4 / 6 / 12 -> LSD passes 26.0 / 24.7 / 24.7 / 23.9 ms)
#endif
#ifndef IPK_PS_PREFETCH
#define IPK_PS_PREFETCH 0          // tiles ahead whose records a tile asks the L2 for (0 = off; r04g: 512 / 1024 / 2048 ahead made the LSD passes 0.7 - 1.6 ms SLOWER)
#endif
#ifndef IPK_MP_PREFETCH
#define IPK_MP_PREFETCH 1024     // r04g: merge 11.54 ms without, 10.76 / 11.07 / 11.48 with 1024 / 2048 / 4096 tiles ahead
#endif
#ifndef IPK_PS_POLL
#define IPK_PS_POLL 4
#endif
#ifndef IPK_PS_SLEEP
#define IPK_PS_SLEEP 64
#endif
    constexpr int PS_LSD_ITEMS = IPK_LSD_ITEMS, PS_LSD_MINB = IPK_LSD_MINB, PS_LSD_TILE = PS_THREADS * PS_LSD_ITEMS;
    constexpr int PS_POLL = IPK_PS_POLL;

    /// asks the L2 for records [first, first + count) of `base` (one bulk prefetch; SASS: UBLKPF): a tile that starts a tile
    /// lifetime later then loads from L2 instead of waiting for DRAM
    __device__ __forceinline__ void ps_prefetch_l2(const uint64_t* base, uint64_t first, uint64_t end, uint32_t count)
    {
        if (first >= end) return;
        const uint64_t lo = first & ~1ull;                                       // 16-byte aligned start
        const uint64_t hi = min(end, first + count) & ~1ull;                    // whole 16-byte units only
        if (hi <= lo) return;
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + lo), "r"(uint32_t((hi - lo) * 8)) : "memory");
    }

    /// bucket geometry of one segmented launch (kernel parameter, 2 KB)
    struct ps_table
    {
        uint32_t start[PS_BUCKETS + 1];        // first record of every bucket in the packed array
        uint32_t tile_first[PS_BUCKETS + 1];   // id of the first tile of every bucket of THIS launch (tiles cut at bucket boundaries)
        uint32_t d0, d1;                       // buckets [d0, d1) take part
    };

    /// destinations of the MSD pass: bucket d goes to out[owner[d]] (this rank's buffer or a peer's, mapped through CUDA IPC)
    struct ps_dest
    {
        uint64_t* out[RS_MAX_PEERS];
        uint8_t owner[PS_BUCKETS];
    };

    /// The emission stream read in branch-group order: run r (one commit of the join kernel, enumerate.cuh: emit_run) holds
    /// the logical positions [lstart[r], lstart[r + 1]) and lies at off[r] of the physical stream; runs ascend by group.
    struct ps_runs
    {
        const uint32_t* lstart;                // [n_runs + 1], lstart[n_runs] = n; null: the stream is read as it lies
        const uint32_t* off;                   // [n_runs]
        const uint32_t* tile_first;            // [tiles of the MSD pass] run that holds the tile's first logical position
        uint32_t n_runs;
    };
    constexpr int PS_RUN_WIN = 64;             // run boundaries a tile stages in shared memory (runs average thousands of records)

    /// logical -> physical position for ascending logical positions of one thread; the staged window covers the runs
    /// first .. first + PS_RUN_WIN - 1 of the tile, anything beyond is looked up in global memory
    struct ps_run_cursor
    {
        uint32_t j, hi, delta;
        __device__ __forceinline__ void init(const uint32_t* s_rs, const uint32_t* s_ro) { j = 0; hi = s_rs[1]; delta = s_ro[0] - s_rs[0]; }
        __device__ __forceinline__ uint32_t map(uint32_t L, const uint32_t* s_rs, const uint32_t* s_ro, const ps_runs& R, uint32_t first)
        {
            while (L >= hi && j + 1 < uint32_t(PS_RUN_WIN)) { ++j; hi = s_rs[j + 1]; delta = s_ro[j] - s_rs[j]; }
            if (L < hi) return L + delta;
            uint32_t lo = first + PS_RUN_WIN - 1, up = R.n_runs;          // last r in [lo, up) with lstart[r] <= L
            while (up - lo > 1) { const uint32_t mid = (lo + up) >> 1; if (R.lstart[mid] <= L) lo = mid; else up = mid; }
            return R.off[lo] + (L - R.lstart[lo]);
        }
    };

    inline void ps_fill_tiles(ps_table& T, uint32_t d0, uint32_t d1, uint32_t tile)
    {
        for (uint32_t d = 0; d <= PS_BUCKETS; ++d) T.tile_first[d] = 0;
        uint32_t t = 0;
        for (uint32_t d = d0; d < d1; ++d) { T.tile_first[d] = t; t += (T.start[d + 1] - T.start[d] + tile - 1) / tile; }
        for (uint32_t d = d1; d <= PS_BUCKETS; ++d) T.tile_first[d] = t;
        T.d0 = d0; T.d1 = d1;
    }

    /// bucket of a tile: the b in [d0, d1) with tile_first[b] <= tile < tile_first[b + 1]
    __device__ __forceinline__ uint32_t ps_bucket_of_tile(const ps_table& T, uint32_t tile)
    {
        uint32_t lo = T.d0, hi = T.d1;
        while (hi - lo > 1)
        {
            const uint32_t mid = (lo + hi) >> 1;
            if (T.tile_first[mid] <= tile) lo = mid; else hi = mid;
        }
        return lo;
    }

    template<int THREADS, int ITEMS, int RADIX = RS_RADIX> constexpr size_t ps_smem_bytes(bool digits) { return size_t(THREADS / 32) * RADIX * 4 + 2 * RADIX * 4 + 64 + size_t(THREADS) * ITEMS * (digits ? 9 : 8); }

    // -----------------------------------------------------------------------------------------------------------------
    // Ranking core shared by the two sweeps: the tile's items in (warp, round, lane) order carry a digit each; afterwards
    // pos[i] = position of item i in the tile's digit-sorted order, dbase / goff as in rs_onesweep.
    // -----------------------------------------------------------------------------------------------------------------
    template<int THREADS, int ITEMS, int RADIX>
    struct ps_smem_view
    {
        uint64_t* sk;                       // [TILE] staged records
        uint32_t (*wh)[RADIX];              // [WARPS][RADIX]
        uint32_t* dbase;                    // [RADIX]
        uint32_t* goff;                     // [RADIX]
        uint32_t* misc;                     // [16]
        uint8_t* sd;                        // [TILE] staged digits (MSD pass only)
        __device__ explicit ps_smem_view(unsigned char* raw)
        {
            constexpr int WARPS = THREADS / 32, TILE = THREADS * ITEMS;
            sk = reinterpret_cast<uint64_t*>(raw);
            wh = reinterpret_cast<uint32_t (*)[RADIX]>(sk + TILE);
            dbase = &wh[0][0] + WARPS * RADIX;
            goff = dbase + RADIX;
            misc = goff + RADIX;
            sd = reinterpret_cast<uint8_t*>(misc + 16);
        }
    };

    /// One pass over one tile.  MSD = true: (keys_in, vals_in) pairs -> packed records, digit = top byte of the composite key,
    /// destination per digit from D, tiles are plain runs of THREADS * ITEMS pairs.  MSD = false: packed -> packed, digit =
    /// byte `shift / 8` of the record, tiles cut at bucket boundaries (T), digit bases per bucket.
    /// BITS = width of the digit (LSD passes of 4 .. 8 bits: everything that is per digit -- the counters' reset and scan,
    /// the status words, the look-back -- shrinks with the radix, the per-record work does not).
    template<int THREADS, int ITEMS, bool MSD, int MINB, bool RUNS = false, int BITS = 8>
    __global__ void __launch_bounds__(THREADS, MINB) ps_sweep(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                              uint64_t* __restrict__ out, const uint32_t* __restrict__ digit_base, uint64_t n,
                                                              uint32_t shift /* MSD: low_bits; else: bit of the record where the digit starts */,
                                                              uint32_t base_stride /* LSD: words between two buckets' digit bases */,
                                                              uint64_t* status, uint32_t* ticket, uint32_t epoch,
                                                              const __grid_constant__ ps_table T, const __grid_constant__ ps_dest D,
                                                              const __grid_constant__ ps_runs R)
    {
        constexpr int WARPS = THREADS / 32, TILE = THREADS * ITEMS;
        constexpr int RADIX = 1 << BITS, DWARPS = RADIX >= 32 ? RADIX / 32 : 1;        // DWARPS: warps that hold one digit per lane
        constexpr uint32_t DMASK = RADIX - 1;
        static_assert(THREADS >= RADIX && THREADS % 32 == 0 && ITEMS % 4 == 0 && BITS >= 1 && BITS <= 8 && (!MSD || BITS == 8), "tile shape");
        extern __shared__ __align__(16) unsigned char ps_smem[];
        const ps_smem_view<THREADS, ITEMS, RADIX> S(ps_smem);
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const unsigned full = 0xffffffffu;
        if (tid == 0)
        {
            const uint32_t t = atomicAdd(ticket, 1u);
            S.misc[15] = t;
            if (!MSD) S.misc[14] = ps_bucket_of_tile(T, t);
        }
        for (int i = tid; i < WARPS * RADIX; i += THREADS) (&S.wh[0][0])[i] = 0;
        __syncthreads();
        const uint32_t tile = S.misc[15];
        uint64_t tile_base;
        uint32_t tile_n, chain_pos;                     // chain_pos: tiles before this one in its look-back chain
        bool first_of_chain;
        const uint32_t* dbase_g = digit_base;
        if (MSD)
        {
            tile_base = uint64_t(tile) * TILE;
            tile_n = uint32_t(min(uint64_t(TILE), n - tile_base));
            first_of_chain = tile == 0;
            chain_pos = tile;
        }
        else
        {
            const uint32_t b = S.misc[14];
            const uint32_t lt = tile - T.tile_first[b];
            tile_base = uint64_t(T.start[b]) + uint64_t(lt) * TILE;
            tile_n = uint32_t(min(uint64_t(TILE), uint64_t(T.start[b + 1]) - tile_base));
            first_of_chain = lt == 0;
            chain_pos = lt;
            dbase_g = digit_base + size_t(b) * base_stride;
            // (bucket boundaries shift the tiles ahead by less than a tile each: close enough for a prefetch)
            if (IPK_PS_PREFETCH > 0 && tid == 32) ps_prefetch_l2(keys_in, tile_base + uint64_t(IPK_PS_PREFETCH) * TILE, T.start[T.d1], TILE);
        }

        uint64_t rec[ITEMS];
        uint32_t dpk[ITEMS / 4];            // the items' digits, four per word
        uint32_t rkp[ITEMS / 2];            // their ranks among the warp's earlier items of the same digit, two per word
#pragma unroll
        for (int i = 0; i < ITEMS / 4; ++i) dpk[i] = 0;
#pragma unroll
        for (int i = 0; i < ITEMS / 2; ++i) rkp[i] = 0;
        if (MSD)
        {
            const uint64_t low_mask = shift >= 64 ? ~0ull : ((1ull << shift) - 1);
            if (RUNS)
            {
                // the tile's logical positions, mapped run by run to where the enumeration put them
                __shared__ uint32_t s_rs[PS_RUN_WIN + 1], s_ro[PS_RUN_WIN];
                const uint32_t first = R.tile_first[tile];
                if (tid <= PS_RUN_WIN)
                {
                    const uint32_t r = first + tid;
                    s_rs[tid] = r <= R.n_runs ? R.lstart[r] : 0xffffffffu;
                    if (tid < PS_RUN_WIN) s_ro[tid] = r < R.n_runs ? R.off[r] : 0u;
                }
                __syncthreads();
                ps_run_cursor c;
                c.init(s_rs, s_ro);
#pragma unroll
                for (int i = 0; i < ITEMS; ++i)
                {
                    const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                    rec[i] = li < tile_n ? keys_in[c.map(uint32_t(tile_base) + li, s_rs, s_ro, R, first)] : 0ull;
                }
                c.init(s_rs, s_ro);
#pragma unroll
                for (int i = 0; i < ITEMS; ++i)
                {
                    const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                    const uint32_t v = li < tile_n ? vals_in[c.map(uint32_t(tile_base) + li, s_rs, s_ro, R, first)] : 0u;
                    dpk[i / 4] |= (uint32_t(rec[i] >> shift) & 0xffu) << (8 * (i & 3));
                    rec[i] = ((rec[i] & low_mask) << 32) | v;
                }
            }
            else
            {
#pragma unroll
                for (int i = 0; i < ITEMS; ++i)
                {
                    const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                    rec[i] = li < tile_n ? keys_in[tile_base + li] : 0ull;
                }
#pragma unroll
                for (int i = 0; i < ITEMS; ++i)
                {
                    const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                    const uint32_t v = li < tile_n ? vals_in[tile_base + li] : 0u;
                    dpk[i / 4] |= (uint32_t(rec[i] >> shift) & 0xffu) << (8 * (i & 3));
                    rec[i] = ((rec[i] & low_mask) << 32) | v;
                }
            }
        }
        else
        {
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
            {
                const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                rec[i] = li < tile_n ? keys_in[tile_base + li] : 0ull;
                dpk[i / 4] |= (uint32_t(rec[i] >> shift) & DMASK) << (8 * (i & 3));
            }
        }
        // The matches of a batch of items are issued back to back (MATCH has a long latency and does not depend on the
        // counters); then, item by item, every lane reads its digit's warp counter and the lowest lane of each digit adds the
        // digit's count to it (no leader election, no shuffle).
        const unsigned lt_mask = (1u << lane) - 1;
#pragma unroll
        for (int i0 = 0; i0 < ITEMS; i0 += IPK_PS_RANKB)
        {
            unsigned pe[IPK_PS_RANKB];
#pragma unroll
            for (int j = 0; j < IPK_PS_RANKB; ++j)
            {
                const int i = i0 + j;
                if (i >= ITEMS) continue;
                const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                const uint32_t d = li < tile_n ? (dpk[i / 4] >> (8 * (i & 3))) & 0xffu : uint32_t(RADIX);
                pe[j] = __match_any_sync(full, d);
            }
#pragma unroll
            for (int j = 0; j < IPK_PS_RANKB; ++j)
            {
                const int i = i0 + j;
                if (i >= ITEMS) continue;
                const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
                const bool valid = li < tile_n;
                const uint32_t d = (dpk[i / 4] >> (8 * (i & 3))) & 0xffu;
                const unsigned below = pe[j] & lt_mask;
                const uint32_t old = valid ? S.wh[warp][d] : 0u;
                __syncwarp();
                if (valid && below == 0) S.wh[warp][d] = old + __popc(pe[j]);
                __syncwarp();
                rkp[i / 2] |= ((old + __popc(below)) & 0xffffu) << (16 * (i & 1));
            }
        }
        __syncthreads();
        const bool is_digit = tid < RADIX;
        uint32_t cnt = 0;
        if (is_digit)
        {
#pragma unroll
            for (int w = 0; w < WARPS; ++w)
            {
                const uint32_t t = S.wh[w][tid];
                S.wh[w][tid] = cnt;
                cnt += t;
            }
        }
        uint64_t* my_status = status + uint64_t(tile) * RADIX + (is_digit ? tid : 0);
        if (is_digit)
            *reinterpret_cast<volatile uint64_t*>(my_status) = rs_status(epoch, first_of_chain ? RS_FLAG_INCL : RS_FLAG_LOCAL, cnt);
        const uint32_t inc = warp_incl_scan(cnt, lane);                 // (cnt = 0 in the lanes beyond the radix)
        if (warp < DWARPS && lane == 31) S.misc[warp] = inc;
        __syncthreads();
        if (warp == 0)
        {
            const uint32_t w = lane < DWARPS ? S.misc[lane] : 0;
            const uint32_t winc = warp_incl_scan(w, lane);
            if (lane < DWARPS) S.misc[lane] = winc - w;
        }
        __syncthreads();
        // The records are placed into the tile's digit order in shared memory FIRST: that only needs the local scan, and it
        // gives the preceding tiles time to publish before the look-back below starts walking (ncu r03p: 30 % of the
        // instructions the kernel executed were polls of words that were not there yet).
        const uint32_t ex_local = is_digit ? S.misc[warp] + inc - cnt : 0u;
        if (is_digit) S.dbase[tid] = ex_local;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            if (li < tile_n)
            {
                const uint32_t d = (dpk[i / 4] >> (8 * (i & 3))) & 0xffu;
                const uint32_t pos = S.dbase[d] + S.wh[warp][d] + ((rkp[i / 2] >> (16 * (i & 1))) & 0xffffu);
                S.sk[pos] = rec[i];
                if (MSD) S.sd[pos] = uint8_t(d);
            }
        }
        if (is_digit)
        {
            uint32_t before = 0;
            if (!first_of_chain)
            {
                // the chain ends at the first tile of the bucket (MSD: of the launch), which published an inclusive word.
                // Four predecessors are polled at a time: their loads overlap, and a walk over LOCAL words costs one
                // memory latency per four tiles instead of one per tile; a word that is not there yet is waited for with a
                // short sleep instead of a hot spin.
                const volatile uint64_t* p = status + uint64_t(tile - 1) * RADIX + tid;
                uint32_t left = chain_pos;                      // predecessors in the chain, nearest first
                bool done = false;
                while (!done)
                {
                    uint64_t w[PS_POLL];
#pragma unroll
                    for (int j = 0; j < PS_POLL; ++j) w[j] = uint32_t(j) < left ? p[-int64_t(j) * RADIX] : 0ull;
                    int used = 0;
#pragma unroll
                    for (int j = 0; j < PS_POLL; ++j)
                    {
                        if (done || uint32_t(j) >= left || used != j) continue;
                        if (uint32_t(w[j] >> 34) != epoch || ((w[j] >> 32) & 3) == 0) continue;      // not published yet: stop consuming here
                        before += uint32_t(w[j]);
                        used = j + 1;
                        if (((w[j] >> 32) & 3) == RS_FLAG_INCL) done = true;
                    }
                    left -= used; p -= int64_t(used) * RADIX;
                    if (IPK_PS_SLEEP > 0 && !done && used < PS_POLL) __nanosleep(IPK_PS_SLEEP);
                }
                *reinterpret_cast<volatile uint64_t*>(my_status) = rs_status(epoch, RS_FLAG_INCL, before + cnt);
            }
            S.goff[tid] = dbase_g[tid] + before - ex_local;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t q = uint32_t(i) * THREADS + tid;
            if (q < tile_n)
            {
                const uint64_t r = S.sk[q];
                if (MSD)
                {
                    const uint32_t d = S.sd[q];
                    D.out[D.owner[d]][S.goff[d] + q] = r;            // a store over NVLink when the owner is a peer
                }
                else
                {
                    const uint32_t d = uint32_t(r >> shift) & DMASK;
                    out[S.goff[d] + q] = r;
                }
            }
        }
    }

    /// Digit widths of the LSD passes over `total` key bits: as few passes as 8-bit digits need, the bits spread evenly over
    /// them (12 bits = 6 + 6, not 8 + 4: the per-digit work of a pass -- counters, status words, look-back -- goes with
    /// 2^width).  IPKGPU_LSD_WIDTHS="8,4" overrides (tuning).
    inline uint32_t ps_pass_widths(uint32_t total, uint32_t* width /*[8]*/)
    {
        uint32_t passes = div_up<uint32_t>(total, 8);
        if (const char* e = getenv("IPKGPU_LSD_WIDTHS"))
        {
            uint32_t w[8], nw = 0, sum = 0;
            for (const char* q = e; *q && nw < 8;)
            {
                const uint32_t v = uint32_t(strtoul(q, const_cast<char**>(&q), 10));
                if (v < 1 || v > 8) { nw = 0; break; }
                w[nw++] = v; sum += v;
                if (*q == ',') ++q;
            }
            if (nw && sum >= total) { for (uint32_t i = 0; i < nw; ++i) width[i] = w[i]; return nw; }
        }
        uint32_t left = total;
        for (uint32_t p = 0; p < passes; ++p) { width[p] = div_up<uint32_t>(left, passes - p); left -= width[p]; }
        return passes;
    }

    /// Digit histograms of every LSD pass and bucket from ONE read of the packed records: hist[b][p][d] += ...
    /// CTA c walks records [c * chunk, (c + 1) * chunk) of the launch's bucket range and flushes its shared counters
    /// whenever it crosses a bucket boundary.
    constexpr int PS_HIST_THREADS = 256;
    constexpr int PS_HIST_ITEMS = 16;
    __global__ void __launch_bounds__(PS_HIST_THREADS) ps_seg_hist(const uint64_t* __restrict__ recs, uint32_t n_passes, uint64_t pass_shift, uint64_t pass_bits, uint64_t chunk,
                                                                  uint32_t* __restrict__ hist, const __grid_constant__ ps_table T)
    {
        extern __shared__ uint32_t ps_sh[];              // [2 copies][n_passes][256]
        const uint32_t per_copy = n_passes * RS_RADIX;
        uint32_t* mine = ps_sh + ((threadIdx.x >> 5) & 1) * per_copy;
        const uint64_t lo = uint64_t(T.start[T.d0]) + uint64_t(blockIdx.x) * chunk;
        const uint64_t hi = min(lo + chunk, uint64_t(T.start[T.d1]));
        if (lo >= hi) return;
        // bucket of the first record: largest b in [d0, d1) with start[b] <= lo
        uint32_t b = T.d0;
        {
            uint32_t l = T.d0, h = T.d1;
            while (h - l > 1) { const uint32_t m = (l + h) >> 1; if (uint64_t(T.start[m]) <= lo) l = m; else h = m; }
            b = l;
        }
        uint64_t pos = lo;
        while (pos < hi)
        {
            while (uint64_t(T.start[b + 1]) <= pos) ++b;         // skips empty buckets
            const uint64_t seg_end = min(hi, uint64_t(T.start[b + 1]));
            for (uint32_t i = threadIdx.x; i < 2 * per_copy; i += PS_HIST_THREADS) ps_sh[i] = 0;
            __syncthreads();
            for (uint64_t base = pos; base < seg_end; base += uint64_t(PS_HIST_THREADS) * PS_HIST_ITEMS)
            {
                uint64_t r[PS_HIST_ITEMS];
#pragma unroll
                for (int i = 0; i < PS_HIST_ITEMS; ++i)
                {
                    const uint64_t idx = base + uint64_t(i) * PS_HIST_THREADS + threadIdx.x;
                    r[i] = idx < seg_end ? recs[idx] : 0;
                }
#pragma unroll
                for (int i = 0; i < PS_HIST_ITEMS; ++i)
                {
                    const uint64_t idx = base + uint64_t(i) * PS_HIST_THREADS + threadIdx.x;
                    if (idx < seg_end)
                    {
                        const uint32_t kk = uint32_t(r[i] >> 32);
                        for (uint32_t p = 0; p < n_passes; ++p)           // pass p: bits [shift_p, shift_p + bits_p) of the key field
                            atomicAdd(&mine[p * RS_RADIX + ((kk >> uint32_t((pass_shift >> (8 * p)) & 0xff)) & ((1u << uint32_t((pass_bits >> (8 * p)) & 0xff)) - 1))], 1u);
                    }
                }
            }
            __syncthreads();
            uint32_t* g = hist + size_t(b) * per_copy;
            for (uint32_t i = threadIdx.x; i < per_copy; i += PS_HIST_THREADS)
            {
                const uint32_t v = ps_sh[i] + ps_sh[per_copy + i];
                if (v) atomicAdd(&g[i], v);
            }
            __syncthreads();
            pos = seg_end;
        }
    }

    /// in place: hist[b][p][0..256) -> start[b] + exclusive scan; grid = (n_passes, buckets of the launch)
    __global__ void __launch_bounds__(RS_RADIX) ps_seg_scan(uint32_t* __restrict__ hist, uint32_t n_passes, const __grid_constant__ ps_table T)
    {
        __shared__ uint32_t s_warp[9];
        const uint32_t b = T.d0 + blockIdx.y;
        uint32_t* h = hist + (size_t(b) * n_passes + blockIdx.x) * RS_RADIX;
        const uint32_t v = h[threadIdx.x];
        h[threadIdx.x] = T.start[b] + block_excl_scan_256(v, s_warp, nullptr);
    }

    // -----------------------------------------------------------------------------------------------------------------
    // max-merge + CSR by k-mer over sorted PACKED records (see merge.cuh for the semantics); tiles cut at bucket boundaries
    // -----------------------------------------------------------------------------------------------------------------
    constexpr int MP_THREADS = 256;
    constexpr int MP_ITEMS = 8;
    constexpr int MP_MINB = 6;
    constexpr int MP_TILE = MP_THREADS * MP_ITEMS;
    constexpr int MP_WARPS = MP_THREADS / 32;

    struct mergep_args
    {
        const uint64_t* recs;              // packed, sorted inside every bucket
        uint32_t group_bits;               // width of the group field inside the packed key
        uint32_t kmer_low_bits;            // k-mer bits below the bucket byte (key_bits - 8)
        uint32_t group_add;
        const uint32_t* group_branch;
        uint64_t* kmer_keys; uint64_t* kmer_offs; uint32_t* ent_branch; float* ent_score;
        uint64_t* status; uint32_t* ticket; unsigned long long* totals;
        uint64_t offs_add;
        uint32_t epoch, n_tiles;
    };

    __global__ void __launch_bounds__(MP_THREADS, MP_MINB) mg_merge_packed(const __grid_constant__ mergep_args A, const __grid_constant__ ps_table T)
    {
        __shared__ uint64_t sk[MP_TILE + 1];            // sk[0] = the record before the tile (same bucket)
        __shared__ uint32_t s_cnt[2][MP_ITEMS][MP_WARPS];
        __shared__ uint32_t s_tile, s_bucket, s_before[2];

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const unsigned full = 0xffffffffu;
        if (tid == 0)
        {
            const uint32_t t = atomicAdd(A.ticket, 1u);
            s_tile = t;
            s_bucket = ps_bucket_of_tile(T, t);
        }
        __syncthreads();
        const uint32_t tile = s_tile, bucket = s_bucket;
        const uint32_t lt = tile - T.tile_first[bucket];
        const uint64_t base = uint64_t(T.start[bucket]) + uint64_t(lt) * MP_TILE, bucket_end = T.start[bucket + 1];
        const uint32_t tile_n = uint32_t(min(uint64_t(MP_TILE), bucket_end - base));
        const bool first_in_bucket = lt == 0;
        if (IPK_MP_PREFETCH > 0 && tid == 32) ps_prefetch_l2(A.recs, base + uint64_t(IPK_MP_PREFETCH) * MP_TILE, T.start[T.d1], MP_TILE);
        {
            const uint64_t* __restrict__ g = A.recs + base;
            uint64_t r[MP_ITEMS];
#pragma unroll
            for (int j = 0; j < MP_ITEMS; ++j)
            {
                const uint32_t q = uint32_t(j) * MP_THREADS + tid;
                r[j] = q < tile_n ? __ldg(g + q) : 0ull;
            }
#pragma unroll
            for (int j = 0; j < MP_ITEMS; ++j)
            {
                const uint32_t q = uint32_t(j) * MP_THREADS + tid;
                if (q < tile_n) sk[q + 1] = r[j];
            }
            if (tid == 0) sk[0] = first_in_bucket ? 0ull : A.recs[base - 1];
        }
        __syncthreads();
        const uint32_t kshift = 32 + A.group_bits;
        unsigned hb[MP_ITEMS], kb[MP_ITEMS];
        uint32_t own[MP_ITEMS];             // the record's own score bits
#pragma unroll
        for (int j = 0; j < MP_ITEMS; ++j)
        {
            const uint32_t q = uint32_t(j) * MP_THREADS + tid;
            bool head = false, khead = false;
            own[j] = 0;
            if (q < tile_n)
            {
                const uint64_t cur = sk[q + 1], prev = sk[q];
                own[j] = uint32_t(cur);
                const bool very_first = q == 0 && first_in_bucket;
                head = (cur >> 32) != (prev >> 32) || very_first;
                khead = (cur >> kshift) != (prev >> kshift) || very_first;
            }
            hb[j] = __ballot_sync(full, head);
            kb[j] = __ballot_sync(full, khead);
            if (lane == 0) { s_cnt[0][j][warp] = __popc(hb[j]); s_cnt[1][j][warp] = __popc(kb[j]); }
        }
        __syncthreads();
        uint32_t total = 0;
        uint64_t* mine = A.status + uint64_t(tile) * 2 + (warp < 2 ? warp : 0);
        if (warp < 2)
        {
            uint32_t* c = &s_cnt[warp][0][0];
            constexpr int N = MP_ITEMS * MP_WARPS, PER = (N + 31) / 32;
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
        // maximum score of every run that starts in this tile (a run may leave the tile, never the bucket)
        float best[MP_ITEMS];
#pragma unroll
        for (int j = 0; j < MP_ITEMS; ++j)
        {
            best[j] = __uint_as_float(own[j]);
            if (!((hb[j] >> lane) & 1)) continue;
            // nine runs in ten are one record long: the next record of the warp's row is a head too (ncu r03p: this loop was 30 % of
            // the kernel's instructions, 16 of 32 lanes active)
            if (lane < 31 && ((hb[j] >> (lane + 1)) & 1)) continue;
            const uint32_t q = uint32_t(j) * MP_THREADS + tid;
            const uint64_t cur = sk[q + 1];
            const uint32_t ck = uint32_t(cur >> 32);
            float m = __uint_as_float(uint32_t(cur));
            uint32_t r = q + 1;
            for (; r < tile_n; ++r)
            {
                const uint64_t o = sk[r + 1];
                if (uint32_t(o >> 32) != ck) break;
                const float v = __uint_as_float(uint32_t(o));
                if (m < v) m = v;                                   // branch_group.cpp:92 `it->second < kmer.score`
            }
            if (r == tile_n)
                for (uint64_t g = base + tile_n; g < bucket_end; ++g)
                {
                    const uint64_t o = A.recs[g];
                    if (uint32_t(o >> 32) != ck) break;
                    const float v = __uint_as_float(uint32_t(o));
                    if (m < v) m = v;
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
                if (tile + 1 == A.n_tiles) A.totals[warp] = (unsigned long long)(before) + total;
            }
        }
        __syncthreads();
        const uint32_t group_mask = A.group_bits >= 32 ? ~0u : ((1u << A.group_bits) - 1);
        const uint64_t kmer_top = uint64_t(bucket) << A.kmer_low_bits;
        const uint32_t lt_mask = (1u << lane) - 1;
#pragma unroll
        for (int j = 0; j < MP_ITEMS; ++j)
        {
            if (!((hb[j] >> lane) & 1)) continue;
            const uint32_t q = uint32_t(j) * MP_THREADS + tid;
            const uint32_t ck = uint32_t(sk[q + 1] >> 32);
            const uint64_t r1 = uint64_t(s_before[0]) + s_cnt[0][j][warp] + __popc(hb[j] & lt_mask);
            const uint32_t group = (ck & group_mask) + A.group_add;
            A.ent_branch[r1] = A.group_branch[group];
            A.ent_score[r1] = best[j];
            if ((kb[j] >> lane) & 1)
            {
                const uint64_t r2 = uint64_t(s_before[1]) + s_cnt[1][j][warp] + __popc(kb[j] & lt_mask);
                A.kmer_keys[r2] = kmer_top | uint64_t(A.group_bits >= 32 ? 0u : (ck >> A.group_bits));
                A.kmer_offs[r2] = r1 + A.offs_add;
            }
        }
    }

    // -----------------------------------------------------------------------------------------------------------------
    // host side
    // -----------------------------------------------------------------------------------------------------------------
    /// true when the packed path applies to composite keys of key_bits + group_bits
    inline bool ps_eligible(uint32_t key_bits, uint32_t group_bits, uint64_t n)
    {
        static const bool off = getenv("IPKGPU_PACKED") && atoi(getenv("IPKGPU_PACKED")) == 0;
        return !off && n > 0 && n <= 0xffffffffull && key_bits >= 8 && key_bits + group_bits - 8 <= 32;
    }

    /// status words + tickets for launches of up to n_tiles tiles (shares ctx->rs_status with radix_sort.cuh)
    inline void ps_reserve_status(ipkgpu_ctx* ctx, uint64_t n_tiles)
    {
        const size_t want = n_tiles * RS_RADIX * 8;
        if (want > ctx->rs_status.cap)
        {
            ctx->reserve(ctx->rs_status, want);
            IPK_CUDA(cudaMemsetAsync(ctx->rs_status.p, 0, ctx->rs_status.cap, ctx->stream));
            ctx->rs_epoch = 0;
        }
    }

    __global__ void ps_runs_keys(const emit_run* __restrict__ runs, uint32_t n_runs, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n_runs) { keys[i] = runs[i].group; vals[i] = i; }
    }
    __global__ void ps_runs_gather(const emit_run* __restrict__ runs, const uint32_t* __restrict__ perm, uint32_t n_runs, uint32_t n_total,
                                   uint32_t* __restrict__ cnt, uint32_t* __restrict__ off, uint32_t* __restrict__ lstart)
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i == 0) lstart[n_runs] = n_total;
        if (i < n_runs) { const emit_run r = runs[perm[i]]; cnt[i] = r.count; off[i] = uint32_t(r.offset); }
    }

    /// tile_first[t] = last run r with lstart[r] <= t * tile (one thread per tile of the MSD pass)
    __global__ void ps_runs_tile_first(const uint32_t* __restrict__ lstart, uint32_t n_runs, uint32_t n_tiles, uint32_t tile, uint32_t* __restrict__ tile_first)
    {
        const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
        if (t >= n_tiles) return;
        const uint32_t target = t * tile;
        uint32_t lo = 0, up = n_runs;
        while (up - lo > 1) { const uint32_t mid = (lo + up) >> 1; if (lstart[mid] <= target) lo = mid; else up = mid; }
        tile_first[t] = lo;
    }

    /// ctx->runs (ctx->runs_n commits of the enumeration, any order) -> the group-ordered index the MSD pass reads through.
    /// A stable sort of the run ids by group, then a scan of the counts.  Uses ctx->hist / ctx->scan_tmp as sort scratch.
    inline ps_runs ps_order_runs(ipkgpu_ctx* ctx, uint64_t n, uint32_t group_bits)
    {
        const uint32_t nr = uint32_t(ctx->runs_n);
        const size_t a8 = (size_t(nr) * 8 + 255) & ~size_t(255), a4 = (size_t(nr + 1) * 4 + 255) & ~size_t(255);
        const uint32_t n_tiles = uint32_t(div_up<uint64_t>(n, PS_MSD_TILE));
        ctx->reserve(ctx->runs_scratch, 2 * a8 + 5 * a4 + size_t(n_tiles) * 4);
        char* base = ctx->runs_scratch.as<char>();
        uint64_t* ka = reinterpret_cast<uint64_t*>(base); uint64_t* kb = reinterpret_cast<uint64_t*>(base + a8);
        uint32_t* va = reinterpret_cast<uint32_t*>(base + 2 * a8); uint32_t* vb = va + a4 / 4;
        uint32_t* cnt = vb + a4 / 4; uint32_t* off = cnt + a4 / 4; uint32_t* lstart = off + a4 / 4;
        const uint32_t grid = div_up<uint32_t>(nr, 256);
        IPK_LAUNCH(ctx, "ps_runs_keys", uint64_t(nr) * 28, ps_runs_keys<<<grid, 256, 0, ctx->stream>>>(ctx->runs.as<emit_run>(), nr, ka, va));
        const bool in_b = radix_sort_pairs(ctx, ka, va, kb, vb, nr, group_bits, ctx->hist, ctx->scan_tmp);
        IPK_LAUNCH(ctx, "ps_runs_gather", uint64_t(nr) * 28,
                   ps_runs_gather<<<grid, 256, 0, ctx->stream>>>(ctx->runs.as<emit_run>(), in_b ? vb : va, nr, uint32_t(n), cnt, off, lstart));
        exclusive_scan_u32(ctx, cnt, lstart, nr, ctx->scan_tmp);
        uint32_t* tile_first = lstart + a4 / 4;
        IPK_LAUNCH(ctx, "ps_runs_tile_first", uint64_t(n_tiles) * 80,
                   ps_runs_tile_first<<<div_up<uint32_t>(n_tiles, 256), 256, 0, ctx->stream>>>(lstart, nr, n_tiles, PS_MSD_TILE, tile_first));
        return ps_runs{ lstart, off, tile_first, nr };
    }

    /// The emission copied into branch-group order through the run index (tiles of PS_MSD_TILE logical positions): for
    /// records that do not pack (amino acids) the pair sort then only has the k-mer bits to sort, stably -- one 24-byte copy
    /// instead of the passes on the group bits.
    __global__ void __launch_bounds__(PS_THREADS) ps_gather_runs(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                                 uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint64_t n,
                                                                 const __grid_constant__ ps_runs R)
    {
        __shared__ uint32_t s_rs[PS_RUN_WIN + 1], s_ro[PS_RUN_WIN];
        const uint32_t tile = blockIdx.x, tid = threadIdx.x;
        const uint64_t tile_base = uint64_t(tile) * PS_MSD_TILE;
        const uint32_t tile_n = uint32_t(min(uint64_t(PS_MSD_TILE), n - tile_base));
        const uint32_t first = R.tile_first[tile];
        if (tid <= uint32_t(PS_RUN_WIN))
        {
            const uint32_t r = first + tid;
            s_rs[tid] = r <= R.n_runs ? R.lstart[r] : 0xffffffffu;
            if (tid < uint32_t(PS_RUN_WIN)) s_ro[tid] = r < R.n_runs ? R.off[r] : 0u;
        }
        __syncthreads();
        ps_run_cursor c;
        c.init(s_rs, s_ro);
        for (uint32_t li = tid; li < tile_n; li += PS_THREADS)
        {
            const uint32_t L = uint32_t(tile_base) + li;
            const uint32_t p = c.map(L, s_rs, s_ro, R, first);
            keys_out[L] = keys_in[p];
            vals_out[L] = vals_in[p];
        }
    }

    inline void ps_gather_in_group_order(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout, uint64_t n, uint32_t group_bits)
    {
        const ps_runs Rn = ps_order_runs(ctx, n, group_bits);
        const uint32_t n_tiles = uint32_t(div_up<uint64_t>(n, PS_MSD_TILE));
        IPK_LAUNCH(ctx, "ps_gather_runs", n * 24, ps_gather_runs<<<n_tiles, PS_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, n, Rn));
    }

    /// top-byte histogram of the composite keys -> hist[0..256) (device, in ctx->hist) ; not yet scanned
    inline void ps_top_histogram(ipkgpu_ctx* ctx, const uint64_t* keys, uint64_t n, uint32_t low_bits)
    {
        ctx->reserve(ctx->hist, (RS_MAX_PASSES * RS_RADIX + 16) * 4);
        IPK_CUDA(cudaMemsetAsync(ctx->hist.p, 0, (RS_MAX_PASSES * RS_RADIX + 16) * 4, ctx->stream));
        if (n == 0) return;
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, RS_HIST_THREADS * RS_HIST_ITEMS), uint64_t(ctx->num_sms) * 8));
        IPK_LAUNCH(ctx, "rs_hist", n * 8, rs_hist_bits<<<hgrid, RS_HIST_THREADS, 2 * RS_RADIX * 4, ctx->stream>>>(keys, n, low_bits, 8, 1, ctx->hist.as<uint32_t>()));
    }

    /// hist[d] = sum over the slots of partial[slot][d]: the join kernel's per-SM counts of the emission by top k-mer byte
    __global__ void __launch_bounds__(RS_RADIX) ps_reduce_top_hist(const uint32_t* __restrict__ partial, uint32_t slots, uint32_t* __restrict__ hist)
    {
        uint32_t s = 0;
        for (uint32_t i = 0; i < slots; ++i) s += partial[size_t(i) * RS_RADIX + threadIdx.x];
        hist[threadIdx.x] = s;
    }

    /// the top-byte histogram from the join kernel's own counts (enum_args::top_hist) instead of a pass over the keys
    inline void ps_top_histogram_from_join(ipkgpu_ctx* ctx)
    {
        ctx->reserve(ctx->hist, (RS_MAX_PASSES * RS_RADIX + 16) * 4);
        IPK_CUDA(cudaMemsetAsync(ctx->hist.p, 0, (RS_MAX_PASSES * RS_RADIX + 16) * 4, ctx->stream));
        IPK_LAUNCH(ctx, "ps_reduce_top_hist", uint64_t(JOIN_HIST_SLOTS) * RS_RADIX * 4,
                   ps_reduce_top_hist<<<1, RS_RADIX, 0, ctx->stream>>>(ctx->join_hist.as<uint32_t>(), JOIN_HIST_SLOTS, ctx->hist.as<uint32_t>()));
    }

    /// MSD pass: n (key, score) pairs -> packed records in D's buffers; d_base[256] (device) = where this rank's records of
    /// each bucket start inside the owner's buffer.  Uses tickets[15] of ctx->hist.
    inline void ps_msd_pack(ipkgpu_ctx* ctx, const uint64_t* keys, const uint32_t* vals, uint64_t n, uint32_t low_bits,
                            const uint32_t* d_base, const ps_dest& D, bool peers, const ps_runs& Rn)
    {
        if (n == 0) return;
        const uint32_t n_tiles = uint32_t(div_up<uint64_t>(n, PS_MSD_TILE));
        ps_reserve_status(ctx, n_tiles);
        uint32_t* ticket = ctx->hist.as<uint32_t>() + RS_MAX_PASSES * RS_RADIX + 15;
        IPK_CUDA(cudaMemsetAsync(ticket, 0, 4, ctx->stream));
        ps_table none;
        memset(&none, 0, sizeof none);
        constexpr size_t smem = ps_smem_bytes<PS_THREADS, PS_MSD_ITEMS>(true);
        auto kern = Rn.lstart ? ps_sweep<PS_THREADS, PS_MSD_ITEMS, true, PS_MSD_MINB, true> : ps_sweep<PS_THREADS, PS_MSD_ITEMS, true, PS_MSD_MINB, false>;
        IPK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        IPK_LAUNCH(ctx, peers ? "ps_msd_scatter_peers" : "ps_msd_pack", n * 20 + uint64_t(n_tiles) * RS_RADIX * 8,
                   kern<<<n_tiles, PS_THREADS, smem, ctx->stream>>>(keys, vals, nullptr, d_base, n, low_bits, 0, ctx->rs_status.as<uint64_t>(), ticket,
                                                                    rs_next_epoch(ctx), none, D, Rn));
    }

    /// LSD passes on the packed records of buckets [d0, d1) (T.start filled by the caller): ping-pongs between p0 and p1,
    /// returns true when the sorted records end up in p1.  ctx->ps_hist holds the per-bucket digit bases.
    /// skip_bits: low bits of the key field the input is already ordered on (the group field, when the MSD pass read the
    /// emission in group order): the passes start above them.
    inline bool ps_sort_buckets(ipkgpu_ctx* ctx, uint64_t* p0, uint64_t* p1, ps_table& T, uint32_t d0, uint32_t d1, uint32_t low_bits, uint32_t skip_bits = 0)
    {
        if (low_bits <= skip_bits) return false;
        uint32_t width[8];
        const uint32_t passes = ps_pass_widths(low_bits - skip_bits, width);
        const uint64_t n = uint64_t(T.start[d1]) - T.start[d0];
        if (n == 0 || passes == 0) return false;
        uint64_t pass_shift = 0, pass_bits = 0;
        for (uint32_t p = 0, at = skip_bits; p < passes; at += width[p], ++p) { pass_shift |= uint64_t(at) << (8 * p); pass_bits |= uint64_t(width[p]) << (8 * p); }
        ps_fill_tiles(T, d0, d1, PS_LSD_TILE);
        const uint32_t n_tiles = T.tile_first[d1];
        ps_reserve_status(ctx, n_tiles);
        const size_t hist_words = size_t(PS_BUCKETS) * passes * RS_RADIX;
        ctx->reserve(ctx->ps_hist, hist_words * 4 + 64);
        uint32_t* hist = ctx->ps_hist.as<uint32_t>();
        uint32_t* tickets = hist + hist_words;
        IPK_CUDA(cudaMemsetAsync(hist + size_t(d0) * passes * RS_RADIX, 0, size_t(d1 - d0) * passes * RS_RADIX * 4, ctx->stream));
        IPK_CUDA(cudaMemsetAsync(tickets, 0, 64, ctx->stream));
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, PS_HIST_THREADS * PS_HIST_ITEMS), uint64_t(ctx->num_sms) * 8));
        uint64_t chunk = div_up<uint64_t>(n, hgrid);
        chunk = div_up<uint64_t>(chunk, PS_HIST_THREADS * PS_HIST_ITEMS) * (PS_HIST_THREADS * PS_HIST_ITEMS);
        IPK_LAUNCH(ctx, "ps_seg_hist", n * 8,
                   ps_seg_hist<<<uint32_t(div_up<uint64_t>(n, chunk)), PS_HIST_THREADS, 2 * passes * RS_RADIX * 4, ctx->stream>>>(p0, passes, pass_shift, pass_bits, chunk, hist, T));
        IPK_LAUNCH(ctx, "ps_seg_scan", uint64_t(d1 - d0) * passes * 2048,
                   ps_seg_scan<<<dim3(passes, d1 - d0), RS_RADIX, 0, ctx->stream>>>(hist, passes, T));
        ps_dest none;
        memset(&none, 0, sizeof none);
        ps_runs no_runs;
        memset(&no_runs, 0, sizeof no_runs);
        bool in1 = false;
        for (uint32_t p = 0; p < passes; ++p)
        {
            // digits narrower than 4 bits run as 4-bit digits: the bits above the key field of a packed record are zero
            const uint32_t bits = std::max<uint32_t>(width[p], 4);
            void (*kern)(const uint64_t*, const uint32_t*, uint64_t*, const uint32_t*, uint64_t, uint32_t, uint32_t, uint64_t*, uint32_t*, uint32_t, const ps_table, const ps_dest,
                         const ps_runs) = nullptr;
            size_t smem = 0;
            switch (bits)
            {
            case 4: kern = ps_sweep<PS_THREADS, PS_LSD_ITEMS, false, PS_LSD_MINB, false, 4>; smem = ps_smem_bytes<PS_THREADS, PS_LSD_ITEMS, 16>(false); break;
            case 5: kern = ps_sweep<PS_THREADS, PS_LSD_ITEMS, false, PS_LSD_MINB, false, 5>; smem = ps_smem_bytes<PS_THREADS, PS_LSD_ITEMS, 32>(false); break;
            case 6: kern = ps_sweep<PS_THREADS, PS_LSD_ITEMS, false, PS_LSD_MINB, false, 6>; smem = ps_smem_bytes<PS_THREADS, PS_LSD_ITEMS, 64>(false); break;
            case 7: kern = ps_sweep<PS_THREADS, PS_LSD_ITEMS, false, PS_LSD_MINB, false, 7>; smem = ps_smem_bytes<PS_THREADS, PS_LSD_ITEMS, 128>(false); break;
            default: kern = ps_sweep<PS_THREADS, PS_LSD_ITEMS, false, PS_LSD_MINB, false, 8>; smem = ps_smem_bytes<PS_THREADS, PS_LSD_ITEMS, 256>(false); break;
            }
            IPK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            IPK_LAUNCH(ctx, "ps_sweep", n * 16 + uint64_t(n_tiles) * (1u << bits) * 8,
                       kern<<<n_tiles, PS_THREADS, smem, ctx->stream>>>(in1 ? p1 : p0, nullptr, in1 ? p0 : p1, hist + size_t(p) * RS_RADIX, n,
                                                                        32 + uint32_t((pass_shift >> (8 * p)) & 0xff), passes * RS_RADIX, ctx->rs_status.as<uint64_t>(),
                                                                        tickets + p, rs_next_epoch(ctx), T, none, no_runs));
            in1 = !in1;
        }
        return in1;
    }
}
