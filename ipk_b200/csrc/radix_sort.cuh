// radix_sort.cuh -- stable LSD radix sort of (u64 key, u32 payload) pairs + device-wide exclusive scan, sm_100a.
//
// Replaces the two hash-map probes the reference pays per phylo-k-mer (ipk::put, branch_group.cpp:88-101, and
// phylo_kmer_db::unsafe_insert, db_builder.cpp:685-693): records are brought together by sorting on the composite
// key (kmer << group_bits | group) instead of being looked up one by one.
//
// Per 8-bit pass:  rs_histogram (per-tile digit counts, digit-major table)  ->  exclusive scan of the table
//                  ->  rs_scatter (stable in-tile ranking by warp match + staged, run-coalesced scatter).
// Only the bits that can differ are sorted: passes = ceil((key_bits + group_bits) / 8).
#pragma once
#include "common.cuh"

namespace ipkgpu
{
    constexpr int RS_THREADS = 256;
    constexpr int RS_ITEMS = 12;
    constexpr int RS_TILE = RS_THREADS * RS_ITEMS;      // 3072 pairs per CTA: 46 KB of static shared memory, 4 CTAs / SM
    constexpr int RS_WARPS = RS_THREADS / 32;
    constexpr int RS_RADIX = 256;

    // ------------------------------------------------------------------------------------------------------------
    // exclusive scan of u32 -> u32 (three-phase: tile sums, scan of tile sums, downsweep); in place allowed
    // ------------------------------------------------------------------------------------------------------------
    constexpr int SCAN_THREADS = 256;
    constexpr int SCAN_ITEMS = 16;
    constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

    __device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane)
    {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        return v;
    }

    /// Block-wide exclusive scan of one value per thread (blockDim.x == 256); returns exclusive prefix, total in *total.
    __device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* s_warp /*[9]*/, uint32_t* total)
    {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const uint32_t inc = warp_incl_scan(v, lane);
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0)
        {
            uint32_t w = lane < 8 ? s_warp[lane] : 0;
            const uint32_t winc = warp_incl_scan(w, lane);
            if (lane < 8) s_warp[lane] = winc - w;
            if (lane == 7) s_warp[8] = winc;
        }
        __syncthreads();
        const uint32_t r = s_warp[warp] + inc - v;
        if (total) *total = s_warp[8];
        __syncthreads();
        return r;
    }

    __global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums(const uint32_t* __restrict__ in, uint32_t* __restrict__ sums, uint64_t n)
    {
        __shared__ uint32_t s_warp[9];
        const uint64_t base = uint64_t(blockIdx.x) * SCAN_TILE;
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i)
        {
            const uint64_t idx = base + uint64_t(i) * SCAN_THREADS + threadIdx.x;
            if (idx < n) acc += in[idx];
        }
        uint32_t total;
        block_excl_scan_256(acc, s_warp, &total);
        if (threadIdx.x == 0) sums[blockIdx.x] = total;
    }

    /// single CTA: exclusive scan of m values in place, grand total to *total_out (may be null)
    __global__ void __launch_bounds__(SCAN_THREADS) scan_small(uint32_t* __restrict__ v, uint32_t m, unsigned long long* total_out)
    {
        __shared__ uint32_t s_warp[9];
        uint32_t carry = 0;
        for (uint32_t base = 0; base < m; base += SCAN_THREADS)
        {
            const uint32_t i = base + threadIdx.x;
            const uint32_t x = i < m ? v[i] : 0;
            uint32_t total;
            const uint32_t ex = block_excl_scan_256(x, s_warp, &total);
            if (i < m) v[i] = carry + ex;
            carry += total;
        }
        if (total_out && threadIdx.x == 0) *total_out = carry;
    }

    __global__ void __launch_bounds__(SCAN_THREADS) scan_downsweep(const uint32_t* in, uint32_t* out,     /* in == out allowed */
                                                                   const uint32_t* __restrict__ sums, uint64_t n)
    {
        __shared__ uint32_t s_warp[9];
        const uint64_t base = uint64_t(blockIdx.x) * SCAN_TILE + uint64_t(threadIdx.x) * SCAN_ITEMS;   // blocked: thread owns SCAN_ITEMS consecutive
        uint32_t x[SCAN_ITEMS];
        uint32_t acc = 0;
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i)
        {
            x[i] = (base + i < n) ? in[base + i] : 0;
            acc += x[i];
        }
        uint32_t ex = block_excl_scan_256(acc, s_warp, nullptr) + sums[blockIdx.x];
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; ++i)
        {
            if (base + i < n) out[base + i] = ex;
            ex += x[i];
        }
    }

    /// out[i] = sum_{j<i} in[j]; n up to 2^32 * SCAN_TILE / ... (tile sums are scanned by one CTA, recursively tiled)
    inline void exclusive_scan_u32(ipkgpu_ctx* ctx, const uint32_t* in, uint32_t* out, uint64_t n, dbuf& tmp,
                                   unsigned long long* d_total = nullptr)
    {
        if (n == 0)
        {
            if (d_total) IPK_CUDA(cudaMemsetAsync(d_total, 0, 8, ctx->stream));
            return;
        }
        const uint64_t tiles = div_up<uint64_t>(n, SCAN_TILE);
        if (tiles > 0x7fffffffull) fail(IPKGPU_EINVAL, "scan of %llu elements is too large", (unsigned long long)n);
        ctx->reserve(tmp, tiles * 4);
        uint32_t* sums = tmp.as<uint32_t>();
        IPK_LAUNCH(ctx, "scan_tile_sums", n * 4, scan_tile_sums<<<uint32_t(tiles), SCAN_THREADS, 0, ctx->stream>>>(in, sums, n));
        IPK_LAUNCH(ctx, "scan_small", tiles * 8, scan_small<<<1, SCAN_THREADS, 0, ctx->stream>>>(sums, uint32_t(tiles), d_total));
        IPK_LAUNCH(ctx, "scan_downsweep", n * 8, scan_downsweep<<<uint32_t(tiles), SCAN_THREADS, 0, ctx->stream>>>(in, out, sums, n));
    }

    // ------------------------------------------------------------------------------------------------------------
    // radix sort
    // ------------------------------------------------------------------------------------------------------------

    /// digit of a key in one LSD pass: bits [shift, shift+8) under mask
    struct digit_bits
    {
        uint32_t shift, mask;
        __device__ __forceinline__ uint32_t operator()(uint64_t key) const { return uint32_t(key >> shift) & mask; }
    };

    __host__ __device__ __forceinline__ uint64_t mix64(uint64_t x)
    {
        x += 0x9E3779B97F4A7C15ull;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
        x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
        return x ^ (x >> 31);
    }

    /// "digit" = owner rank of a k-mer in the multi-GPU exchange: mix64(kmer) % n_parts (n_parts <= 256).
    /// The same kernels then bucket the records by owner, stably.
    struct digit_owner
    {
        uint32_t group_bits, n_parts;
        __device__ __forceinline__ uint32_t operator()(uint64_t ckey) const { return uint32_t(mix64(ckey >> group_bits) % n_parts); }
    };

    /// table[d * n_tiles + tile] = number of keys of this tile whose digit is d
    template<class DigitFn>
    __global__ void __launch_bounds__(RS_THREADS) rs_histogram(const uint64_t* __restrict__ keys, uint32_t* __restrict__ table,
                                                               uint64_t n, uint32_t n_tiles, DigitFn digit)
    {
        __shared__ uint32_t sh[RS_RADIX];
        sh[threadIdx.x] = 0;
        __syncthreads();
        const uint64_t base = uint64_t(blockIdx.x) * RS_TILE;
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i)
        {
            const uint64_t idx = base + uint64_t(i) * RS_THREADS + threadIdx.x;
            if (idx < n) atomicAdd(&sh[digit(keys[idx])], 1u);
        }
        __syncthreads();
        table[uint64_t(threadIdx.x) * n_tiles + blockIdx.x] = sh[threadIdx.x];
    }

    /// Stable scatter of one tile.  offs = exclusive scan of the digit-major table.
    template<class DigitFn>
    __global__ void __launch_bounds__(RS_THREADS) rs_scatter(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                             uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                             const uint32_t* __restrict__ offs, uint64_t n, uint32_t n_tiles,
                                                             DigitFn digit)
    {
        __shared__ uint32_t wh[RS_WARPS][RS_RADIX];     // per-warp digit counts, then exclusive offsets over warps
        __shared__ uint32_t dbase[RS_RADIX];            // tile-local position of the first key with digit d
        __shared__ uint32_t goff[RS_RADIX];             // global position of the first key of this tile with digit d
        __shared__ uint32_t s_warp[9];
        __shared__ uint64_t sk[RS_TILE];
        __shared__ uint32_t sv[RS_TILE];

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const unsigned full = 0xffffffffu;
        for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&wh[0][0])[i] = 0;
        __syncthreads();

        const uint64_t tile_base = uint64_t(blockIdx.x) * RS_TILE;
        const uint32_t tile_n = uint32_t(min(uint64_t(RS_TILE), n - tile_base));
        uint64_t key[RS_ITEMS];
        uint32_t val[RS_ITEMS];
        uint16_t rank[RS_ITEMS];
        // warp w owns items [w*512, (w+1)*512) of the tile, 32 consecutive items per round: order = (warp, round, lane)
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * RS_ITEMS) + i * 32 + lane;
            const bool valid = li < tile_n;
            key[i] = valid ? keys_in[tile_base + li] : ~0ull;
            val[i] = valid ? vals_in[tile_base + li] : 0u;
        }
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * RS_ITEMS) + i * 32 + lane;
            const bool valid = li < tile_n;
            const uint32_t d = valid ? digit(key[i]) : RS_RADIX;   // invalid lanes form their own group
            const unsigned peers = __match_any_sync(full, d);
            const int leader = __ffs(peers) - 1;
            uint32_t old = 0;
            if (lane == leader && valid)
            {
                old = wh[warp][d];
                wh[warp][d] = old + __popc(peers);
            }
            old = __shfl_sync(full, old, leader);
            rank[i] = uint16_t(old + __popc(peers & ((1u << lane) - 1)));
            __syncwarp();
        }
        __syncthreads();
        // digit d = tid: exclusive prefix over warps, tile total, then exclusive scan over digits
        uint32_t cnt = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w)
        {
            const uint32_t t = wh[w][tid];
            wh[w][tid] = cnt;
            cnt += t;
        }
        const uint32_t ex = block_excl_scan_256(cnt, s_warp, nullptr);
        dbase[tid] = ex;
        goff[tid] = offs[uint64_t(tid) * n_tiles + blockIdx.x];
        __syncthreads();
#pragma unroll
        for (int i = 0; i < RS_ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * RS_ITEMS) + i * 32 + lane;
            if (li < tile_n)
            {
                const uint32_t d = digit(key[i]);
                const uint32_t pos = dbase[d] + wh[warp][d] + rank[i];
                sk[pos] = key[i];
                sv[pos] = val[i];
            }
        }
        __syncthreads();
        for (uint32_t i = tid; i < tile_n; i += RS_THREADS)
        {
            const uint64_t kk = sk[i];
            const uint32_t d = digit(kk);
            const uint32_t g = goff[d] + (i - dbase[d]);
            keys_out[g] = kk;
            vals_out[g] = sv[i];
        }
    }

    /// One stable bucketing pass (histogram -> scan -> scatter) under an arbitrary digit function.
    template<class DigitFn>
    inline void radix_pass(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout, uint64_t n,
                           DigitFn digit, dbuf& table, dbuf& scan_tmp)
    {
        const uint32_t n_tiles = uint32_t(div_up<uint64_t>(n, RS_TILE));
        const uint64_t table_n = uint64_t(n_tiles) * RS_RADIX;
        ctx->reserve(table, table_n * 4);
        uint32_t* tab = table.as<uint32_t>();
        IPK_LAUNCH(ctx, "rs_histogram", n * 8 + table_n * 4, rs_histogram<DigitFn><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, tab, n, n_tiles, digit));
        exclusive_scan_u32(ctx, tab, tab, table_n, scan_tmp);
        IPK_LAUNCH(ctx, "rs_scatter", n * 24 + table_n * 4, rs_scatter<DigitFn><<<n_tiles, RS_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, tab, n, n_tiles, digit));
    }

    /// Sorts n pairs on key bits [bit_begin, bit_begin + n_bits), stably.  Ping-pongs between (ka,va) and (kb,vb);
    /// returns true if the sorted result is in (kb,vb).  n < 2^32.
    inline bool radix_sort_pairs(ipkgpu_ctx* ctx, uint64_t* ka, uint32_t* va, uint64_t* kb, uint32_t* vb, uint64_t n,
                                 uint32_t n_bits, dbuf& table, dbuf& scan_tmp, uint32_t* passes_out = nullptr,
                                 uint32_t bit_begin = 0)
    {
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "radix sort of %llu records exceeds 2^32-1", (unsigned long long)n);
        const uint32_t passes = n_bits ? div_up<uint32_t>(n_bits, 8) : 0;
        if (passes_out) *passes_out = passes;
        if (n == 0 || passes == 0) return false;
        bool in_b = false;
        for (uint32_t p = 0; p < passes; ++p)
        {
            const uint32_t shift = p * 8;
            const uint32_t width = std::min<uint32_t>(8, n_bits - shift);
            const uint32_t mask = (1u << width) - 1;
            const uint64_t* kin = in_b ? kb : ka;
            const uint32_t* vin = in_b ? vb : va;
            uint64_t* kout = in_b ? ka : kb;
            uint32_t* vout = in_b ? va : vb;
            radix_pass(ctx, kin, vin, kout, vout, n, digit_bits{ bit_begin + shift, mask }, table, scan_tmp);
            in_b = !in_b;
        }
        return in_b;
    }
}
