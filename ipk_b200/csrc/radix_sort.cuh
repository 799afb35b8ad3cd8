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
    // radix sort ("one sweep" per pass): the digit histograms of ALL passes come from one read of the keys; every
    // pass is then a single kernel that ranks a tile in shared memory, obtains the tile's global digit offsets by a
    // decoupled look-back over the preceding tiles' status words, and scatters.  Traffic per pass: 12 B read + 12 B
    // written per pair (+ one 8-byte status word per tile and digit).
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

    /// owner rank of a k-mer in the multi-GPU exchange: the high word of mix64(kmer) scaled to [0, n_parts) by a
    /// multiply-shift (a 64-bit `%` by a run-time divisor cost the r01g partition pass 3x its memory time)
    __host__ __device__ __forceinline__ uint32_t kmer_owner(uint64_t kmer, uint32_t n_parts)
    {
        return uint32_t(((mix64(kmer) >> 32) * uint64_t(n_parts)) >> 32);
    }

    /// "digit" = owner rank (n_parts <= 256).  The same kernels then bucket the records by owner, stably.
    struct digit_owner
    {
        uint32_t group_bits, n_parts;
        __device__ __forceinline__ uint32_t operator()(uint64_t ckey) const { return kmer_owner(ckey >> group_bits, n_parts); }
    };

    /// destinations of the fused partition + exchange: bucket d is written straight into rank d's receive buffers
    /// (peer device memory mapped through CUDA IPC, or this rank's own) starting at the element offset digit_base[d]
    constexpr int RS_MAX_PEERS = 16;
    struct rs_peer_table
    {
        uint64_t* keys[RS_MAX_PEERS];
        uint32_t* vals[RS_MAX_PEERS];
    };

    constexpr int RS_MAX_PASSES = 8;
    constexpr int RS_HIST_THREADS = 256;
    constexpr int RS_HIST_ITEMS = 16;

    /// hist[p * 256 + d] += number of keys whose p-th digit (bits [bit_begin + 8p, +8) under the pass's mask) is d.
    /// One read of the keys serves every pass.  Per-warp replicated shared histograms keep same-address atomics apart.
    __global__ void __launch_bounds__(RS_HIST_THREADS) rs_hist_bits(const uint64_t* __restrict__ keys, uint64_t n, uint32_t bit_begin,
                                                                    uint32_t n_bits, uint32_t n_passes, uint32_t* __restrict__ hist)
    {
        extern __shared__ uint32_t sh[];                 // [2 copies][n_passes][256]
        const uint32_t per_copy = n_passes * RS_RADIX;
        for (uint32_t i = threadIdx.x; i < 2 * per_copy; i += RS_HIST_THREADS) sh[i] = 0;
        __syncthreads();
        uint32_t* mine = sh + ((threadIdx.x >> 5) & 1) * per_copy;
        const uint64_t chunk = uint64_t(RS_HIST_THREADS) * RS_HIST_ITEMS;
        for (uint64_t base = blockIdx.x * chunk; base < n; base += uint64_t(gridDim.x) * chunk)
        {
            uint64_t k[RS_HIST_ITEMS];
#pragma unroll
            for (int i = 0; i < RS_HIST_ITEMS; ++i)
            {
                const uint64_t idx = base + uint64_t(i) * RS_HIST_THREADS + threadIdx.x;
                k[i] = idx < n ? keys[idx] : 0;
            }
#pragma unroll
            for (int i = 0; i < RS_HIST_ITEMS; ++i)
            {
                const uint64_t idx = base + uint64_t(i) * RS_HIST_THREADS + threadIdx.x;
                if (idx < n)
                {
                    const uint64_t kk = k[i] >> bit_begin;
                    for (uint32_t p = 0; p < n_passes; ++p)
                    {
                        const uint32_t width = min(8u, n_bits - 8 * p);
                        atomicAdd(&mine[p * RS_RADIX + (uint32_t(kk >> (8 * p)) & ((1u << width) - 1))], 1u);
                    }
                }
            }
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < per_copy; i += RS_HIST_THREADS)
        {
            const uint32_t v = sh[i] + sh[per_copy + i];
            if (v) atomicAdd(&hist[i], v);
        }
    }

    /// single-pass histogram under an arbitrary digit function (the owner partition)
    template<class DigitFn>
    __global__ void __launch_bounds__(RS_HIST_THREADS) rs_hist_fn(const uint64_t* __restrict__ keys, uint64_t n, DigitFn digit, uint32_t* __restrict__ hist)
    {
        __shared__ uint32_t sh[RS_RADIX];
        sh[threadIdx.x] = 0;
        __syncthreads();
        for (uint64_t idx = blockIdx.x * uint64_t(RS_HIST_THREADS) + threadIdx.x; idx < n; idx += uint64_t(gridDim.x) * RS_HIST_THREADS)
            atomicAdd(&sh[digit(keys[idx])], 1u);
        __syncthreads();
        if (sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
    }

    /// in place, one CTA per pass: hist[p][0..256) -> exclusive scan (the global start of every digit in that pass)
    __global__ void __launch_bounds__(RS_RADIX) rs_scan_hist(uint32_t* __restrict__ hist)
    {
        __shared__ uint32_t s_warp[9];
        uint32_t* h = hist + blockIdx.x * RS_RADIX;
        const uint32_t v = h[threadIdx.x];
        h[threadIdx.x] = block_excl_scan_256(v, s_warp, nullptr);
    }

    // status word of (tile, digit): epoch << 34 | flag << 32 | value.  The epoch changes with every pass, so words left by
    // earlier passes read as "not ready" and the table is zeroed only when it is (re)allocated.
    constexpr uint64_t RS_FLAG_LOCAL = 1, RS_FLAG_INCL = 2;
    __device__ __forceinline__ uint64_t rs_status(uint32_t epoch, uint64_t flag, uint32_t value) { return uint64_t(epoch) << 34 | flag << 32 | value; }

    /// shared memory of rs_onesweep<THREADS, ITEMS>: wh u32[WARPS][256], dbase u32[256], goff u32[256], misc u32[16], sk u64[TILE]
    template<int THREADS, int ITEMS> constexpr size_t rs_smem_bytes() { return size_t(THREADS / 32) * RS_RADIX * 4 + 2 * RS_RADIX * 4 + 64 + size_t(THREADS) * ITEMS * 8; }

    /// One pass over one tile of THREADS * ITEMS pairs (tile ids are handed out by an atomic ticket, so every predecessor a
    /// tile waits for is already running).  digit_base = exclusive scan of this pass's global digit histogram.
    template<int THREADS, int ITEMS, class DigitFn, bool PEERS, int MINB = 1>
    __global__ void __launch_bounds__(THREADS, MINB) rs_onesweep(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                           uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                           const uint32_t* __restrict__ digit_base, uint64_t n,
                                                           uint64_t* status, uint32_t* ticket, uint32_t epoch, DigitFn digit,
                                                           const __grid_constant__ rs_peer_table PT)
    {
        __shared__ uint64_t* s_pk[PEERS ? RS_MAX_PEERS : 1];
        __shared__ uint32_t* s_pv[PEERS ? RS_MAX_PEERS : 1];
        if (PEERS && threadIdx.x < RS_MAX_PEERS) { s_pk[threadIdx.x] = PT.keys[threadIdx.x]; s_pv[threadIdx.x] = PT.vals[threadIdx.x]; }
        constexpr int WARPS = THREADS / 32, TILE = THREADS * ITEMS;
        static_assert(THREADS >= RS_RADIX && THREADS % 32 == 0, "one thread per digit is needed");
        extern __shared__ __align__(16) unsigned char rs_smem[];
        uint64_t* sk = reinterpret_cast<uint64_t*>(rs_smem);                        // [TILE]; the values reuse it once the keys are out
        uint32_t* sv = reinterpret_cast<uint32_t*>(sk);
        uint32_t (*wh)[RS_RADIX] = reinterpret_cast<uint32_t (*)[RS_RADIX]>(sk + TILE);   // per-warp digit counts, then exclusive offsets over warps
        uint32_t* dbase = &wh[0][0] + WARPS * RS_RADIX;                             // tile-local position of the first key with digit d
        uint32_t* goff = dbase + RS_RADIX;                                          // global position of the tile's first key with digit d, minus dbase
        uint32_t* s_misc = goff + RS_RADIX;                                         // [0..8] warp sums of the digit scan, [15] tile id

        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const unsigned full = 0xffffffffu;
        if (tid == 0) s_misc[15] = atomicAdd(ticket, 1u);
        for (int i = tid; i < WARPS * RS_RADIX; i += THREADS) (&wh[0][0])[i] = 0;
        __syncthreads();
        const uint32_t tile = s_misc[15];
        const uint64_t tile_base = uint64_t(tile) * TILE;
        const uint32_t tile_n = uint32_t(min(uint64_t(TILE), n - tile_base));

        uint64_t key[ITEMS];
        uint32_t val[ITEMS];
        uint16_t rank[ITEMS];
        // warp w owns items [w*32*ITEMS, (w+1)*32*ITEMS) of the tile, 32 consecutive items per round: order = (warp, round, lane)
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            key[i] = li < tile_n ? keys_in[tile_base + li] : ~0ull;
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            val[i] = li < tile_n ? vals_in[tile_base + li] : 0u;
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            const bool valid = li < tile_n;
            const uint32_t d = valid ? digit(key[i]) : RS_RADIX;   // invalid lanes form their own group
            // (eight ballots, one per digit bit, instead of match.any measured 6 % SLOWER per pass on cfg3 -- r01f)
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
        // digit d = tid (< 256): exclusive prefix over warps, the tile's count, its status word and the look-back
        const bool is_digit = tid < RS_RADIX;
        uint32_t cnt = 0;
        if (is_digit)
        {
#pragma unroll
            for (int w = 0; w < WARPS; ++w)
            {
                const uint32_t t = wh[w][tid];
                wh[w][tid] = cnt;
                cnt += t;
            }
        }
        uint64_t* my_status = status + uint64_t(tile) * RS_RADIX + (is_digit ? tid : 0);
        if (is_digit)
            *reinterpret_cast<volatile uint64_t*>(my_status) = rs_status(epoch, tile == 0 ? RS_FLAG_INCL : RS_FLAG_LOCAL, cnt);
        // exclusive scan of cnt over the 256 digits (threads >= 256 only take part in the barriers)
        const uint32_t inc = warp_incl_scan(cnt, lane);
        if (is_digit && lane == 31) s_misc[warp] = inc;
        __syncthreads();
        if (warp == 0)
        {
            const uint32_t w = lane < 8 ? s_misc[lane] : 0;
            const uint32_t winc = warp_incl_scan(w, lane);
            if (lane < 8) s_misc[lane] = winc - w;
        }
        __syncthreads();
        if (is_digit)
        {
            const uint32_t ex_local = s_misc[warp] + inc - cnt;
            uint32_t before = 0;
            if (tile > 0)
            {
                const volatile uint64_t* p = status + uint64_t(tile - 1) * RS_RADIX + tid;
                while (true)
                {
                    const uint64_t w = *p;
                    if (uint32_t(w >> 34) != epoch || ((w >> 32) & 3) == 0) continue;      // not published yet in this pass
                    before += uint32_t(w);
                    if (((w >> 32) & 3) == RS_FLAG_INCL) break;
                    p -= RS_RADIX;
                }
                *reinterpret_cast<volatile uint64_t*>(my_status) = rs_status(epoch, RS_FLAG_INCL, before + cnt);
            }
            dbase[tid] = ex_local;
            goff[tid] = digit_base[tid] + before - ex_local;
        }
        __syncthreads();
        // keys: tile-local order -> shared memory -> runs of equal digits to global memory
        uint32_t pos[ITEMS];
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            pos[i] = 0;
            if (li < tile_n)
            {
                const uint32_t d = digit(key[i]);
                pos[i] = dbase[d] + wh[warp][d] + rank[i];
                sk[pos[i]] = key[i];
            }
        }
        __syncthreads();
        uint32_t gpos[ITEMS];
        uint8_t gdst[PEERS ? ITEMS : 1];
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t q = uint32_t(i) * THREADS + tid;
            gpos[i] = 0;
            if (PEERS) gdst[i] = 0;
            if (q < tile_n)
            {
                const uint64_t kk = sk[q];
                const uint32_t d = digit(kk);
                gpos[i] = goff[d] + q;
                if (PEERS) { gdst[i] = uint8_t(d); s_pk[d][gpos[i]] = kk; }          // a store over NVLink when d is a peer
                else keys_out[gpos[i]] = kk;
            }
        }
        // values: same route through the (reused) shared buffer
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t li = warp * (32 * ITEMS) + i * 32 + lane;
            if (li < tile_n) sv[pos[i]] = val[i];
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
        {
            const uint32_t q = uint32_t(i) * THREADS + tid;
            if (q < tile_n)
            {
                if (PEERS) s_pv[gdst[i]][gpos[i]] = sv[q];
                else vals_out[gpos[i]] = sv[q];
            }
        }
    }

    /// tile shapes the sweep kernel is instantiated for; ctx->rs_shape picks one (tuning knob; 2 = default).  r01r-t sweeps on
    /// cfg3 (ms of sort per step): 256x12 77.9, 256x16 81.5, 384x16 63.0, 512x12 91.7, 512x16 86.2, 384x20 92.4; forcing 3 CTAs/SM
    /// (384x12 at 56 registers) 66.0; the 384x16 shape must keep 2 CTAs/SM (80 registers): __launch_bounds__(384, 2)
    struct rs_shape { int threads, items; };
    constexpr rs_shape RS_SHAPES[] = { { 256, 12 }, { 256, 16 }, { 384, 16 }, { 512, 12 }, { 512, 16 }, { 384, 20 } };
    constexpr int RS_N_SHAPES = int(sizeof(RS_SHAPES) / sizeof(RS_SHAPES[0]));
    inline rs_shape rs_current_shape(const ipkgpu_ctx* ctx) { return RS_SHAPES[ctx->rs_shape >= 0 && ctx->rs_shape < RS_N_SHAPES ? ctx->rs_shape : 0]; }
    inline uint32_t rs_tile(const ipkgpu_ctx* ctx) { const rs_shape s = rs_current_shape(ctx); return uint32_t(s.threads * s.items); }

    /// scratch of the sort inside `table`: [0, 8*256) u32 digit histograms / bases of up to 8 passes, then 16 u32 tickets.
    /// The status words live in ctx->rs_status (zeroed when it grows; epochs keep passes apart afterwards).
    struct rs_scratch
    {
        uint32_t* hist;
        uint32_t* tickets;
        uint64_t* status;
    };

    inline rs_scratch rs_prepare(ipkgpu_ctx* ctx, uint64_t n, dbuf& table)
    {
        const uint64_t n_tiles = div_up<uint64_t>(n, std::min<uint32_t>(rs_tile(ctx), 3072u));      // enough for every tile shape
        ctx->reserve(table, (RS_MAX_PASSES * RS_RADIX + 16) * 4);
        const size_t want = n_tiles * RS_RADIX * 8;
        if (want > ctx->rs_status.cap)
        {
            ctx->reserve(ctx->rs_status, want);
            IPK_CUDA(cudaMemsetAsync(ctx->rs_status.p, 0, ctx->rs_status.cap, ctx->stream));
            ctx->rs_epoch = 0;
        }
        IPK_CUDA(cudaMemsetAsync(table.p, 0, (RS_MAX_PASSES * RS_RADIX + 16) * 4, ctx->stream));
        return { table.as<uint32_t>(), table.as<uint32_t>() + RS_MAX_PASSES * RS_RADIX, ctx->rs_status.as<uint64_t>() };
    }

    inline uint32_t rs_next_epoch(ipkgpu_ctx* ctx)
    {
        if (++ctx->rs_epoch >= (1u << 30))
        {
            IPK_CUDA(cudaMemsetAsync(ctx->rs_status.p, 0, ctx->rs_status.cap, ctx->stream));
            ctx->rs_epoch = 1;
        }
        return ctx->rs_epoch;
    }

    template<int THREADS, int ITEMS, class DigitFn, int MINB = 1>
    inline void rs_launch_shape(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout,
                                const uint32_t* digit_base, uint64_t n, uint64_t* status, uint32_t* ticket, DigitFn digit,
                                const rs_peer_table* peers = nullptr)
    {
        constexpr size_t smem = rs_smem_bytes<THREADS, ITEMS>();
        const uint32_t n_tiles = uint32_t(div_up<uint64_t>(n, uint64_t(THREADS) * ITEMS));
        rs_peer_table none;
        memset(&none, 0, sizeof none);
        if (peers)
        {
            IPK_CUDA(cudaFuncSetAttribute(rs_onesweep<THREADS, ITEMS, DigitFn, true, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            IPK_LAUNCH(ctx, "rs_scatter_peers", n * 24 + uint64_t(n_tiles) * RS_RADIX * 8,
                       rs_onesweep<THREADS, ITEMS, DigitFn, true, MINB><<<n_tiles, THREADS, smem, ctx->stream>>>(kin, vin, nullptr, nullptr, digit_base, n, status,
                                                                                                           ticket, rs_next_epoch(ctx), digit, *peers));
            return;
        }
        IPK_CUDA(cudaFuncSetAttribute(rs_onesweep<THREADS, ITEMS, DigitFn, false, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        IPK_LAUNCH(ctx, "rs_onesweep", n * 24 + uint64_t(n_tiles) * RS_RADIX * 8,
                   rs_onesweep<THREADS, ITEMS, DigitFn, false, MINB><<<n_tiles, THREADS, smem, ctx->stream>>>(kin, vin, kout, vout, digit_base, n, status, ticket,
                                                                                                        rs_next_epoch(ctx), digit, none));
    }

    template<class DigitFn>
    inline void rs_launch_onesweep(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout,
                                   const uint32_t* digit_base, uint64_t n, uint64_t* status, uint32_t* ticket, DigitFn digit)
    {
        switch (ctx->rs_shape)
        {
        case 1: rs_launch_shape<256, 16>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        case 2: rs_launch_shape<384, 16, DigitFn, 2>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        case 3: rs_launch_shape<512, 12>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        case 4: rs_launch_shape<512, 16>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        case 5: rs_launch_shape<384, 20>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        default: rs_launch_shape<256, 12>(ctx, kin, vin, kout, vout, digit_base, n, status, ticket, digit); break;
        }
    }

    /// One stable bucketing pass (histogram -> scan -> one sweep) under an arbitrary digit function.  Leaves the
    /// exclusive digit starts in table[0..256) (device) for the caller.
    template<class DigitFn>
    inline void radix_pass(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t* kout, uint32_t* vout, uint64_t n,
                           DigitFn digit, dbuf& table, dbuf& scan_tmp)
    {
        (void)scan_tmp;
        if (n == 0) return;
        const rs_scratch S = rs_prepare(ctx, n, table);
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, RS_HIST_THREADS * 8), uint64_t(ctx->num_sms) * 8));
        IPK_LAUNCH(ctx, "rs_hist", n * 8, rs_hist_fn<DigitFn><<<hgrid, RS_HIST_THREADS, 0, ctx->stream>>>(kin, n, digit, S.hist));
        IPK_LAUNCH(ctx, "rs_scan_hist", 2048, rs_scan_hist<<<1, RS_RADIX, 0, ctx->stream>>>(S.hist));
        rs_launch_onesweep(ctx, kin, vin, kout, vout, S.hist, n, S.status, S.tickets, digit);
    }

    /// Owner histogram of n records: table[0..n_parts) (device) = number of records per owner rank.
    inline void owner_histogram(ipkgpu_ctx* ctx, const uint64_t* kin, uint64_t n, digit_owner own, dbuf& table)
    {
        const rs_scratch S = rs_prepare(ctx, std::max<uint64_t>(n, 1), table);
        if (n == 0) return;
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, RS_HIST_THREADS * 8), uint64_t(ctx->num_sms) * 8));
        IPK_LAUNCH(ctx, "rs_hist", n * 8, rs_hist_fn<digit_owner><<<hgrid, RS_HIST_THREADS, 0, ctx->stream>>>(kin, n, own, S.hist));
    }

    /// The fused partition + exchange pass: bucket d of the n records goes to peers.keys[d] / peers.vals[d] starting at
    /// element d_base[d] (device table, 256 words; the caller computed it from the all-gathered owner histograms), in
    /// stable order.  Remote buckets are plain stores into peer memory over NVLink -- no staging buffer, no collective.
    inline void scatter_to_peers(ipkgpu_ctx* ctx, const uint64_t* kin, const uint32_t* vin, uint64_t n, digit_owner own,
                                 const uint32_t* d_base, const rs_peer_table& peers, dbuf& table)
    {
        if (n == 0) return;
        const rs_scratch S = rs_prepare(ctx, n, table);
        rs_launch_shape<384, 16, digit_owner, 2>(ctx, kin, vin, nullptr, nullptr, d_base, n, S.status, S.tickets, own, &peers);
    }

    /// One most-significant-digit pass: (ka, va) -> (kb, vb) by the 8 key bits from bit_begin, stably.  start[0..256) (host) =
    /// where each digit's range begins in the output; the caller sorts the ranges on the bits below independently.
    inline void radix_partition_top(ipkgpu_ctx* ctx, const uint64_t* ka, const uint32_t* va, uint64_t* kb, uint32_t* vb, uint64_t n,
                                    uint32_t bit_begin, dbuf& table, uint32_t* start)
    {
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "radix partition of %llu records exceeds 2^32-1", (unsigned long long)n);
        const rs_scratch S = rs_prepare(ctx, n, table);
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, RS_HIST_THREADS * RS_HIST_ITEMS), uint64_t(ctx->num_sms) * 8));
        IPK_LAUNCH(ctx, "rs_hist", n * 8, rs_hist_bits<<<hgrid, RS_HIST_THREADS, 2 * RS_RADIX * 4, ctx->stream>>>(ka, n, bit_begin, 8, 1, S.hist));
        IPK_LAUNCH(ctx, "rs_scan_hist", 2048, rs_scan_hist<<<1, RS_RADIX, 0, ctx->stream>>>(S.hist));
        IPK_CUDA(cudaMemcpyAsync(start, S.hist, RS_RADIX * 4, cudaMemcpyDeviceToHost, ctx->stream));
        rs_launch_onesweep(ctx, ka, va, kb, vb, S.hist, n, S.status, S.tickets, digit_bits{ bit_begin, 0xffu });
        IPK_CUDA(cudaStreamSynchronize(ctx->stream));
    }

    /// Sorts n pairs on key bits [bit_begin, bit_begin + n_bits), stably.  Ping-pongs between (ka,va) and (kb,vb);
    /// returns true if the sorted result is in (kb,vb).  n < 2^32.
    inline bool radix_sort_pairs(ipkgpu_ctx* ctx, uint64_t* ka, uint32_t* va, uint64_t* kb, uint32_t* vb, uint64_t n,
                                 uint32_t n_bits, dbuf& table, dbuf& scan_tmp, uint32_t* passes_out = nullptr,
                                 uint32_t bit_begin = 0)
    {
        (void)scan_tmp;
        if (n > 0xffffffffull) fail(IPKGPU_EINVAL, "radix sort of %llu records exceeds 2^32-1", (unsigned long long)n);
        const uint32_t passes = n_bits ? div_up<uint32_t>(n_bits, 8) : 0;
        if (passes > uint32_t(RS_MAX_PASSES)) fail(IPKGPU_EINVAL, "radix sort on %u bits needs more than %d passes", n_bits, RS_MAX_PASSES);
        if (passes_out) *passes_out = passes;
        if (n == 0 || passes == 0) return false;
        const rs_scratch S = rs_prepare(ctx, n, table);
        const uint32_t hgrid = uint32_t(std::min<uint64_t>(div_up<uint64_t>(n, RS_HIST_THREADS * RS_HIST_ITEMS), uint64_t(ctx->num_sms) * 8));
        IPK_LAUNCH(ctx, "rs_hist", n * 8,
                   rs_hist_bits<<<hgrid, RS_HIST_THREADS, 2 * passes * RS_RADIX * 4, ctx->stream>>>(ka, n, bit_begin, n_bits, passes, S.hist));
        IPK_LAUNCH(ctx, "rs_scan_hist", uint64_t(passes) * 2048, rs_scan_hist<<<passes, RS_RADIX, 0, ctx->stream>>>(S.hist));
        bool in_b = false;
        for (uint32_t p = 0; p < passes; ++p)
        {
            const uint32_t shift = p * 8;
            const uint32_t width = std::min<uint32_t>(8, n_bits - shift);
            const uint32_t mask = (1u << width) - 1;
            rs_launch_onesweep(ctx, in_b ? kb : ka, in_b ? vb : va, in_b ? ka : kb, in_b ? va : vb, S.hist + p * RS_RADIX, n, S.status, S.tickets + p,
                               digit_bits{ bit_begin + shift, mask });
            in_b = !in_b;
        }
        return in_b;
    }
}
