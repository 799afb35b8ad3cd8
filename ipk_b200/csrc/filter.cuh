// filter.cuh -- filter values, ordering by filter value and the final gather, sm_100a.
//
// Replaces   mif0_filter::calc_filter_values   ipk/src/filter.cpp:60-119   (+ logscore_to_score :20-23, shannon :55-58)
//            random_filter::calc_filter_values ipk/src/filter.cpp:133-146
//            std::sort(kmer_order) + the db.at(kmer) walk of filter_in_ram   ipk/src/db_builder.cpp:284,323-332
#pragma once
#include "common.cuh"
#include "mif_math.cuh"

namespace ipkgpu
{
    /// filter.cpp:20-23: std::min(std::pow(10, log_score), 1.0) evaluated in double, returned as score_type (float).
    /// mif_exp10 (mif_math.cuh; relative error < 1e-14) stands in for pow(10, x): after the rounding to float the two agree.
    __device__ __forceinline__ float logscore_to_score(float log_score)
    {
        const double p = (log_score > -300.0f && log_score <= 0.0f) ? mif_exp10(log_score) : exp10(double(log_score));
        return float(p < 1.0 ? p : 1.0);
    }
    __device__ __forceinline__ double shfl_xor_f64(unsigned mask, double v, int lane_mask)
    {
        int lo = __double2loint(v), hi = __double2hiint(v);
        lo = __shfl_xor_sync(mask, lo, lane_mask);
        hi = __shfl_xor_sync(mask, hi, lane_mask);
        return __hiloint2double(hi, lo);
    }

    // ----------------------------------------------------------------------------------------------------------
    // K4 MIF0.  filter.cpp:60-119 per k-mer with entries x_1..x_n (x = logscore_to_score), T groups, threshold t:
    //     ss = sum x_i + (T - n) t;   H = T h(t/ss) + sum (h(x_i/ss) - h(t/ss)),  h(w) = -w log2 w;   fv = ss (H - log2 T)
    // Multiplied out, every division by ss cancels:
    //     fv = ss (log2 ss - log2 T)  -  sum x_i log2 x_i  -  (T - n) t log2 t
    // so an entry contributes two numbers that depend on its score alone, x_i and x_i log2 x_i, and a k-mer needs ONE pass
    // over its entries (the reference's two: the sum, then the weighted terms) and no division.  Same value up to fp64
    // rounding (~1e-15 relative); the rounding to fp32 differs from the reference's in ~1e-7 of the k-mers, as any other
    // libm would (tests/parity.py: FV_RTOL).
    // A CTA owns the k-mers whose first entry lies in its chunk of MIF_CHUNK entries, i.e. a contiguous run of entries.
    // The transcendental work is done one ENTRY per thread into shared memory; the two fp64 sums are then taken one K-MER
    // per thread in stored order, or -- from MIF_LONG entries on -- by a warp in a fixed lane order (lane l adds entries
    // l, l+32, ..., then the xor butterfly): a fixed association per k-mer, whichever path, chunk or GPU it falls into.
    // ----------------------------------------------------------------------------------------------------------
    constexpr int MIF_THREADS = 256;
    constexpr int MIF_CHUNK = 1024;          // entries per CTA chunk (k-mers are assigned by their first entry)
    constexpr int MIF_LONG = 16;             // k-mers with at least this many entries are summed by 8 lanes (fixed lane order) ...
    constexpr int MIF_WIDE = 256;            // ... and from this many on by a whole warp
    constexpr int MIF_CAP = 2048;            // entries a CTA can stage; longer runs take the streaming path
    constexpr size_t MIF_SMEM = size_t(MIF_CAP) * 8 + size_t(MIF_CAP) * 4 + size_t(MIF_CHUNK + 2) * 4;

    /// chunk_first[b] = first k-mer whose entries start at or after b * MIF_CHUNK  (b = 0 .. n_chunks; last = n_kmers)
    __global__ void k4_chunk_first(const uint64_t* __restrict__ offs, uint64_t n_kmers, uint64_t n_chunks, uint32_t* __restrict__ chunk_first)
    {
        const uint64_t b = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (b > n_chunks) return;
        if (b == n_chunks) { chunk_first[b] = uint32_t(n_kmers); return; }
        const uint64_t target = b * MIF_CHUNK;
        uint64_t lo = 0, hi = n_kmers;           // first q in [0, n_kmers] with offs[q] >= target
        while (lo < hi)
        {
            const uint64_t mid = (lo + hi) >> 1;
            if (offs[mid] >= target) hi = mid; else lo = mid + 1;
        }
        chunk_first[b] = uint32_t(lo);
    }

    struct mif0_consts
    {
        double threshold, Hc, T, tlog;       // tlog = threshold * log2(threshold)
    };

    /// x log2 x of an entry's score x in [0, 1] (a float widened to double)
    __device__ __forceinline__ double mif_xlog2x(double x, const mif_log2_entry* tab)
    {
        return x >= 2.3e-308 ? x * mif_log2(x, tab) : x * log2(x);      // x == 0: 0 * -inf = NaN, as the reference's shannon(0)
    }
    /// the filter value from the two sums of a k-mer with n entries
    __device__ __forceinline__ float mif0_value(double S, double A, uint64_t n, const mif0_consts& C, const mif_log2_entry* tab)
    {
        const double rest = C.T - double(n);
        const double ss = __dadd_rn(S, __dmul_rn(rest, C.threshold));                                       // filter.cpp:86
        const double l = (ss >= 2.3e-308 && ss < 1.0e300) ? mif_log2(ss, tab) : log2(ss);
        return float(__dsub_rn(__dsub_rn(__dmul_rn(ss, __dsub_rn(l, C.Hc)), A), __dmul_rn(rest, C.tlog)));
    }

    /// the constants of a build, evaluated once on the host (the two log2 calls per THREAD were 17 % of the kernel's instructions, ncu r03p)
    inline mif0_consts mif0_make_consts(float threshold, uint32_t total_num_groups)
    {
        const double thr = double(threshold);
        return mif0_consts{ thr, std::log2(double(total_num_groups)), double(total_num_groups), thr * std::log2(thr) };
    }

    __global__ void __launch_bounds__(MIF_THREADS) k4_mif0(const uint64_t* __restrict__ offs, const float* __restrict__ scores,
                                                           const uint32_t* __restrict__ chunk_first, const __grid_constant__ mif0_consts C,
                                                           float* __restrict__ fv, const mif_log2_entry* __restrict__ log2_tab)
    {
        __shared__ mif_log2_entry s_tab[MIF_LOG2_BUCKETS];
        extern __shared__ __align__(16) unsigned char mif_smem[];
        double* st = reinterpret_cast<double*>(mif_smem);                       // [MIF_CAP] x log2 x per entry
        float* sx = reinterpret_cast<float*>(st + MIF_CAP);                     // [MIF_CAP] x per entry
        uint32_t* koff = reinterpret_cast<uint32_t*>(sx + MIF_CAP);             // [MIF_CHUNK + 2] per k-mer: first entry, relative to the run
        const uint32_t q0 = chunk_first[blockIdx.x], q1 = chunk_first[blockIdx.x + 1];
        if (q0 == q1) return;
        const int tid = threadIdx.x;
        if (tid < MIF_LOG2_BUCKETS) s_tab[tid] = log2_tab[tid];
        const uint32_t nk = q1 - q0;
        const uint64_t e0 = offs[q0], e1 = offs[q1];
        const uint64_t span = e1 - e0;
        const int lane = tid & 31, warp = tid >> 5;
        constexpr int MIF_WARPS = MIF_THREADS / 32;
        const unsigned full = 0xffffffffu;
        __syncthreads();                                           // s_tab
        if (span <= uint64_t(MIF_CAP))
        {
            for (uint32_t j = tid; j <= nk; j += MIF_THREADS) koff[j] = uint32_t(offs[q0 + j] - e0);
            for (uint32_t i = tid; i < uint32_t(span); i += MIF_THREADS)
            {
                const float x = logscore_to_score(scores[e0 + i]);
                sx[i] = x;
                st[i] = mif_xlog2x(double(x), s_tab);
            }
            __syncthreads();
            for (uint32_t j = tid; j < nk; j += MIF_THREADS)
            {
                const uint32_t a = koff[j], b = koff[j + 1];
                if (b - a >= uint32_t(MIF_LONG)) continue;
                double S = 0, A = 0;
                for (uint32_t i = a; i < b; ++i) { S = __dadd_rn(S, double(sx[i])); A = __dadd_rn(A, st[i]); }       // stored order (filter.cpp:78-82)
                fv[q0 + j] = mif0_value(S, A, b - a, C, s_tab);
            }
            // MIF_LONG .. MIF_WIDE - 1 entries: 8 lanes, lane l adds entries l, l + 8, ..., then the xor butterfly over 4, 2, 1
            {
                const int gl = tid & 7;
                const unsigned gmask = 0xffu << (lane & ~7);
                for (uint32_t j = tid >> 3; j < nk; j += MIF_THREADS / 8)
                {
                    const uint32_t a = koff[j], b = koff[j + 1];
                    if (b - a < uint32_t(MIF_LONG) || b - a >= uint32_t(MIF_WIDE)) continue;      // uniform within the 8 lanes
                    double S = 0, A = 0;
                    for (uint32_t i = a + gl; i < b; i += 8) { S = __dadd_rn(S, double(sx[i])); A = __dadd_rn(A, st[i]); }
#pragma unroll
                    for (int o = 4; o >= 1; o >>= 1) { S = __dadd_rn(S, shfl_xor_f64(gmask, S, o)); A = __dadd_rn(A, shfl_xor_f64(gmask, A, o)); }
                    if (gl == 0) fv[q0 + j] = mif0_value(S, A, b - a, C, s_tab);
                }
            }
            // MIF_WIDE entries and more: a warp, lane l adds entries l, l + 32, ..., butterfly over 16 .. 1
            for (uint32_t j = warp; j < nk; j += MIF_WARPS)
            {
                const uint32_t a = koff[j], b = koff[j + 1];
                if (b - a < uint32_t(MIF_WIDE)) continue;
                double S = 0, A = 0;
                for (uint32_t i = a + lane; i < b; i += 32) { S = __dadd_rn(S, double(sx[i])); A = __dadd_rn(A, st[i]); }
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) { S = __dadd_rn(S, shfl_xor_f64(full, S, o)); A = __dadd_rn(A, shfl_xor_f64(full, A, o)); }
                if (lane == 0) fv[q0 + j] = mif0_value(S, A, b - a, C, s_tab);
            }
            return;
        }
        // streaming path: a run longer than the staging buffer (a k-mer present on thousands of branches).  Straight from
        // global memory, one warp per k-mer, in the same two orders.
        for (uint32_t j = warp; j < nk; j += MIF_WARPS)
        {
            const uint64_t a = offs[q0 + j], b = offs[q0 + j + 1];
            double S = 0, A = 0;
            if (b - a >= uint64_t(MIF_LONG))
            {
                // the association of the staged path: 8 summing lanes below MIF_WIDE entries (the other lanes add exact zeros
                // in the first two butterfly steps), 32 from there on
                const int width = b - a >= uint64_t(MIF_WIDE) ? 32 : 8;
                if (lane < width)
                    for (uint64_t i = a + lane; i < b; i += width)
                    {
                        const double x = double(logscore_to_score(scores[i]));
                        S = __dadd_rn(S, x); A = __dadd_rn(A, mif_xlog2x(x, s_tab));
                    }
#pragma unroll
                for (int o = 16; o >= 1; o >>= 1) { S = __dadd_rn(S, shfl_xor_f64(full, S, o)); A = __dadd_rn(A, shfl_xor_f64(full, A, o)); }
            }
            else if (lane == 0)
                for (uint64_t i = a; i < b; ++i)
                {
                    const double x = double(logscore_to_score(scores[i]));
                    S = __dadd_rn(S, x); A = __dadd_rn(A, mif_xlog2x(x, s_tab));
                }
            if (lane == 0) fv[q0 + j] = mif0_value(S, A, b - a, C, s_tab);
        }
    }

    __device__ __forceinline__ uint64_t mulmod31(uint64_t a, uint64_t b) { return (a * b) % 2147483647ull; }

    /// filter.cpp:137-144 with libstdc++'s std::default_random_engine (minstd_rand0: x <- 16807 x mod 2^31-1, seed 42) and
    /// std::uniform_real_distribution<double>(0,1) (generate_canonical: two draws, (g1 + g2*R) / R^2, R = 2^31-2).
    /// The reference draws in hash-map iteration order, which is unknowable here (parity unpinned); the canonical order of
    /// this repo is a function of the k-mer alone, so that key shards on several GPUs draw what one GPU draws: the k-mer
    /// with key K takes draws 2e+1 and 2e+2 of the stream, e = K mod (2^31-2) (the period).  The LCG is jumped ahead by
    /// modular exponentiation.
    __global__ void k4_random(uint64_t n_kmers, const uint64_t* __restrict__ kmer_keys, float* __restrict__ fv)
    {
        const uint64_t q = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (q >= n_kmers) return;
        uint64_t e = 2 * (kmer_keys[q] % 2147483646ull) + 1, base = 16807, acc = 1;
        while (e)
        {
            if (e & 1) acc = mulmod31(acc, base);
            base = mulmod31(base, base);
            e >>= 1;
        }
        const uint64_t x1 = mulmod31(42, acc), x2 = mulmod31(x1, 16807);
        const double R = 2147483646.0;
        double r = __ddiv_rn(__dadd_rn(double(x1 - 1), __dmul_rn(double(x2 - 1), R)), __dmul_rn(R, R));
        if (r >= 1.0) r = 0.99999999999999988897769753748;      // nextafter(1.0, 0.0)
        fv[q] = float(r);
    }

    /// ascending-sortable image of a float, and the original index as payload
    __global__ void k5_order_keys(const float* __restrict__ fv, uint64_t n, uint64_t* __restrict__ okeys, uint32_t* __restrict__ ovals)
    {
        const uint64_t q = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (q >= n) return;
        const uint32_t b = __float_as_uint(fv[q]);
        okeys[q] = b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
        ovals[q] = uint32_t(q);
    }

    /// out position i takes k-mer perm[i]: its key, filter value and entry count
    __global__ void k6_gather_heads(const uint32_t* __restrict__ perm, const uint64_t* __restrict__ kmer_keys,
                                    const uint64_t* __restrict__ kmer_offs, const float* __restrict__ fv, uint64_t n,
                                    uint64_t* __restrict__ out_keys, float* __restrict__ out_fv, uint32_t* __restrict__ out_cnt,
                                    uint64_t* __restrict__ out_offs /* indexed results: where the k-mer's entries start, or null */, uint64_t n_entries)
    {
        const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (out_offs && i == 0) out_offs[n] = n_entries;
        if (i >= n) return;
        const uint32_t src = perm[i];
        out_keys[i] = kmer_keys[src];
        out_fv[i] = fv[src];
        const uint64_t a = kmer_offs[src];
        out_cnt[i] = uint32_t(kmer_offs[src + 1] - a);
        if (out_offs) out_offs[i] = a;
    }

    /// u32 -> u16 branch ids (IPKGPU_RESULT_BRANCH16) of the entries [lo, hi), four per thread and step between the
    /// multiples of four (both arrays are indexed from their 16-byte aligned starts, so one alignment serves both)
    __global__ void __launch_bounds__(256) k6_narrow_branch(const uint32_t* __restrict__ in, uint64_t lo, uint64_t hi, uint16_t* __restrict__ out)
    {
        const uint64_t v0 = min(hi, (lo + 3) & ~uint64_t(3)), v1 = max(v0, hi & ~uint64_t(3));
        for (uint64_t i = v0 / 4 + blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < v1 / 4; i += uint64_t(gridDim.x) * blockDim.x)
        {
            const uint4 v = reinterpret_cast<const uint4*>(in)[i];
            reinterpret_cast<uint2*>(out)[i] = make_uint2(v.x | v.y << 16, v.z | v.w << 16);
        }
        if (blockIdx.x == 0)
        {
            if (lo + threadIdx.x < v0) out[lo + threadIdx.x] = uint16_t(in[lo + threadIdx.x]);
            if (v1 + threadIdx.x < hi) out[v1 + threadIdx.x] = uint16_t(in[v1 + threadIdx.x]);
        }
    }

    /// entries of k-mer perm[i] move to [out_offs32[i], +cnt), G lanes per k-mer; also widens the offsets to u64 (+ the final sentinel)
    template<int G>
    __global__ void __launch_bounds__(256) k6_gather_entries(const uint32_t* __restrict__ perm, const uint64_t* __restrict__ kmer_offs,
                                                             const uint32_t* __restrict__ out_offs32, uint64_t n, uint64_t n_entries,
                                                             const uint32_t* __restrict__ ent_branch, const float* __restrict__ ent_score,
                                                             uint64_t* __restrict__ out_offs, uint32_t* __restrict__ out_branch, float* __restrict__ out_score)
    {
        const uint64_t t = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        const uint64_t i = t / G;
        const int gl = int(threadIdx.x) & (G - 1);
        if (t == 0) out_offs[n] = n_entries;
        if (i >= n) return;
        const uint32_t src = perm[i];
        const uint64_t a = kmer_offs[src], cnt = kmer_offs[src + 1] - a;
        const uint64_t o = out_offs32[i];
        if (gl == 0) out_offs[i] = o;
        for (uint64_t j = gl; j < cnt; j += G)
        {
            out_branch[o + j] = ent_branch[a + j];
            out_score[o + j] = ent_score[a + j];
        }
    }
}
