// filter.cuh -- filter values, ordering by filter value and the final gather, sm_100a.
//
// Replaces   mif0_filter::calc_filter_values   ipk/src/filter.cpp:60-119   (+ logscore_to_score :20-23, shannon :55-58)
//            random_filter::calc_filter_values ipk/src/filter.cpp:133-146
//            std::sort(kmer_order) + the db.at(kmer) walk of filter_in_ram   ipk/src/db_builder.cpp:284,323-332
#pragma once
#include "common.cuh"

namespace ipkgpu
{
    /// filter.cpp:20-23: std::min(std::pow(10, log_score), 1.0) evaluated in double, returned as score_type (float).
    /// exp10 (<= 1 ulp in fp64) stands in for pow(10, x); after the rounding to float the two agree (FV_RTOL covers the rest).
    __device__ __forceinline__ float logscore_to_score(float log_score)
    {
        const double p = exp10(double(log_score));
        return float(p < 1.0 ? p : 1.0);
    }
    __device__ __forceinline__ double shannon(double x) { return -x * log2(x); }    // filter.cpp:55-58

    __device__ __forceinline__ double shfl_f64(unsigned mask, double v, int src, int width)
    {
        int lo = __double2loint(v), hi = __double2hiint(v);
        lo = __shfl_sync(mask, lo, src, width);
        hi = __shfl_sync(mask, hi, src, width);
        return __hiloint2double(hi, lo);
    }

    /// G lanes per k-mer (G = 1, 4, 8 or 32, chosen by the host from the mean entries per k-mer).  The transcendental
    /// work (exp10, log2, the divisions) is spread over the lanes, G entries at a time; the two fp64 running sums are then
    /// advanced in the reference's order -- entries are visited in stored (group processing) order, as
    /// filter.cpp:78-82,101-108 does -- by broadcasting the lanes' terms one after the other.
    template<int G>
    __global__ void __launch_bounds__(256) k4_mif0(const uint64_t* __restrict__ offs, const float* __restrict__ scores, uint64_t n_kmers,
                                                   uint32_t total_num_groups, float threshold, float* __restrict__ fv)
    {
        const uint64_t q = (blockIdx.x * uint64_t(blockDim.x) + threadIdx.x) / G;
        if (q >= n_kmers) return;                                    // whole groups leave together
        const int gl = int(threadIdx.x) & (G - 1), lane = int(threadIdx.x) & 31;
        const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (lane & ~(G - 1)));
        const uint64_t a = offs[q], b = offs[q + 1];
        double score_sum = 0;
        for (uint64_t base = a; base < b; base += G)
        {
            const uint64_t i = base + gl;
            const float x = i < b ? logscore_to_score(scores[i]) : 0.0f;
            const int cnt = int(min(uint64_t(G), b - base));
            if (G == 1) score_sum = __dadd_rn(score_sum, double(x));
            else for (int j = 0; j < cnt; ++j) score_sum = __dadd_rn(score_sum, double(__shfl_sync(gmask, x, j, G)));
        }
        score_sum = __dadd_rn(score_sum, __dmul_rn(double(uint64_t(total_num_groups) - (b - a)), double(threshold)));   // :86
        const double weighted_threshold = double(threshold) / score_sum;
        const double target_threshold = shannon(weighted_threshold);
        double HcBw1 = __dmul_rn(double(total_num_groups), target_threshold);                                            // :93
        for (uint64_t base = a; base < b; base += G)
        {
            const uint64_t i = base + gl;
            double target_value = 0;
            if (i < b)
            {
                const double weighted_score = double(logscore_to_score(scores[i])) / score_sum;
                target_value = shannon(weighted_score);
            }
            const int cnt = int(min(uint64_t(G), b - base));
            if (G == 1) HcBw1 = __dadd_rn(__dsub_rn(HcBw1, target_threshold), target_value);                             // :107
            else for (int j = 0; j < cnt; ++j) HcBw1 = __dadd_rn(__dsub_rn(HcBw1, target_threshold), shfl_f64(gmask, target_value, j, G));
        }
        const double Hc = log2(double(total_num_groups));
        if (gl == 0) fv[q] = float(__dmul_rn(score_sum, __dsub_rn(HcBw1, Hc)));                                          // :115
    }

    __device__ __forceinline__ uint64_t mulmod31(uint64_t a, uint64_t b) { return (a * b) % 2147483647ull; }

    /// filter.cpp:137-144 with libstdc++'s std::default_random_engine (minstd_rand0: x <- 16807 x mod 2^31-1, seed 42) and
    /// std::uniform_real_distribution<double>(0,1) (generate_canonical: two draws, (g1 + g2*R) / R^2, R = 2^31-2).
    /// The i-th k-mer IN ASCENDING KEY ORDER takes draws 2i+1 and 2i+2 (the reference draws in hash-map iteration
    /// order, which is unknowable here: parity unpinned).  The LCG is jumped ahead by modular exponentiation.
    __global__ void k4_random(uint64_t n_kmers, float* __restrict__ fv)
    {
        const uint64_t q = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (q >= n_kmers) return;
        uint64_t e = 2 * q + 1, base = 16807, acc = 1;
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
                                    uint64_t* __restrict__ out_keys, float* __restrict__ out_fv, uint32_t* __restrict__ out_cnt)
    {
        const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (i >= n) return;
        const uint32_t src = perm[i];
        out_keys[i] = kmer_keys[src];
        out_fv[i] = fv[src];
        out_cnt[i] = uint32_t(kmer_offs[src + 1] - kmer_offs[src]);
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
