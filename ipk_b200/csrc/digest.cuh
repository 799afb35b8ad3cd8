// digest.cuh -- order-independent digest of the database a build left on the device, sm_100a.
//
// No reference counterpart: the reference compares databases entry by entry with tools/src/diff.cpp:129-295.  A digest is
// what lets two builds that never meet in one memory (device-resident vs host-bound, 1 GPU vs the shards of G processes)
// be compared entry for entry: sums and xors of a 64-bit mix of every (k-mer, branch, score bits), every (k-mer, entry
// count) and every (k-mer, filter value bits).  Sums / xors over shards add / xor up to the digest of the whole database.
// tests/parity.py holds the same definition in numpy.
#pragma once
#include "common.cuh"

namespace ipkgpu
{
    __host__ __device__ __forceinline__ uint64_t dg_mix64(uint64_t x)
    {
        x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
        x ^= x >> 27; x *= 0x94d049bb133111ebull;
        x ^= x >> 31;
        return x;
    }
    __host__ __device__ __forceinline__ uint64_t dg_key(uint64_t kmer) { return dg_mix64(kmer + 0x9e3779b97f4a7c15ull); }
    __host__ __device__ __forceinline__ uint64_t dg_entry(uint64_t hk, uint32_t branch, uint32_t score_bits) { return dg_mix64(hk ^ (uint64_t(branch) << 32 | score_bits)); }
    __host__ __device__ __forceinline__ uint64_t dg_count(uint64_t hk, uint64_t count) { return dg_mix64(hk ^ (count + 0x632be59bd9b4e019ull)); }
    __host__ __device__ __forceinline__ uint64_t dg_fv(uint64_t hk, uint32_t fv_bits) { return dg_mix64(hk ^ (uint64_t(fv_bits) | 0x5bd1e99500000000ull)); }

    constexpr int DG_THREADS = 256;
    constexpr int DG_LANES = 8;             // lanes that share one k-mer

    /// acc[0..5] = entries sum / xor, k-mers sum / xor, filter values sum / xor; acc[6] = entries whose score lies
    /// within `band` above log_threshold (the emitted side of the north star's boundary cases)
    __global__ void __launch_bounds__(DG_THREADS) k7_digest(const uint64_t* __restrict__ kmer_keys, const uint64_t* __restrict__ kmer_offs,
                                                            const uint32_t* __restrict__ ent_branch, const float* __restrict__ ent_score,
                                                            const float* __restrict__ fv, uint64_t n_kmers, float log_threshold, float band,
                                                            unsigned long long* __restrict__ acc)
    {
        uint64_t es = 0, ex = 0, ks = 0, kx = 0, fs = 0, fx = 0, nb = 0;
        const int gl = threadIdx.x & (DG_LANES - 1);
        const float hi = __fadd_rn(log_threshold, band);
        for (uint64_t q = (blockIdx.x * uint64_t(DG_THREADS) + threadIdx.x) / DG_LANES; q < n_kmers; q += uint64_t(gridDim.x) * (DG_THREADS / DG_LANES))
        {
            const uint64_t hk = dg_key(kmer_keys[q]);
            const uint64_t a = kmer_offs[q], b = kmer_offs[q + 1];
            if (gl == 0)
            {
                const uint64_t h = dg_count(hk, b - a);
                ks += h; kx ^= h;
                const uint64_t g = dg_fv(hk, __float_as_uint(fv[q]));
                fs += g; fx ^= g;
            }
            for (uint64_t i = a + gl; i < b; i += DG_LANES)
            {
                const float s = ent_score[i];
                const uint64_t h = dg_entry(hk, ent_branch[i], __float_as_uint(s));
                es += h; ex ^= h;
                nb += s < hi ? 1 : 0;
            }
        }
        uint64_t v[7] = { es, ex, ks, kx, fs, fx, nb };
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1)
#pragma unroll
            for (int j = 0; j < 7; ++j)
            {
                const uint64_t t = __shfl_xor_sync(0xffffffffu, v[j], o);
                if (j == 1 || j == 3 || j == 5) v[j] ^= t; else v[j] += t;
            }
        if ((threadIdx.x & 31) == 0)
        {
            atomicAdd(acc + 0, (unsigned long long)v[0]); atomicXor(acc + 1, (unsigned long long)v[1]);
            atomicAdd(acc + 2, (unsigned long long)v[2]); atomicXor(acc + 3, (unsigned long long)v[3]);
            atomicAdd(acc + 4, (unsigned long long)v[4]); atomicXor(acc + 5, (unsigned long long)v[5]);
            atomicAdd(acc + 6, (unsigned long long)v[6]);
        }
    }

    // ---- the N-way merge order of key shards (merge_stage2, ipk/src/db_builder.cpp:392-458) as one stable sort ---------------
    /// keys[i] = ascending-sortable image of fv[src], vals[i] = src, with src = perm[i] (or i without a permutation)
    __global__ void __launch_bounds__(256) k9_image_keys(const float* __restrict__ fv, const uint32_t* __restrict__ perm, uint64_t n,
                                                         uint64_t* __restrict__ okeys, uint32_t* __restrict__ ovals)
    {
        const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (i >= n) return;
        const uint32_t src = perm ? perm[i] : uint32_t(i);
        const uint32_t b = __float_as_uint(fv[src]);
        okeys[i] = b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u);
        ovals[i] = src;
    }

    __global__ void __launch_bounds__(256) k9_iota(uint32_t* __restrict__ v, uint64_t n)
    {
        const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (i < n) v[i] = uint32_t(i);
    }

    struct mo_table { uint64_t prefix[257]; uint32_t n_shards; };

    /// global position (shards laid end to end) -> (shard, index inside the shard)
    __global__ void __launch_bounds__(256) k9_split_positions(const uint32_t* __restrict__ g, uint64_t n, const __grid_constant__ mo_table T,
                                                              uint8_t* __restrict__ shard, uint32_t* __restrict__ index)
    {
        const uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        if (i >= n) return;
        const uint64_t p = g[i];
        uint32_t lo = 0, hi = T.n_shards;              // last s with prefix[s] <= p
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (T.prefix[mid] <= p) lo = mid; else hi = mid; }
        shard[i] = uint8_t(lo);
        index[i] = uint32_t(p - T.prefix[lo]);
    }

    /// Boundary cases (BASELINE.json: k-mers whose score lies within a band of the threshold are listed separately): the
    /// unique (k-mer, group) records of a build at threshold - band (`loose`, sorted by composite key) against those of a
    /// build at threshold + band (`tight`, a subset as a set of keys).  A loose record that the tight build lacks, or holds
    /// with another score (the per-branch maximum was taken over a different set of windows), depends on which side of the
    /// band the threshold falls: it is appended to the list.
    __global__ void __launch_bounds__(256) k8_boundary_diff(const uint64_t* __restrict__ loose_keys, const float* __restrict__ loose_scores, uint64_t n_loose,
                                                            const uint64_t* __restrict__ tight_keys, const float* __restrict__ tight_scores, uint64_t n_tight,
                                                            uint32_t group_bits, const uint32_t* __restrict__ group_branch, uint64_t cap,
                                                            uint64_t* __restrict__ out_kmer, uint32_t* __restrict__ out_branch, float* __restrict__ out_loose,
                                                            float* __restrict__ out_tight, unsigned long long* __restrict__ counter)
    {
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n_loose; i += uint64_t(gridDim.x) * blockDim.x)
        {
            const uint64_t key = loose_keys[i];
            uint64_t lo = 0, hi = n_tight;                 // first j with tight_keys[j] >= key
            while (lo < hi) { const uint64_t mid = (lo + hi) >> 1; if (tight_keys[mid] < key) lo = mid + 1; else hi = mid; }
            const bool present = lo < n_tight && tight_keys[lo] == key;
            const float sl = loose_scores[i];
            const float st = present ? tight_scores[lo] : __uint_as_float(0x7fc00000u);
            if (present && __float_as_uint(st) == __float_as_uint(sl)) continue;
            const unsigned long long o = atomicAdd(counter, 1ull);
            if (o < cap)
            {
                out_kmer[o] = key >> group_bits;
                out_branch[o] = group_branch[uint32_t(key & ((1ull << group_bits) - 1))];
                out_loose[o] = sl;
                out_tight[o] = st;
            }
        }
    }
}
