// merge.cuh -- per-branch max-merge and regrouping by k-mer over SORTED composite keys, sm_100a.
//
// Input: records sorted by  ckey = (kmer << group_bits) | group  (group = branch group in processing order).
//   * equal ckey  = the same k-mer emitted again for the same branch (another window, or the branch's other ghost
//     node): keep the maximum score                      -- ipk::put, ipk/src/branch_group.cpp:88-101
//   * equal kmer  = one database row: its (branch, score) entries in group processing order
//                                                        -- unsafe_insert loop, ipk/src/db_builder.cpp:685-693
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace ipkgpu
{
    constexpr int MG_THREADS = 256;
    constexpr int MG_ITEMS = 8;
    constexpr int MG_TILE = MG_THREADS * MG_ITEMS;

    /// per tile: number of record heads (ckey differs from predecessor) and of k-mer heads
    __global__ void __launch_bounds__(MG_THREADS) mg_count(const uint64_t* __restrict__ ckeys, uint64_t n, uint32_t group_bits,
                                                           uint32_t* __restrict__ rec_counts, uint32_t* __restrict__ kmer_counts)
    {
        __shared__ uint32_t s_warp[9];
        const uint64_t base = uint64_t(blockIdx.x) * MG_TILE + uint64_t(threadIdx.x) * MG_ITEMS;
        uint32_t c1 = 0, c2 = 0;
        if (base < n)
        {
            uint64_t prev = base ? ckeys[base - 1] : ~ckeys[0];
#pragma unroll
            for (int i = 0; i < MG_ITEMS; ++i)
            {
                if (base + i < n)
                {
                    const uint64_t cur = ckeys[base + i];
                    c1 += cur != prev;
                    c2 += (cur >> group_bits) != (prev >> group_bits) || (base + i == 0);
                    prev = cur;
                }
            }
        }
        uint32_t t1, t2;
        block_excl_scan_256(c1, s_warp, &t1);
        block_excl_scan_256(c2, s_warp, &t2);
        if (threadIdx.x == 0) { rec_counts[blockIdx.x] = t1; kmer_counts[blockIdx.x] = t2; }
    }

    /// Writes the compacted records (and, when kmer_keys != null, the CSR by k-mer).
    /// rec_offs / kmer_offs_tile: exclusive scans of mg_count's outputs.
    /// The group field of the input keys is batch-local (group_bits wide, offset group_add); compacted records are
    /// re-encoded as (kmer << out_group_bits) | (group + group_add), the global form stage 2 and the exchange use.
    __global__ void __launch_bounds__(MG_THREADS) mg_write(const uint64_t* __restrict__ ckeys, const float* __restrict__ scores, uint64_t n,
                                                           uint32_t group_bits, uint32_t out_group_bits, uint32_t group_add,
                                                           const uint32_t* __restrict__ rec_offs,
                                                           const uint32_t* __restrict__ kmer_offs_tile,
                                                           uint64_t* __restrict__ u_keys, float* __restrict__ u_scores,      // may be null
                                                           const uint32_t* __restrict__ group_branch,                        // CSR outputs below may be null
                                                           uint64_t* __restrict__ kmer_keys, uint64_t* __restrict__ kmer_offs,
                                                           uint32_t* __restrict__ ent_branch, float* __restrict__ ent_score)
    {
        __shared__ uint32_t s_warp[9];
        const uint64_t base = uint64_t(blockIdx.x) * MG_TILE + uint64_t(threadIdx.x) * MG_ITEMS;
        uint64_t k[MG_ITEMS];
        uint32_t c1 = 0, c2 = 0;
        uint64_t prev0 = 0;
        if (base < n)
        {
            prev0 = base ? ckeys[base - 1] : ~ckeys[0];
            uint64_t prev = prev0;
#pragma unroll
            for (int i = 0; i < MG_ITEMS; ++i)
            {
                if (base + i < n)
                {
                    k[i] = ckeys[base + i];
                    c1 += k[i] != prev;
                    c2 += (k[i] >> group_bits) != (prev >> group_bits) || (base + i == 0);
                    prev = k[i];
                }
            }
        }
        uint64_t r1 = uint64_t(rec_offs[blockIdx.x]) + block_excl_scan_256(c1, s_warp, nullptr);
        uint64_t r2 = uint64_t(kmer_offs_tile[blockIdx.x]) + block_excl_scan_256(c2, s_warp, nullptr);
        if (base >= n) return;
        const uint64_t group_mask = (1ull << group_bits) - 1;
        uint64_t prev = prev0;
#pragma unroll
        for (int i = 0; i < MG_ITEMS; ++i)
        {
            if (base + i < n)
            {
                const uint64_t cur = k[i];
                if (cur != prev)
                {
                    // head of a run of equal (kmer, group): maximum score of the run (runs may cross tiles)
                    float m = scores[base + i];
                    for (uint64_t j = base + i + 1; j < n && ckeys[j] == cur; ++j)
                    {
                        const float v = scores[j];
                        if (m < v) m = v;                           // branch_group.cpp:92 `it->second < kmer.score`
                    }
                    const uint32_t group = uint32_t(cur & group_mask) + group_add;
                    if (u_keys) { u_keys[r1] = ((cur >> group_bits) << out_group_bits) | group; u_scores[r1] = m; }
                    if (kmer_keys)
                    {
                        ent_branch[r1] = group_branch[group];
                        ent_score[r1] = m;
                        if ((cur >> group_bits) != (prev >> group_bits) || (base + i == 0))
                        {
                            kmer_keys[r2] = cur >> group_bits;
                            kmer_offs[r2] = r1;
                            ++r2;
                        }
                    }
                    ++r1;
                }
                prev = cur;
            }
        }
    }
}
