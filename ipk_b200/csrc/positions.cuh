// positions.cuh -- the KEEP_POSITIONS build (ipk-aa-pos): every (k-mer, branch) entry also carries the alignment position of
// the window that gave it its score.  sm_100a.
//
// Reference: ipk::put under KEEP_POSITIONS (ipk/src/branch_group.cpp:72-86): `if (it->second.score < kmer.score)` -- a strict
// comparison, so of several windows with the maximal score the FIRST in processing order keeps its position; the order is the
// group's ghost nodes one after the other, windows ascending (db_builder.cpp:642-665, position = window.get_position()).
// The entry then is {branch, score, position} (db_builder.cpp:686-690).  Filter values ignore the position (filter.cpp:69-72).
//
// This is the reference's rarely used variant (amino acids only, ipk.py:275-283), so it is built for clarity, not for the
// roofline: the enumeration runs with its per-record (node, window) output, a sequence number  rank_in_group * windows + window
// is packed BELOW the group field of the composite key, the pair sort orders (k-mer, group), and one thread per run takes the
// maximum with the smaller sequence number winning ties.
#pragma once
#include "common.cuh"

namespace ipkgpu
{
    /// keys[i] (plain k-mer) -> ((kmer << group_bits | group) << seq_bits) | rank_in_group * nwin + window
    __global__ void __launch_bounds__(256) kp_compose(uint64_t* __restrict__ keys, const uint32_t* __restrict__ node, const uint32_t* __restrict__ win,
                                                      const uint32_t* __restrict__ node_group, const uint32_t* __restrict__ node_rank, uint64_t n,
                                                      uint32_t group_bits, uint32_t seq_bits, uint32_t nwin)
    {
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
        {
            const uint32_t nd = node[i];
            const uint64_t seq = uint64_t(node_rank[nd]) * nwin + win[i];
            keys[i] = (((keys[i] << group_bits) | node_group[nd]) << seq_bits) | seq;
        }
    }

    /// head_e[i] = 1 where a (k-mer, group) run starts, head_k[i] = 1 where a k-mer starts (records sorted on the bits above seq_bits)
    __global__ void __launch_bounds__(256) kp_flags(const uint64_t* __restrict__ keys, uint64_t n, uint32_t group_bits, uint32_t seq_bits,
                                                    uint32_t* __restrict__ head_e, uint32_t* __restrict__ head_k)
    {
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
        {
            const uint64_t cur = keys[i] >> seq_bits;
            const uint64_t prev = i ? keys[i - 1] >> seq_bits : ~cur;
            head_e[i] = cur != prev ? 1u : 0u;
            head_k[i] = (cur >> group_bits) != (prev >> group_bits) ? 1u : 0u;
        }
    }

    /// one thread per run: maximum score, the smallest sequence number among the maxima (= the first window in processing order,
    /// branch_group.cpp:76 strict `<`), written to the CSR at the scanned positions
    __global__ void __launch_bounds__(256) kp_write(const uint64_t* __restrict__ keys, const float* __restrict__ scores, uint64_t n, uint32_t group_bits,
                                                    uint32_t seq_bits, uint32_t nwin, const uint32_t* __restrict__ head_e, const uint32_t* __restrict__ idx_e,
                                                    const uint32_t* __restrict__ head_k, const uint32_t* __restrict__ idx_k, const uint32_t* __restrict__ group_branch,
                                                    uint64_t* __restrict__ kmer_keys, uint64_t* __restrict__ kmer_offs, uint32_t* __restrict__ ent_branch,
                                                    float* __restrict__ ent_score, uint32_t* __restrict__ ent_pos)
    {
        const uint64_t seq_mask = seq_bits >= 64 ? ~0ull : ((1ull << seq_bits) - 1);
        for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x)
        {
            if (!head_e[i]) continue;
            const uint64_t run = keys[i] >> seq_bits;
            float best = scores[i];
            uint64_t best_seq = keys[i] & seq_mask;
            for (uint64_t j = i + 1; j < n && (keys[j] >> seq_bits) == run; ++j)
            {
                const float s = scores[j];
                const uint64_t q = keys[j] & seq_mask;
                if (best < s || (best == s && q < best_seq)) { best = s; best_seq = q; }
            }
            const uint32_t e = idx_e[i];
            ent_branch[e] = group_branch[uint32_t(run & ((1ull << group_bits) - 1))];
            ent_score[e] = best;
            ent_pos[e] = uint32_t(best_seq % nwin);
            if (head_k[i])
            {
                const uint32_t q = idx_k[i];
                kmer_keys[q] = run >> group_bits;
                kmer_offs[q] = e;
            }
        }
    }

    /// positions of a contiguous (gathered) result: out position i takes k-mer perm[i], G lanes per k-mer
    template<int G>
    __global__ void __launch_bounds__(256) kp_gather_pos(const uint32_t* __restrict__ perm, const uint64_t* __restrict__ kmer_offs,
                                                         const uint64_t* __restrict__ out_offs, uint64_t n_kmers, const uint32_t* __restrict__ ent_pos,
                                                         uint32_t* __restrict__ out_pos)
    {
        const uint64_t t = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x;
        const uint64_t i = t / G;
        const int gl = int(threadIdx.x) & (G - 1);
        if (i >= n_kmers) return;
        const uint32_t src = perm[i];
        const uint64_t a = kmer_offs[src], cnt = kmer_offs[src + 1] - a, o = out_offs[i];
        for (uint64_t j = gl; j < cnt; j += G) out_pos[o + j] = ent_pos[a + j];
    }
}
