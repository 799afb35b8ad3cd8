// TEST INFRASTRUCTURE (oracle/_ref) -- NOT product code.  Nothing under ipk_b200/ may link or call this.
//
// C-ABI driver around the reference's OWN object code: /root/reference/ipk/src/pk_compute.cpp and
// /root/reference/ipk/src/window.cpp are compiled UNMODIFIED, where they lie, against oracle/shim/i2l/*.h
// (the i2l submodule is absent from the checkout).  The reference's db_builder.cpp,
// branch_group.cpp and filter.cpp need Boost + the i2l database and cannot be compiled here, so the few
// dozen lines of glue around DCLA are restated below, each block citing the lines it follows.
//
// Built by oracle/Makefile into oracle/_ref/libipkref_{dna,aa}.so (alphabet is a compile-time constant
// in the reference: i2l::seq_traits::alphabet_size).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <unordered_map>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "window.h"       // reference: ipk/include/window.h
#include "pk_compute.h"   // reference: ipk/include/pk_compute.h

namespace
{
    constexpr size_t SIGMA = i2l::seq_traits::alphabet_size;
    using key_type = i2l::phylo_kmer::key_type;
    using score_type = i2l::phylo_kmer::score_type;

    /// packed [W][sigma] fp32 log10 -> ipk::matrix (ctor calls preprocess(), window.cpp:10-14)
    ipk::matrix make_matrix(const float* logp, uint32_t W)
    {
        std::vector<ipk::matrix::column> data(W);
        for (uint32_t j = 0; j < W; ++j)
            for (size_t i = 0; i < SIGMA; ++i)
                data[j][i] = logp[size_t(j) * SIGMA + i];
        return ipk::matrix(std::move(data), "n");
    }

    struct pkdb_value { uint32_t branch; float score; };
    using group_hash_map = std::unordered_map<key_type, score_type>;   // i2l::hash_map stand-in (branch_group.h:23)

    /// branch_group.cpp:88-101 (non-positional put): insert, or replace when stored < new
    inline void put(group_hash_map& map, const ipk::uphylo_kmer& kmer)
    {
        if (auto it = map.find(kmer.key); it != map.end())
        {
            if (it->second < kmer.score) map[kmer.key] = kmer.score;
        }
        else map[kmer.key] = kmer.score;
    }

    /// filter.cpp:20-23 -- note the float return type: the double result is rounded to score_type
    score_type logscore_to_score(score_type log_score)
    {
        return std::min(std::pow(10, log_score), 1.0);
    }
    double shannon(double x) { return -x * std::log2(x); }   // filter.cpp:55-58
}

extern "C"
{
    int ipkref_sigma() { return int(SIGMA); }

    /// matrix::preprocess (window.cpp:16-27): B[0..W], fp32 sequential
    void ipkref_best_prefix(const float* logp, uint32_t W, float* B)
    {
        auto m = make_matrix(logp, W);
        for (uint32_t j = 0; j <= W; ++j) B[j] = m.range_max_sum(0, j);   // B[j]-B[0], B[0]=0
    }

    /// Runs `for window in to_windows(matrix,k): DCLA(window,k).run(eps)` (db_builder.cpp:645-649) over one
    /// matrix and returns every emitted (window start, key, score) in the reference's own emission order.
    /// Returns the number emitted; writes at most cap of them.
    int64_t ipkref_enumerate(const float* logp, uint32_t W, uint32_t k, float eps,
                             uint32_t* wpos, uint64_t* keys, float* scores, int64_t cap)
    {
        auto m = make_matrix(logp, W);
        int64_t n = 0;
        for (const auto& window : ipk::to_windows(&m, k))
        {
            auto alg = ipk::DCLA(window, k);
            alg.run(eps);
            for (const auto& kmer : alg.get_result())
            {
                if (n < cap)
                {
                    if (wpos) wpos[n] = uint32_t(window.get_position());
                    keys[n] = kmer.key;
                    scores[n] = kmer.score;
                }
                ++n;
            }
        }
        return n;
    }

    struct ipkref_db
    {
        std::vector<uint64_t> keys;       // fv-ascending, ties by key ascending (canonical order of this repo)
        std::vector<float> fv;
        std::vector<uint64_t> offsets;
        std::vector<uint32_t> branch;
        std::vector<float> score;
        uint64_t n_emitted = 0;
        double ms_stage1 = 0, ms_stage2 = 0;
    };

    /// Stage 1 + stage 2 of db_builder::run for the in-RAM path (db_builder.cpp:182-338), serialization excluded.
    ///   logp        [n_nodes][W][sigma] fp32 log10 posteriors, nodes in processing order
    ///   node_branch [n_nodes] original post-order id; equal ids = one branch group (db_builder.cpp:524-553)
    ///   filter      0 = mif0, 1 = random (filter.h:12-18)
    ///   threads     1 = the reference's real behaviour (OpenMP commented out, db_builder.cpp:602-605);
    ///               >1 parallelises the group loop the reference itself once marked `omp parallel for`.
    ipkref_db* ipkref_build(const float* logp, const uint32_t* node_branch, uint32_t n_nodes, uint32_t W, uint32_t k,
                            float log_threshold, float threshold, uint32_t total_num_groups, int filter, int threads)
    {
        auto* out = new ipkref_db();
        const auto t0 = std::chrono::steady_clock::now();

        // group_ghost_ids (db_builder.cpp:524-553): groups in order of first appearance
        std::vector<std::vector<uint32_t>> groups;
        std::vector<uint32_t> group_branch;
        {
            std::unordered_map<uint32_t, size_t> index_mapping;
            for (uint32_t n = 0; n < n_nodes; ++n)
            {
                auto it = index_mapping.find(node_branch[n]);
                if (it != index_mapping.end()) groups[it->second].push_back(n);
                else
                {
                    index_mapping[node_branch[n]] = groups.size();
                    groups.push_back({ n });
                    group_branch.push_back(node_branch[n]);
                }
            }
        }

        // explore_kmers / explore_group (db_builder.cpp:576-698), in-RAM branch
        std::vector<group_hash_map> group_maps(groups.size());
        uint64_t n_emitted = 0;
        (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) reduction(+: n_emitted) num_threads(threads > 0 ? threads : 1)
#endif
        for (size_t g = 0; g < groups.size(); ++g)
        {
            auto& group_map = group_maps[g];
            for (uint32_t node : groups[g])
            {
                auto node_matrix = make_matrix(logp + size_t(node) * W * SIGMA, W);
                for (const auto& window : ipk::to_windows(&node_matrix, k))
                {
                    auto alg = ipk::DCLA(window, k);
                    alg.run(log_threshold);
                    for (const auto& kmer : alg.get_result())
                    {
                        put(group_map, kmer);
                        ++n_emitted;
                    }
                }
                node_matrix.clear();
            }
        }
        // db_builder.cpp:685-693: unsafe_insert(kmer, {branch, score}) = append to db[kmer], groups in order
        std::unordered_map<key_type, std::vector<pkdb_value>> db;
        for (size_t g = 0; g < groups.size(); ++g)
        {
            for (const auto& [kmer, score] : group_maps[g]) db[kmer].push_back({ group_branch[g], score });
            group_hash_map().swap(group_maps[g]);
        }
        out->n_emitted = n_emitted;
        const auto t1 = std::chrono::steady_clock::now();

        // filter_in_ram (db_builder.cpp:254-290)
        struct kmer_fv { key_type key; float filter_value; };
        std::vector<kmer_fv> order;
        order.reserve(db.size());
        if (filter == 0)
        {
            // mif0_filter::calc_filter_values (filter.cpp:60-119); _threshold is score_type (filter.h:37)
            const size_t T = total_num_groups;
            for (const auto& [key, entries] : db)
            {
                double score_sum = 0;
                for (const auto& e : entries) score_sum += logscore_to_score(e.score);
                score_sum += static_cast<double>(T - entries.size()) * threshold;
                const auto weighted_threshold = threshold / score_sum;
                const auto target_threshold = shannon(weighted_threshold);
                auto HcBw1 = static_cast<double>(T) * target_threshold;
                for (const auto& e : entries)
                {
                    const auto weighted_score = logscore_to_score(e.score) / score_sum;
                    const auto target_value = shannon(weighted_score);
                    HcBw1 = HcBw1 - target_threshold + target_value;
                }
                const auto Hc = std::log2(T);
                order.push_back({ key, float(score_sum * (HcBw1 - Hc)) });
            }
        }
        else
        {
            // random_filter (filter.cpp:133-146).  The reference draws in hash-map iteration order, which is
            // unknowable without i2l::hash_map: PARITY UNPINNED.  This repo's canonical draw is a function of the k-mer:
            // key K takes draws 2e+1, 2e+2 of std::default_random_engine(42), e = K mod (2^31-2); the engine is jumped
            // there (x <- 42 * 16807^(2e) mod 2^31-1) and std::uniform_real_distribution does the rest.
            for (const auto& kv : db)
            {
                const uint64_t M = 2147483647ull;
                uint64_t e = 2 * (uint64_t(kv.first) % (M - 1)), base = 16807, acc = 42;
                for (; e; e >>= 1) { if (e & 1) acc = acc * base % M; base = base * base % M; }
                std::default_random_engine generator(static_cast<std::default_random_engine::result_type>(acc));
                std::uniform_real_distribution<double> distribution(0, 1);
                order.push_back({ kv.first, (float)distribution(generator) });
            }
        }
        // db_builder.cpp:284 std::sort(kmer_order) -- tie order unspecified there; canonical here: key ascending
        std::sort(order.begin(), order.end(), [](const kmer_fv& a, const kmer_fv& b) {
            return a.filter_value < b.filter_value || (a.filter_value == b.filter_value && a.key < b.key);
        });
        const auto t2 = std::chrono::steady_clock::now();

        // what save_phylo_kmer would be handed, k-mer by k-mer (db_builder.cpp:323-332)
        out->offsets.push_back(0);
        for (const auto& o : order)
        {
            const auto& entries = db.at(o.key);
            out->keys.push_back(o.key);
            out->fv.push_back(o.filter_value);
            for (const auto& e : entries) { out->branch.push_back(e.branch); out->score.push_back(e.score); }
            out->offsets.push_back(out->branch.size());
        }
        out->ms_stage1 = std::chrono::duration<double, std::milli>(t1 - t0).count();
        out->ms_stage2 = std::chrono::duration<double, std::milli>(t2 - t1).count();
        return out;
    }

    uint64_t ipkref_db_num_kmers(const ipkref_db* d) { return d->keys.size(); }
    uint64_t ipkref_db_num_entries(const ipkref_db* d) { return d->branch.size(); }
    uint64_t ipkref_db_num_emitted(const ipkref_db* d) { return d->n_emitted; }
    double ipkref_db_ms(const ipkref_db* d, int stage) { return stage == 1 ? d->ms_stage1 : d->ms_stage2; }
    void ipkref_db_copy(const ipkref_db* d, uint64_t* keys, float* fv, uint64_t* offsets, uint32_t* branch, float* score)
    {
        std::memcpy(keys, d->keys.data(), d->keys.size() * 8);
        std::memcpy(fv, d->fv.data(), d->fv.size() * 4);
        std::memcpy(offsets, d->offsets.data(), d->offsets.size() * 8);
        std::memcpy(branch, d->branch.data(), d->branch.size() * 4);
        std::memcpy(score, d->score.data(), d->score.size() * 4);
    }
    void ipkref_db_free(ipkref_db* d) { delete d; }
}
