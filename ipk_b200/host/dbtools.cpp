// dbtools.cpp -- ipk-b200-db: the two database tools of the reference over this repo's .ipk reader.
//
//   ipk-b200-db diff A.ipk B.ipk [-v] [--eps E] [--boundary E]     tools/src/diff.cpp:24-295 (same checks, same report lines)
//   ipk-b200-db dump DB.ipk                                         tools/src/dump.cpp:18-33  (k-mer, then `10^score <TAB> pre-order id`)
//
// diff compares what the reference's differ compares -- sequence type, k, omega / threshold, tree, number of k-mers and of
// phylo-k-mers, and every (k-mer, branch) score within EPS (1e-2 there; --eps changes it, --eps 0 asks for bit equality).
// --boundary E is BASELINE.json's rule on top (diff-plain-text.py:36-45): a difference where a score lies within E of the
// threshold is listed separately as a boundary case and does not fail the comparison.  Exit status: 0 equal, 2 different,
// 1 error.
#include "ipkhost.h"
#include "tree.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace
{
    struct database
    {
        ipkhost_db_file* file = nullptr;
        ipkhost_db v;
        explicit database(const char* path)
        {
            memset(&v, 0, sizeof v);
            const int rc = ipkhost_read_ipk(path, &file, &v);
            if (rc != IPKHOST_OK) throw std::runtime_error(std::string("cannot read ") + path + (rc == IPKHOST_EFORMAT ? " (not a database this tool knows)" : ""));
        }
        ~database() { ipkhost_free_ipk(file); }
        database(const database&) = delete;
    };

    const char* ok(bool x) { return x ? "OK" : "DIFF"; }

    /// i2l::decode_kmer: most significant digits first; state order of the alphabets as ipkhost.h lists them
    std::string decode_kmer(uint64_t key, uint64_t k, bool dna)
    {
        static const char* DNA = "ACGT";
        static const char* AA = "RHKDESTNQCGPAILMFWYV";
        const unsigned bits = dna ? 2 : 5;
        std::string s(k, '?');
        for (uint64_t i = 0; i < k; ++i)
        {
            const unsigned c = unsigned(key >> (bits * (k - 1 - i))) & ((1u << bits) - 1);
            s[i] = dna ? DNA[c] : (c < 20 ? AA[c] : '?');
        }
        return s;
    }

    float log_threshold(float omega, uint64_t k, bool dna)
    {
        return log10f(powf(omega / (dna ? 4.0f : 20.0f), float(k)));          // as ipkgpu_score_threshold (include/ipkgpu.h)
    }

    struct pk_diff { uint64_t kmer; uint32_t branch; float a, b; };

    int run_diff(int argc, char** argv)
    {
        bool verbose = false;
        double eps = 1e-2, boundary = -1.0;
        std::vector<const char*> files;
        for (int i = 2; i < argc; ++i)
        {
            const std::string a = argv[i];
            if (a == "-v" || a == "--verbose") verbose = true;
            else if (a == "--eps" && i + 1 < argc) eps = atof(argv[++i]);
            else if (a == "--boundary" && i + 1 < argc) boundary = atof(argv[++i]);
            else files.push_back(argv[i]);
        }
        if (files.size() != 2) { fprintf(stderr, "Usage: ipk-b200-db diff DATABASE_A DATABASE_B [-v] [--eps E] [--boundary E]\n"); return 1; }
        const database A(files[0]), B(files[1]);
        const ipkhost_db& a = A.v; const ipkhost_db& b = B.v;
        bool all_ok = true;
        auto report = [&](const char* what, bool match, const std::string& va, const std::string& vb) {
            all_ok = all_ok && match;
            printf("%s:\t%s\t%s\t%s\n", what, ok(match), va.c_str(), vb.c_str());
        };
        report("Sequence type", !strcmp(a.sequence_type, b.sequence_type), a.sequence_type, b.sequence_type);
        report("k-mer size", a.kmer_size == b.kmer_size, std::to_string(a.kmer_size), std::to_string(b.kmer_size));
        report("Omega", a.omega == b.omega, std::to_string(a.omega), std::to_string(b.omega));
        const bool dna_a = !strcmp(a.sequence_type, "DNA"), dna_b = !strcmp(b.sequence_type, "DNA");
        const float thr_a = log_threshold(a.omega, a.kmer_size, dna_a), thr_b = log_threshold(b.omega, b.kmer_size, dna_b);
        report("Threshold", a.omega == b.omega, std::to_string(thr_a), std::to_string(thr_b));
        report("Reference tree", !strcmp(a.tree, b.tree), " ", " ");
        {
            bool same = a.n_index == b.n_index;
            for (uint64_t i = 0; same && i < a.n_index; ++i)
                same = a.index_num_nodes[i] == b.index_num_nodes[i] && a.index_subtree_branch_length[i] == b.index_subtree_branch_length[i];
            report("Tree index", same, std::to_string(a.n_index), std::to_string(b.n_index));        // ("???" in the reference: not implemented there)
        }
        report("Number of k-mers", a.n_kmers == b.n_kmers, std::to_string(a.n_kmers), std::to_string(b.n_kmers));
        report("Number of phylo-k-mers", a.n_entries == b.n_entries, std::to_string(a.n_entries), std::to_string(b.n_entries));

        // check_phylo_kmers (diff.cpp:203-289): every (k-mer, branch) of either database
        std::unordered_map<uint64_t, uint64_t> where_b;
        where_b.reserve(size_t(b.n_kmers) * 2 + 1);
        for (uint64_t q = 0; q < b.n_kmers; ++q) where_b[b.keys[q]] = q;
        std::vector<pk_diff> diffs, boundary_cases;
        const float na = NAN;
        auto differs = [&](float x, float y) {
            if (std::isnan(x) || std::isnan(y)) return true;
            return eps > 0 ? !(std::fabs(double(x) - double(y)) < eps) : memcmp(&x, &y, 4) != 0;
        };
        auto add = [&](uint64_t kmer, uint32_t branch, float x, float y) {
            const bool near = boundary >= 0 && ((!std::isnan(x) && std::fabs(double(x) - double(thr_a)) < boundary) ||
                                                (!std::isnan(y) && std::fabs(double(y) - double(thr_b)) < boundary));
            (near ? boundary_cases : diffs).push_back({ kmer, branch, x, y });
        };
        std::vector<char> seen_b(b.n_kmers, 0);
        std::unordered_map<uint32_t, float> vb;
        for (uint64_t q = 0; q < a.n_kmers; ++q)
        {
            const auto it = where_b.find(a.keys[q]);
            if (it == where_b.end())
            {
                for (uint64_t e = a.offsets[q]; e < a.offsets[q + 1]; ++e) add(a.keys[q], a.branch[e], a.score[e], na);
                continue;
            }
            const uint64_t p = it->second;
            seen_b[p] = 1;
            vb.clear();
            for (uint64_t e = b.offsets[p]; e < b.offsets[p + 1]; ++e) vb[b.branch[e]] = b.score[e];
            for (uint64_t e = a.offsets[q]; e < a.offsets[q + 1]; ++e)
            {
                const auto s = vb.find(a.branch[e]);
                if (s == vb.end()) add(a.keys[q], a.branch[e], a.score[e], na);
                else { if (differs(a.score[e], s->second)) add(a.keys[q], a.branch[e], a.score[e], s->second); vb.erase(s); }
            }
            for (const auto& [branch, score] : vb) add(a.keys[q], branch, na, score);        // branches only B scores
        }
        for (uint64_t p = 0; p < b.n_kmers; ++p)
            if (!seen_b[p])
                for (uint64_t e = b.offsets[p]; e < b.offsets[p + 1]; ++e) add(b.keys[p], b.branch[e], na, b.score[e]);
        const bool match = diffs.empty();
        all_ok = all_ok && match;
        printf("Phylo-k-mer scores:\t%s\t%zu\n", ok(match), diffs.size());
        if (boundary >= 0) printf("Boundary cases (a score within %g of the threshold):\t%zu\n", boundary, boundary_cases.size());
        if (verbose)
        {
            auto list = [&](const std::vector<pk_diff>& v) {
                printf("\t\tcode\tk-mer\tbranch\tA score\tB score\n");
                for (const auto& d : v)
                    printf("\t\t%llu\t%s\t%u\t%g\t%g\t\n", (unsigned long long)d.kmer, decode_kmer(d.kmer, a.kmer_size, dna_a).c_str(), d.branch,
                           std::pow(10.0, double(d.a)), std::pow(10.0, double(d.b)));
            };
            list(diffs);
            if (boundary >= 0) { printf("\tboundary cases:\n"); list(boundary_cases); }
        }
        return all_ok ? 0 : 2;
    }

    int run_dump(int argc, char** argv)
    {
        if (argc != 3) { fprintf(stderr, "Usage: ipk-b200-db dump DATABASE\n"); return 1; }
        const database D(argv[2]);
        const ipkhost_db& d = D.v;
        const bool dna = !strcmp(d.sequence_type, "DNA");
        // dump.cpp:20,31: branch ids are printed as PRE-order ids of the header's tree
        std::vector<uint32_t> pre_of_post;
        if (d.tree && *d.tree)
        {
            const ipkhost::ptree t = ipkhost::parse_newick(d.tree);
            pre_of_post.resize(t.node_count());
            for (size_t i = 0; i < t.postorder.size(); ++i) pre_of_post[i] = t.nodes[t.postorder[i]].preorder_id;
        }
        for (uint64_t q = 0; q < d.n_kmers; ++q)
        {
            printf("%s\n", decode_kmer(d.keys[q], d.kmer_size, dna).c_str());
            for (uint64_t e = d.offsets[q]; e < d.offsets[q + 1]; ++e)
            {
                const uint32_t br = d.branch[e];
                printf("\t%g\t%u\n", std::pow(10.0, double(d.score[e])), br < pre_of_post.size() ? pre_of_post[br] : br);
            }
        }
        return 0;
    }
}

int main(int argc, char** argv)
{
    try
    {
        if (argc >= 2 && !strcmp(argv[1], "diff")) return run_diff(argc, argv);
        if (argc >= 2 && !strcmp(argv[1], "dump")) return run_dump(argc, argv);
        fprintf(stderr, "Usage: ipk-b200-db diff DATABASE_A DATABASE_B [-v] [--eps E] [--boundary E]\n       ipk-b200-db dump DATABASE\n");
        return 1;
    }
    catch (const std::exception& e)
    {
        fprintf(stderr, "Exception: %s\n", e.what());
        return 1;
    }
}
