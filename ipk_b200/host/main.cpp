// main.cpp -- ipk-b200: the database-build half of `ipk-dna` / `ipk-aa` (ipk/src/main.cpp:76-246, ipk::build) on a B200.
//
// It starts where the reference's hot path starts -- ancestral reconstruction done, ghost nodes mapped to branches --
// and ends where it ends, with the serialized database:
//     raxml-ng ancestralProbs (--ar-probs)  +  ghost node -> branch table (--node-map)
//         -> libipkgpu (include/ipkgpu.h): enumerate, per-branch max, regroup, filter values, order
//         -> DB.ipk (--output)
// Options keep the reference's names and defaults where it has them (ipk/src/command_line.cpp:83-147): -k (8), --omega
// (1.5), --mif0 | --random, -o/--output (<workdir>/DB.ipk), -w/--workdir.  Tree extension, alignment reduction and the
// raxml-ng run itself stay upstream (out of scope); --node-map is what extended_tree + ar::mapping boil
// down to for this path: one line `AR-node-label <TAB> post-order id of the original branch`, in processing order.
// Errors: message on stderr, exit status 1 (main.cpp:225-246).
#include "ipkhost.h"
#include "../../include/ipkgpu.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <memory>
#include <vector>

namespace
{
    struct options
    {
        std::string ar_probs, node_map, tree_file, output, workdir = ".";
        std::string states = "dna";
        uint32_t k = 8;
        float omega = 1.5f;
        int filter = IPKGPU_FILTER_MIF0;
        uint32_t node_count = 0;
        int device = 0, threads = 0;
        bool dump = false;
    };

    [[noreturn]] void usage(const char* why)
    {
        if (why) fprintf(stderr, "ipk-b200: %s\n", why);
        fprintf(stderr,
                "usage: ipk-b200 --ar-probs FILE --node-map FILE [--states dna|aa] [-k K] [--omega W] [--mif0|--random]\n"
                "                [--tree NEWICK] [--node-count N] [-w DIR] [-o DB.ipk] [--device N] [-j THREADS] [--dump]\n");
        exit(1);
    }

    options parse(int argc, char** argv)
    {
        options o;
        auto need = [&](int& i) -> const char* { if (i + 1 >= argc) usage((std::string("missing value for ") + argv[i]).c_str()); return argv[++i]; };
        for (int i = 1; i < argc; ++i)
        {
            const std::string a = argv[i];
            if (a == "--ar-probs") o.ar_probs = need(i);
            else if (a == "--node-map") o.node_map = need(i);
            else if (a == "--tree" || a == "-t" || a == "--reftree") o.tree_file = need(i);
            else if (a == "--states") o.states = need(i);
            else if (a == "-k" || a == "--k") o.k = uint32_t(atoi(need(i)));
            else if (a == "--omega") o.omega = float(atof(need(i)));
            else if (a == "--mif0") o.filter = IPKGPU_FILTER_MIF0;
            else if (a == "--random") o.filter = IPKGPU_FILTER_RANDOM;
            else if (a == "--node-count") o.node_count = uint32_t(atoll(need(i)));
            else if (a == "-w" || a == "--workdir") o.workdir = need(i);
            else if (a == "-o" || a == "--output") o.output = need(i);
            else if (a == "--device") o.device = atoi(need(i));
            else if (a == "-j" || a == "--num-threads") o.threads = atoi(need(i));
            else if (a == "--dump") o.dump = true;
            else if (a == "-u" || a == "--mu") need(i);                 // accepted and ignored, as the reference build does
            else if (a == "--help" || a == "-h") usage(nullptr);
            else usage(("unknown option " + a).c_str());
        }
        if (o.ar_probs.empty() || o.node_map.empty()) usage("--ar-probs and --node-map are required");
        if (o.states != "dna" && o.states != "aa") usage("--states must be dna or aa");
        if (o.output.empty()) o.output = o.workdir + "/DB.ipk";         // command_line.cpp:259-263
        return o;
    }

    double seconds_since(std::chrono::steady_clock::time_point t0)
    {
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
}

int main(int argc, char** argv)
{
    try
    {
        const options o = parse(argc, argv);
        const uint32_t sigma = o.states == "dna" ? 4 : 20, bits = o.states == "dna" ? 2 : 5;

        // ---- node map: processing order of the ghost nodes and the branch each belongs to (db_builder.cpp:495-553)
        std::vector<std::string> labels;
        std::vector<uint32_t> node_branch;
        {
            std::ifstream in(o.node_map);
            if (!in) throw std::runtime_error("cannot open " + o.node_map);
            std::string line;
            while (std::getline(in, line))
            {
                if (line.empty() || line[0] == '#') continue;
                const size_t tab = line.find('\t');
                if (tab == std::string::npos) throw std::runtime_error("node map line without a TAB: " + line);
                labels.push_back(line.substr(0, tab));
                node_branch.push_back(uint32_t(std::stoul(line.substr(tab + 1))));
            }
        }
        uint32_t max_branch = 0;
        for (uint32_t b : node_branch) max_branch = std::max(max_branch, b);
        const uint32_t node_count = o.node_count ? o.node_count : max_branch + 2;      // original_tree.get_node_count(): branches + root

        // ---- posteriors
        auto t0 = std::chrono::steady_clock::now();
        ipkhost_ar* ar = nullptr;
        if (ipkhost_ar_open(o.ar_probs.c_str(), sigma, o.threads, &ar) != IPKHOST_OK) throw std::runtime_error(ipkhost_last_error());
        const uint32_t width = ipkhost_ar_width(ar);
        std::vector<const char*> label_ptrs;
        for (auto& s : labels) label_ptrs.push_back(s.c_str());
        // page-locked, so that the library uploads in node chunks under its first kernels
        void* logp_mem = nullptr;
        if (ipkgpu_host_alloc(o.device, size_t(labels.size()) * width * sigma * sizeof(float), &logp_mem) != IPKGPU_OK)
            throw std::runtime_error("cannot allocate page-locked memory for the posteriors (no CUDA device?)");
        std::unique_ptr<float, void (*)(float*)> logp(static_cast<float*>(logp_mem), [](float* p) { ipkgpu_host_free(p); });
        if (ipkhost_ar_read_nodes(ar, label_ptrs.data(), uint32_t(labels.size()), logp.get()) != IPKHOST_OK)
            throw std::runtime_error(ipkhost_last_error());
        ipkhost_ar_close(ar);
        fprintf(stderr, "Loaded %zu ghost-node matrices x %u sites x %u states in %.3f s\n", labels.size(), width, sigma, seconds_since(t0));

        // ---- the device path
        t0 = std::chrono::steady_clock::now();
        ipkgpu_ctx* ctx = nullptr;
        if (ipkgpu_create(&ctx, o.device, nullptr) != IPKGPU_OK) throw std::runtime_error(ipkgpu_last_error(nullptr));
        ipkgpu_input in;
        memset(&in, 0, sizeof in);
        in.logp = logp.get();
        in.node_branch = node_branch.data();
        in.n_nodes = uint32_t(labels.size());
        in.width = width; in.sigma = sigma; in.bits = bits; in.k = o.k;
        ipkgpu_score_threshold(o.omega, sigma, o.k, &in.threshold, &in.log_threshold);
        in.total_num_groups = node_count;
        in.filter = o.filter;
        ipkgpu_result res;
        if (ipkgpu_build(ctx, &in, IPKGPU_RESULT_HOST_INDEXED | IPKGPU_RESULT_BRANCH16, &res) != IPKGPU_OK)
        {
            const std::string msg = ipkgpu_last_error(ctx);
            ipkgpu_destroy(ctx);
            throw std::runtime_error(msg);
        }
        fprintf(stderr, "Built %llu k-mers, %llu phylo-k-mers (%llu emitted over %llu windows) in %.3f s (device %.3f ms)\n",
                (unsigned long long)res.n_kmers, (unsigned long long)res.n_entries, (unsigned long long)res.stats.n_emitted,
                (unsigned long long)res.stats.n_windows, seconds_since(t0), double(res.stats.ms_total));

        // ---- serialize
        t0 = std::chrono::steady_clock::now();
        std::string tree;
        if (!o.tree_file.empty())
        {
            std::ifstream tf(o.tree_file);
            if (!tf) throw std::runtime_error("cannot open " + o.tree_file);
            std::stringstream ss;
            ss << tf.rdbuf();
            tree = ss.str();
            while (!tree.empty() && (tree.back() == '\n' || tree.back() == '\r')) tree.pop_back();
        }
        ipkhost_db db;
        memset(&db, 0, sizeof db);
        db.sequence_type = sigma == 4 ? "DNA" : "Proteins";
        db.tree = tree.c_str();
        db.kmer_size = o.k;
        db.omega = o.omega;
        db.n_kmers = res.n_kmers; db.n_entries = res.n_entries;
        db.keys = res.keys; db.fv = res.fv; db.offsets = res.offsets; db.counts = res.counts; db.branch = res.branch; db.branch16 = res.branch16; db.score = res.score;
        if (ipkhost_write_ipk(o.output.c_str(), &db) != IPKHOST_OK) throw std::runtime_error("cannot write " + o.output);
        fprintf(stderr, "Wrote %s in %.3f s\n", o.output.c_str(), seconds_since(t0));
        if (o.dump)
            for (uint64_t q = 0; q < res.n_kmers; ++q)
                for (uint64_t e = res.offsets[q]; e < (res.counts ? res.offsets[q] + res.counts[q] : res.offsets[q + 1]); ++e)
                    printf("%llu\t%.9g\t%u\t%.9g\n", (unsigned long long)res.keys[q], double(res.fv[q]), res.branch16 ? uint32_t(res.branch16[e]) : res.branch[e], double(res.score[e]));
        ipkgpu_free_result(ctx, &res);
        ipkgpu_destroy(ctx);
        return 0;
    }
    catch (const std::exception& e)
    {
        fprintf(stderr, "Exception: %s\n", e.what());
        return 1;
    }
}
