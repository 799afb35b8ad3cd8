// main.cpp -- ipk-b200: the database build of `ipk-dna` / `ipk-aa` (ipk/src/main.cpp:76-246, ipk::build) on a B200.
//
// Two ways in:
//  (1) the reference's own command line (ipk/src/command_line.cpp:83-147), exactly as ipk.py:292-329 spawns it:
//          ipk-dna --ar-binary X --refalign A.fasta -t TREE -w DIR -k K --model M --alpha A --categories C
//                  --reduction-ratio R --omega W --mif0|--random --both|--inner-only|--outer-only -u MU -j N -o OUT -v V
//                  [--ar-dir DIR] [--ar-only] [--ar-parameters "..."] [--no-reduction] [--use-unrooted] [--uncompressed] [--on-disk]
//      Ancestral reconstruction stays the upstream CPU stage it is in BASELINE.json: --ar-dir must point at raxml-ng's
//      `*.raxml.ancestralProbs` / `*.raxml.ancestralTree` (the reference's own resume point, ar.cpp:610-640); everything
//      after it is done here: original tree -> ghost nodes -> extended/AR label mapping -> ghost selection and grouping ->
//      posteriors -> libipkgpu -> tree index + DB.ipk.  Options that only steer raxml-ng or the alignment reduction are
//      parsed and ignored (the alignment is only needed to RUN the reconstruction); -u/--mu and --BB/--DC/--DCLA/--DCCW are
//      parsed and ignored as the reference build itself ignores them; --on-disk needs no flag here (the device
//      path batches by itself when the emitted volume exceeds its memory) and is accepted.
//      The alphabet comes from the executable's name like the reference's three binaries (ipk-dna, ipk-aa), or --states.
//  (2) the build half alone: --ar-probs FILE --node-map FILE (`AR-node-label <TAB> post-order id of the branch`, in
//      processing order), for callers that have done the tree work themselves.
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
#include <condition_variable>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <dirent.h>
#include <sys/stat.h>

namespace
{
    struct options
    {
        std::string ar_probs, node_map, tree_file, output, workdir = ".", ar_dir, refalign;
        std::string states;
        uint32_t k = 8;
        float omega = 1.5f;
        int filter = IPKGPU_FILTER_MIF0;
        uint32_t node_count = 0;
        int device = 0, threads = 0, verbosity = 1;
        std::string logp_f32;              // --logp-f32 FILE --width W: the matrices as raw float32 [node][site][state] in node-map order (benchmarks: no text to parse)
        uint32_t width = 0;
        std::string boundary_file;         // --boundary-list FILE: the entries that depend on which side of +-band the threshold falls
        float boundary_band = 1e-5f;
        std::vector<int> devices;          // --devices 0,1,...: one rank per device, the multi-GPU build of include/ipkgpu.h
        int ghosts = 0;                    // 0 both, 1 inner only, 2 outer only
        bool dump = false, use_unrooted = false, ar_only = false, merge_branches = false, keep_positions = false;
    };

    [[noreturn]] void usage(const char* why)
    {
        if (why) fprintf(stderr, "ipk-b200: %s\n", why);
        fprintf(stderr,
                "usage: ipk-b200 --refalign FASTA -t|--reftree NEWICK --ar-dir DIR [-k K] [--omega W] [--mif0|--random]\n"
                "                [--both|--inner-only|--outer-only] [--use-unrooted] [-w DIR] [-o DB.ipk] [--states dna|aa]\n"
                "                [-j THREADS] [-v N] [--device N | --devices 0,1,...] [--dump]   (and every other ipk-dna / ipk-aa option, see main.cpp)\n"
                "   or: ipk-b200 --ar-probs FILE --node-map FILE [--tree NEWICK] [--node-count N] [...]\n");
        exit(1);
    }

    options parse(int argc, char** argv)
    {
        options o;
        {
            // ipk-dna / ipk-aa / ipk-aa-pos pick the alphabet at compile time (ipk/CMakeLists.txt); here by the name invoked
            std::string self = argc ? argv[0] : "";
            const size_t slash = self.find_last_of('/');
            if (slash != std::string::npos) self = self.substr(slash + 1);
            o.keep_positions = self.find("-pos") != std::string::npos;       // ipk-aa-pos: the KEEP_POSITIONS binary (ipk/CMakeLists.txt; ipk.py:275-283)
            o.states = self.find("aa") != std::string::npos ? "aa" : "dna";
        }
        auto need = [&](int& i) -> const char* { if (i + 1 >= argc) usage((std::string("missing value for ") + argv[i]).c_str()); return argv[++i]; };
        for (int i = 1; i < argc; ++i)
        {
            const std::string a = argv[i];
            if (a == "--ar-probs") o.ar_probs = need(i);
            else if (a == "--node-map") o.node_map = need(i);
            else if (a == "--tree" || a == "-t" || a == "--reftree") o.tree_file = need(i);
            else if (a == "--refalign") o.refalign = need(i);
            else if (a == "--ar-dir") o.ar_dir = need(i);
            else if (a == "--states") o.states = need(i);
            else if (a == "-k" || a == "--k") o.k = uint32_t(atoi(need(i)));
            else if (a == "--omega") o.omega = float(atof(need(i)));
            else if (a == "--mif0") o.filter = IPKGPU_FILTER_MIF0;
            else if (a == "--random") o.filter = IPKGPU_FILTER_RANDOM;
            else if (a == "--both") o.ghosts = 0;
            else if (a == "--inner-only") o.ghosts = 1;
            else if (a == "--outer-only") o.ghosts = 2;
            else if (a == "--use-unrooted") o.use_unrooted = true;
            else if (a == "--ar-only") o.ar_only = true;
            else if (a == "--merge-branches") o.merge_branches = true;
            else if (a == "--keep-positions") o.keep_positions = true;
            else if (a == "--node-count") o.node_count = uint32_t(atoll(need(i)));
            else if (a == "-w" || a == "--workdir") o.workdir = need(i);
            else if (a == "-o" || a == "--output") o.output = need(i);
            else if (a == "--device") o.device = atoi(need(i));
            else if (a == "--devices")
            {
                const std::string v = need(i);
                for (size_t p = 0; p < v.size();) { size_t q = v.find(',', p); if (q == std::string::npos) q = v.size(); o.devices.push_back(atoi(v.substr(p, q - p).c_str())); p = q + 1; }
                if (o.devices.empty() || o.devices.size() > 16) usage("--devices takes 1 to 16 comma-separated device ordinals");
                o.device = o.devices[0];
            }
            else if (a == "-j" || a == "--num-threads") o.threads = atoi(need(i));
            else if (a == "-v" || a == "--verbosity") o.verbosity = atoi(need(i));
            else if (a == "--dump") o.dump = true;
            else if (a == "--logp-f32") o.logp_f32 = need(i);
            else if (a == "--width") o.width = uint32_t(atoll(need(i)));
            else if (a == "--boundary-list") o.boundary_file = need(i);
            else if (a == "--boundary-band") o.boundary_band = float(atof(need(i)));
            // parsed and ignored: they steer raxml-ng / the alignment reduction (upstream), or nothing at all in the reference build
            else if (a == "--ar-binary" || a == "-m" || a == "--model" || a == "-a" || a == "--alpha" || a == "--categories" ||
                     a == "--reduction-ratio" || a == "-u" || a == "--mu") need(i);
            else if (a == "--ar-parameters")
            {
                // ipk.py passes the value inside literal quotes, possibly split by the shell: swallow up to the closing quote (command_line.cpp:170-207)
                std::string v = need(i);
                if (!v.empty() && (v[0] == '"' || v[0] == '\''))
                {
                    const char q = v[0];
                    while ((v.size() < 2 || v.back() != q) && i + 1 < argc) v = argv[++i];
                }
            }
            else if (a == "--no-reduction" || a == "--uncompressed" || a == "--on-disk" || a == "--BB" || a == "--DC" || a == "--DCLA" || a == "--DCCW") {}
            else if (a == "--help" || a == "-h") usage(nullptr);
            else usage(("unknown option " + a).c_str());
        }
        if (o.states != "dna" && o.states != "aa") usage("--states must be dna or aa");
        if (!o.logp_f32.empty() && (o.node_map.empty() || o.width == 0)) usage("--logp-f32 needs --node-map and --width");
        if (!o.logp_f32.empty()) o.ar_probs = o.logp_f32;
        const bool reference_mode = o.ar_probs.empty() && o.node_map.empty();
        if (reference_mode)
        {
            if (o.tree_file.empty()) usage("the option '--reftree' is required but missing");       // boost::program_options' wording
            if (o.ar_dir.empty())
                usage("--ar-dir is required: ancestral reconstruction is the upstream CPU stage (run raxml-ng through the reference's --ar-only, "
                      "then point --ar-dir at its *.raxml.ancestralProbs / *.raxml.ancestralTree)");
        }
        else if (o.ar_probs.empty() || o.node_map.empty()) usage("--ar-probs and --node-map go together");
        // main.cpp:31-37 (check_parameters): ipk-dna / ipk-aa refuse it with exactly this message; only the KEEP_POSITIONS binary takes it
        if (o.merge_branches && !o.keep_positions) { fprintf(stderr, "Exception: --merge-branches is only supported for IPK compiled with the KEEP_POSITIONS flag.\n"); exit(1); }
        // (with positions the reference accepts the flag and then never reads it: db_builder.cpp stores _merge_branches and that is all)
        if (o.keep_positions && o.devices.size() > 1) { fprintf(stderr, "Exception: a KEEP_POSITIONS build runs on one device\n"); exit(1); }
        if (o.output.empty()) o.output = o.workdir + "/DB.ipk";         // command_line.cpp:259-263
        return o;
    }

    std::string read_text(const std::string& path)
    {
        std::ifstream in(path);
        if (!in) throw std::runtime_error("cannot open " + path);
        std::stringstream ss;
        ss << in.rdbuf();
        return ss.str();
    }

    /// find_file_by_suffix (ar.cpp:594-608); the lexicographically first match, so that the choice does not depend on readdir order
    std::string find_by_suffix(const std::string& dir, const std::string& suffix)
    {
        DIR* d = opendir(dir.c_str());
        if (!d) throw std::runtime_error("Error! No such directory: " + dir);
        std::string best;
        while (const dirent* e = readdir(d))
        {
            const std::string name = e->d_name;
            if (name.size() >= suffix.size() && name.compare(name.size() - suffix.size(), suffix.size(), suffix) == 0 && (best.empty() || name < best)) best = name;
        }
        closedir(d);
        if (best.empty()) throw std::runtime_error("Could not find \"*" + suffix + "\" in the folder provided by --ar-dir: " + dir);
        return dir + "/" + best;
    }

    double seconds_since(std::chrono::steady_clock::time_point t0)
    {
        return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
}

namespace
{
    /// all threads wait until `n` have arrived (C++17: no std::barrier)
    struct thread_barrier
    {
        std::mutex m; std::condition_variable cv; unsigned n, waiting = 0, phase = 0;
        explicit thread_barrier(unsigned n_) : n(n_) {}
        void wait()
        {
            std::unique_lock<std::mutex> l(m);
            const unsigned p = phase;
            if (++waiting == n) { waiting = 0; ++phase; cv.notify_all(); }
            else cv.wait(l, [&] { return phase != p; });
        }
    };

    struct shard_result
    {
        ipkgpu_ctx* ctx = nullptr;
        ipkgpu_result res;
        std::string error;
        ipkgpu_stats stage1;
    };

    /// The multi-GPU build of include/ipkgpu.h with the ranks as threads of this process, one per device: branch groups in
    /// contiguous blocks -> stage 1 -> the MSD pass of the merge sort stores each k-mer range into its owner's buffer over
    /// NVLink (direct peer access) -> every owner sorts / merges / filters its range -> shards in pinned host memory.
    /// The caller merges them by (filter value, key) while it serializes (merge_stage2, db_builder.cpp:392-458).
    void build_on_devices(const std::vector<int>& devices, const float* logp, const std::vector<uint32_t>& node_branch, uint32_t width, uint32_t sigma,
                          uint32_t bits, uint32_t k, float omega, uint32_t node_count, int filter, std::vector<shard_result>& out)
    {
        const uint32_t G = uint32_t(devices.size()), n_nodes = uint32_t(node_branch.size());
        // groups by first appearance; the nodes must come group after group so that a rank's matrices are one slice
        std::vector<uint32_t> group_branch, group_first_node;
        {
            std::unordered_map<uint32_t, uint32_t> seen;
            for (uint32_t i = 0; i < n_nodes; ++i)
            {
                const auto it = seen.find(node_branch[i]);
                if (it == seen.end()) { seen[node_branch[i]] = uint32_t(group_branch.size()); group_branch.push_back(node_branch[i]); group_first_node.push_back(i); }
                else if (it->second + 1 != group_branch.size()) throw std::runtime_error("--devices needs the ghost nodes listed branch group after branch group");
            }
            group_first_node.push_back(n_nodes);
        }
        const uint32_t n_groups = uint32_t(group_branch.size());
        uint32_t group_bits = 0;
        while ((1ull << group_bits) < n_groups) ++group_bits;
        const uint32_t key_bits = k * bits;
        if (key_bits < 8 || key_bits + group_bits - 8 > 32)
            throw std::runtime_error("--devices: this alphabet / k / tree size does not fit the packed exchange (k*bits >= 8, k*bits + group bits <= 40)");
        float threshold = 0, log_threshold = 0;
        ipkgpu_score_threshold(omega, sigma, k, &threshold, &log_threshold);
        out.assign(G, shard_result());
        std::vector<uint64_t> counts(size_t(G) * 256, 0);
        std::vector<uint64_t*> recv(G, nullptr);
        std::vector<int32_t> ordered(G, 0);
        thread_barrier bar(G);
        std::vector<std::thread> th;
        for (uint32_t r = 0; r < G; ++r)
            th.emplace_back([&, r] {
                shard_result& S = out[r];
                auto check = [&](int rc, const char* what) {
                    if (rc != IPKGPU_OK && S.error.empty()) S.error = std::string(what) + ": " + ipkgpu_last_error(S.ctx);
                    return rc == IPKGPU_OK;
                };
                bool ok = true;
                if (ipkgpu_create(&S.ctx, devices[r], nullptr) != IPKGPU_OK) { S.error = ipkgpu_last_error(nullptr); ok = false; }
                for (uint32_t p = 0; ok && p < G; ++p) ok = check(ipkgpu_enable_peer(S.ctx, devices[p]), "peer access");
                const uint32_t g0 = uint32_t(uint64_t(n_groups) * r / G), g1 = uint32_t(uint64_t(n_groups) * (r + 1) / G);
                const uint32_t n0 = group_first_node[g0], n1 = group_first_node[g1];
                ipkgpu_input in;
                memset(&in, 0, sizeof in);
                in.logp = logp + size_t(n0) * width * sigma;
                in.node_branch = node_branch.data() + n0;
                in.n_nodes = n1 - n0;
                in.width = width; in.sigma = sigma; in.bits = bits; in.k = k;
                in.threshold = threshold; in.log_threshold = log_threshold;
                in.total_num_groups = node_count; in.filter = filter;
                uint64_t n_raw = 0;
                memset(&S.stage1, 0, sizeof S.stage1);
                if (ok) ok = check(ipkgpu_stage1_raw(S.ctx, &in, g0, group_bits, &n_raw, &S.stage1), "stage 1");
                if (ok) ok = check(ipkgpu_stage1_bucket_counts(S.ctx, counts.data() + size_t(r) * 256), "bucket counts");
                bar.wait();                                    // the G x 256 matrix is complete (a failed rank contributed zeros)
                uint8_t owner[256]; uint64_t base[256], start[257], n_recv = 0;
                if (ipkgpu_exchange_plan(G, r, counts.data(), owner, base, start, &n_recv) != IPKGPU_OK && S.error.empty()) { S.error = "a key range exceeds 2^32-1 records"; ok = false; }
                if (ok) ok = check(ipkgpu_recv_packed(S.ctx, n_recv ? n_recv : 1, &recv[r], nullptr), "receive buffer");
                bar.wait();                                    // every rank's buffer exists
                bool all_ok = true;
                for (uint32_t p = 0; p < G; ++p) all_ok = all_ok && out[p].error.empty();
                if (ok && all_ok) ok = check(ipkgpu_stage1_scatter_packed(S.ctx, G, recv.data(), owner, base, &ordered[r]), "exchange");
                bar.wait();                                    // every rank's stores have landed
                int32_t all_ordered = 1;
                for (uint32_t p = 0; p < G; ++p) { all_ordered &= ordered[p]; all_ok = all_ok && out[p].error.empty(); }
                if (ok && all_ok)
                    check(ipkgpu_stage2_packed(S.ctx, n_recv, start, group_bits, all_ordered, group_branch.data(), n_groups, key_bits, threshold, node_count, filter,
                                               IPKGPU_RESULT_HOST_INDEXED | IPKGPU_RESULT_BRANCH16, &S.res), "stage 2");
            });
        for (auto& t : th) t.join();
        for (uint32_t r = 0; r < G; ++r)
            if (!out[r].error.empty())
            {
                const std::string msg = "device " + std::to_string(devices[r]) + ": " + out[r].error;
                for (auto& s : out) if (s.ctx) ipkgpu_destroy(s.ctx);
                throw std::runtime_error(msg);
            }
    }
}

int main(int argc, char** argv)
{
    try
    {
        const options o = parse(argc, argv);
        const uint32_t sigma = o.states == "dna" ? 4 : 20, bits = o.states == "dna" ? 2 : 5;

        std::vector<std::string> labels;
        std::vector<uint32_t> node_branch;
        std::vector<uint64_t> index_num_nodes;
        std::vector<double> index_lengths;
        std::string tree, ar_probs = o.ar_probs;
        uint32_t node_count = 0;
        if (o.node_map.empty())
        {
            // ---- the reference's path from the tree to ipk::build's arguments (main.cpp:145-198)
            ar_probs = find_by_suffix(o.ar_dir, ".raxml.ancestralProbs");
            const std::string ar_tree_file = find_by_suffix(o.ar_dir, ".raxml.ancestralTree");
            if (o.verbosity > 0)
                fprintf(stderr, "Ancestral reconstruction results have been found:\n\t%s\n\t%s\n", ar_probs.c_str(), ar_tree_file.c_str());
            if (o.ar_only)
            {
                if (o.verbosity > 0) fprintf(stderr, "--ar-only requested. Finishing after ancestral reconstruction.\n");
                return 0;
            }
            ipkhost_plan* plan = nullptr;
            if (ipkhost_plan_create(read_text(o.tree_file).c_str(), read_text(ar_tree_file).c_str(), o.ghosts, o.use_unrooted ? 1 : 0, &plan) != IPKHOST_OK)
                throw std::runtime_error(ipkhost_last_error());
            const uint32_t n = ipkhost_plan_num_ghosts(plan);
            for (uint32_t i = 0; i < n; ++i) labels.push_back(ipkhost_plan_ar_label(plan, i));
            node_branch.assign(ipkhost_plan_node_branch(plan), ipkhost_plan_node_branch(plan) + n);
            node_count = ipkhost_plan_node_count(plan);
            index_num_nodes.assign(ipkhost_plan_index_num_nodes(plan), ipkhost_plan_index_num_nodes(plan) + node_count);
            index_lengths.assign(ipkhost_plan_index_lengths(plan), ipkhost_plan_index_lengths(plan) + node_count);
            tree = ipkhost_plan_newick(plan);
            // save_extended_tree (main.cpp:39-46)
            const std::string ext_dir = o.workdir + "/extended_trees";
            mkdir(ext_dir.c_str(), 0777);
            std::ofstream(ext_dir + "/extended_tree.newick") << ipkhost_plan_extended_newick(plan) << "\n";
            if (o.verbosity > 0) fprintf(stderr, "%u ghost nodes in %u branch groups of a %u-node tree\n", n, ipkhost_plan_num_groups(plan), node_count);
            ipkhost_plan_free(plan);
        }
        else
        {
            // ---- node map: processing order of the ghost nodes and the branch each belongs to (db_builder.cpp:495-553)
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
            uint32_t max_branch = 0;
            for (uint32_t b : node_branch) max_branch = std::max(max_branch, b);
            node_count = o.node_count ? o.node_count : max_branch + 2;      // original_tree.get_node_count(): branches + root
            if (!o.tree_file.empty())
            {
                // with a tree the header gets its index and the canonical newick; without, both stay empty
                tree = read_text(o.tree_file);
                ipkhost_plan* plan = nullptr;
                if (ipkhost_plan_create(tree.c_str(), nullptr, 0, 1, &plan) == IPKHOST_OK)
                {
                    const uint32_t nc = ipkhost_plan_node_count(plan);
                    index_num_nodes.assign(ipkhost_plan_index_num_nodes(plan), ipkhost_plan_index_num_nodes(plan) + nc);
                    index_lengths.assign(ipkhost_plan_index_lengths(plan), ipkhost_plan_index_lengths(plan) + nc);
                    tree = ipkhost_plan_newick(plan);
                    if (!o.node_count) node_count = nc;
                    ipkhost_plan_free(plan);
                }
                else while (!tree.empty() && (tree.back() == '\n' || tree.back() == '\r')) tree.pop_back();
            }
        }

        // ---- posteriors
        auto t0 = std::chrono::steady_clock::now();
        uint32_t width = o.width;
        void* logp_mem = nullptr;
        std::unique_ptr<float, void (*)(float*)> logp(nullptr, [](float* p) { ipkgpu_host_free(p); });
        if (!o.logp_f32.empty())
        {
            const size_t bytes = size_t(labels.size()) * width * sigma * sizeof(float);
            if (ipkgpu_host_alloc(o.device, bytes, &logp_mem) != IPKGPU_OK)
                throw std::runtime_error("cannot allocate page-locked memory for the posteriors (no CUDA device?)");
            logp.reset(static_cast<float*>(logp_mem));
            FILE* f = fopen(o.logp_f32.c_str(), "rb");
            if (!f || fread(logp.get(), 1, bytes, f) != bytes) { if (f) fclose(f); throw std::runtime_error("cannot read " + std::to_string(bytes) + " bytes from " + o.logp_f32); }
            fclose(f);
        }
        else
        {
            ipkhost_ar* ar = nullptr;
            if (ipkhost_ar_open(ar_probs.c_str(), sigma, o.threads, &ar) != IPKHOST_OK) throw std::runtime_error(ipkhost_last_error());
            width = ipkhost_ar_width(ar);
            std::vector<const char*> label_ptrs;
            for (auto& s : labels) label_ptrs.push_back(s.c_str());
            // page-locked, so that the library uploads in node chunks under its first kernels
            if (ipkgpu_host_alloc(o.device, size_t(labels.size()) * width * sigma * sizeof(float), &logp_mem) != IPKGPU_OK)
                throw std::runtime_error("cannot allocate page-locked memory for the posteriors (no CUDA device?)");
            logp.reset(static_cast<float*>(logp_mem));
            if (ipkhost_ar_read_nodes(ar, label_ptrs.data(), uint32_t(labels.size()), logp.get()) != IPKHOST_OK)
                throw std::runtime_error(ipkhost_last_error());
            ipkhost_ar_close(ar);
        }
        fprintf(stderr, "Loaded %zu ghost-node matrices x %u sites x %u states in %.3f s\n", labels.size(), width, sigma, seconds_since(t0));

        // ---- the device path
        t0 = std::chrono::steady_clock::now();
        ipkhost_db header;
        memset(&header, 0, sizeof header);
        header.sequence_type = sigma == 4 ? "DNA" : "Proteins";
        header.tree = tree.c_str();
        header.n_index = index_num_nodes.size();
        header.index_num_nodes = index_num_nodes.data();
        header.index_subtree_branch_length = index_lengths.data();
        header.kmer_size = o.k;
        header.omega = o.omega;
        const char* const layout_note = "(record layout as documented in ipkhost.h: UNVERIFIED against i2l's serializer, whose source is not available)";
        if (o.devices.size() > 1)
        {
            std::vector<shard_result> shards;
            build_on_devices(o.devices, logp.get(), node_branch, width, sigma, bits, o.k, o.omega, node_count, o.filter, shards);
            std::vector<ipkhost_db> views(shards.size());
            uint64_t n_kmers = 0, n_entries = 0, n_emitted = 0;
            for (size_t r = 0; r < shards.size(); ++r)
            {
                const ipkgpu_result& res = shards[r].res;
                ipkhost_db& v = views[r];
                memset(&v, 0, sizeof v);
                v.n_kmers = res.n_kmers; v.n_entries = res.n_entries;
                v.keys = res.keys; v.fv = res.fv; v.offsets = res.offsets; v.counts = res.counts; v.branch = res.branch; v.branch16 = res.branch16; v.score = res.score;
                n_kmers += res.n_kmers; n_entries += res.n_entries; n_emitted += shards[r].stage1.n_emitted;
            }
            fprintf(stderr, "Built %llu k-mers, %llu phylo-k-mers (%llu emitted) on %zu devices in %.3f s\n", (unsigned long long)n_kmers,
                    (unsigned long long)n_entries, (unsigned long long)n_emitted, shards.size(), seconds_since(t0));
            // ---- N-way merge by (filter value, key) while serializing (merge_stage2, db_builder.cpp:392-458)
            t0 = std::chrono::steady_clock::now();
            std::vector<uint8_t> order_shard(n_kmers);
            std::vector<uint32_t> order_index(n_kmers);
            // the shards are ascending k-mer ranges (packed exchange): one stable device sort of the filter values gives the order
            // (ipkgpu_merge_order on the first device)
            {
                std::vector<const float*> fvs(views.size());
                std::vector<const uint64_t*> kys(views.size());
                std::vector<uint64_t> cnt(views.size());
                for (size_t r = 0; r < views.size(); ++r) { fvs[r] = views[r].fv; kys[r] = views[r].keys; cnt[r] = views[r].n_kmers; }
                if (ipkgpu_merge_order(shards[0].ctx, uint32_t(views.size()), fvs.data(), kys.data(), cnt.data(), 1, o.k * bits,
                                       order_shard.data(), order_index.data()) != IPKGPU_OK)
                    throw std::runtime_error(std::string("cannot merge the shards: ") + ipkgpu_last_error(shards[0].ctx));
            }
            if (ipkhost_write_ipk_sharded(o.output.c_str(), &header, uint32_t(views.size()), views.data(), order_shard.data(), order_index.data(), o.threads) != IPKHOST_OK)
                throw std::runtime_error("cannot write " + o.output);
            fprintf(stderr, "Wrote %s in %.3f s %s\n", o.output.c_str(), seconds_since(t0), layout_note);
            if (o.dump)
                for (uint64_t i = 0; i < n_kmers; ++i)
                {
                    const ipkhost_db& v = views[order_shard[i]];
                    const uint64_t q = order_index[i];
                    for (uint64_t e = v.offsets[q]; e < v.offsets[q] + v.counts[q]; ++e)
                        printf("%llu\t%.9g\t%u\t%.9g\n", (unsigned long long)v.keys[q], double(v.fv[q]), v.branch16 ? uint32_t(v.branch16[e]) : v.branch[e], double(v.score[e]));
                }
            for (auto& s : shards) ipkgpu_destroy(s.ctx);
            return 0;
        }
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
        if (o.keep_positions) ipkgpu_set_keep_positions(ctx, 1);
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
        ipkhost_db db = header;
        db.n_kmers = res.n_kmers; db.n_entries = res.n_entries;
        db.keys = res.keys; db.fv = res.fv; db.offsets = res.offsets; db.counts = res.counts; db.branch = res.branch; db.branch16 = res.branch16; db.score = res.score;
        if (o.keep_positions)
        {
            int32_t pos_on_device = 0;
            ipkgpu_result_positions(ctx, &db.position, &pos_on_device);
            if (res.n_entries && (!db.position || pos_on_device)) throw std::runtime_error("the build kept no positions");
        }
        if (ipkhost_write_ipk_mt(o.output.c_str(), &db, o.threads) != IPKHOST_OK) throw std::runtime_error("cannot write " + o.output);
        fprintf(stderr, "Wrote %s in %.3f s %s\n", o.output.c_str(), seconds_since(t0), layout_note);
        if (o.dump)
            for (uint64_t q = 0; q < res.n_kmers; ++q)
                for (uint64_t e = res.offsets[q]; e < (res.counts ? res.offsets[q] + res.counts[q] : res.offsets[q + 1]); ++e)
                    printf("%llu\t%.9g\t%u\t%.9g\n", (unsigned long long)res.keys[q], double(res.fv[q]), res.branch16 ? uint32_t(res.branch16[e]) : res.branch[e], double(res.score[e]));
        if (!o.boundary_file.empty())
        {
            // BASELINE.json: k-mers within 1e-5 of the threshold are listed separately (the threshold's rounding is i2l's, unpinned here)
            const uint64_t cap = 1ull << 22;
            std::vector<uint64_t> bk(cap);
            std::vector<uint32_t> bb(cap);
            std::vector<float> bl(cap), bt(cap);
            uint64_t nb = 0;
            if (ipkgpu_boundary_list(ctx, &in, o.boundary_band, cap, bk.data(), bb.data(), bl.data(), bt.data(), &nb) != IPKGPU_OK)
                throw std::runtime_error(std::string("boundary list: ") + ipkgpu_last_error(ctx));
            FILE* bf = fopen(o.boundary_file.c_str(), "w");
            if (!bf) throw std::runtime_error("cannot write " + o.boundary_file);
            fprintf(bf, "# entries whose presence or score changes when log10(threshold) = %.9g moves by less than %g: %llu\n# k-mer code\tbranch\tscore at threshold-band\tscore at threshold+band (nan: absent)\n",
                    double(in.log_threshold), double(o.boundary_band), (unsigned long long)nb);
            for (uint64_t i = 0; i < std::min(nb, cap); ++i) fprintf(bf, "%llu\t%u\t%.9g\t%.9g\n", (unsigned long long)bk[i], bb[i], double(bl[i]), double(bt[i]));
            fclose(bf);
            fprintf(stderr, "Boundary cases (within %g of the threshold): %llu, listed in %s\n", double(o.boundary_band), (unsigned long long)nb, o.boundary_file.c_str());
        }
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
