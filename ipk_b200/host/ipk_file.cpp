// ipk_file.cpp -- .ipk database writer / reader of the B200 build's host side.
// Writer = the serialization walk of db_builder::filter_in_ram (ipk/src/db_builder.cpp:297-306 header, :323-332 one
// save_phylo_kmer per k-mer in filter-value order).  The byte layout of i2l::save_header / save_phylo_kmer is
// UNVERIFIED here (the i2l submodule is absent); ipkhost.h documents the layout adopted.  Output is
// buffered in 8 MiB blocks: the reference's per-field Boost archive calls are its stage-3 bottleneck.
#include "ipkhost.h"

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace
{

    const char PREAMBLE[] = "serialization::archive";
    const uint16_t ARCHIVE_VERSION = 19;

    struct writer
    {
        FILE* f;
        std::vector<unsigned char> buf;
        bool ok = true;
        explicit writer(FILE* f_) : f(f_) { buf.reserve(8u << 20); }
        void flush()
        {
            if (!buf.empty() && fwrite(buf.data(), 1, buf.size(), f) != buf.size()) ok = false;
            buf.clear();
        }
        void raw(const void* p, size_t n)
        {
            if (buf.size() + n > buf.capacity()) flush();
            if (n > buf.capacity()) { if (fwrite(p, 1, n, f) != n) ok = false; return; }
            const unsigned char* b = static_cast<const unsigned char*>(p);
            buf.insert(buf.end(), b, b + n);
        }
        template<class T> void put(T v) { raw(&v, sizeof v); }
        void str(const char* s) { const uint64_t n = s ? strlen(s) : 0; put(n); if (n) raw(s, n); }
    };

    struct reader
    {
        FILE* f;
        bool ok = true;
        template<class T> T get() { T v{}; if (fread(&v, sizeof v, 1, f) != 1) ok = false; return v; }
        std::string str()
        {
            const uint64_t n = get<uint64_t>();
            if (!ok || n > (1ull << 32)) { ok = false; return {}; }
            std::string s(n, '\0');
            if (n && fread(&s[0], 1, n, f) != n) ok = false;
            return s;
        }
    };
}

struct ipkhost_db_file
{
    std::string sequence_type, tree;
    std::vector<uint64_t> index_num_nodes, keys, offsets;
    std::vector<double> index_len;
    std::vector<float> fv, score;
    std::vector<uint32_t> branch;
};

extern "C"
{

int ipkhost_write_ipk(const char* path, const ipkhost_db* db)
{
    if (!path || !db) return IPKHOST_EINVAL;
    if (db->n_kmers && (!db->keys || !db->fv || !db->offsets)) return IPKHOST_EINVAL;
    if (db->n_entries && ((!db->branch && !db->branch16) || !db->score)) return IPKHOST_EINVAL;
    FILE* f = fopen(path, "wb");
    if (!f) return IPKHOST_EIO;
    writer w(f);
    w.put<uint64_t>(sizeof PREAMBLE - 1);
    w.raw(PREAMBLE, sizeof PREAMBLE - 1);
    w.put<uint16_t>(ARCHIVE_VERSION);
    // ipk_header (db_builder.cpp:297-305): sequence_type, tree_index, tree, kmer_size, omega, num_kmers, num_entries
    w.str(db->sequence_type);
    w.put<uint64_t>(db->n_index);
    for (uint64_t i = 0; i < db->n_index; ++i)
    {
        w.put<uint64_t>(db->index_num_nodes[i]);
        w.put<double>(db->index_subtree_branch_length[i]);
    }
    w.str(db->tree);
    w.put<uint64_t>(db->kmer_size);
    w.put<float>(db->omega);
    w.put<uint64_t>(db->n_kmers);
    w.put<uint64_t>(db->n_entries);
    // save_phylo_kmer(kmer, filter_value, entries) in the order handed over (db_builder.cpp:323-332)
    for (uint64_t q = 0; q < db->n_kmers; ++q)
    {
        const uint64_t a = db->offsets[q], b = db->counts ? a + db->counts[q] : db->offsets[q + 1];
        w.put<uint64_t>(db->keys[q]);
        w.put<float>(db->fv[q]);
        w.put<uint64_t>(b - a);
        for (uint64_t e = a; e < b; ++e) { w.put<uint32_t>(db->branch16 ? uint32_t(db->branch16[e]) : db->branch[e]); w.put<float>(db->score[e]); }
    }
    w.flush();
    const bool ok = w.ok && fclose(f) == 0;
    return ok ? IPKHOST_OK : IPKHOST_EIO;
}

int ipkhost_read_ipk(const char* path, ipkhost_db_file** out, ipkhost_db* view)
{
    if (!path || !out || !view) return IPKHOST_EINVAL;
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return IPKHOST_EIO;
    setvbuf(f, nullptr, _IOFBF, 8u << 20);
    reader r{ f };
    auto d = new ipkhost_db_file();
    const std::string pre = r.str();
    const uint16_t ver = r.get<uint16_t>();
    if (!r.ok || pre != PREAMBLE || ver != ARCHIVE_VERSION) { fclose(f); delete d; return IPKHOST_EFORMAT; }
    d->sequence_type = r.str();
    const uint64_t n_index = r.get<uint64_t>();
    if (!r.ok || n_index > (1ull << 32)) { fclose(f); delete d; return IPKHOST_EFORMAT; }
    d->index_num_nodes.resize(n_index);
    d->index_len.resize(n_index);
    for (uint64_t i = 0; i < n_index; ++i) { d->index_num_nodes[i] = r.get<uint64_t>(); d->index_len[i] = r.get<double>(); }
    d->tree = r.str();
    const uint64_t kmer_size = r.get<uint64_t>();
    const float omega = r.get<float>();
    const uint64_t n_kmers = r.get<uint64_t>(), n_entries = r.get<uint64_t>();
    if (!r.ok || n_kmers > (1ull << 40) || n_entries > (1ull << 40)) { fclose(f); delete d; return IPKHOST_EFORMAT; }
    d->keys.resize(n_kmers); d->fv.resize(n_kmers); d->offsets.resize(n_kmers + 1);
    d->branch.resize(n_entries); d->score.resize(n_entries);
    uint64_t e = 0;
    for (uint64_t q = 0; q < n_kmers && r.ok; ++q)
    {
        d->keys[q] = r.get<uint64_t>();
        d->fv[q] = r.get<float>();
        const uint64_t n = r.get<uint64_t>();
        d->offsets[q] = e;
        if (!r.ok || e + n > n_entries) { r.ok = false; break; }
        for (uint64_t i = 0; i < n; ++i, ++e) { d->branch[e] = r.get<uint32_t>(); d->score[e] = r.get<float>(); }
    }
    d->offsets[n_kmers] = e;
    fclose(f);
    if (!r.ok || e != n_entries) { delete d; return IPKHOST_EFORMAT; }
    view->sequence_type = d->sequence_type.c_str();
    view->tree = d->tree.c_str();
    view->n_index = n_index;
    view->index_num_nodes = d->index_num_nodes.data();
    view->index_subtree_branch_length = d->index_len.data();
    view->kmer_size = kmer_size;
    view->omega = omega;
    view->n_kmers = n_kmers; view->n_entries = n_entries;
    view->keys = d->keys.data(); view->fv = d->fv.data(); view->offsets = d->offsets.data();
    view->branch = d->branch.data(); view->score = d->score.data();
    view->counts = nullptr;
    view->branch16 = nullptr;
    *out = d;
    return IPKHOST_OK;
}

void ipkhost_free_ipk(ipkhost_db_file* f) { delete f; }

}   // extern "C"
