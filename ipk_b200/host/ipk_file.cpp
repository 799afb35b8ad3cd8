// ipk_file.cpp -- .ipk database writer / reader of the B200 build's host side.
// Writer = the serialization walk of db_builder::filter_in_ram (ipk/src/db_builder.cpp:297-306 header, :323-332 one
// save_phylo_kmer per k-mer in filter-value order).  The byte layout of i2l::save_header / save_phylo_kmer is
// UNVERIFIED here (the i2l submodule is absent); ipkhost.h documents the layout adopted.  Output is
// buffered in 8 MiB blocks: the reference's per-field Boost archive calls are its stage-3 bottleneck.
#include "ipkhost.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <unistd.h>
#include <sys/mman.h>

namespace
{

    const char PREAMBLE[] = "serialization::archive";
    const uint16_t ARCHIVE_VERSION = 19;

    struct writer
    {
        FILE* f;
        std::vector<unsigned char> buf;
        bool ok = true;
        uint64_t total = 0;               // bytes flushed so far
        explicit writer(FILE* f_) : f(f_) { buf.reserve(8u << 20); }
        void flush()
        {
            if (!buf.empty() && fwrite(buf.data(), 1, buf.size(), f) != buf.size()) ok = false;
            total += buf.size();
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
    std::vector<uint32_t> branch, position;
};

namespace
{
    /// bytes of the part of the file before the first k-mer record
    void put_header(writer& w, const ipkhost_db* db)
    {
        // boost::archive::binary_oarchive with default flags (db_builder.cpp:176-177): basic_binary_oarchive::init writes the
        // signature string and the library version, basic_binary_oprimitive::init the sizes of int / long / float / double as
        // one byte each and int(1) for the endianness check.
        w.put<uint64_t>(sizeof PREAMBLE - 1);
        w.raw(PREAMBLE, sizeof PREAMBLE - 1);
        w.put<uint16_t>(ARCHIVE_VERSION);
        w.put<uint8_t>(sizeof(int)); w.put<uint8_t>(sizeof(long)); w.put<uint8_t>(sizeof(float)); w.put<uint8_t>(sizeof(double));
        w.put<int32_t>(1);
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
    }

    inline uint64_t entries_of(const ipkhost_db* db, uint64_t q) { return db->counts ? db->counts[q] : db->offsets[q + 1] - db->offsets[q]; }
    inline size_t entry_bytes(const ipkhost_db* db) { return db->position ? 10 : 8; }

    /// where the k-mer at output position q lives: one database, or k-mer order_index[q] of shard order_shard[q]
    struct source
    {
        const ipkhost_db* dbs;
        const uint8_t* order_shard;        // null: dbs[0], position q itself
        const uint32_t* order_index;
        const ipkhost_db* db(uint64_t q) const { return order_shard ? dbs + order_shard[q] : dbs; }
        uint64_t at(uint64_t q) const { return order_shard ? order_index[q] : q; }
    };

    /// output positions [q0, q1) formatted into private blocks and written at file offset `at` (record = 20 + 8 n bytes)
    bool write_range(int fd, const source& S, uint64_t q0, uint64_t q1, uint64_t at)
    {
        const size_t BLOCK = 8u << 20;
        std::vector<unsigned char> buf(BLOCK + 64);
        size_t fill = 0;
        auto flush = [&]() -> bool {
            size_t done = 0;
            while (done < fill)
            {
                const ssize_t r = pwrite(fd, buf.data() + done, fill - done, off_t(at + done));
                if (r <= 0) return false;
                done += size_t(r);
            }
            at += fill; fill = 0;
            return true;
        };
        for (uint64_t q = q0; q < q1; ++q)
        {
            const ipkhost_db* db = S.db(q);
            const uint64_t i = S.at(q);
            const uint64_t a = db->offsets[i], n = entries_of(db, i);
            if (fill + 20 > BLOCK && !flush()) return false;
            unsigned char* p = buf.data() + fill;
            memcpy(p, &db->keys[i], 8); memcpy(p + 8, &db->fv[i], 4); memcpy(p + 12, &n, 8);
            fill += 20;
            for (uint64_t e = a; e < a + n;)
            {
                if (fill + 16 > BLOCK && !flush()) return false;
                const size_t eb = entry_bytes(db);
                const uint64_t room = (BLOCK - fill) / eb, m = std::min<uint64_t>(room, a + n - e);
                unsigned char* o = buf.data() + fill;
                for (uint64_t j = 0; j < m; ++j)
                {
                    const uint32_t b = db->branch16 ? uint32_t(db->branch16[e + j]) : db->branch[e + j];
                    memcpy(o + eb * j, &b, 4); memcpy(o + eb * j + 4, &db->score[e + j], 4);
                    if (db->position) { const uint16_t p16 = uint16_t(db->position[e + j]); memcpy(o + eb * j + 8, &p16, 2); }
                }
                fill += size_t(m) * eb;
                e += m;
            }
        }
        return flush();
    }

    /// output positions [q0, q1) formatted straight into the mapped file at `dst`
    void format_range(unsigned char* dst, const source& S, uint64_t q0, uint64_t q1)
    {
        for (uint64_t q = q0; q < q1; ++q)
        {
            const ipkhost_db* db = S.db(q);
            const uint64_t i = S.at(q);
            const uint64_t a = db->offsets[i], n = entries_of(db, i);
            memcpy(dst, &db->keys[i], 8); memcpy(dst + 8, &db->fv[i], 4); memcpy(dst + 12, &n, 8);
            dst += 20;
            const float* sc = db->score + a;
            if (db->position)
            {
                // KEEP_POSITIONS entries: {branch u32, score f32, position u16} (i2l::phylo_kmer::pos_type; UNVERIFIED like the rest)
                for (uint64_t j = 0; j < n; ++j)
                {
                    const uint32_t b = db->branch16 ? uint32_t(db->branch16[a + j]) : db->branch[a + j];
                    const uint16_t p16 = uint16_t(db->position[a + j]);
                    memcpy(dst + 10 * j, &b, 4); memcpy(dst + 10 * j + 4, &sc[j], 4); memcpy(dst + 10 * j + 8, &p16, 2);
                }
                dst += 10 * n;
                continue;
            }
            if (db->branch16)
            {
                const uint16_t* br = db->branch16 + a;
                for (uint64_t j = 0; j < n; ++j) { const uint32_t b = br[j]; memcpy(dst + 8 * j, &b, 4); memcpy(dst + 8 * j + 4, &sc[j], 4); }
            }
            else
            {
                const uint32_t* br = db->branch + a;
                for (uint64_t j = 0; j < n; ++j) { memcpy(dst + 8 * j, &br[j], 4); memcpy(dst + 8 * j + 4, &sc[j], 4); }
            }
            dst += 8 * n;
        }
    }

    /// header + all records of `S` (n_kmers output positions, n_entries entries in total) with T threads
    int write_all(const char* path, const ipkhost_db* header, const source& S, uint64_t n_kmers, uint64_t n_entries, int n_threads)
    {
        FILE* f = fopen(path, "w+b");           // read-write: the records are written through a shared mapping
        if (!f) return IPKHOST_EIO;
        uint64_t header_bytes = 0;
        {
            ipkhost_db h = *header;
            h.n_kmers = n_kmers; h.n_entries = n_entries;
            writer w(f);
            put_header(w, &h);
            header_bytes = w.total + w.buf.size();
            w.flush();
            if (!w.ok || fflush(f) != 0) { fclose(f); return IPKHOST_EIO; }
        }
        // save_phylo_kmer(kmer, filter_value, entries) in the order handed over (db_builder.cpp:323-332): the k-mers are cut
        // into runs of about equal bytes, one thread formats and pwrites each run at its prefix-summed offset
        unsigned T = n_threads > 0 ? unsigned(n_threads) : std::max(1u, std::thread::hardware_concurrency());
        T = unsigned(std::min<uint64_t>(T, std::max<uint64_t>(1, n_kmers / 4096)));
        const size_t eb = entry_bytes(S.dbs);
        const uint64_t total = n_kmers * 20 + n_entries * eb;
        std::vector<uint64_t> cut(T + 1, n_kmers), at(T + 1, total);
        cut[0] = 0; at[0] = 0;
        {
            uint64_t acc = 0, next = 1;
            for (uint64_t q = 0; q < n_kmers && next < T; ++q)
            {
                while (next < T && acc >= total / T * next) { cut[next] = q; at[next] = acc; ++next; }
                acc += 20 + eb * entries_of(S.db(q), S.at(q));
            }
            // cuts never reached (the last k-mers are long) stay at the end: empty ranges
        }
        const int fd = fileno(f);
        std::vector<char> ok(T, 1);
        std::vector<std::thread> th;
        // Buffered writes to ONE file serialize on its inode lock, whatever the number of threads (measured: 3.3 GB/s on tmpfs
        // with 16 threads).  The file is therefore sized up front and mapped, and every thread formats its run straight into the
        // mapping: page faults are per page and run in parallel, and the private blocks are gone.  Files that cannot be mapped
        // (pipes, some network mounts) take the pwrite path.
        void* map = MAP_FAILED;
        if (total && ftruncate(fd, off_t(header_bytes + total)) == 0)
            map = mmap(nullptr, size_t(header_bytes + total), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (map != MAP_FAILED)
        {
            unsigned char* base = static_cast<unsigned char*>(map) + header_bytes;
            for (unsigned t = 0; t < T; ++t)
                th.emplace_back([&, t] { format_range(base + at[t], S, cut[t], cut[t + 1]); });
            for (auto& x : th) x.join();
            if (munmap(map, size_t(header_bytes + total)) != 0) ok[0] = 0;
        }
        else
        {
            for (unsigned t = 0; t < T; ++t)
                th.emplace_back([&, t] { ok[t] = write_range(fd, S, cut[t], cut[t + 1], header_bytes + at[t]) ? 1 : 0; });
            for (auto& x : th) x.join();
        }
        bool good = true;
        for (char c : ok) good = good && c;
        good = fclose(f) == 0 && good;
        return good ? IPKHOST_OK : IPKHOST_EIO;
    }

    /// the order the device sorts filter values in (filter.cuh: k5_order_keys): sign-flipped bit image, ties by key
    inline uint32_t fv_image(float f) { uint32_t b; memcpy(&b, &f, 4); return b ^ ((b >> 31) ? 0xffffffffu : 0x80000000u); }
    inline bool head_less(const ipkhost_db& a, uint64_t i, const ipkhost_db& b, uint64_t j)
    {
        const uint32_t x = fv_image(a.fv[i]), y = fv_image(b.fv[j]);
        return x != y ? x < y : a.keys[i] < b.keys[j];
    }
}

extern "C"
{

int ipkhost_write_ipk_mt(const char* path, const ipkhost_db* db, int n_threads)
{
    if (!path || !db) return IPKHOST_EINVAL;
    if (db->n_kmers && (!db->keys || !db->fv || !db->offsets)) return IPKHOST_EINVAL;
    if (db->n_entries && ((!db->branch && !db->branch16) || !db->score)) return IPKHOST_EINVAL;
    const source S{ db, nullptr, nullptr };
    return write_all(path, db, S, db->n_kmers, db->n_entries, n_threads);
}

int ipkhost_merge_order(uint32_t n_shards, const ipkhost_db* shards, int n_threads, uint8_t* out_shard, uint32_t* out_index)
{
    if (!shards || n_shards < 1 || n_shards > 255) return IPKHOST_EINVAL;
    uint64_t total = 0;
    uint32_t biggest = 0;
    for (uint32_t s = 0; s < n_shards; ++s)
    {
        if (shards[s].n_kmers > 0xffffffffull) return IPKHOST_EINVAL;
        total += shards[s].n_kmers;
        if (shards[s].n_kmers > shards[biggest].n_kmers) biggest = s;
    }
    if (total == 0) return IPKHOST_OK;
    if (!out_shard || !out_index) return IPKHOST_EINVAL;
    unsigned T = n_threads > 0 ? unsigned(n_threads) : std::max(1u, std::thread::hardware_concurrency());
    T = unsigned(std::min<uint64_t>(T, std::max<uint64_t>(1, total / 65536)));
    // splitters: T - 1 evenly spaced k-mers of the largest shard; lower bounds of each in every shard give the threads
    // disjoint input ranges and their output offsets (keys are unique across shards, so no tie straddles a cut)
    std::vector<std::vector<uint64_t>> lo(T + 1, std::vector<uint64_t>(n_shards, 0));
    for (uint32_t s = 0; s < n_shards; ++s) lo[T][s] = shards[s].n_kmers;
    const ipkhost_db& B = shards[biggest];
    for (unsigned t = 1; t < T; ++t)
    {
        const uint64_t pivot = B.n_kmers * t / T;
        for (uint32_t s = 0; s < n_shards; ++s)
        {
            uint64_t a = 0, b = shards[s].n_kmers;           // first i with !(shard[i] < pivot)
            while (a < b) { const uint64_t m = (a + b) / 2; if (head_less(shards[s], m, B, pivot)) a = m + 1; else b = m; }
            lo[t][s] = a;
        }
    }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < T; ++t)
        th.emplace_back([&, t] {
            uint64_t out = 0;
            for (uint32_t s = 0; s < n_shards; ++s) out += lo[t][s];
            std::vector<uint64_t> cur(lo[t]);
            const std::vector<uint64_t>& end = lo[t + 1];
            while (true)
            {
                int best = -1;
                for (uint32_t s = 0; s < n_shards; ++s)
                    if (cur[s] < end[s] && (best < 0 || head_less(shards[s], cur[s], shards[best], cur[best]))) best = int(s);
                if (best < 0) break;
                out_shard[out] = uint8_t(best);
                out_index[out] = uint32_t(cur[best]++);
                ++out;
            }
        });
    for (auto& x : th) x.join();
    return IPKHOST_OK;
}

int ipkhost_digest(uint32_t n_shards, const ipkhost_db* shards, int n_threads, uint64_t* out)
{
    // the digest of ipk_b200/csrc/digest.cuh over host databases: out[0..5] = entries sum / xor, k-mers sum / xor, fv sum / xor
    if (!shards || !out) return IPKHOST_EINVAL;
    auto mix = [](uint64_t x) { x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x; };
    unsigned T = n_threads > 0 ? unsigned(n_threads) : std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::vector<uint64_t>> acc(T, std::vector<uint64_t>(6, 0));
    for (uint32_t s = 0; s < n_shards; ++s)
    {
        const ipkhost_db* d = shards + s;
        std::vector<std::thread> th;
        for (unsigned t = 0; t < T; ++t)
            th.emplace_back([&, t] {
                std::vector<uint64_t>& a = acc[t];
                for (uint64_t q = d->n_kmers * t / T; q < d->n_kmers * (t + 1) / T; ++q)
                {
                    const uint64_t hk = mix(d->keys[q] + 0x9e3779b97f4a7c15ull);
                    const uint64_t b = d->offsets[q], n = entries_of(d, q);
                    const uint64_t hc = mix(hk ^ (n + 0x632be59bd9b4e019ull));
                    a[2] += hc; a[3] ^= hc;
                    uint32_t fb; memcpy(&fb, &d->fv[q], 4);
                    const uint64_t hf = mix(hk ^ (uint64_t(fb) | 0x5bd1e99500000000ull));
                    a[4] += hf; a[5] ^= hf;
                    for (uint64_t e = b; e < b + n; ++e)
                    {
                        uint32_t sb; memcpy(&sb, &d->score[e], 4);
                        const uint32_t br = d->branch16 ? uint32_t(d->branch16[e]) : d->branch[e];
                        const uint64_t he = mix(hk ^ (uint64_t(br) << 32 | sb));
                        a[0] += he; a[1] ^= he;
                    }
                }
            });
        for (auto& x : th) x.join();
    }
    for (int i = 0; i < 6; ++i) out[i] = 0;
    for (unsigned t = 0; t < T; ++t)
        for (int i = 0; i < 6; ++i) { if (i & 1) out[i] ^= acc[t][i]; else out[i] += acc[t][i]; }
    return IPKHOST_OK;
}

int ipkhost_write_ipk_sharded(const char* path, const ipkhost_db* header, uint32_t n_shards, const ipkhost_db* shards,
                              const uint8_t* order_shard, const uint32_t* order_index, int n_threads)
{
    if (!path || !header || !shards || n_shards < 1) return IPKHOST_EINVAL;
    uint64_t n_kmers = 0, n_entries = 0;
    for (uint32_t s = 0; s < n_shards; ++s) { n_kmers += shards[s].n_kmers; n_entries += shards[s].n_entries; }
    if (n_kmers && (!order_shard || !order_index)) return IPKHOST_EINVAL;
    const source S{ shards, order_shard, order_index };
    return write_all(path, header, S, n_kmers, n_entries, n_threads);
}

int ipkhost_write_ipk(const char* path, const ipkhost_db* db) { return ipkhost_write_ipk_mt(path, db, 0); }

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
    const uint8_t s_int = r.get<uint8_t>(), s_long = r.get<uint8_t>(), s_float = r.get<uint8_t>(), s_double = r.get<uint8_t>();
    const int32_t one = r.get<int32_t>();
    if (!r.ok || pre != PREAMBLE || ver != ARCHIVE_VERSION || s_int != sizeof(int) || s_long != sizeof(long) || s_float != sizeof(float) ||
        s_double != sizeof(double) || one != 1) { fclose(f); delete d; return IPKHOST_EFORMAT; }
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
    // entries with positions are 10 bytes instead of 8: the header does not say which (i2l: a property of the binary that wrote
    // the file), the length of the file does
    bool with_pos = false;
    {
        const long here = ftell(f);
        if (here >= 0 && fseek(f, 0, SEEK_END) == 0)
        {
            const long end = ftell(f);
            with_pos = n_entries && uint64_t(end - here) == n_kmers * 20 + n_entries * 10;
            fseek(f, here, SEEK_SET);
        }
    }
    if (with_pos) d->position.resize(n_entries);
    uint64_t e = 0;
    for (uint64_t q = 0; q < n_kmers && r.ok; ++q)
    {
        d->keys[q] = r.get<uint64_t>();
        d->fv[q] = r.get<float>();
        const uint64_t n = r.get<uint64_t>();
        d->offsets[q] = e;
        if (!r.ok || e + n > n_entries) { r.ok = false; break; }
        for (uint64_t i = 0; i < n; ++i, ++e)
        {
            d->branch[e] = r.get<uint32_t>(); d->score[e] = r.get<float>();
            if (with_pos) d->position[e] = r.get<uint16_t>();
        }
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
    view->position = with_pos ? d->position.data() : nullptr;
    *out = d;
    return IPKHOST_OK;
}

void ipkhost_free_ipk(ipkhost_db_file* f) { delete f; }

}   // extern "C"
