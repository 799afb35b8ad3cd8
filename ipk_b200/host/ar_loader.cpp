// ar_loader.cpp -- raxml-ng ancestral-probability loader of the B200 build's host side.
// Behaviour follows ar::raxmlng_reader (ipk/src/ar.cpp:150-270) and the csv reader it instantiates
// (CSVReader<3+sigma, trim_chars<' '>, no_quote_escape<'\t'>, throw_on_overflow, single_and_empty_line_comment<'.'>>):
// the file is mapped once, indexed by node label in one pass, and any set of nodes is parsed in parallel straight into
// the packed [node][site][sigma] float layout the device path consumes.
#include "ipkhost.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cmath>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace
{
    thread_local std::string g_error;
}
namespace ipkhost { void set_last_error(const std::string& s) { g_error = s; } }      // plan.cpp, ipk_file.cpp
namespace
{

    int fail(int code, const std::string& what)
    {
        g_error = what;
        return code;
    }

    // csv.h:1024-1085 parse_float<float>: digit-by-digit accumulation IN FLOAT (not strtof): the integer part by
    // x = x*10 + y, the fraction by pos /= 10; x += y*pos, the exponent by square-and-multiply on base 0.1f / 10.
    bool parse_float_ref(const char* col, const char* end, float& x)
    {
        bool is_neg = false;
        if (col < end && *col == '-') { is_neg = true; ++col; }
        else if (col < end && *col == '+') ++col;
        x = 0;
        while (col < end && '0' <= *col && *col <= '9')
        {
            const int y = *col - '0';
            x *= 10;
            x += y;
            ++col;
        }
        if (col < end && (*col == '.' || *col == ','))
        {
            ++col;
            float pos = 1;
            while (col < end && '0' <= *col && *col <= '9')
            {
                pos /= 10;
                const int y = *col - '0';
                ++col;
                x += y * pos;
            }
        }
        if (col < end && (*col == 'e' || *col == 'E'))
        {
            ++col;
            // parse_signed_integer<set_to_max_on_overflow>(col, e)  (csv.h:975-1008)
            bool eneg = false;
            if (col < end && *col == '-') { eneg = true; ++col; }
            else if (col < end && *col == '+') ++col;
            if (col >= end) return false;
            long long ev = 0;
            while (col < end)
            {
                if (*col < '0' || *col > '9') return false;
                ev = ev * 10 + (*col - '0');
                if (ev > 2147483647LL) ev = 2147483647LL;
                ++col;
            }
            int e = int(ev);
            if (e != 0)
            {
                float base;
                if (eneg) base = 0.1;           // the double literal 0.1 rounded to float, as in the reference
                else base = 10;
                while (e != 1)
                {
                    if ((e & 1) == 0) { base = base * base; e >>= 1; }
                    else { x *= base; --e; }
                }
                x *= base;
            }
        }
        else if (col != end) return false;   // error::no_digit
        if (is_neg) x = -x;
        return true;
    }

    // RAxML-ng prints amino acids as A R N D C Q E G H I L K M F P S T W Y V; IPK encodes them as
    // R H K D E S T N Q C G P A I L M F W Y V (ar.cpp:232-234): dst[j] = src[AA_FROM[j]]
    const int AA_FROM[20] = { 1, 8, 11, 3, 6, 15, 16, 2, 5, 4, 7, 14, 0, 9, 10, 12, 13, 17, 18, 19 };
}

struct ipkhost_ar
{
    int fd = -1;
    const char* data = nullptr;
    size_t size = 0;
    uint32_t sigma = 0, width = 0;
    int n_threads = 1;
    struct block { size_t begin, end; uint32_t rows; };
    std::vector<std::string> labels;
    std::vector<block> blocks;
    std::unordered_map<std::string, uint32_t> index;

    ~ipkhost_ar()
    {
        if (data) munmap(const_cast<char*>(data), size);
        if (fd >= 0) close(fd);
    }

    /// one node block -> out[rows][sigma]; returns an error string or empty
    std::string parse_block(const block& b, const std::string& label, float* out) const
    {
        const char* p = data + b.begin;
        const char* const stop = data + b.end;
        uint32_t row = 0;
        float vals[32];
        while (p < stop)
        {
            const char* eol = static_cast<const char*>(memchr(p, '\n', size_t(stop - p)));
            const char* line_end = eol ? eol : stop;
            const char* next = eol ? eol + 1 : stop;
            if (line_end > p && line_end[-1] == '\r') --line_end;
            if (line_end == p || *p == '.') { p = next; continue; }       // single_and_empty_line_comment<'.'>
            // fields: Node, Site, State, then sigma probabilities; '\t' separated, spaces trimmed
            const char* f = p;
            for (uint32_t col = 0; col < 3 + sigma; ++col)
            {
                if (f > line_end) return "too few columns in a row of node " + label;
                const char* fe = static_cast<const char*>(memchr(f, '\t', size_t(line_end - f)));
                if (!fe) fe = line_end;
                if (col >= 3)
                {
                    const char* a = f;
                    const char* z = fe;
                    while (a < z && *a == ' ') ++a;
                    while (z > a && z[-1] == ' ') --z;
                    if (!parse_float_ref(a, z, vals[col - 3])) return "not a number in a row of node " + label;
                }
                if (col + 1 == 3 + sigma && fe != line_end) return "too many columns in a row of node " + label;
                f = fe + 1;
            }
            if (row >= b.rows) return "row count changed while reading node " + label;
            float* dst = out + size_t(row) * sigma;
            if (sigma == 20) for (int j = 0; j < 20; ++j) dst[j] = std::log10(vals[AA_FROM[j]]);      // ar.cpp:258-259
            else for (uint32_t j = 0; j < sigma; ++j) dst[j] = std::log10(vals[j]);
            ++row;
            p = next;
        }
        if (row != b.rows) return "row count changed while reading node " + label;
        return {};
    }
};

extern "C"
{

const char* ipkhost_last_error(void) { return g_error.c_str(); }

int ipkhost_parse_float(const char* text, float* out)
{
    if (!text || !out) return fail(IPKHOST_EINVAL, "null argument");
    return parse_float_ref(text, text + strlen(text), *out) ? IPKHOST_OK : fail(IPKHOST_EFORMAT, "not a number");
}

int ipkhost_ar_open(const char* path, uint32_t sigma, int n_threads, ipkhost_ar** out)
{
    if (!path || !out || (sigma != 4 && sigma != 20)) return fail(IPKHOST_EINVAL, "path / out null or sigma not 4 | 20");
    *out = nullptr;
    auto ar = new ipkhost_ar();
    ar->sigma = sigma;
    ar->n_threads = n_threads > 0 ? n_threads : int(std::max(1u, std::thread::hardware_concurrency()));
    ar->fd = open(path, O_RDONLY);
    struct stat st;
    if (ar->fd < 0 || fstat(ar->fd, &st) != 0) { delete ar; return fail(IPKHOST_EIO, std::string("cannot open ") + path); }
    ar->size = size_t(st.st_size);
    if (ar->size)
    {
        void* m = mmap(nullptr, ar->size, PROT_READ, MAP_PRIVATE, ar->fd, 0);
        if (m == MAP_FAILED) { delete ar; return fail(IPKHOST_EIO, std::string("cannot map ") + path); }
        ar->data = static_cast<const char*>(m);
    }
    // build_index (ar.cpp:150-188): skip the header line; a block starts wherever the first field changes
    const char* p = ar->data;
    const char* const stop = ar->data + ar->size;
    if (p < stop)
    {
        const char* eol = static_cast<const char*>(memchr(p, '\n', size_t(stop - p)));
        p = eol ? eol + 1 : stop;
    }
    std::string current;
    while (p < stop)
    {
        const char* eol = static_cast<const char*>(memchr(p, '\n', size_t(stop - p)));
        const char* line_end = eol ? eol : stop;
        const char* next = eol ? eol + 1 : stop;
        if (line_end > p && *p != '.' && !(line_end == p + 1 && *p == '\r'))
        {
            const char* tab = static_cast<const char*>(memchr(p, '\t', size_t(line_end - p)));
            const size_t len = size_t((tab ? tab : line_end) - p);
            if (len != current.size() || memcmp(p, current.data(), len) != 0)
            {
                if (!ar->blocks.empty()) ar->blocks.back().end = size_t(p - ar->data);
                current.assign(p, len);
                if (ar->index.count(current)) { delete ar; return fail(IPKHOST_EFORMAT, "rows of node " + current + " are not consecutive"); }
                ar->index.emplace(current, uint32_t(ar->labels.size()));
                ar->labels.push_back(current);
                ar->blocks.push_back({ size_t(p - ar->data), ar->size, 0 });
            }
            ar->blocks.back().rows++;
        }
        p = next;
    }
    ar->width = ar->blocks.empty() ? 0 : ar->blocks.front().rows;
    *out = ar;
    return IPKHOST_OK;
}

void ipkhost_ar_close(ipkhost_ar* ar) { delete ar; }
uint32_t ipkhost_ar_num_nodes(const ipkhost_ar* ar) { return ar ? uint32_t(ar->labels.size()) : 0; }
uint32_t ipkhost_ar_width(const ipkhost_ar* ar) { return ar ? ar->width : 0; }
const char* ipkhost_ar_label(const ipkhost_ar* ar, uint32_t i) { return ar && i < ar->labels.size() ? ar->labels[i].c_str() : nullptr; }

int ipkhost_ar_read_nodes(ipkhost_ar* ar, const char* const* labels, uint32_t n, float* out)
{
    if (!ar || (n && (!labels || !out))) return fail(IPKHOST_EINVAL, "null argument");
    std::vector<uint32_t> which(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        auto it = ar->index.find(labels[i]);
        if (it == ar->index.end()) return fail(IPKHOST_ENOENT, std::string("Could not read the AR matrix for the node ") + labels[i]);
        which[i] = it->second;
        if (ar->blocks[it->second].rows != ar->width)
            return fail(IPKHOST_EFORMAT, std::string("node ") + labels[i] + " has a different number of sites");
    }
    std::atomic<uint32_t> next{ 0 };
    std::atomic<bool> failed{ false };
    std::string first_error;
    std::atomic_flag error_lock = ATOMIC_FLAG_INIT;
    const size_t per_node = size_t(ar->width) * ar->sigma;
    auto work = [&] {
        for (uint32_t i = next++; i < n && !failed; i = next++)
        {
            const std::string err = ar->parse_block(ar->blocks[which[i]], ar->labels[which[i]], out + size_t(i) * per_node);
            if (!err.empty() && !failed.exchange(true))
            {
                while (error_lock.test_and_set()) {}
                first_error = err;
                error_lock.clear();
            }
        }
    };
    const int nt = int(std::min<uint32_t>(uint32_t(ar->n_threads), std::max(n, 1u)));
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    if (failed) return fail(IPKHOST_EFORMAT, first_error);
    return IPKHOST_OK;
}

}   // extern "C"
