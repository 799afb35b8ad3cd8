/* TEST INFRASTRUCTURE (oracle, kind "port") -- NOT product code.
 *
 * Plain-C restatement of the reference's phylo-k-mer database build hot path (phylo42/IPK @ 84b65e9), used
 * ONLY by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the checker
 * for the CUDA path.  Nothing under ipk_b200/ may import, link or call it.
 *
 * Parity status: the enumeration + max-merge arithmetic is PINNED -- tests/test_oracle.py checks it against
 * fixtures produced by the reference's own object code (oracle/_ref, see oracle/ref_driver.cpp and
 * tests/golden/make_golden.py).  UNPINNED (i2l submodule absent): the rounding of
 * score_threshold(), the tie order of equal filter values, the random filter's draw order, the .ipk bytes.
 *
 * Each function cites the reference file:line it follows.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { uint64_t key; float score; } pk_t;                 /* i2l::unpositioned_phylo_kmer */
typedef struct { pk_t* v; size_t n, cap; } pkvec;

static void pv_push(pkvec* a, uint64_t key, float score)
{
    if (a->n == a->cap)
    {
        a->cap = a->cap ? a->cap * 2 : 16;
        a->v = (pk_t*)realloc(a->v, a->cap * sizeof(pk_t));
    }
    a->v[a->n].key = key;
    a->v[a->n].score = score;
    a->n++;
}

/* matrix::preprocess, ipk/src/window.cpp:16-27: B[j+1] = B[j] + max_i M[j][i], sequential fp32 */
void ipko_best_prefix(const float* logp, uint32_t W, uint32_t sigma, float* B)
{
    volatile float product = 0.0f;      /* volatile: forbid any widening / reassociation */
    B[0] = 0.0f;
    for (uint32_t j = 0; j < W; ++j)
    {
        const float* col = logp + (size_t)j * sigma;
        float best = col[0];
        for (uint32_t i = 1; i < sigma; ++i) if (best < col[i]) best = col[i];   /* std::max_element */
        product = product + best;
        B[j + 1] = product;
    }
}

/* matrix::range_max_sum, window.cpp:69-72 (window::range_max_product adds the window start, :134-137) */
static inline float rms(const float* B, size_t start, size_t len)
{
    volatile float d = B[start + len] - B[start];
    return d;
}

/* kmer_score_comparator, pk_compute.cpp:8-11: descending score */
static int cmp_desc(const void* a, const void* b)
{
    const float x = ((const pk_t*)a)->score, y = ((const pk_t*)b)->score;
    return (x > y) ? -1 : (x < y) ? 1 : 0;
}

typedef struct { const float* M; const float* B; uint32_t sigma, bits; size_t s; } dc_ctx;

/* DCLA::DC, pk_compute.cpp:42-114.  j = offset in the window, h = length, eps = threshold of this node. */
static pkvec dc(const dc_ctx* c, size_t j, size_t h, float eps)
{
    pkvec result = { 0, 0, 0 };
    if (h == 1)
    {
        /* as_column, pk_compute.cpp:14-26: strict > */
        const float* col = c->M + (c->s + j) * c->sigma;
        for (uint32_t i = 0; i < c->sigma; ++i)
            if (col[i] > eps) pv_push(&result, i, col[i]);
        return result;
    }
    /* :54-55 -- two separate fp32 operations each: (B[a+len]-B[a]) then eps - that */
    volatile float eps_l = eps - rms(c->B, c->s + j + h / 2, h - h / 2);
    volatile float eps_r = eps - rms(c->B, c->s + j, h / 2);
    pkvec l = dc(c, j, h / 2, eps_l);
    pkvec r = dc(c, j + h / 2, h - h / 2, eps_r);

    const int prefix_sort = l.n < r.n;                 /* :61 */
    pkvec* mn = prefix_sort ? &l : &r;
    pkvec* mx = prefix_sort ? &r : &l;
    const float eps_min = prefix_sort ? eps_l : eps_r;
    const float eps_max = prefix_sort ? eps_r : eps_l;
    if (mn->n)
    {
        qsort(mn->v, mn->n, sizeof(pk_t), cmp_desc);   /* :70 */
        const unsigned shift = (unsigned)((h - h / 2) * c->bits);
        for (size_t i = 0; i < mx->n; ++i)
        {
            const uint64_t a = mx->v[i].key;
            const float a_score = mx->v[i].score;
            if (a_score < eps_max) break;              /* :75-79 (never fires) */
            for (size_t i2 = 0; i2 < mn->n; ++i2)
            {
                const uint64_t b = mn->v[i2].key;
                const float b_score = mn->v[i2].score;
                if (b_score < eps_min) break;          /* :84-88 (never fires) */
                volatile float score = a_score + b_score;
                if (score <= eps) break;               /* :90-94 */
                const uint64_t kmer = prefix_sort ? ((b << shift) | a) : ((a << shift) | b);   /* :96-104 */
                pv_push(&result, kmer, score);
            }
        }
    }
    free(l.v);
    free(r.v);
    return result;
}

/* `for window in to_windows(matrix,k): DCLA(window,k).run(eps)`  (db_builder.cpp:645-649; window.cpp:159-197) */
int64_t ipko_enumerate(const float* logp, uint32_t W, uint32_t sigma, uint32_t bits, uint32_t k, float eps,
                       uint32_t* wpos, uint64_t* keys, float* scores, int64_t cap)
{
    if (W < k) return 0;
    float* B = (float*)malloc(((size_t)W + 1) * sizeof(float));
    ipko_best_prefix(logp, W, sigma, B);
    int64_t n = 0;
    for (size_t s = 0; s + k <= W; ++s)
    {
        dc_ctx c = { logp, B, sigma, bits, s };
        pkvec r = dc(&c, 0, k, eps);
        for (size_t i = 0; i < r.n; ++i, ++n)
            if (n < cap)
            {
                if (wpos) wpos[n] = (uint32_t)s;
                keys[n] = r.v[i].key;
                scores[n] = r.v[i].score;
            }
        free(r.v);
    }
    free(B);
    return n;
}

/* Order-free specification of the same output (SURVEY.md §3.3 "parity predicate"): a k-mer is emitted iff every
 * node v of the floor(h/2) | h-floor(h/2) split tree has fp32 partial sum S(v) > eps(v).  Brute force over
 * sigma^k codes; for cross-checking dc() on small k only. */
static int pred(const dc_ctx* c, const uint8_t* code, size_t j, size_t h, float eps, float* S)
{
    if (h == 1)
    {
        *S = c->M[(c->s + j) * c->sigma + code[j]];
        return *S > eps;
    }
    volatile float eps_l = eps - rms(c->B, c->s + j + h / 2, h - h / 2);
    volatile float eps_r = eps - rms(c->B, c->s + j, h / 2);
    float sl, sr;
    if (!pred(c, code, j, h / 2, eps_l, &sl)) return 0;
    if (!pred(c, code, j + h / 2, h - h / 2, eps_r, &sr)) return 0;
    volatile float sum = sl + sr;
    *S = sum;
    return sum > eps;
}

int64_t ipko_enumerate_predicate(const float* logp, uint32_t W, uint32_t sigma, uint32_t bits, uint32_t k, float eps,
                                 uint32_t* wpos, uint64_t* keys, float* scores, int64_t cap)
{
    if (W < k) return 0;
    float* B = (float*)malloc(((size_t)W + 1) * sizeof(float));
    ipko_best_prefix(logp, W, sigma, B);
    uint64_t total = 1;
    for (uint32_t i = 0; i < k; ++i) total *= sigma;
    int64_t n = 0;
    uint8_t code[64];
    for (size_t s = 0; s + k <= W; ++s)
    {
        dc_ctx c = { logp, B, sigma, bits, s };
        for (uint64_t x = 0; x < total; ++x)
        {
            uint64_t t = x, key = 0;
            for (int p = (int)k - 1; p >= 0; --p) { code[p] = (uint8_t)(t % sigma); t /= sigma; }
            for (uint32_t p = 0; p < k; ++p) key = (key << bits) | code[p];
            float S;
            if (pred(&c, code, 0, k, eps, &S))
            {
                if (n < cap) { if (wpos) wpos[n] = (uint32_t)s; keys[n] = key; scores[n] = S; }
                ++n;
            }
        }
    }
    free(B);
    return n;
}

/* ---------------------------------------------------------------------------------------------------------------
 * Per-branch max-merge, regroup by k-mer, filter values, order.
 * ------------------------------------------------------------------------------------------------------------- */

/* group_hash_map (branch_group.h:23) as an open-addressing table: key -> score */
typedef struct { uint64_t* k; float* s; size_t cap, n; } gmap;
#define GMAP_EMPTY UINT64_MAX

static void gmap_init(gmap* m, size_t cap)
{
    m->cap = cap; m->n = 0;
    m->k = (uint64_t*)malloc(cap * sizeof(uint64_t));
    m->s = (float*)malloc(cap * sizeof(float));
    memset(m->k, 0xff, cap * sizeof(uint64_t));
}
static inline size_t gmap_slot(const gmap* m, uint64_t key)
{
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    size_t i = (size_t)(h >> 17) & (m->cap - 1);
    while (m->k[i] != GMAP_EMPTY && m->k[i] != key) i = (i + 1) & (m->cap - 1);
    return i;
}
static void gmap_put(gmap* m, uint64_t key, float score);
static void gmap_grow(gmap* m)
{
    gmap o = *m;
    gmap_init(m, o.cap * 2);
    for (size_t i = 0; i < o.cap; ++i) if (o.k[i] != GMAP_EMPTY) gmap_put(m, o.k[i], o.s[i]);
    free(o.k); free(o.s);
}
/* ipk::put, branch_group.cpp:88-101: insert, or overwrite when stored < new (i.e. keep the maximum) */
static void gmap_put(gmap* m, uint64_t key, float score)
{
    size_t i = gmap_slot(m, key);
    if (m->k[i] == key) { if (m->s[i] < score) m->s[i] = score; return; }
    m->k[i] = key; m->s[i] = score;
    if (++m->n * 2 > m->cap) gmap_grow(m);
}

typedef struct { uint64_t key; uint32_t group; float score; } rec_t;
static int cmp_rec(const void* a, const void* b)
{
    const rec_t* x = (const rec_t*)a; const rec_t* y = (const rec_t*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->group < y->group ? -1 : x->group > y->group ? 1 : 0;
}

/* logscore_to_score, filter.cpp:20-23: min(pow(10, x), 1.0) in double, RETURNED AS float (score_type) */
static inline float logscore_to_score(float log_score)
{
    double p = pow(10.0, (double)log_score);
    return (float)(p < 1.0 ? p : 1.0);
}
static inline double shannon(double x) { return -x * log2(x); }     /* filter.cpp:55-58 */

/* mif0_filter::calc_filter_values body for one k-mer, filter.cpp:63-116 */
float ipko_mif0(const float* log_scores, size_t n, uint32_t total_num_groups, float threshold)
{
    double score_sum = 0;
    for (size_t i = 0; i < n; ++i) score_sum += logscore_to_score(log_scores[i]);
    score_sum += (double)(total_num_groups - n) * threshold;
    const double weighted_threshold = threshold / score_sum;
    const double target_threshold = shannon(weighted_threshold);
    double HcBw1 = (double)total_num_groups * target_threshold;
    for (size_t i = 0; i < n; ++i)
    {
        const double weighted_score = logscore_to_score(log_scores[i]) / score_sum;
        const double target_value = shannon(weighted_score);
        HcBw1 = HcBw1 - target_threshold + target_value;
    }
    const double Hc = log2((double)total_num_groups);
    return (float)(score_sum * (HcBw1 - Hc));
}

/* random_filter, filter.cpp:133-146: std::default_random_engine(42) (libstdc++: minstd_rand0, x <- 16807 x mod 2^31-1)
 * feeding std::uniform_real_distribution<double>(0,1) (libstdc++ generate_canonical<double,53>: two draws,
 * ((g1-1) + (g2-1)*R) / R^2 with R = 2^31-2, result 1.0 mapped to nextafter(1,0)), cast to float.
 * Draw order: the reference iterates its hash map (unknowable, PARITY UNPINNED); canonical here: the k-mer with key K takes
 * draws 2e+1 and 2e+2 of the stream, e = K mod (2^31-2) -- a function of the k-mer alone, so key shards agree with the whole. */
static uint64_t mulmod31(uint64_t a, uint64_t b) { return (a * b) % 2147483647ull; }
void ipko_random_fv(size_t n, const uint64_t* keys, float* fv)
{
    const double R = 2147483646.0;
    for (size_t i = 0; i < n; ++i)
    {
        uint64_t e = 2 * (keys[i] % 2147483646ull) + 1, base = 16807, acc = 1;
        for (; e; e >>= 1) { if (e & 1) acc = mulmod31(acc, base); base = mulmod31(base, base); }
        const uint64_t x1 = mulmod31(42, acc), x2 = mulmod31(x1, 16807);
        double r = ((double)(x1 - 1) + (double)(x2 - 1) * R) / (R * R);
        if (r >= 1.0) r = nextafter(1.0, 0.0);
        fv[i] = (float)r;
    }
}

typedef struct
{
    uint64_t n_kmers, n_entries, n_emitted;
    uint64_t* keys; float* fv; uint64_t* offsets; uint32_t* branch; float* score;
    double ms_stage1, ms_stage2;
} ipko_db;

typedef struct { float fv; uint64_t key; uint64_t idx; } ord_t;
static int cmp_ord(const void* a, const void* b)
{
    const ord_t* x = (const ord_t*)a; const ord_t* y = (const ord_t*)b;
    if (x->fv != y->fv) return x->fv < y->fv ? -1 : 1;
    return x->key < y->key ? -1 : x->key > y->key ? 1 : 0;
}
static double now_ms(void)
{
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

/* db_builder::run stages 1+2, in-RAM (db_builder.cpp:182-290, 576-698).  Arguments as ipkgpu_input (include/ipkgpu.h). */
ipko_db* ipko_build(const float* logp, const uint32_t* node_branch, uint32_t n_nodes, uint32_t W, uint32_t sigma,
                    uint32_t bits, uint32_t k, float log_threshold, float threshold, uint32_t total_num_groups,
                    int filter, int threads)
{
    const double t0 = now_ms();
    /* group_ghost_ids, db_builder.cpp:524-553: group index = order of first appearance of the branch id */
    uint32_t* node_group = (uint32_t*)malloc((size_t)n_nodes * 4 + 4);
    uint32_t* group_branch = (uint32_t*)malloc((size_t)n_nodes * 4 + 4);
    uint32_t n_groups = 0;
    {
        size_t cap = 16;
        while (cap < (size_t)n_nodes * 2 + 2) cap *= 2;
        uint32_t* tk = (uint32_t*)malloc(cap * 4);      /* branch id   */
        uint32_t* tv = (uint32_t*)malloc(cap * 4);      /* group index */
        uint8_t* used = (uint8_t*)calloc(cap, 1);
        for (uint32_t n = 0; n < n_nodes; ++n)
        {
            size_t i = ((size_t)node_branch[n] * 2654435761u) & (cap - 1);
            while (used[i] && tk[i] != node_branch[n]) i = (i + 1) & (cap - 1);
            if (!used[i])
            {
                used[i] = 1; tk[i] = node_branch[n]; tv[i] = n_groups;
                group_branch[n_groups++] = node_branch[n];
            }
            node_group[n] = tv[i];
        }
        free(tk); free(tv); free(used);
    }
    rec_t** grecs = (rec_t**)calloc(n_groups ? n_groups : 1, sizeof(rec_t*));
    size_t* gcount = (size_t*)calloc(n_groups ? n_groups : 1, sizeof(size_t));
    uint64_t n_emitted = 0;
    (void)threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) reduction(+: n_emitted) num_threads(threads > 0 ? threads : 1)
#endif
    for (uint32_t g = 0; g < n_groups; ++g)
    {
        /* explore_group, db_builder.cpp:629-698 */
        gmap map; gmap_init(&map, 1024);
        float* B = (float*)malloc(((size_t)W + 1) * sizeof(float));
        for (uint32_t n = 0; n < n_nodes; ++n)
        {
            if (node_group[n] != g) continue;
            const float* M = logp + (size_t)n * W * sigma;
            ipko_best_prefix(M, W, sigma, B);
            for (size_t s = 0; s + k <= W; ++s)
            {
                dc_ctx c = { M, B, sigma, bits, s };
                pkvec r = dc(&c, 0, k, log_threshold);
                for (size_t i = 0; i < r.n; ++i) gmap_put(&map, r.v[i].key, r.v[i].score);
                n_emitted += r.n;
                free(r.v);
            }
        }
        free(B);
        rec_t* out = (rec_t*)malloc((map.n ? map.n : 1) * sizeof(rec_t));
        size_t m = 0;
        for (size_t i = 0; i < map.cap; ++i)
            if (map.k[i] != GMAP_EMPTY) { out[m].key = map.k[i]; out[m].group = g; out[m].score = map.s[i]; ++m; }
        grecs[g] = out; gcount[g] = m;
        free(map.k); free(map.s);
    }
    /* unsafe_insert(kmer, {branch, score}) group after group (db_builder.cpp:685-693) == records grouped by
     * key with entries in group order */
    size_t n_entries = 0;
    for (uint32_t g = 0; g < n_groups; ++g) n_entries += gcount[g];
    rec_t* recs = (rec_t*)malloc((n_entries ? n_entries : 1) * sizeof(rec_t));
    for (size_t g = 0, o = 0; g < n_groups; ++g) { memcpy(recs + o, grecs[g], gcount[g] * sizeof(rec_t)); o += gcount[g]; free(grecs[g]); }
    free(grecs); free(gcount);
    qsort(recs, n_entries, sizeof(rec_t), cmp_rec);
    size_t n_kmers = 0;
    for (size_t i = 0; i < n_entries; ++i) if (i == 0 || recs[i].key != recs[i - 1].key) ++n_kmers;
    const double t1 = now_ms();

    /* filter_in_ram, db_builder.cpp:254-290 */
    size_t* start = (size_t*)malloc((n_kmers + 1) * sizeof(size_t));
    for (size_t i = 0, q = 0; i < n_entries; ++i) if (i == 0 || recs[i].key != recs[i - 1].key) start[q++] = i;
    start[n_kmers] = n_entries;
    ord_t* ord = (ord_t*)malloc((n_kmers ? n_kmers : 1) * sizeof(ord_t));
    float* rnd = NULL;
    if (filter != 0)
    {
        rnd = (float*)malloc((n_kmers ? n_kmers : 1) * sizeof(float));
        uint64_t* ks = (uint64_t*)malloc((n_kmers ? n_kmers : 1) * sizeof(uint64_t));
        for (size_t q = 0; q < n_kmers; ++q) ks[q] = recs[start[q]].key;
        ipko_random_fv(n_kmers, ks, rnd);
        free(ks);
    }
#ifdef _OPENMP
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
#endif
    {
        float* tmp = (float*)malloc(((size_t)n_groups + 1) * sizeof(float));
#ifdef _OPENMP
#pragma omp for schedule(static)
#endif
        for (size_t q = 0; q < n_kmers; ++q)
        {
            const size_t a = start[q], n = start[q + 1] - a;
            ord[q].key = recs[a].key; ord[q].idx = q;
            if (filter == 0)
            {
                for (size_t i = 0; i < n; ++i) tmp[i] = recs[a + i].score;
                ord[q].fv = ipko_mif0(tmp, n, total_num_groups, threshold);
            }
            else ord[q].fv = rnd[q];
        }
        free(tmp);
    }
    free(rnd);
    qsort(ord, n_kmers, sizeof(ord_t), cmp_ord);      /* db_builder.cpp:284; ties: key ascending (canonical here) */
    const double t2 = now_ms();

    ipko_db* db = (ipko_db*)calloc(1, sizeof(ipko_db));
    db->n_kmers = n_kmers; db->n_entries = n_entries; db->n_emitted = n_emitted;
    db->keys = (uint64_t*)malloc((n_kmers + 1) * 8); db->fv = (float*)malloc((n_kmers + 1) * 4);
    db->offsets = (uint64_t*)malloc((n_kmers + 1) * 8);
    db->branch = (uint32_t*)malloc((n_entries + 1) * 4); db->score = (float*)malloc((n_entries + 1) * 4);
    size_t o = 0;
    for (size_t q = 0; q < n_kmers; ++q)
    {
        const size_t a = start[ord[q].idx], n = start[ord[q].idx + 1] - a;
        db->keys[q] = ord[q].key; db->fv[q] = ord[q].fv; db->offsets[q] = o;
        for (size_t i = 0; i < n; ++i, ++o) { db->branch[o] = group_branch[recs[a + i].group]; db->score[o] = recs[a + i].score; }
    }
    db->offsets[n_kmers] = o;
    db->ms_stage1 = t1 - t0; db->ms_stage2 = t2 - t1;
    free(ord); free(start); free(recs); free(node_group); free(group_branch);
    return db;
}

void ipko_db_free(ipko_db* db)
{
    if (!db) return;
    free(db->keys); free(db->fv); free(db->offsets); free(db->branch); free(db->score); free(db);
}
