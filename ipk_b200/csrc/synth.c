/* Synthetic input generator for the phylo-k-mer database build (BASELINE.json configs 2-5).
 *
 * Host-only tooling shared by tests/ and bench.py (both arms): a seeded, counter-based generator of
 *   - a random rooted binary tree with n_taxa leaves (post-order ids, as db_builder.cpp:191-197 indexes it),
 *   - ancestral states evolved down that tree,
 *   - the per-ghost-node posterior matrices raxml-ng would hand to `ipk build --ar-dir` (ar.cpp:200-270): for
 *     every non-root node two ghost nodes X0 (mid-branch) and X1 (its sister ghost), extended_tree.cpp:103-150,
 *     each a [W][sigma] fp32 matrix of log10 posteriors, 9-decimal rounded like `raxml-ng --precision 9`
 *     (ar.cpp:670) so that exact zeros (-inf) occur.
 * Every value is a pure function of (seed, node, site), so any node range can be produced in isolation
 * (one rank per GPU generates only its own branch groups).
 *
 * Not part of the timed path and not part of the oracle.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t mix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
static inline uint64_t h4(uint64_t seed, uint64_t a, uint64_t b, uint64_t c)
{
    return mix64(mix64(mix64(mix64(seed) ^ a) ^ (b * 0xD1B54A32D192ED03ull)) ^ (c * 0x8CB92BA72F3D8DD7ull));
}
static inline double u01(uint64_t h) { return ((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0); }

/* Random rooted binary tree by random pairwise joins; returns parent[] indexed by POST-ORDER id.
 * n_nodes = 2*n_taxa-1, root id = n_nodes-1, parent[root] = UINT32_MAX.  branch_len may be NULL. */
int ipk_synth_tree(uint32_t n_taxa, uint64_t seed, uint32_t* parent, float* branch_len)
{
    if (n_taxa < 2) return -1;
    const uint32_t n_nodes = 2 * n_taxa - 1;
    uint32_t* left = (uint32_t*)malloc(n_nodes * 4);
    uint32_t* right = (uint32_t*)malloc(n_nodes * 4);
    uint32_t* active = (uint32_t*)malloc(n_taxa * 4);
    for (uint32_t i = 0; i < n_taxa; ++i) { active[i] = i; left[i] = right[i] = UINT32_MAX; }
    uint32_t n_active = n_taxa, next = n_taxa;
    while (n_active > 1)
    {
        uint32_t a = (uint32_t)(h4(seed, 1, next, 0) % n_active);
        uint32_t b = (uint32_t)(h4(seed, 1, next, 1) % (n_active - 1));
        if (b >= a) ++b;
        left[next] = active[a]; right[next] = active[b];
        const uint32_t lo = a < b ? a : b, hi = a < b ? b : a;
        active[lo] = next; active[hi] = active[n_active - 1];
        --n_active; ++next;
    }
    /* iterative post-order from the root (creation id n_nodes-1) */
    uint32_t* post = (uint32_t*)malloc(n_nodes * 4);       /* creation id -> post-order id */
    uint32_t* stack = (uint32_t*)malloc(n_nodes * 4);
    uint8_t* state = (uint8_t*)calloc(n_nodes, 1);
    uint32_t sp = 0, counter = 0;
    stack[sp++] = n_nodes - 1;
    while (sp)
    {
        const uint32_t v = stack[sp - 1];
        if (left[v] == UINT32_MAX || state[v] == 2) { post[v] = counter++; --sp; continue; }
        if (state[v] == 0) { state[v] = 1; stack[sp++] = left[v]; }
        else { state[v] = 2; stack[sp++] = right[v]; }
    }
    for (uint32_t v = 0; v < n_nodes; ++v)
    {
        if (left[v] != UINT32_MAX) { parent[post[left[v]]] = post[v]; parent[post[right[v]]] = post[v]; }
    }
    parent[n_nodes - 1] = UINT32_MAX;
    if (branch_len)
        for (uint32_t v = 0; v < n_nodes; ++v)
            branch_len[v] = (v == n_nodes - 1) ? 0.0f : (float)(-0.05 * log(u01(h4(seed, 2, v, 0))));
    free(left); free(right); free(active); free(post); free(stack); free(state);
    return 0;
}

typedef struct
{
    float p_peak_x0, p_peak_x1;   /* probability that a (ghost, site) posterior is sharply peaked */
    float mut_mean;               /* mean per-branch substitution probability (exponential, clipped at 0.5) */
    float x1_mut;                 /* extra substitution probability of the sister ghost X1 */
    float diffuse_lo, diffuse_hi; /* p_max range of a non-peaked column */
    float peak_exp_lo, peak_exp_hi; /* a peaked column has p_max = 1 - 10^-U(peak_exp_lo, peak_exp_hi) */
} ipk_synth_params;

/* Ghost matrices for branches [branch_begin, branch_end) (post-order ids of non-root nodes), two per branch in the
 * order X0, X1.  logp: [(branch_end-branch_begin)*2][W][sigma]; node_branch: [(branch_end-branch_begin)*2]. */
int ipk_synth_posteriors(uint32_t n_taxa, uint32_t W, uint32_t sigma, uint64_t seed, const ipk_synth_params* prm,
                         uint32_t branch_begin, uint32_t branch_end, float* logp, uint32_t* node_branch)
{
    const uint32_t n_nodes = 2 * n_taxa - 1;
    if (branch_end > n_nodes - 1 || branch_begin > branch_end || sigma < 2 || sigma > 32) return -1;
    uint32_t* parent = (uint32_t*)malloc(n_nodes * 4);
    if (ipk_synth_tree(n_taxa, seed, parent, NULL)) { free(parent); return -1; }
    /* ancestral states, root downwards: in post-order a parent has a larger id than its children */
    uint8_t* st = (uint8_t*)malloc((size_t)n_nodes * W);
    for (uint32_t j = 0; j < W; ++j) st[(size_t)(n_nodes - 1) * W + j] = (uint8_t)(h4(seed, 3, n_nodes - 1, j) % sigma);
    for (int64_t v = (int64_t)n_nodes - 2; v >= 0; --v)
    {
        double m = -prm->mut_mean * log(u01(h4(seed, 4, (uint64_t)v, 0)));
        if (m > 0.5) m = 0.5;
        const uint8_t* ps = st + (size_t)parent[v] * W;
        uint8_t* cs = st + (size_t)v * W;
        for (uint32_t j = 0; j < W; ++j)
        {
            const uint64_t h = h4(seed, 5, (uint64_t)v, j);
            cs[j] = (u01(h) < m) ? (uint8_t)(mix64(h) % sigma) : ps[j];
        }
    }
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (int64_t g = 0; g < (int64_t)(branch_end - branch_begin) * 2; ++g)
    {
        const uint32_t v = branch_begin + (uint32_t)(g / 2);
        const int x1 = (int)(g & 1);
        const uint64_t ghost = (uint64_t)v * 2 + x1;
        node_branch[g] = v;
        float* out = logp + (size_t)g * W * sigma;
        const uint8_t* cs = st + (size_t)v * W;
        const uint8_t* ps = st + (size_t)parent[v] * W;
        const double p_peak = x1 ? prm->p_peak_x1 : prm->p_peak_x0;
        for (uint32_t j = 0; j < W; ++j)
        {
            /* X0 sits mid-branch: each site follows the child or the parent; X1 hangs off X0 */
            uint32_t c = (h4(seed, 6, v, j) & 1) ? cs[j] : ps[j];
            if (x1)
            {
                const uint64_t h = h4(seed, 7, v, j);
                if (u01(h) < prm->x1_mut) c = (uint32_t)(mix64(h) % sigma);
            }
            const uint64_t hp = h4(seed, 8, ghost, j);
            double pmax;
            if (u01(hp) < p_peak) pmax = 1.0 - pow(10.0, -(prm->peak_exp_lo + (prm->peak_exp_hi - prm->peak_exp_lo) * u01(mix64(hp))));
            else pmax = prm->diffuse_lo + (prm->diffuse_hi - prm->diffuse_lo) * u01(mix64(hp));
            double share[32], tot = 0;
            for (uint32_t i = 0; i < sigma; ++i)
            {
                share[i] = (i == c) ? 0.0 : -log(u01(h4(seed, 9 + i, ghost, j)));
                tot += share[i];
            }
            for (uint32_t i = 0; i < sigma; ++i)
            {
                double p = (i == c) ? pmax : (1.0 - pmax) * share[i] / tot;
                p = floor(p * 1e9 + 0.5) * 1e-9;                  /* --precision 9 */
                out[(size_t)j * sigma + i] = log10f((float)p);     /* ar.cpp:258-259 */
            }
        }
    }
    free(st); free(parent);
    return 0;
}

/* i2l::score_threshold(omega, k) and its log10 as the call sites' types imply (float powf / log10f; db_builder.cpp:260,640) --
 * the same two lines as ipkgpu_score_threshold (include/ipkgpu.h), so that the CPU arm of bench.py and the fixture
 * generators get the constants without mapping the GPU library. */
void ipk_synth_score_threshold(float omega, uint32_t sigma, uint32_t k, float* threshold, float* log_threshold)
{
    const float t = powf(omega / (float)sigma, (float)k);
    if (threshold) *threshold = t;
    if (log_threshold) *log_threshold = log10f(t);
}
