/* ipkhost.h -- host side of the B200 phylo-k-mer database build above the device C ABI (include/ipkgpu.h).
 *
 * Mirrors, for the hot path only, what ipk::build's callers hand it and what it hands the serializer:
 *   posterior loader   ar::raxmlng_reader::build_index / read_node      ipk/src/ar.cpp:150-270
 *                      (csv float parsing: strasser csv.h parse_float, csv.h:1024-1085; AA column order ar.cpp:232-234)
 *   database writer    i2l::save_header + i2l::save_phylo_kmer walk     ipk/src/db_builder.cpp:297-306,323-332
 *   database reader    what tools/src/diff.cpp:129-295 and dump.cpp:18-33 consume
 * Plain C entry points so that tests bind them with ctypes; the ipk-b200 executable (main.cpp) is built on them.
 */
#ifndef IPKHOST_H
#define IPKHOST_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define IPKHOST_OK       0
#define IPKHOST_EIO     -1    /* file cannot be opened / read / written */
#define IPKHOST_EFORMAT -2    /* malformed row, wrong column count, not a number (csv.h error::no_digit) */
#define IPKHOST_ENOENT  -3    /* node label not in the file ("Could not read the AR matrix for the node", ar.cpp:262) */
#define IPKHOST_EINVAL  -4

typedef struct ipkhost_ar ipkhost_ar;

/* Index a raxml-ng `*.ancestralProbs` file (ar.cpp:150-188): header line skipped, one block of consecutive rows per
 * node label.  sigma = 4 (DNA: p_A p_C p_G p_T) or 20 (AA, raxml-ng column order).  n_threads <= 0: all cores. */
int  ipkhost_ar_open(const char* path, uint32_t sigma, int n_threads, ipkhost_ar** out);
void ipkhost_ar_close(ipkhost_ar* ar);
const char* ipkhost_last_error(void);
uint32_t ipkhost_ar_num_nodes(const ipkhost_ar* ar);
uint32_t ipkhost_ar_width(const ipkhost_ar* ar);                 /* sites of the first node; every node must agree */
const char* ipkhost_ar_label(const ipkhost_ar* ar, uint32_t i);  /* labels in file order */
/* read_node (ar.cpp:200-270) for n labels at once: out[n][width][sigma] log10 posteriors in IPK's state order
 * (DNA A,C,G,T; AA R,H,K,D,E,S,T,N,Q,C,G,P,A,I,L,M,F,W,Y,V), float arithmetic as the reference's csv reader. */
int  ipkhost_ar_read_nodes(ipkhost_ar* ar, const char* const* labels, uint32_t n, float* out);
/* the csv reader's float parser alone (known-answer tests against the reference's own csv.h) */
int  ipkhost_parse_float(const char* text, float* out);

/* ---- build plan: everything ipk::build is handed besides the posteriors (ipk/src/main.cpp:145-198) -----------------
 * original tree (newick text) -> ghost-node injection (extended_tree.cpp:76-155) -> ghost selection and grouping
 * (db_builder.cpp:495-553) -> extended label -> AR label through the AR tree (ar::map_nodes, ar.cpp:790-834; the AR tree
 * is re-rooted first when raxml-ng unrooted it, extended_tree.cpp:186-205) -> tree index (db_builder.cpp:191-197).
 * ar_tree_newick may be NULL: the posteriors are then keyed by the extended labels themselves.
 * ghost_strategy: 0 both (default), 1 inner only (_X0), 2 outer only (_X1).  An unrooted original tree is IPKHOST_EINVAL
 * unless use_unrooted (extended_tree.cpp:169-177). */
typedef struct ipkhost_plan ipkhost_plan;
int  ipkhost_plan_create(const char* original_newick, const char* ar_tree_newick, int ghost_strategy, int use_unrooted, ipkhost_plan** out);
void ipkhost_plan_free(ipkhost_plan* p);
uint32_t ipkhost_plan_num_ghosts(const ipkhost_plan* p);                  /* ghost matrices to load, in processing order */
uint32_t ipkhost_plan_num_groups(const ipkhost_plan* p);
const char* ipkhost_plan_ext_label(const ipkhost_plan* p, uint32_t i);    /* label in the extended tree ("17_X1") */
const char* ipkhost_plan_ar_label(const ipkhost_plan* p, uint32_t i);     /* label of the same node in the AR output */
const uint32_t* ipkhost_plan_node_branch(const ipkhost_plan* p);          /* [num_ghosts] post-order id of the ORIGINAL node */
uint32_t ipkhost_plan_node_count(const ipkhost_plan* p);                  /* original_tree.get_node_count() */
const uint64_t* ipkhost_plan_index_num_nodes(const ipkhost_plan* p);      /* [node_count] tree index, post-order */
const double* ipkhost_plan_index_lengths(const ipkhost_plan* p);
const char* ipkhost_plan_newick(const ipkhost_plan* p);                   /* to_newick(original_tree): the header's tree */
const char* ipkhost_plan_extended_newick(const ipkhost_plan* p);          /* what save_extended_tree writes (main.cpp:39-46) */

/* ---- .ipk database file ------------------------------------------------------------------------------------
 * Layout (UNVERIFIED against i2l, whose source is absent here -- SURVEY.md 8f1; everything below is written and read
 * through this one module so that a correction is local):
 *   preamble of a boost::archive::binary_oarchive with default flags (db_builder.cpp:176-177): u64 length 22, the 22 bytes
 *   "serialization::archive", u16 library version 19 (BOOST_ARCHIVE_VERSION of Boost 1.76 - 1.8x), then what
 *   basic_binary_oprimitive::init adds: u8 sizeof(int), u8 sizeof(long), u8 sizeof(float), u8 sizeof(double), i32 1;
 *   then raw little-endian primitives:
 *   header   string sequence_type ("DNA" | "Proteins"); u64 n_index, n_index x { u64 num_nodes; f64 subtree_branch_length };
 *            string tree (newick); u64 kmer_size; f32 omega; u64 num_kmers; u64 num_entries
 *   k-mers   num_kmers x { u64 key; f32 filter_value; u64 n; n x { u32 branch; f32 score } }   in the order given
 *   (strings: u64 length + bytes). */
typedef struct ipkhost_db
{
    const char* sequence_type;
    const char* tree;
    uint64_t n_index;
    const uint64_t* index_num_nodes;           /* [n_index] */
    const double* index_subtree_branch_length; /* [n_index] */
    uint64_t kmer_size;
    float omega;
    uint64_t n_kmers, n_entries;
    const uint64_t* keys;                      /* [n_kmers] */
    const float* fv;                           /* [n_kmers] */
    const uint64_t* offsets;                   /* [n_kmers + 1] */
    const uint32_t* counts;                    /* NULL (contiguous: k-mer q owns [offsets[q], offsets[q+1])) or [n_kmers]: k-mer q
                                                * owns [offsets[q], offsets[q] + counts[q]) -- ipkgpu's indexed results */
    const uint32_t* branch;                    /* [n_entries], or NULL when branch16 is set */
    const float* score;                        /* [n_entries] */
    const uint16_t* branch16;                  /* NULL, or the branch ids as 16-bit values (ipkgpu's IPKGPU_RESULT_BRANCH16) */
    const uint32_t* position;                  /* NULL, or [n_entries] alignment positions of a KEEP_POSITIONS build (ipk-aa-pos): entries
                                                * are then written as { u32 branch; f32 score; u16 position } (pos_type; UNVERIFIED(i2l)) */
} ipkhost_db;

int  ipkhost_write_ipk(const char* path, const ipkhost_db* db);                 /* all cores */
/* n_threads <= 0: all cores.  The k-mers are cut into runs of equal bytes (a record is 20 + 8 n bytes), every thread
 * formats its run into private 8 MiB blocks and pwrites them at the prefix-summed offset. */
int  ipkhost_write_ipk_mt(const char* path, const ipkhost_db* db, int n_threads);
/* The shards of a multi-GPU build (one per rank: disjoint k-mer sets, each in ascending (filter value, key)) as ONE database:
 * ipkhost_merge_order is the N-way merge by (filter value, key) of merge_stage2 (db_builder.cpp:392-458) over the k-mer
 * heads only -- the k-mer at merged position i is k-mer out_index[i] of shard out_shard[i], its entries stay where they are;
 * n_threads workers own disjoint filter-value ranges.  ipkhost_write_ipk_sharded streams that order into one file (header
 * fields from `header`, counts summed over the shards), as merge_stage2 does while it merges. */
int  ipkhost_merge_order(uint32_t n_shards, const ipkhost_db* shards, int n_threads, uint8_t* out_shard, uint32_t* out_index);
int  ipkhost_write_ipk_sharded(const char* path, const ipkhost_db* header, uint32_t n_shards, const ipkhost_db* shards,
                               const uint8_t* order_shard, const uint32_t* order_index, int n_threads);
/* the order-independent digest of include/ipkgpu.h (ipkgpu_result_digest) over databases in HOST memory:
 * out[0..5] = entries sum / xor, k-mers sum / xor, filter values sum / xor -- what arrived on the host is what the device built */
int  ipkhost_digest(uint32_t n_shards, const ipkhost_db* shards, int n_threads, uint64_t* out /*[6]*/);
/* reads a file written by ipkhost_write_ipk; arrays are owned by the returned handle */
typedef struct ipkhost_db_file ipkhost_db_file;
int  ipkhost_read_ipk(const char* path, ipkhost_db_file** out, ipkhost_db* view);
void ipkhost_free_ipk(ipkhost_db_file* f);

#ifdef __cplusplus
}
#endif
#endif
