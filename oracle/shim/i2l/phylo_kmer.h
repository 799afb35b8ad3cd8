// TEST INFRASTRUCTURE (oracle) -- minimal stand-in for i2l/phylo_kmer.h (submodule absent).
// Types recovered from call sites: key u64, score float, branch u32
// (ipk/include/row.h:11-17, ipk/include/pk_compute.h:32, ipk/src/branch_group.cpp:88-101).
#ifndef IPK_ORACLE_SHIM_I2L_PHYLO_KMER_H
#define IPK_ORACLE_SHIM_I2L_PHYLO_KMER_H
#include <cstdint>
#include <vector>
#include <string>
#include <stdexcept>
#include <tuple>
#include "seq.h"

namespace i2l
{
    struct unpositioned_phylo_kmer
    {
        using key_type = uint64_t;
        using score_type = float;
        using branch_type = uint32_t;
        using pos_type = int32_t;

        key_type key;
        score_type score;
    };
    using phylo_kmer = unpositioned_phylo_kmer;
}
#endif
