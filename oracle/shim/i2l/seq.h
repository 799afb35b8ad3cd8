// TEST INFRASTRUCTURE (oracle) -- minimal stand-in for the absent phylo42/i2l submodule
// (pinned 553629790985196adca69970f19313a032ec50a8, see /root/reference/.SUBMODULES.json).
// Only the symbols that the reference's ipk/src/pk_compute.cpp and ipk/src/window.cpp touch:
//   seq_traits::alphabet_size   (ipk/include/window.h:13, ipk/src/pk_compute.cpp:17)
//   seq_type, bit_length<>()    (ipk/src/pk_compute.cpp:99,103)
// Values: DNA 4 states / 2 bits, AA 20 states / 5 bits (UNVERIFIED(i2l); consistent with
// ipk/src/ar.cpp:225,232-234 and CHANGELOG.txt:23).
#ifndef IPK_ORACLE_SHIM_I2L_SEQ_H
#define IPK_ORACLE_SHIM_I2L_SEQ_H
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include <tuple>

namespace i2l
{
    struct dna { static constexpr const char* name = "DNA"; };
    struct aa  { static constexpr const char* name = "Proteins"; };

#ifdef SEQ_TYPE_AA
    using seq_type = aa;
#else
    using seq_type = dna;
#endif

    template<class Seq> struct seq_traits_impl;
    template<> struct seq_traits_impl<dna> { static constexpr size_t alphabet_size = 4; };
    template<> struct seq_traits_impl<aa>  { static constexpr size_t alphabet_size = 20; };
    using seq_traits = seq_traits_impl<seq_type>;

    template<class Seq> constexpr uint64_t bit_length();
    template<> constexpr uint64_t bit_length<dna>() { return 2; }
    template<> constexpr uint64_t bit_length<aa>()  { return 5; }
}
#endif
