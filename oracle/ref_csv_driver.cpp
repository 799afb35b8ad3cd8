// TEST INFRASTRUCTURE.  The reference's own csv reader (strasser/csv-parser csv.h, compiled unmodified from
// /root/reference) instantiated exactly as ar::raxmlng_reader::read_node instantiates it (ipk/src/ar.cpp:200-259),
// followed by the statements of that function that turn a row into a matrix column (AA reorder :232-234, log10
// :258-259).  Used to pin ipk_b200/host/ar_loader.cpp bit-for-bit.   usage: ref_csv_{dna,aa} in.tsv out.f32
#include <csv-parser/csv.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <string>

using score_type = float;      // i2l::phylo_kmer::score_type (shim/i2l/phylo_kmer.h)

int main(int argc, char** argv)
{
    if (argc != 3) return 2;
#if defined(SEQ_TYPE_AA)
    constexpr size_t sigma = 20;
#else
    constexpr size_t sigma = 4;
#endif
    std::ifstream file_stream(argv[1]);
    std::string header;
    std::getline(file_stream, header);                       // build_index skips the header line (ar.cpp:162)
    ::io::CSVReader<3 + sigma, ::io::trim_chars<' '>, ::io::no_quote_escape<'\t'>, ::io::throw_on_overflow,
                    ::io::single_and_empty_line_comment<'.'>> in(argv[1], file_stream);
    FILE* out = fopen(argv[2], "wb");
    if (!out) return 3;
    std::string node_label, site, state;
    try
    {
#if defined(SEQ_TYPE_AA)
        score_type a, r, n, d, c, q, e, g, h, i, l, k, m, f, p, s, t, w, y, v;
        while (in.read_row(node_label, site, state, a, r, n, d, c, q, e, g, h, i, l, k, m, f, p, s, t, w, y, v))
        {
            auto new_column = std::array<score_type, 20>{ r, h, k, d, e, s, t, n, q, c, g, p, a, i, l, m, f, w, y, v };
#else
        score_type a, c, g, t;
        while (in.read_row(node_label, site, state, a, c, g, t))
        {
            auto new_column = std::array<score_type, 4>{ a, c, g, t };
#endif
            auto log = [](auto value) { return std::log10(value); };
            std::transform(begin(new_column), end(new_column), begin(new_column), log);
            fwrite(new_column.data(), sizeof(score_type), new_column.size(), out);
        }
    }
    catch (const std::exception& ex)
    {
        fprintf(stderr, "%s\n", ex.what());
        fclose(out);
        return 1;
    }
    fclose(out);
    return 0;
}
