// TEST INFRASTRUCTURE ONLY — never linked into or called from the product path.
//
// Tiny driver around the *reference's own* in-memory SBWT constructor
// (/root/reference/SBWT/include/sbwt/NodeBOSSInMemoryConstructor.hh:166-190, the constructor
// Finito's tests use at src/tests.cpp:44-46).  It writes the plain-matrix ".sbwt" file that
// `build-fmin -i` expects: variant string + SBWT::serialize (SBWT/include/sbwt/SBWT.hh:395-424,
// header written the way SBWT/src/CLI/sbwt_build.cpp does it), optionally with the k-mer prefix
// precalc table (`sbwt build` default is 8, sbwt_build.cpp:44).  Used to (a) produce golden
// fixtures and (b) check finito_b200's own SBWT construction tool bit for bit.
//
// usage: sbwt_inmem <unitigs.fna> <out.sbwt> <k> [precalc_k=0]
#include <fstream>
#include <iostream>
#include <string>
#include <vector>
#include "globals.hh"
#include "SBWT.hh"
#include "SeqIO.hh"
#include "variants.hh"
#include "NodeBOSSInMemoryConstructor.hh"

int main(int argc, char** argv) {
    if (argc < 4) {
        std::cerr << "usage: " << argv[0] << " <unitigs.fna> <out.sbwt> <k> [precalc_k]\n";
        return 2;
    }
    const std::string in_path = argv[1], out_path = argv[2];
    const int64_t k = std::stoll(argv[3]);
    int64_t precalc_k = argc > 4 ? std::stoll(argv[4]) : 0;
    if (precalc_k > k) precalc_k = k;

    std::vector<std::string> seqs;
    sbwt::SeqIO::Reader<> reader(in_path);
    while (true) {
        int64_t len = reader.get_next_read_to_buffer();
        if (len == 0) break;
        seqs.emplace_back(reader.read_buf);
    }
    sbwt::plain_matrix_sbwt_t index;
    sbwt::NodeBOSSInMemoryConstructor<sbwt::plain_matrix_sbwt_t> ctor;
    ctor.build(seqs, index, k, true);
    if (precalc_k > 0) index.do_kmer_prefix_precalc(precalc_k);

    sbwt::throwing_ofstream out(out_path, std::ios::binary);
    sbwt::serialize_string("plain-matrix", out.stream);
    index.serialize(out.stream);
    std::cerr << "n_nodes " << index.number_of_subsets() << " n_kmers " << index.number_of_kmers() << "\n";
    return 0;
}
