// sviper_b200/host/msa_harness.cpp -- case-file driver for build_consensus (include/sviper/basics.h:484-520) as the GPU
// programs run it: the reference's unchanged function over the shim (host/shim/sviper/basics.h), i.e. the pairwise DP fills
// on the device (svp_msa_*) and everything else SeqAn's own host code. Reads the "MSA n / seq mapq" cases of a case file and
// prints "MSA len=.. seq=.." lines in the format of oracle/ref_harness, so that tests/test_gpu_msa.py can compare them with
// what the unmodified reference printed (tests/golden/msa_expected.txt).
//   msa_harness cases.txt [repeat]      repeat > 1: run the file several times and print timing to stderr
// Built by sviper_b200/csrc/Makefile (needs the reference tree); the binary travels to the GPU box.
#include <thread>
#include <omp.h>
#include <chrono>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <sviper/config.h>
#include <sviper/basics.h>

int main(int argc, char ** argv)
{
    if (argc < 2) { std::cerr << "usage: msa_harness cases.txt [repeat]\n"; return 2; }
    int const repeat = argc > 2 ? std::atoi(argv[2]) : 1;
    std::vector<seqan::StringSet<seqan::Dna5String> > sets;
    std::vector<std::vector<double> > quals;
    {
        std::ifstream in(argv[1]);
        std::string line, cmd;
        while (in >> cmd)
        {
            if (cmd != "MSA") { std::getline(in, line); continue; }
            unsigned n; in >> n;
            sets.emplace_back(); quals.emplace_back();
            for (unsigned i = 0; i < n; ++i)
            {
                std::string sq; double q; in >> sq >> q;
                seqan::appendValue(sets.back(), seqan::Dna5String(sq)); quals.back().push_back(q);
            }
        }
    }
    for (int r = 0; r < repeat; ++r)
    {
        auto t0 = std::chrono::steady_clock::now();
        for (size_t c = 0; c < sets.size(); ++c)
        {
            seqan::Dna5String cns = sviper::build_consensus(sets[c], quals[c]);
            if (r) continue;
            std::string out(seqan::length(cns), 'N');
            for (size_t i = 0; i < out.size(); ++i) out[i] = (char)cns[i];
            std::cout << "MSA len=" << out.size() << " seq=" << (out.empty() ? std::string("*") : out) << "\n";
        }
        if (repeat > 1)
            std::cerr << "[msa_harness] pass " << r << ": " << sets.size() << " consensus calls in "
                      << std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() << " s\n";
    }
    return 0;
}
