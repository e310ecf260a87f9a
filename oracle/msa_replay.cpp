// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Exercises the SeqAn plumbing of sviper_b200/host/shim/sviper_b200_msa.hpp WITHOUT a GPU: the staged globalMsaAlignment
// (trace matrices handed to SeqAn's own traceback, the four local rounds with recorded forbidden cells, fragment order,
// distance matrix) with the fills supplied by the oracle's C restatement (oracle_msa.c) instead of the device. Reads the MSA
// cases of a case file (tests/golden/msa_cases.txt), runs the reference's UNCHANGED build_consensus over the shim's
// globalMsaAlignment overload and prints "MSA len=.. seq=.." lines in ref_harness's format; tests/test_msa_golden.py compares
// them with what the unmodified reference printed. The GPU counterpart is sviper_b200/host/msa_harness.cpp.
//
// build: oracle/Makefile (needs /root/reference; links oracle_all.c's fills statically).
#include <thread>
#include <omp.h>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "../sviper_b200/host/shim/sviper_b200_msa.hpp"

extern "C" void orc_msa_sw_fill(const uint8_t *h, int n, const uint8_t *v, int m, int mt, int mm, int ge, int go,
                                const uint8_t *forbidden, uint8_t *trace, int *out4);
extern "C" int orc_msa_global_fill(const uint8_t *h, int n, const uint8_t *v, int m, int mt, int mm, int ge, int go,
                                   uint8_t *trace, int *end3);

struct OracleFills
{
    std::vector<std::vector<uint8_t> > h, v, forb;
    void begin(std::vector<std::vector<uint8_t> > const & hh, std::vector<std::vector<uint8_t> > const & vv)
    {
        h = hh; v = vv; forb.assign(h.size(), std::vector<uint8_t>());
    }
    void global(std::vector<int32_t> const & idx, std::vector<uint8_t *> const & trace, std::vector<svp_msa_fill_result> & res)
    {
        res.resize(idx.size());
        for (size_t i = 0; i < idx.size(); ++i)
        {
            int p = idx[i], end3[3];
            res[i].score = orc_msa_global_fill(h[p].data(), (int)h[p].size(), v[p].data(), (int)v[p].size(), 5, -3, -1, -3, trace[i], end3);
            res[i].S = end3[0]; res[i].H = end3[1]; res[i].V = end3[2]; res[i].row = (int)v[p].size(); res[i].col = (int)h[p].size(); res[i].dir = 0;
        }
    }
    void local(std::vector<int32_t> const & idx, std::vector<std::vector<int32_t> > const & new_cells,
               std::vector<uint8_t *> const & trace, std::vector<svp_msa_fill_result> & res)
    {
        res.resize(idx.size());
        for (size_t i = 0; i < idx.size(); ++i)
        {
            int p = idx[i], out4[4];
            size_t n = h[p].size(), m = v[p].size();
            if (!new_cells[i].empty() && forb[p].empty()) forb[p].assign(n * m, 0);
            for (size_t c = 0; c + 1 < new_cells[i].size(); c += 2)
                forb[p][(size_t)(new_cells[i][c] - 1) * m + (size_t)(new_cells[i][c + 1] - 1)] = 1;
            orc_msa_sw_fill(h[p].data(), (int)n, v[p].data(), (int)m, 5, -3, -1, -3, forb[p].empty() ? nullptr : forb[p].data(), trace[i], out4);
            res[i].score = out4[0]; res[i].row = out4[1]; res[i].col = out4[2]; res[i].dir = out4[3]; res[i].S = res[i].H = res[i].V = 0;
        }
    }
};

// the overload build_consensus resolves to in THIS test program: the shim's staged alignment over the oracle's fills
namespace seqan
{
inline void globalMsaAlignment(Align<Dna5String, ArrayGaps> & align, Score<int, Simple> const & scoreObject)
{
    OracleFills fills;
    sviper_b200::shim::global_msa_alignment(fills, align, scoreObject);
}
}  // namespace seqan

#include <sviper/config.h>
#include <sviper/basics.h>

int main(int argc, char ** argv)
{
    if (argc < 2) { std::cerr << "usage: msa_replay cases.txt\n"; return 2; }
    std::ifstream in(argv[1]);
    std::string line, cmd;
    while (in >> cmd)
    {
        if (cmd != "MSA") { std::getline(in, line); continue; }
        unsigned n; in >> n;
        seqan::StringSet<seqan::Dna5String> seqs;
        std::vector<double> quals;
        for (unsigned i = 0; i < n; ++i)
        {
            std::string sq; double q; in >> sq >> q;
            seqan::appendValue(seqs, seqan::Dna5String(sq)); quals.push_back(q);
        }
        seqan::Dna5String cns = sviper::build_consensus(seqs, quals);
        std::string out(seqan::length(cns), 'N');
        for (size_t i = 0; i < out.size(); ++i) out[i] = (char)cns[i];
        std::cout << "MSA len=" << out.size() << " seq=" << (out.empty() ? std::string("*") : out) << "\n";
    }
    return 0;
}
