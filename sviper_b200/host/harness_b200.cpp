// Case-file driver over sviper_b200/host/polisher.hpp: reads the same case files as the reference harness
// (oracle/ref_harness.cpp, tests/golden/cases.txt) and prints the POLISH and FINAL result lines in the same
// text format, so the two outputs can be compared with diff. Everything is computed on the GPU through the C ABI.
//
//   POLISH  flank ins_mean ins_sd n ref      then n lines "r1 q1 r2 q2"   -> Polisher / BatchScheduler
//   FINAL   ref_part polished                                              -> Polisher::final_alignment
// Other commands of the case format (ALIGN, ALIGNX, ROUND) are consumed and answered with "SKIP".
//
// usage: harness_b200 cases.txt [n_gpus]
#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>

#include "polisher.hpp"
#include "prepare.hpp"

using namespace sviper_b200;

static std::string hexd(double d)
{
    char buf[64];
    snprintf(buf, sizeof buf, "%a", d);
    return buf;
}

static QRead mkread(std::string const & s, std::string const & q)
{
    QRead r;
    if (s != "*") { r.seq = s; r.qual = q; }
    return r;
}

struct PolishCase
{
    Config cfg;
    std::future<std::unique_ptr<ResultBuffers>> fut;
};

// EVAL lines (tests/golden/eval_cases.txt): the host arithmetic after the final alignment, no GPU involved.
// Prints what oracle/ref_harness.cpp prints for the same line.
static int eval_cases(char const * path)
{
    std::ifstream in(path);
    if (!in) { std::cerr << "cannot open " << path << "\n"; return 2; }
    std::string cmd;
    while (in >> cmd)
    {
        if (cmd != "EVAL") { std::cerr << "unknown command " << cmd << "\n"; return 2; }
        std::string type, cig, sq; int k, begin;
        VariantFields var;
        in >> type >> var.ref_pos >> var.ref_pos_end >> var.sv_length >> k >> begin >> var.info >> cig >> sq;
        var.sv_type = type == "DEL" ? SVP_SV_DEL : SVP_SV_INS;
        var.filter = "PASS";
        std::vector<CigarElement> cigar;
        for (size_t i = 0; i < cig.size();)
        {
            uint32_t c = 0;
            while (i < cig.size() && cig[i] >= '0' && cig[i] <= '9') c = c * 10 + (uint32_t)(cig[i++] - '0');
            cigar.push_back(CigarElement{cig[i++], c});
        }
        bool has = evaluate_final_mapping(var, cigar, sq == "*" ? "" : sq, begin, k);
        std::cout << "EVAL has=" << has << " pos=" << var.ref_pos << " end=" << var.ref_pos_end << " svlen=" << var.sv_length
                  << " qual=" << hexd(var.quality) << " qualstr=" << var.quality_string() << " filter=" << var.filter
                  << " info=" << var.info << "\n";
    }
    return 0;
}

static std::vector<CigarElement> parse_cigar(std::string const & cig)
{
    std::vector<CigarElement> cigar;
    for (size_t i = 0; i < cig.size();)
    {
        uint32_t c = 0;
        while (i < cig.size() && cig[i] >= '0' && cig[i] <= '9') c = c * 10 + (uint32_t)(cig[i++] - '0');
        cigar.push_back(CigarElement{cig[i++], c});
    }
    return cigar;
}

// SUPPORT / REGION / CUTCOV lines (tests/golden/prep_cases.txt): the CIGAR walks before the hot path, no GPU involved.
// Prints what oracle/ref_harness.cpp prints for the same lines.
static int prep_cases(char const * path)
{
    std::ifstream in(path);
    if (!in) { std::cerr << "cannot open " << path << "\n"; return 2; }
    std::string cmd;
    while (in >> cmd)
    {
        if (cmd == "SUPPORT")
        {
            std::string type, cig; AlignedRecord rec; VariantFields var;
            in >> type >> var.ref_pos >> var.sv_length >> rec.beginPos >> cig;
            var.sv_type = type == "DEL" ? SVP_SV_DEL : SVP_SV_INS;
            var.ref_pos_end = type == "DEL" ? var.ref_pos + var.sv_length : var.ref_pos;
            rec.cigar = parse_cigar(cig);
            std::cout << "SUPPORT supports=" << record_supports_variant(rec, var) << "\n";
        }
        else if (cmd == "REGION")
        {
            std::string cig; AlignedRecord rec; int rb, re;
            in >> rec.beginPos >> cig >> rb >> re;
            rec.qName = "r";
            rec.cigar = parse_cigar(cig);
            auto t = get_read_region_boundaries(rec, rb, re);
            std::cout << "REGION begin=" << std::get<0>(t) << " end=" << std::get<1>(t) << "\n";
        }
        else if (cmd == "CUTCOV")
        {
            int cov; unsigned n;
            in >> cov >> n;
            std::vector<AlignedRecord> recs(n);
            for (unsigned i = 0; i < n; ++i)
            {
                std::string cig; int mq;
                in >> recs[i].qName >> recs[i].beginPos >> mq >> cig;
                recs[i].mapQ = (uint8_t)mq;
                recs[i].cigar = parse_cigar(cig);
            }
            cut_down_high_coverage(recs, (int16_t)cov);
            std::cout << "CUTCOV n=" << recs.size() << " keep=";
            for (size_t i = 0; i < recs.size(); ++i) std::cout << (i ? "," : "") << recs[i].qName;
            std::cout << "\n";
        }
        else if (cmd == "MERGE")
        {
            unsigned n;
            in >> n;
            std::vector<AlignedRecord> recs(n);
            for (unsigned i = 0; i < n; ++i)
            {
                std::string cig; int flag, mq;
                in >> recs[i].qName >> flag >> recs[i].rID >> recs[i].beginPos >> mq >> cig >> recs[i].seq;
                recs[i].flag = (uint16_t)flag; recs[i].mapQ = (uint8_t)mq;
                recs[i].cigar = parse_cigar(cig);
            }
            std::vector<AlignedRecord> out;
            try { out = merge_alignments(recs); }
            catch (std::logic_error const &) { std::cout << "MERGE undefined\n"; continue; }   // out-of-bounds access in the reference
            std::cout << "MERGE n=" << out.size();
            for (AlignedRecord const & r : out)
            {
                std::cout << " " << r.qName << ":" << r.flag << ":" << r.beginPos << ":" << (int)r.mapQ << ":";
                for (CigarElement const & e : r.cigar) std::cout << e.count << e.operation;
                std::cout << ":" << r.seq;
            }
            std::cout << "\n";
        }
        else { std::cerr << "unknown command " << cmd << "\n"; return 2; }
    }
    return 0;
}

int main(int argc, char ** argv)
{
    if (argc < 2) { std::cerr << "usage: harness_b200 cases.txt [n_gpus] | harness_b200 --eval eval_cases.txt | harness_b200 --prep prep_cases.txt\n"; return 2; }
    if (std::string(argv[1]) == "--eval") return argc > 2 ? eval_cases(argv[2]) : 2;
    if (std::string(argv[1]) == "--prep") return argc > 2 ? prep_cases(argv[2]) : 2;
    const int n_gpus = argc > 2 ? atoi(argv[2]) : 1;
    std::ifstream in(argv[1]);
    if (!in) { std::cerr << "cannot open " << argv[1] << "\n"; return 2; }
    std::vector<int> devs;
    for (int g = 0; g < n_gpus; ++g) devs.push_back(g);

    std::vector<std::string> lines;          // output, in input order
    std::vector<std::pair<size_t, PolishCase>> pending;
    try
    {
        BatchScheduler sched(devs);
        Polisher single(0);
        std::string cmd;
        while (in >> cmd)
        {
            if (cmd == "ALIGNX") { std::string a; for (int i = 0; i < 6; ++i) in >> a; lines.push_back("SKIP"); }
            else if (cmd == "ALIGN") { std::string a; for (int i = 0; i < 3; ++i) in >> a; lines.push_back("SKIP"); }
            else if (cmd == "ROUND" || cmd == "POLISH")
            {
                int fix = 0; int flank; double im, isd; unsigned n; std::string ref;
                if (cmd == "ROUND") in >> fix;
                in >> flank >> im >> isd >> n >> ref;
                std::vector<QRead> r1, r2;
                for (unsigned i = 0; i < n; ++i)
                {
                    std::string a, qa, b, qb;
                    in >> a >> qa >> b >> qb;
                    r1.push_back(mkread(a, qa)); r2.push_back(mkread(b, qb));
                }
                if (cmd == "ROUND") { lines.push_back("SKIP"); continue; }
                PolishCase pc;
                pc.cfg.mean_insert_size = im; pc.cfg.stdev_insert_size = isd; pc.cfg.ref_flank_length = flank;
                compute_baseQ_stats(pc.cfg, r1, r2);
                std::unique_ptr<PackedVariant> pv(new PackedVariant(r1, r2, ref, pc.cfg));
                pc.fut = sched.submit(std::move(pv));
                lines.push_back("");
                pending.emplace_back(lines.size() - 1, std::move(pc));
            }
            else if (cmd == "FINAL")
            {
                std::string a, b; in >> a >> b;
                double s = 0;
                std::vector<CigarElement> cigar = single.final_alignment(a == "*" ? "" : a, b == "*" ? "" : b, &s);
                std::ostringstream o;
                o << "FINAL score=" << hexd(s) << " ncigar=" << cigar.size() << " cigar=";
                for (CigarElement const & c : cigar) o << c.count << c.operation;
                lines.push_back(o.str());
            }
            else { std::cerr << "unknown command " << cmd << "\n"; return 2; }
        }
        for (auto & p : pending)
        {
            std::unique_ptr<ResultBuffers> rb = p.second.fut.get();
            if (rb->res.status != SVP_OK) { std::cerr << "variant failed with status " << rb->res.status << "\n"; return 1; }
            Config & cfg = p.second.cfg;
            rb->to_config(cfg);
            std::ostringstream o;
            o << "POLISH rounds=" << cfg.rounds << " subst=" << cfg.substituted_bases << " ins=" << cfg.inserted_bases
              << " del=" << cfg.deleted_bases << " baseQ_mean=" << hexd(cfg.baseQ_mean) << " mappQ_mean="
              << hexd(cfg.mappQ_mean) << " seq=" << rb->sequence() << " cov=";
            for (size_t i = 0; i < cfg.cov_profile.size(); ++i) o << (i ? "," : "") << cfg.cov_profile[i];
            lines[p.first] = o.str();
        }
    }
    catch (std::exception const & e)
    {
        std::cerr << "harness_b200: " << e.what() << "\n";
        return 1;
    }
    for (std::string const & l : lines) std::cout << l << "\n";
    return 0;
}
