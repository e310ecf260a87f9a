// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Thin driver around the UNMODIFIED reference headers (compiled from where they
// lie under /root/reference, see oracle/Makefile; output goes to oracle/_ref/).
// It reads hot-path cases from a line-based text file and prints what the
// reference computes, so the C restatement (oracle/sviper_oracle.c) and the CUDA
// path can be pinned against the real SeqAn/SViper code.
//
// Commands (one per line, whitespace separated; '*' = empty string):
//   ALIGNX  mm mx ge go  ref read          generic semi-global affine (SeqAn golden vectors)
//   ALIGN   ref read qual                  sviper::map_single_read      (mapping.h:84-137)
//   ROUND   fix flank ins_mean ins_sd n ref  then n lines "r1 q1 r2 q2"   one polish() round with profile dump
//   POLISH  flank ins_mean ins_sd n ref      then n lines "r1 q1 r2 q2"   polish_to_perfection (polishing.h:319)
//   FINAL   ref_part polished              convex-gap global alignment + CIGAR (sviper.h:452-466); polished = '@': the
//                                          sequence the preceding POLISH returned
//   EVAL    type pos end svlen k begin info cigar seq   assign_quality + refine_variant on the final record
//                                                  (sviper.h:468-489, evaluate_final_mapping.h:17-28, variant.h:277-353)
//   SUPPORT type pos svlen begin cigar         record_supports_variant on a long-read record (variant.h:239-275)
//   REGION  begin cigar ref_begin ref_end      get_read_region_boundaries (basics.h:189-241)
//   CUTCOV  cov n  then n lines "name begin mapq cigar"   cut_down_high_coverage (basics.h:403-442)
//   MSA     n  then n lines "seq mapq"       sviper::build_consensus (basics.h:484-520: globalMsaAlignment + column vote)
//   SWFILL  a b k  then k pairs "col row"    seqan::_alignSmithWaterman (graph_msa/graph_align_tcoffee_library.h:252-390): the
//                                          legacy clumping Smith-Waterman fill of the local pairwise library, forbidden cells given
//   GFILL   a b                            the trace matrix of globalAlignment(String<Fragment>, ..., Gotoh()) as the global pairwise
//                                          library computes it (global_alignment_unbanded.h:386-407 -> dp_algorithm_impl.h:1393-1469)
//   MERGE   n  then n lines "name flag rid begin mapq cigar seq" (sorted by name)   merge_alignments
//                                                  (merge_split_alignments.h:378-561, called at sviper.h:266)
#include <thread>
#include <omp.h>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include <seqan/align.h>
#include <seqan/bam_io.h>
#include <seqan/graph_msa.h>
#include <sviper/config.h>
#include <sviper/polishing.h>
#include <sviper/basics.h>
#include <sviper/variant.h>
#include <sviper/merge_split_alignments.h>
#include <sviper/evaluate_final_mapping.h>

using namespace seqan;

static std::string hexd(double d)
{
    char buf[64];
    snprintf(buf, sizeof buf, "%a", d);
    return buf;
}

template <typename TGaps>
static std::string row(TGaps const & g)
{
    std::string s;
    for (unsigned i = 0; i < length(g); ++i)
        s += isGap(g, i) ? '-' : (char)Dna5(g[i]);
    return s.empty() ? "*" : s;
}

static unsigned long long fnv1a(unsigned char const * p, size_t n)
{
    unsigned long long h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}

static Dna5QString mkread(std::string const & s, std::string const & q)
{
    Dna5QString r;
    if (s == "*") return r;
    for (size_t i = 0; i < s.size(); ++i)
    {
        Dna5Q c = Dna5Q(s[i]);
        assignQualityValue(c, (int)q[i] - 33);
        appendValue(r, c);
    }
    return r;
}

static void read_pairs(std::istream & in, unsigned n, StringSet<Dna5QString> & r1, StringSet<Dna5QString> & r2)
{
    for (unsigned i = 0; i < n; ++i)
    {
        std::string a, qa, b, qb;
        in >> a >> qa >> b >> qb;
        appendValue(r1, mkread(a, qa));
        appendValue(r2, mkread(b, qb));
    }
}

static void parse_cigar(std::string const & cig, String<CigarElement<char, unsigned>> & out)
{
    for (size_t i = 0; i < cig.size();)
    {
        unsigned c = 0;
        while (i < cig.size() && cig[i] >= '0' && cig[i] <= '9') c = c * 10 + (cig[i++] - '0');
        appendValue(out, CigarElement<char, unsigned>(cig[i++], c));
    }
}

static std::string dna(Dna5String const & s)
{
    std::string o;
    for (unsigned i = 0; i < length(s); ++i) o += (char)s[i];
    return o.empty() ? "*" : o;
}

int main(int argc, char ** argv)
{
    if (argc < 2) { std::cerr << "usage: ref_harness cases.txt\n"; return 2; }
    std::ifstream in(argv[1]);
    std::string cmd;
    Dna5String last_polished;
    while (in >> cmd)
    {
        if (cmd == "ALIGNX")
        {
            int mm, mx, ge, go; std::string ref, rd;
            in >> mm >> mx >> ge >> go >> ref >> rd;
            Dna5String h = ref; Dna5String v = rd;
            Gaps<Dna5String, ArrayGaps> gh(h), gv(v);
            int s = globalAlignment(gh, gv, Score<int, Simple>(mm, mx, ge, go),
                                    AlignConfig<true, false, false, true>(), AffineGaps());
            std::cout << "ALIGNX score=" << s << " ref=" << row(gh) << " read=" << row(gv) << "\n";
        }
        else if (cmd == "ALIGN")
        {
            std::string ref, rd, q;
            in >> ref >> rd >> q;
            Dna5String h = ref; Dna5QString r = mkread(rd, q);
            Gaps<Dna5QString, ArrayGaps> gr; Gaps<Dna5String, ArrayGaps> gh;
            assignSource(gr, r); assignSource(gh, h);
            double mapQ = 0;
            bool rc = sviper::map_single_read(gr, gh, r, h, mapQ);
            std::cout << "ALIGN rc=" << rc << " mapQ=" << (int)mapQ
                      << " lead=" << sviper::gapsBeginPos(gr) << " end=" << sviper::gapsEndPos(gr)
                      << " ref=" << row(gh) << " read=" << row(gr) << "\n";
        }
        else if (cmd == "ROUND" || cmd == "POLISH")
        {
            int fix = 0; unsigned flank; double im, isd; unsigned n; std::string ref;
            if (cmd == "ROUND") in >> fix;
            in >> flank >> im >> isd >> n >> ref;
            StringSet<Dna5QString> r1, r2;
            read_pairs(in, n, r1, r2);
            sviper::CmdOptions opt;
            opt.mean_insert_size_of_short_reads = im;
            opt.stdev_insert_size_of_short_reads = isd;
            sviper::SViperConfig config{opt};
            config.ref_flank_length = flank;
            sviper::compute_baseQ_stats(config, r1, r2);
            Dna5String h = ref;
            if (cmd == "POLISH")
            {
                Dna5String out = sviper::polish_to_perfection(r1, r2, h, config);
                last_polished = out;
                std::cout << "POLISH rounds=" << config.rounds << " subst=" << config.substituted_bases
                          << " ins=" << config.inserted_bases << " del=" << config.deleted_bases
                          << " baseQ_mean=" << hexd(config.baseQ_mean) << " mappQ_mean=" << hexd(config.mappQ_mean)
                          << " seq=" << dna(out) << " cov=";
                for (size_t i = 0; i < config.cov_profile.size(); ++i) std::cout << (i ? "," : "") << config.cov_profile[i];
                std::cout << "\n";
            }
            else
            {
                // body of sviper::polish (polishing.h:295-317) with the profile exposed
                config.fix_indels = fix;
                Gaps<Dna5String, ArrayGaps> contigGaps(h);
                std::vector<sviper::Mapping_object> mobs = sviper::mapping(r1, r2, h, config);
                sviper::compute_mappQ_stats(config, mobs);
                String<ProfileChar<Dna5, double> > profile;
                resize(profile, length(h));
                config.cov_profile.clear();
                resize(config.cov_profile, length(h));
                sviper::fill_profile(profile, config.cov_profile, contigGaps, mobs, config);
                Dna5String out = sviper::consensus_from_profile(profile, contigGaps, config);
                std::cout << "ROUND subst=" << config.substituted_bases << " ins=" << config.inserted_bases
                          << " del=" << config.deleted_bases << " baseQ_mean=" << hexd(config.baseQ_mean)
                          << " mappQ_mean=" << hexd(config.mappQ_mean) << " seq=" << dna(out) << " proper=";
                for (size_t i = 0; i < mobs.size(); ++i) std::cout << (mobs[i].proper_pair ? '1' : '0');
                std::cout << " gaps=" << row(contigGaps) << " cov=";
                for (size_t i = 0; i < config.cov_profile.size(); ++i) std::cout << (i ? "," : "") << config.cov_profile[i];
                std::cout << " profile=";
                for (unsigned i = 0; i < length(profile); ++i)
                    for (int k = 0; k < 6; ++k) std::cout << ((i || k) ? "," : "") << hexd(profile[i].count[k]);
                std::cout << "\n";
            }
        }
        else if (cmd == "FINAL")
        {
            std::string a, b; in >> a >> b;
            Dna5String ref_part = a, polished = b;
            if (b == "@") polished = last_polished;   // the output of the preceding POLISH (sviper.h:431 -> :457)
            sviper::CmdOptions opt; sviper::SViperConfig config{opt};
            Gaps<Dna5String, ArrayGaps> gapsRef(ref_part), gapsSeq(polished);
            double s = globalAlignment(gapsRef, gapsSeq, Score<double, Simple>(config.MM, config.MX, config.GE, config.GO),
                                       AlignConfig<false, false, false, false>(), ConvexGaps());
            String<CigarElement<char, unsigned> > cigar;
            getCigarString(cigar, gapsRef, gapsSeq, std::numeric_limits<unsigned>::max());
            std::cout << "FINAL score=" << hexd(s) << " ncigar=" << length(cigar) << " cigar=";
            for (unsigned i = 0; i < length(cigar); ++i) std::cout << cigar[i].count << cigar[i].operation;
            std::cout << "\n";
        }
        else if (cmd == "EVAL")
        {
            std::string type, info, cig, sq; int pos, end, svlen, k, begin;
            in >> type >> pos >> end >> svlen >> k >> begin >> info >> cig >> sq;
            sviper::Variant var;
            var.sv_type = type == "DEL" ? sviper::SV_TYPE::DEL : sviper::SV_TYPE::INS;
            var.ref_pos = pos; var.ref_pos_end = end; var.sv_length = svlen; var.info = info; var.filter = "PASS";
            BamAlignmentRecord rec;
            rec.beginPos = begin;
            if (sq != "*") rec.seq = sq;
            for (size_t i = 0; i < cig.size();)
            {
                unsigned c = 0;
                while (i < cig.size() && cig[i] >= '0' && cig[i] <= '9') c = c * 10 + (cig[i++] - '0');
                appendValue(rec.cigar, CigarElement<char, unsigned>(cig[i++], c));
            }
            sviper::CmdOptions opt; opt.flanking_region = k;
            sviper::SViperConfig config{opt};
            sviper::assign_quality(rec, var, config);
            bool has = sviper::refine_variant(rec, var);
            if (!has) var.filter = "FAIL5";                       // sviper.h:476-478
            std::ostringstream qs; qs << var.quality;             // variant.h:142 (default ostream precision)
            std::cout << "EVAL has=" << has << " pos=" << var.ref_pos << " end=" << var.ref_pos_end
                      << " svlen=" << var.sv_length << " qual=" << hexd(var.quality) << " qualstr=" << qs.str()
                      << " filter=" << var.filter << " info=" << var.info << "\n";
        }
        else if (cmd == "SUPPORT")
        {
            std::string type, cig; int pos, svlen, begin;
            in >> type >> pos >> svlen >> begin >> cig;
            sviper::Variant var;
            var.sv_type = type == "DEL" ? sviper::SV_TYPE::DEL : sviper::SV_TYPE::INS;
            var.ref_pos = pos; var.sv_length = svlen;
            var.ref_pos_end = type == "DEL" ? pos + svlen : pos;
            BamAlignmentRecord rec;
            rec.beginPos = begin;
            parse_cigar(cig, rec.cigar);
            std::cout << "SUPPORT supports=" << sviper::record_supports_variant(rec, var) << "\n";
        }
        else if (cmd == "REGION")
        {
            std::string cig; int begin, rb, re;
            in >> begin >> cig >> rb >> re;
            BamAlignmentRecord rec;
            rec.beginPos = begin; rec.qName = "r";
            parse_cigar(cig, rec.cigar);
            auto t = sviper::get_read_region_boundaries(rec, rb, re);
            std::cout << "REGION begin=" << std::get<0>(t) << " end=" << std::get<1>(t) << "\n";
        }
        else if (cmd == "CUTCOV")
        {
            int cov; unsigned n;
            in >> cov >> n;
            std::vector<BamAlignmentRecord> recs(n);
            for (unsigned i = 0; i < n; ++i)
            {
                std::string name, cig; int begin, mq;
                in >> name >> begin >> mq >> cig;
                recs[i].qName = name; recs[i].beginPos = begin; recs[i].mapQ = mq;
                parse_cigar(cig, recs[i].cigar);
            }
            sviper::cut_down_high_coverage(recs, (int16_t)cov);
            std::cout << "CUTCOV n=" << recs.size() << " keep=";
            for (size_t i = 0; i < recs.size(); ++i) std::cout << (i ? "," : "") << recs[i].qName;
            std::cout << "\n";
        }
        else if (cmd == "MSA")
        {
            unsigned n; in >> n;
            StringSet<Dna5String> seqs;
            std::vector<double> quals;
            for (unsigned i = 0; i < n; ++i)
            {
                std::string sq; double q; in >> sq >> q;
                appendValue(seqs, Dna5String(sq)); quals.push_back(q);
            }
            Dna5String cns = sviper::build_consensus(seqs, quals);
            std::cout << "MSA len=" << length(cns) << " seq=" << (length(cns) ? dna(cns) : std::string("*")) << "\n";
        }
        else if (cmd == "SWFILL")
        {
            std::string a, b; unsigned k; in >> a >> b >> k;
            Dna5String s0 = a, s1 = b;
            StringSet<Dna5String, Dependent<> > str;
            appendValue(str, s0); appendValue(str, s1);
            String<bool> forbidden;
            resize(forbidden, length(s0) * length(s1), false);
            for (unsigned i = 0; i < k; ++i) { size_t col, row; in >> col >> row; forbidden[(col - 1) * length(s1) + (row - 1)] = true; }
            String<unsigned char> trace;
            unsigned char initialDir;
            size_t indexPair[2];
            int sc = _alignSmithWaterman(trace, str, Score<int, Simple>(5, -3, -1, -3), initialDir, indexPair, forbidden);
            std::cout << "SWFILL score=" << sc << " row=" << indexPair[0] << " col=" << indexPair[1] << " dir=" << (int)initialDir
                      << " tracelen=" << length(trace) << " hash=" << fnv1a(begin(trace, Standard()), length(trace)) << "\n";
        }
        else if (cmd == "GFILL")
        {
            std::string a, b; in >> a >> b;
            Dna5String seqH = a, seqV = b;
            Score<int, Simple> sc(5, -3, -1, -3);
            // the set-up of _computeAlignment (dp_algorithm_impl.h:1393-1469) with the fill and the traceback as separate steps
            typedef DPCell_<int, AffineGaps> TCell;
            typedef TraceBitMap_<>::Type TTraceValue;
            typedef SetupAlignmentProfile_<DPGlobal, FreeEndGaps_<False, False, False, False>, AffineGaps,
                                           TracebackOn<TracebackConfig_<SingleTrace, GapsLeft> > >::Type TDPProfile;
            typedef DPContext<TCell, TTraceValue> TContext;
            TContext dpContext;
            typedef std::remove_reference<decltype(getDpScoreMatrix(dpContext))>::type TScoreHost;
            typedef std::remove_reference<decltype(getDpTraceMatrix(dpContext))>::type TTraceHost;
            typedef DPMatrix_<TCell, SparseDPMatrix, TScoreHost> TDPScoreMatrix;
            typedef DPMatrix_<TTraceValue, FullDPMatrix, TTraceHost> TDPTraceMatrix;
            typedef DPMatrixNavigator_<TDPScoreMatrix, DPScoreMatrix, NavigateColumnWise> TScoreNav;
            typedef DPMatrixNavigator_<TDPTraceMatrix, DPTraceMatrix<TracebackOn<TracebackConfig_<SingleTrace, GapsLeft> > >, NavigateColumnWise> TTraceNav;
            typedef DPScout_<TCell, Default> TDPScout;
            DPBandConfig<BandOff> band;
            TDPScoreMatrix dpScoreMatrix; TDPTraceMatrix dpTraceMatrix;
            setLength(dpScoreMatrix, +DPMatrixDimension_::HORIZONTAL, length(seqH) + 1);
            setLength(dpTraceMatrix, +DPMatrixDimension_::HORIZONTAL, length(seqH) + 1);
            setLength(dpScoreMatrix, +DPMatrixDimension_::VERTICAL, length(seqV) + 1);
            setLength(dpTraceMatrix, +DPMatrixDimension_::VERTICAL, length(seqV) + 1);
            setHost(dpScoreMatrix, getDpScoreMatrix(dpContext));
            setHost(dpTraceMatrix, getDpTraceMatrix(dpContext));
            resize(dpScoreMatrix); resize(dpTraceMatrix);
            TScoreNav scoreNav{dpScoreMatrix, band};
            TTraceNav traceNav{dpTraceMatrix, band};
            DPScoutState_<Default> scoutState;
            TDPScout dpScout(scoutState);
            _computeAlignmentImpl(dpScout, scoreNav, traceNav, seqH, seqV, sc, band, TDPProfile(), NavigateColumnWise{});
            TTraceHost & th = getDpTraceMatrix(dpContext);
            std::cout << "GFILL score=" << maxScore(dpScout) << " S=" << _scoreOfCell(dpScout._maxScore) << " H="
                      << _horizontalScoreOfCell(dpScout._maxScore) << " V=" << _verticalScoreOfCell(dpScout._maxScore)
                      << " pos=" << dpScout._maxHostPosition << " tracelen=" << length(th) << " hash="
                      << fnv1a((unsigned char const *)&th[0], length(th) * sizeof(TTraceValue));
            // and the fragments the library call returns
            String<Fragment<> > matches;
            StringSet<Dna5String, Dependent<> > pairSet; appendValue(pairSet, seqH); appendValue(pairSet, seqV);
            int sc2 = globalAlignment(matches, pairSet, sc, AlignConfig<>(), Gotoh());
            std::cout << " libscore=" << sc2 << " frags=";
            for (unsigned i = 0; i < length(matches); ++i)
                std::cout << (i ? "," : "") << matches[i].begin1 << ":" << matches[i].begin2 << ":" << matches[i].len;
            std::cout << "\n";
        }
        else if (cmd == "MERGE")
        {
            unsigned n;
            in >> n;
            std::vector<BamAlignmentRecord> recs(n);
            for (unsigned i = 0; i < n; ++i)
            {
                std::string name, cig, sq; int flag, rid, begin, mq;
                in >> name >> flag >> rid >> begin >> mq >> cig >> sq;
                recs[i].qName = name; recs[i].flag = flag; recs[i].rID = rid; recs[i].beginPos = begin; recs[i].mapQ = mq;
                parse_cigar(cig, recs[i].cigar);
                recs[i].seq = sq;
            }
            std::vector<BamAlignmentRecord> out = sviper::merge_alignments(recs);
            std::cout << "MERGE n=" << out.size();
            for (auto const & r : out)
            {
                std::cout << " " << r.qName << ":" << r.flag << ":" << r.beginPos << ":" << (int)r.mapQ << ":";
                for (unsigned i = 0; i < length(r.cigar); ++i) std::cout << r.cigar[i].count << r.cigar[i].operation;
                std::cout << ":" << r.seq;
            }
            std::cout << "\n";
        }
        else { std::cerr << "unknown command " << cmd << "\n"; return 2; }
    }
    return 0;
}
