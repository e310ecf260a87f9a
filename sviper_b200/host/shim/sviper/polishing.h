// sviper_b200/host/shim/sviper/polishing.h -- the drop-in seam, without touching a line of the reference.
//
// The reference's sviper.h does `#include <sviper/polishing.h>` (include/sviper/sviper.h:14). Putting THIS directory in
// front of the reference's include path makes the unmodified src/sviper.cpp pick up this file instead. It pulls in the
// reference's own polishing.h (compute_baseQ_stats etc. stay the reference's code) and replaces exactly the two calls of
// the hot path (SURVEY 8b seams 1 and 2):
//
//   sviper::polish_to_perfection(reads1, reads2, ref, config)        polishing.h:319-346, called at sviper.h:431
//       -> one variant submitted to the device batch scheduler (sviper_b200::BatchScheduler): the OpenMP threads of
//          src/sviper.cpp:43-47 keep running polish_variant's I/O half, whatever they have submitted at the same time
//          is polished as one device batch (svp_polish_batch), the caller blocks on its own future (seam 3)
//   seqan::globalAlignment(gapsRef, gapsSeq, Score<double>, AlignConfig<0,0,0,0>, ConvexGaps)   sviper.h:457-460
//       -> svp_final_align_batch; the device CIGAR is written back into the two Gaps objects, so the reference's own
//          getCigarString (sviper.h:466) and everything behind it run unchanged
//
// Build (sviper_b200/csrc/Makefile, target ../host/sviper_b200_cli):
//   g++ -std=c++14 -O3 -DNDEBUG -fopenmp -I sviper_b200/host/shim -I $REF/extern/seqan/include -I $REF/include \
//       $REF/src/sviper.cpp -lsviper_b200 -lz
// GPUs: SVIPER_B200_DEVICES="0,1,2,3" (default "0"). There is no CPU fallback: without a GPU the first call throws.
#pragma once

#define polish_to_perfection polish_to_perfection_reference_cpu
#include_next <sviper/polishing.h>
#undef polish_to_perfection

#include <cstdlib>
#include <mutex>

#include <seqan/align.h>

#include "../../polisher.hpp"

namespace sviper_b200
{
namespace shim
{
inline std::vector<int> devices()
{
    std::vector<int> d;
    char const * e = std::getenv("SVIPER_B200_DEVICES");
    std::string s = e ? e : "0";
    size_t p = 0;
    while (p < s.size())
    {
        size_t q = s.find(',', p);
        if (q == std::string::npos) q = s.size();
        if (q > p) d.push_back(std::atoi(s.substr(p, q - p).c_str()));
        p = q + 1;
    }
    if (d.empty()) d.push_back(0);
    return d;
}

inline BatchScheduler & scheduler()
{
    // SVIPER_B200_LINGER_US (default 500): how long a device worker waits for further OpenMP threads to arrive
    char const * e = std::getenv("SVIPER_B200_LINGER_US");
    static BatchScheduler s(devices(), 256, e ? (unsigned)std::atoi(e) : 500u);
    return s;
}

// final alignments are single jobs of a millisecond: one context behind a mutex
inline std::vector<CigarElement> final_alignment(std::string const & ref_part, std::string const & polished, double * score)
{
    static Polisher p(devices()[0]);
    static std::mutex m;
    std::lock_guard<std::mutex> lk(m);
    return p.final_alignment(ref_part, polished, score);
}

template <typename TString>
inline std::string to_std_string(TString const & s)
{
    std::string o(seqan::length(s), 'N');
    for (size_t i = 0; i < o.size(); ++i) o[i] = (char)seqan::Dna5(s[i]);
    return o;
}

inline std::vector<QRead> to_qreads(seqan::StringSet<seqan::Dna5QString> const & reads)
{
    std::vector<QRead> out(seqan::length(reads));
    for (size_t r = 0; r < out.size(); ++r)
    {
        auto const & rd = reads[r];
        out[r].seq = to_std_string(rd);
        out[r].qual.resize(out[r].seq.size());
        for (size_t i = 0; i < out[r].seq.size(); ++i) out[r].qual[i] = (char)(33 + seqan::getQualityValue(rd[i]));
    }
    return out;
}
}  // namespace shim
}  // namespace sviper_b200

namespace sviper
{
inline seqan::Dna5String polish_to_perfection(seqan::StringSet<seqan::Dna5QString> const & reads1,
                                              seqan::StringSet<seqan::Dna5QString> const & reads2,
                                              seqan::Dna5String ref,
                                              SViperConfig & config)
{
    namespace b200 = sviper_b200;
    b200::Config cfg;
    cfg.baseQ_mean = config.baseQ_mean;
    cfg.mean_insert_size = config.mean_insert_size_of_short_reads;
    cfg.stdev_insert_size = config.stdev_insert_size_of_short_reads;
    cfg.ref_flank_length = (int)config.ref_flank_length;
    cfg.min_coverage = config.min_coverage;
    std::unique_ptr<b200::PackedVariant> pv(new b200::PackedVariant(b200::shim::to_qreads(reads1), b200::shim::to_qreads(reads2),
                                                                   b200::shim::to_std_string(ref), cfg));
    std::unique_ptr<b200::ResultBuffers> rb = b200::shim::scheduler().submit(std::move(pv)).get();
    if (rb->res.status != SVP_OK) throw std::runtime_error("sviper_b200: polish_to_perfection failed on the device");
    // the side effects of polishing.h:319-346 on config (config.h:72-90)
    config.rounds += rb->res.rounds;
    config.substituted_bases += rb->res.substituted_bases;
    config.inserted_bases += rb->res.inserted_bases;
    config.deleted_bases += rb->res.deleted_bases;
    config.mappQ_mean = rb->res.mappQ_mean;
    config.alpha = config.mappQ_mean / config.baseQ_mean;
    config.fix_indels = true;
    config.cov_profile.assign(rb->cov.begin(), rb->cov.begin() + rb->res.cov_len);
    seqan::Dna5String out = rb->sequence();
    return out;
}
}  // namespace sviper

namespace seqan
{
// Non-template overload for exactly the call at sviper.h:457-460; preferred over SeqAn's template by overload resolution.
inline double globalAlignment(Gaps<Dna5String, ArrayGaps> & gapsRef, Gaps<Dna5String, ArrayGaps> & gapsSeq,
                              Score<double, Simple> const & scoring, AlignConfig<false, false, false, false> const &,
                              ConvexGaps const &)
{
    if (scoreMatch(scoring) != 90000 || scoreMismatch(scoring) != -80000 || scoreGapExtend(scoring) != -1 ||
        scoreGapOpen(scoring) != -100000)                                     // config.h:66-69: what the device kernel implements
        throw std::logic_error("sviper_b200: final alignment called with other scores than SViperConfig's");
    double score = 0;
    std::vector<sviper_b200::CigarElement> cigar = sviper_b200::shim::final_alignment(
        sviper_b200::shim::to_std_string(source(gapsRef)), sviper_b200::shim::to_std_string(source(gapsSeq)), &score);
    clearGaps(gapsRef); clearGaps(gapsSeq);
    clearClipping(gapsRef); clearClipping(gapsSeq);
    size_t view = 0;                                                          // position in the gapped rows
    for (sviper_b200::CigarElement const & e : cigar)
    {
        if (e.operation == 'I') insertGaps(gapsRef, view, e.count);           // bases of the polished sequence only
        else if (e.operation == 'D') insertGaps(gapsSeq, view, e.count);      // bases of the reference only
        view += e.count;
    }
    return score;
}
}  // namespace seqan
