// sviper_b200/host/shim/sviper/polishing.h -- the drop-in seam, without touching a line of the reference.
//
// The reference's sviper.h does `#include <sviper/polishing.h>` (include/sviper/sviper.h:14). Putting THIS directory in
// front of the reference's include path makes the reference's sources pick up this file instead. It pulls in the
// reference's own polishing.h (compute_baseQ_stats etc. stay the reference's code) and replaces exactly the two calls of
// the hot path (SURVEY 8b seams 1 and 2):
//
//   sviper::polish_to_perfection(reads1, reads2, ref, config)        polishing.h:319-346, called at sviper.h:431
//       -> one variant submitted to the asynchronous device scheduler (Gateway::polish -> svp_submit_polish). With the
//          reference part attached (see the basics.h shim) the final alignment runs in the same device job. The caller is
//          suspended (a fiber of the prepare / device / finish pipeline, pipeline.hpp) or blocks (a plain OpenMP thread
//          of the unmodified src/sviper.cpp) until its result is back (seam 3).
//   seqan::globalAlignment(gapsRef, gapsSeq, Score<double>, AlignConfig<0,0,0,0>, ConvexGaps)   sviper.h:457-460
//       -> the CIGAR the device job already produced (or svp_final_align_batch when called on other sequences) is written
//          back into the two Gaps objects, so the reference's own getCigarString (sviper.h:466) and everything behind it
//          run unchanged
//
// Two programs are built from this (sviper_b200/csrc/Makefile):
//   sviper_b200        host/sviper_b200_main.cpp: the reference's main() sequence with the OpenMP loop replaced by the
//                      fiber pipeline -- prepare threads never wait for the device
//   sviper_b200_cli    the reference's UNMODIFIED src/sviper.cpp: each OpenMP thread blocks on its own variant
// GPUs: SVIPER_B200_DEVICES="0,1,2,3" (default "0"). There is no CPU fallback: without a GPU the first call throws.
#pragma once

#define polish_to_perfection polish_to_perfection_reference_cpu
#include_next <sviper/polishing.h>
#undef polish_to_perfection

#include <cstdlib>
#include <mutex>
#include <sstream>

#include <seqan/align.h>
#include <seqan/seq_io.h>

#include <sviper/basics.h>   // (the shim in this directory: VariantScope, append_ref_flanks hook)

namespace sviper_b200
{
namespace shim
{
template <typename TString>
inline std::string to_std_string(TString const & s)
{
    std::string o(seqan::length(s), 'N');
    for (size_t i = 0; i < o.size(); ++i) o[i] = (char)seqan::Dna5(s[i]);
    return o;
}

inline uint64_t fnv(std::string const & s, uint64_t h = 1469598103934665603ull)
{
    for (unsigned char c : s) { h ^= c; h *= 1099511628211ull; }
    return h;
}

// SVIPER_B200_TRACE=1: one line per variant on stderr with hashes of what went into the device job and what came back
// (tools/gpu_wp_repeat.py compares them between runs)
inline bool trace_on()
{
    static bool const on = [] { char const * e = std::getenv("SVIPER_B200_TRACE"); return e && std::atoi(e) == 1; }();
    return on;
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
    // the reference part of the final alignment: the readRegion call of sviper.h:447-450, same expressions and types
    // (int - unsigned is unsigned arithmetic there), made here so that the final alignment rides in the same device job
    b200::shim::VariantScope & vs = b200::shim::scope();
    std::string ref_part;
    bool with_final = false;
    if (vs.have_region && vs.flank == config.ref_flank_length)
    {
        seqan::Dna5String rp;
        seqan::readRegion(rp, *vs.fai, vs.fai_idx, std::max(1u, vs.start - config.ref_flank_length),
                          std::min(vs.end + config.ref_flank_length, static_cast<unsigned>(vs.ref_length)));
        ref_part = b200::shim::to_std_string(rp);
        with_final = !ref_part.empty();
    }
    vs.have_region = false; vs.have_final = false;
    b200::PackedVariant pv(b200::shim::to_qreads(reads1), b200::shim::to_qreads(reads2), b200::shim::to_std_string(ref), cfg,
                           with_final ? &ref_part : nullptr);
    b200::ResultBuffers rb(seqan::length(ref), (int)(seqan::length(ref) + ref_part.size() + 4096));
    b200::Gateway::instance().polish(pv, rb);   // suspends this fiber / blocks this thread until the device batch is done
    if (rb.res.status == SVP_ERR_UNSUPPORTED || rb.res.status == SVP_ERR_ARG || rb.res.status == SVP_ERR_CAPACITY)
    {
        // outside the device's per-variant limits (include/sviper_b200.h; e.g. a read longer than SVP_MAX_READ_LEN): the
        // other variants are not affected, this one stays unpolished -- said loudly, never silently replaced by CPU code
        std::cerr << "[sviper_b200] variant not polished: input outside the device limits (status " << rb.res.status
                  << "); its consensus is left as built from the long reads" << std::endl;
        config.fix_indels = true;
        return ref;
    }
    if (rb.res.status != SVP_OK) throw std::runtime_error("sviper_b200: polish_to_perfection failed on the device");
    // the side effects of polishing.h:319-346 on config (config.h:72-90)
    config.rounds += rb.res.rounds;
    config.substituted_bases += rb.res.substituted_bases;
    config.inserted_bases += rb.res.inserted_bases;
    config.deleted_bases += rb.res.deleted_bases;
    config.mappQ_mean = rb.res.mappQ_mean;
    config.alpha = config.mappQ_mean / config.baseQ_mean;
    config.fix_indels = true;
    config.cov_profile.assign(rb.cov.begin(), rb.cov.begin() + rb.res.cov_len);
    std::string polished = rb.sequence();
    if (b200::shim::trace_on())
    {
        uint64_t hr = 1469598103934665603ull;
        for (auto const * set : {&reads1, &reads2})
            for (auto const & q : b200::shim::to_qreads(*set)) { hr = b200::shim::fnv(q.seq, hr); hr = b200::shim::fnv(q.qual, hr); }
        std::ostringstream os;
        os << "[trace] refpart=" << b200::shim::fnv(ref_part) << " cons=" << b200::shim::fnv(b200::shim::to_std_string(ref))
           << " cons_len=" << seqan::length(ref) << " reads=" << seqan::length(reads1) << " reads_hash=" << hr
           << " baseQ=" << config.baseQ_mean << " ins=" << config.mean_insert_size_of_short_reads << "/" << config.stdev_insert_size_of_short_reads
           << " polished=" << b200::shim::fnv(polished) << " len=" << polished.size() << " rounds=" << rb.res.rounds
           << " edits=" << rb.res.substituted_bases << "/" << rb.res.inserted_bases << "/" << rb.res.deleted_bases
           << " mappQ=" << rb.res.mappQ_mean << " cigar=";
        for (auto const & e : rb.cigar()) os << e.count << e.operation;
        os << "\n";
        std::cerr << os.str();
    }
    if (with_final)
    {
        vs.have_final = true; vs.ref_part.swap(ref_part); vs.polished = polished; vs.cigar = rb.cigar();
        vs.score = rb.res.final_score;
    }
    seqan::Dna5String out = polished;
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
    std::vector<sviper_b200::CigarElement> cigar;
    sviper_b200::shim::VariantScope & vs = sviper_b200::shim::scope();
    std::string const h = sviper_b200::shim::to_std_string(source(gapsRef)), v = sviper_b200::shim::to_std_string(source(gapsSeq));
    if (vs.have_final && h == vs.ref_part && v == vs.polished)
    {
        cigar.swap(vs.cigar); score = vs.score;      // computed in the variant's polish job
        vs.have_final = false;
    }
    else
        cigar = sviper_b200::Gateway::instance().final_alignment(h, v, &score);
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
