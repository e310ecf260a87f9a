// sviper_b200/host/shim/sviper/basics.h -- shadows the reference's header of the same name (see polishing.h in this
// directory for the mechanism). It includes the reference's own basics.h unchanged and wraps ONE function:
//
//   sviper::append_ref_flanks(cns, faiIndex, fai_idx, ref_length, start, end, flank)     basics.h:269-290, called at
//   sviper.h:419-422 right before polish_to_perfection
//
// The wrapper calls the reference's function and remembers WHERE the flanks came from. With that, the shim of
// polish_to_perfection can read the reference part of the final alignment itself -- the same readRegion call the
// reference makes at sviper.h:447-450 -- and send it along, so that polishing and final alignment of a variant run in
// ONE device job instead of two round trips.
#pragma once

// build_consensus (basics.h:484-520) stays the reference's code; the library call in it,
// seqan::globalMsaAlignment(align, seqan::SimpleScore(5, -3, -1, -3)) (basics.h:494), resolves to the overload below
// (declared before the reference's header is read, preferred over SeqAn's template by overload resolution): the same stages
// with SeqAn's own functions, the pairwise DP fills on the GPU (sviper_b200_msa.hpp). SVIPER_B200_MSA=host keeps SeqAn's
// own globalMsaAlignment (A/B measurements of the whole program; the polishing hot path is on the GPU either way).
#include "../sviper_b200_msa.hpp"

namespace seqan
{
inline void globalMsaAlignment(Align<Dna5String, ArrayGaps> & align, Score<int, Simple> const & scoreObject)
{
    static bool const on_host = [] { char const * e = std::getenv("SVIPER_B200_MSA"); return e && std::string(e) == "host"; }();
    if (on_host)
    {
        globalMsaAlignment<Dna5String, ArrayGaps, Score<int, Simple> >(align, scoreObject);   // SeqAn's template, unchanged
        return;
    }
    sviper_b200::shim::DeviceFills fills;
    sviper_b200::shim::global_msa_alignment(fills, align, scoreObject);
}
}  // namespace seqan

#define append_ref_flanks append_ref_flanks_reference
#include_next <sviper/basics.h>
#undef append_ref_flanks

#include <string>
#include <vector>

#include "../../pipeline.hpp"

namespace sviper_b200
{
namespace shim
{
// per-variant notes of the shims: lives with the fiber that runs polish_variant (Pipeline), else with the thread
struct VariantScope
{
    // set by append_ref_flanks, consumed by polish_to_perfection
    bool have_region = false;
    seqan::FaiIndex const * fai = nullptr;
    unsigned fai_idx = 0;
    int ref_length = 0, start = 0, end = 0;
    unsigned flank = 0;
    // set by polish_to_perfection, consumed by the final globalAlignment
    bool have_final = false;
    std::string ref_part, polished;
    std::vector<CigarElement> cigar;
    double score = 0;
};

inline VariantScope & scope()
{
    Fiber * f = Fiber::current();
    if (f && f->user)
    {
        FiberHome * h = static_cast<FiberHome *>(f->user);
        if (!h->scope) h->scope = new VariantScope();     // freed with the process: one small object per variant
        return *static_cast<VariantScope *>(h->scope);
    }
    static thread_local VariantScope tls;
    return tls;
}
}  // namespace shim
}  // namespace sviper_b200

namespace sviper
{
inline seqan::Dna5String append_ref_flanks(seqan::Dna5String const & seq, seqan::FaiIndex const & fai_index, unsigned fai_idx,
                                           int ref_length, int start, int end, int length)
{
    sviper_b200::shim::VariantScope & vs = sviper_b200::shim::scope();
    vs.have_region = true; vs.have_final = false;
    vs.fai = &fai_index; vs.fai_idx = fai_idx; vs.ref_length = ref_length; vs.start = start; vs.end = end;
    vs.flank = (unsigned)length;
    return append_ref_flanks_reference(seq, fai_index, fai_idx, ref_length, start, end, length);
}
}  // namespace sviper
