// sviper_b200/host/prepare.hpp -- the CIGAR walks of polish_variant's prepare side on plain
// structs, above the C ABI (include/sviper_b200.h). Same names, argument meaning and error behaviour as the reference:
//
//   reference (SeqAn types)                                         here (no SeqAn types)
//   record_supports_variant(record, variant)   variant.h:239-275    sviper_b200::record_supports_variant(record, var)
//   get_read_region_boundaries(record, b, e)   basics.h:189-241     sviper_b200::get_read_region_boundaries(record, b, e)
//   cut_down_high_coverage(short_reads, cov)   basics.h:403-442     sviper_b200::cut_down_high_coverage(short_reads, cov)
//   merge_alignments(records)       merge_split_alignments.h:536    sviper_b200::merge_alignments(records)
//
// AlignedRecord carries the seqan::BamAlignmentRecord fields these functions read. Reading BAM / FASTA files, the mate
// hunt and the long-read MSA stay the reference's own SeqAn code.
#pragma once
#include <iostream>
#include <tuple>

#include "polisher.hpp"

namespace sviper_b200
{
struct AlignedRecord
{
    std::string qName;
    int32_t beginPos = 0;
    uint8_t mapQ = 0;
    std::vector<CigarElement> cigar;
    uint16_t flag = 0;      // only merge_alignments reads flag, rID and seq
    int32_t rID = 0;
    std::string seq;
};

namespace detail
{
inline void split_cigar(std::vector<CigarElement> const & cigar, std::string & ops, std::vector<uint32_t> & cnt)
{
    ops.assign(cigar.size(), 'M');
    cnt.resize(cigar.size());
    for (size_t i = 0; i < cigar.size(); ++i) { ops[i] = cigar[i].operation; cnt[i] = cigar[i].count; }
}
}  // namespace detail

inline bool record_supports_variant(AlignedRecord const & record, VariantFields const & variant)
{
    std::string ops; std::vector<uint32_t> cnt;
    detail::split_cigar(record.cigar, ops, cnt);
    svp_variant_pos vp = {variant.sv_type, variant.ref_pos, variant.ref_pos_end, variant.sv_length};
    int rc = svp_record_supports_variant(record.beginPos, ops.data(), cnt.data(), (int32_t)cnt.size(), &vp);
    if (rc < 0) throw std::logic_error("UNKNOWN CIGAR CHARACTER.");   // basics.h:169
    return rc == 1;
}

inline std::tuple<int, int> get_read_region_boundaries(AlignedRecord const & record, int ref_region_begin,
                                                       int ref_region_end)
{
    std::string ops; std::vector<uint32_t> cnt;
    detail::split_cigar(record.cigar, ops, cnt);
    int32_t b = 0, e = 0;
    int rc = svp_read_region_boundaries(record.beginPos, ops.data(), cnt.data(), (int32_t)cnt.size(), ref_region_begin,
                                        ref_region_end, &b, &e);
    if (rc == 1)                                                      // basics.h:196-203
        std::cerr << "[GET READ REGION ERROR] for record " << record.qName << " Start (" << ref_region_begin
                  << ") >= End (" << ref_region_end << ")" << std::endl;
    if (rc < 0) throw std::logic_error("UNKNOWN CIGAR CHARACTER.");
    return std::make_tuple((int)b, (int)e);
}

inline void cut_down_high_coverage(std::vector<AlignedRecord> & short_reads, int16_t mean_coverage_of_short_reads)
{
    if (short_reads.empty()) return;
    const size_t n = short_reads.size();
    std::vector<int32_t> bp(n), len(n), keep(n);
    std::vector<uint8_t> mq(n);
    std::string ops; std::vector<uint32_t> cnt;
    for (size_t i = 0; i < n; ++i)
    {
        bp[i] = short_reads[i].beginPos; mq[i] = short_reads[i].mapQ;
        detail::split_cigar(short_reads[i].cigar, ops, cnt);
        len[i] = svp_alignment_length_in_ref(ops.data(), cnt.data(), (int32_t)cnt.size());
    }
    int k = svp_cut_down_high_coverage((int32_t)n, bp.data(), mq.data(), len.data(), mean_coverage_of_short_reads,
                                       keep.data());
    if (k < 0) throw std::out_of_range("cut_down_high_coverage: a record reaches beyond the region's end buffer");
    std::vector<AlignedRecord> tmp;
    tmp.reserve((size_t)k);
    for (int i = 0; i < k; ++i) tmp.push_back(short_reads[(size_t)keep[(size_t)i]]);
    short_reads.swap(tmp);
}

// merge_split_alignments.h:536-561: records sorted by read name; one merged record per name, in input order
inline std::vector<AlignedRecord> merge_alignments(std::vector<AlignedRecord> const & records)
{
    std::vector<AlignedRecord> merged;
    size_t b = 0;
    while (b < records.size())
    {
        size_t e = b + 1;
        while (e < records.size() && records[e].qName == records[b].qName) ++e;
        const size_t n = e - b;
        std::vector<std::string> ops(n);
        std::vector<std::vector<uint32_t>> cnt(n);
        std::vector<svp_bam_record> in(n);
        size_t cig_cap = 16, seq_cap = 16;
        for (size_t i = 0; i < n; ++i)
        {
            AlignedRecord const & r = records[b + i];
            detail::split_cigar(r.cigar, ops[i], cnt[i]);
            in[i] = svp_bam_record{r.beginPos, r.rID, r.flag, r.mapQ, ops[i].data(), cnt[i].data(), (int32_t)cnt[i].size(),
                                   r.seq.data(), (int32_t)r.seq.size()};
            cig_cap += r.cigar.size() + 2; seq_cap += r.seq.size();
        }
        std::string out_op(cig_cap, 'M'), out_seq(seq_cap, 'N');
        std::vector<uint32_t> out_cnt(cig_cap);
        int32_t chosen = 0, begin = 0, n_cig = 0, seq_len = 0;
        int rc = svp_merge_record_group((int32_t)n, in.data(), &chosen, &begin, &out_op[0], out_cnt.data(), (int32_t)cig_cap,
                                        &n_cig, &out_seq[0], (int32_t)seq_cap, &seq_len);
        if (rc != SVP_OK) throw std::logic_error("merge_alignments: malformed record group");
        AlignedRecord m = records[b + (size_t)chosen];
        m.beginPos = begin;
        m.cigar.resize((size_t)n_cig);
        for (int32_t i = 0; i < n_cig; ++i) m.cigar[(size_t)i] = CigarElement{out_op[(size_t)i], out_cnt[(size_t)i]};
        m.seq = out_seq.substr(0, (size_t)seq_len);
        merged.push_back(m);
        b = e;
    }
    return merged;
}
}  // namespace sviper_b200
