// sviper_b200/host/shim/sviper_b200_msa.hpp -- build_consensus (include/sviper/basics.h:484-520) with its pairwise DP fills
// on the GPU.
//
// The reference's build_consensus stays UNCHANGED (its column vote included); what is replaced is the one library call in
// it, seqan::globalMsaAlignment(align, SimpleScore(5,-3,-1,-3)) (basics.h:494), by an overload for exactly these argument
// types (see the basics.h shim). The overload runs the stages of globalMsaAlignment
// (graph_msa/graph_align_tcoffee_msa.h:323-548, options of :553-562) with SeqAn's own functions in SeqAn's own order --
// selectPairs, buildAlignmentGraph(ReScore), njTree, tripletLibraryExtension, progressiveAlignment, convertAlignment --
// except for the two stages that are ~90 % of its time (profiles/r2_msa_profile.jsonl):
//
//   global pairwise library   graph_align_tcoffee_library.h:979-1008: per pair a Gotoh alignment. The device fills the
//                             trace matrix (svp_msa_global_fill); SeqAn's own _correctTraceValue / _computeTraceback /
//                             _adaptTraceSegmentsTo turn it into fragments; _setDistanceValue as in the reference.
//   local pairwise library    library.h:431-462, :746-769: per pair up to four rounds of the clumping Smith-Waterman. The
//                             device fills each round's trace (svp_msa_local_fill, all pairs of a round in one launch);
//                             SeqAn's own _alignSmithWatermanTrace follows it, appends the fragments and reports the cells
//                             it forbids for the next round.
//
// The fill provider is a template parameter so that the SeqAn plumbing of this file can be exercised without a GPU by the
// test infrastructure (oracle/msa_replay.cpp plugs the oracle's C fills in); the product only ever instantiates
// DeviceFills -- there is no CPU implementation behind it.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <seqan/align.h>
#include <seqan/graph_msa.h>

#include "../../../include/sviper_b200.h"

namespace sviper_b200
{
namespace shim
{
// stand-in for the String<bool> "forbidden" matrix of the local library: SeqAn's traceback only ever WRITES it
// (_setForbiddenCell, align/local_alignment_banded_waterman_eggert_impl.h:177-186), so the cells are recorded and sent to
// the device, which owns the matrix as a bitmap
struct RecordedCells
{
    std::vector<int32_t> cells;   // (col, row) pairs, 1-based
};

template <typename TSize>
inline void _setForbiddenCell(RecordedCells & f, TSize len1, TSize len2, TSize /*numRows*/)
{
    f.cells.push_back((int32_t)len1);
    f.cells.push_back((int32_t)len2);
}

// the fills on the GPU: one context (own stream) per host thread
class DeviceFills
{
public:
    void begin(std::vector<std::vector<uint8_t> > const & h, std::vector<std::vector<uint8_t> > const & v)
    {
        std::vector<svp_msa_pair> p(h.size());
        for (size_t i = 0; i < h.size(); ++i) p[i] = svp_msa_pair{h[i].data(), (int32_t)h[i].size(), v[i].data(), (int32_t)v[i].size()};
        check(svp_msa_begin(ctx(), (int32_t)p.size(), p.data(), 5, -3, -1, -3), "svp_msa_begin");
    }
    void global(std::vector<int32_t> const & idx, std::vector<uint8_t *> const & trace, std::vector<svp_msa_fill_result> & res)
    {
        res.resize(idx.size());
        check(svp_msa_global_fill(ctx(), (int32_t)idx.size(), idx.data(), trace.data(), res.data()), "svp_msa_global_fill");
    }
    void local(std::vector<int32_t> const & idx, std::vector<std::vector<int32_t> > const & new_cells,
               std::vector<uint8_t *> const & trace, std::vector<svp_msa_fill_result> & res)
    {
        std::vector<const int32_t *> cp(idx.size());
        std::vector<int32_t> cn(idx.size());
        for (size_t i = 0; i < idx.size(); ++i) { cp[i] = new_cells[i].data(); cn[i] = (int32_t)(new_cells[i].size() / 2); }
        res.resize(idx.size());
        check(svp_msa_local_fill(ctx(), (int32_t)idx.size(), idx.data(), cp.data(), cn.data(), trace.data(), res.data()),
              "svp_msa_local_fill");
    }

private:
    struct Holder
    {
        svp_ctx * c = nullptr;
        ~Holder() { if (c) svp_destroy(c); }
    };
    static svp_ctx * ctx()
    {
        static thread_local Holder h;
        if (!h.c)
        {
            // GPUs as for polishing (SVIPER_B200_DEVICES), host threads dealt round-robin
            static std::mutex mu;
            static int next = 0;
            std::vector<int> ids;
            char const * e = std::getenv("SVIPER_B200_DEVICES");
            std::string s = e ? e : "0";
            for (size_t p = 0; p < s.size();)
            {
                size_t q = s.find(',', p);
                if (q == std::string::npos) q = s.size();
                if (q > p) ids.push_back(std::atoi(s.substr(p, q - p).c_str()));
                p = q + 1;
            }
            if (ids.empty()) ids.push_back(0);
            int mine;
            { std::lock_guard<std::mutex> lk(mu); mine = ids[(size_t)(next++) % ids.size()]; }
            h.c = svp_create(mine, nullptr);
            if (!h.c) throw std::runtime_error("sviper_b200: no usable CUDA device for build_consensus (there is no CPU fallback)");
        }
        return h.c;
    }
    static void check(int rc, char const * what)
    {
        if (rc != SVP_OK) throw std::runtime_error(std::string(what) + " failed: " + svp_last_error(ctx()));
    }
};

template <typename TString>
inline std::vector<uint8_t> ord_values(TString const & str)
{
    std::vector<uint8_t> out(seqan::length(str));
    for (size_t i = 0; i < out.size(); ++i) out[i] = (uint8_t)seqan::ordValue(seqan::Dna5(str[i]));
    return out;
}

// Global pairwise library for every pair of pList (library.h:979-1008), fills by TFills. Appends to matches / scores and
// sets the distance matrix exactly as the reference does, pair after pair.
template <typename TFills, typename TStringSet, typename TPairList, typename TMatches, typename TScores, typename TDist>
inline void global_pairwise_library(TFills & fills, TStringSet const & str, TPairList const & pList, TMatches & matches,
                                    TScores & scores, TDist & dist, size_t n_pairs)
{
    using namespace seqan;
    typedef DPCell_<int, AffineGaps> TCell;
    typedef TraceBitMap_<>::Type TTraceValue;
    static_assert(sizeof(TTraceValue) == 1, "the device writes one byte per trace cell");
    typedef TracebackOn<TracebackConfig_<SingleTrace, GapsLeft> > TTraceFlag;
    typedef typename SetupAlignmentProfile_<DPGlobal, FreeEndGaps_<False, False, False, False>, AffineGaps, TTraceFlag>::Type TDPProfile;
    typedef DPContext<TCell, TTraceValue> TContext;
    typedef typename std::remove_reference<decltype(getDpTraceMatrix(std::declval<TContext &>()))>::type TTraceHost;
    typedef DPMatrix_<TTraceValue, FullDPMatrix, TTraceHost> TDPTraceMatrix;
    typedef DPMatrixNavigator_<TDPTraceMatrix, DPTraceMatrix<TTraceFlag>, NavigateColumnWise> TTraceNav;
    typedef typename Value<TMatches>::Type TFragment;
    typedef typename Size<TFragment>::Type TFragSize;
    typedef typename Position<TMatches>::Type TPosition;
    typedef TraceSegment_<TPosition, TFragSize> TTraceSegment;
    typedef typename Size<TStringSet>::Type TSize;

    TSize nseq = length(str);
    _resizeWithRespectToDistance(dist, nseq);
    // the trace matrices: set up as _computeAlignment does (dp_algorithm_impl.h:1424-1447), filled by the device
    std::vector<TContext> contexts(n_pairs);
    std::vector<TDPTraceMatrix> tms(n_pairs);
    std::vector<TStringSet> pairSets;
    std::vector<uint8_t *> bufs(n_pairs);
    std::vector<int32_t> idx(n_pairs);
    DPBandConfig<BandOff> band;
    for (size_t p = 0; p < n_pairs; ++p)
    {
        pairSets.push_back(_makePairSet(str, begin(pList, Standard()) + 2 * p, begin(pList, Standard()) + 2 * p + 1));
        setLength(tms[p], +DPMatrixDimension_::HORIZONTAL, length(value(pairSets[p], 0)) + 1);
        setLength(tms[p], +DPMatrixDimension_::VERTICAL, length(value(pairSets[p], 1)) + 1);
        setHost(tms[p], getDpTraceMatrix(contexts[p]));
        resize(tms[p]);
        bufs[p] = reinterpret_cast<uint8_t *>(&getDpTraceMatrix(contexts[p])[0]);
        idx[p] = (int32_t)p;
    }
    std::vector<svp_msa_fill_result> res;
    fills.global(idx, bufs, res);
    for (size_t p = 0; p < n_pairs; ++p)
    {
        TStringSet & pairSet = pairSets[p];
        TSize from = length(matches);
        // what _computeAlignmentImpl leaves in the scout for plain global alignment: the bottom-right cell
        DPScoutState_<Default> scoutState;
        DPScout_<TCell, Default> dpScout(scoutState);
        _scoreOfCell(dpScout._maxScore) = res[p].S;
        _horizontalScoreOfCell(dpScout._maxScore) = res[p].H;
        _verticalScoreOfCell(dpScout._maxScore) = res[p].V;
        dpScout._maxHostPosition = length(getDpTraceMatrix(contexts[p])) - 1;
        TTraceNav traceNav{tms[p], band};
        String<TTraceSegment> traceSegments;
        int myScore = _finishAlignment(traceSegments, traceNav, dpScout, value(pairSet, 0), value(pairSet, 1), band, TDPProfile());
        _adaptTraceSegmentsTo(matches, positionToId(pairSet, 0), positionToId(pairSet, 1), traceSegments);
        TSize to = length(matches);
        _recordScores(scores, myScore, from, to);
        _setDistanceValue(matches, pairSet, dist, (TSize)pList[2 * p], (TSize)pList[2 * p + 1], (TSize)nseq, (TSize)from);
        clear(getDpTraceMatrix(contexts[p]));
    }
}

// Local pairwise library for every pair (library.h:746-769 -> _multiLocalAlignment(.., 4, SmithWatermanClump) -> :431-462).
// Round r of all pairs that are still searching is one device call; the fragments of a pair are appended in the
// reference's order: pair after pair, round after round.
template <typename TFills, typename TStringSet, typename TPairList, typename TMatches, typename TScores>
inline void local_pairwise_library(TFills & fills, TStringSet const & str, TPairList const & pList, TMatches & matches,
                                   TScores & scores, size_t n_pairs)
{
    using namespace seqan;
    typedef typename Size<TMatches>::Type TSize;
    struct PairState
    {
        TStringSet pairSet;
        TMatches found;
        String<int> found_scores;
        String<unsigned char> trace;
        RecordedCells fresh;          // cells forbidden by the last traceback, not yet sent to the device
        int last_score = 0;
        bool active = true;
    };
    std::vector<PairState> st(n_pairs);
    for (size_t p = 0; p < n_pairs; ++p)
    {
        st[p].pairSet = _makePairSet(str, begin(pList, Standard()) + 2 * p, begin(pList, Standard()) + 2 * p + 1);
        size_t len1 = length(value(st[p].pairSet, 0)), len2 = length(value(st[p].pairSet, 1));
        resize(st[p].trace, len1 * ((len2 >> 1) + (len2 & 1)), 0);       // library.h:299
    }
    for (int count = 0; count < 4; ++count)
    {
        std::vector<int32_t> idx;
        std::vector<uint8_t *> bufs;
        std::vector<std::vector<int32_t> > cells;
        for (size_t p = 0; p < n_pairs; ++p)
            if (st[p].active)
            {
                idx.push_back((int32_t)p);
                bufs.push_back(reinterpret_cast<uint8_t *>(&st[p].trace[0]));
                cells.push_back(st[p].fresh.cells);
                st[p].fresh.cells.clear();
            }
        if (idx.empty()) break;
        std::vector<svp_msa_fill_result> res;
        fills.local(idx, cells, bufs, res);
        for (size_t k = 0; k < idx.size(); ++k)
        {
            PairState & ps = st[(size_t)idx[k]];
            TSize from = length(ps.found);
            size_t indexPair[2] = {(size_t)res[k].row, (size_t)res[k].col};
            unsigned char initialDir = (unsigned char)res[k].dir;
            _alignSmithWatermanTrace(ps.found, ps.pairSet, ps.trace, initialDir, indexPair, ps.fresh);
            int local_score = res[k].score;
            TSize to = length(ps.found);
            if (2 * local_score < ps.last_score)                          // library.h:455-458
            {
                resize(ps.found, from, Generous());
                ps.active = false;
                continue;
            }
            ps.last_score = local_score;
            resize(ps.found_scores, to);
            for (TSize x = from; x < to; ++x) ps.found_scores[x] = local_score;
        }
    }
    for (size_t p = 0; p < n_pairs; ++p)
    {
        append(matches, st[p].found);
        resize(scores, length(matches));
        TSize base = length(matches) - length(st[p].found);
        for (TSize x = 0; x < length(st[p].found); ++x) scores[base + x] = st[p].found_scores[x];
    }
}

// seqan::globalMsaAlignment(align, score) as build_consensus calls it (basics.h:494): graph_align_tcoffee_msa.h:564-578 ->
// :548-562 -> :323-548 with MsaOptions' defaults (rescore = true, build = 0: neighbour joining, methods 0 and 1).
template <typename TFills>
inline void global_msa_alignment(TFills & fills, seqan::Align<seqan::Dna5String, seqan::ArrayGaps> & align,
                                 seqan::Score<int, seqan::Simple> const & sc)
{
    using namespace seqan;
    if (scoreMatch(sc) != 5 || scoreMismatch(sc) != -3 || scoreGapExtend(sc) != -1 || scoreGapOpen(sc) != -3)
        throw std::logic_error("sviper_b200: globalMsaAlignment called with other scores than build_consensus uses");
    typedef StringSet<Dna5String, Dependent<> > TStringSet;
    TStringSet sequenceSet = stringSet(align);
    Graph<Alignment<TStringSet, void, WithoutEdgeId> > gAlign(sequenceSet);
    typedef Size<TStringSet>::Type TSize;
    typedef Graph<Alignment<TStringSet, TSize> > TGraph;
    clear(gAlign);
    assignStringSet(gAlign, sequenceSet);
    TStringSet & seqSet = stringSet(gAlign);
    TSize nSeq = length(seqSet);
    String<TSize> pList;
    selectPairs(seqSet, pList);
    size_t n_pairs = length(pList) / 2;
    String<double> distanceMatrix;
    String<Fragment<> > matches;
    String<int> scores;
    if (n_pairs)
    {
        std::vector<std::vector<uint8_t> > h(n_pairs), v(n_pairs);
        std::vector<std::vector<uint8_t> > ord(nSeq);
        for (TSize i = 0; i < nSeq; ++i) ord[i] = ord_values(seqSet[i]);
        for (size_t p = 0; p < n_pairs; ++p) { h[p] = ord[pList[2 * p]]; v[p] = ord[pList[2 * p + 1]]; }
        fills.begin(h, v);
        global_pairwise_library(fills, seqSet, pList, matches, scores, distanceMatrix, n_pairs);
        local_pairwise_library(fills, seqSet, pList, matches, scores, n_pairs);
    }
    else
        _resizeWithRespectToDistance(distanceMatrix, nSeq);
    TGraph g(seqSet);
    buildAlignmentGraph(matches, scores, g, sc, ReScore());
    clear(matches);
    clear(scores);
    Graph<Tree<double> > guideTree;
    if (empty(distanceMatrix)) getDistanceMatrix(g, distanceMatrix, KmerDistance());
    for (unsigned i = 0; i < length(distanceMatrix); ++i)
        distanceMatrix[i] = static_cast<int64_t>(distanceMatrix[i] * 1e10) / 1e10;
    njTree(distanceMatrix, guideTree);
    clear(distanceMatrix);
    if (nSeq < 10) tripletLibraryExtension(g);
    else tripletLibraryExtension(g, guideTree, 5);
    progressiveAlignment(g, guideTree, gAlign);
    clear(guideTree);
    clear(g);
    convertAlignment(gAlign, align);
}
}  // namespace shim
}  // namespace sviper_b200
