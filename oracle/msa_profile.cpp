// TEST / STUDY INFRASTRUCTURE ONLY -- never linked into the product.
//
// Where does sviper::build_consensus (include/sviper/basics.h:484-520) spend its time? The function is
// seqan::globalMsaAlignment(align, SimpleScore(5,-3,-1,-3)) (graph_msa/graph_align_tcoffee_msa.h:564-578 ->
// :323-548) plus a weighted column vote. This driver replays the SAME sequence of SeqAn calls as
// globalMsaAlignment's body on synthetic long-read crops, with a timer around each stage, and checks that
// the result equals the library call. Output: one JSON line per (length, repeat).
//
//   msa_profile <crop_len> <n_reads> <error_rate> <seed> [repeats]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <random>
#include <string>
#include <vector>

#include <seqan/align.h>
#include <seqan/graph_msa.h>

using namespace seqan;
typedef std::chrono::steady_clock Clock;
static double ms(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); }

static Dna5String noisy(Dna5String const & truth, double err, std::mt19937 & rng)
{
    // long-read error model of SURVEY 8d: 30% substitutions / 35% insertions / 35% deletions of the error budget
    std::uniform_real_distribution<double> U(0, 1);
    std::uniform_int_distribution<int> B(0, 3);
    Dna5String out;
    for (unsigned i = 0; i < length(truth); ++i)
    {
        double u = U(rng);
        if (u < err * 0.30) appendValue(out, Dna5((ordValue(truth[i]) + 1 + B(rng) % 3) % 4));
        else if (u < err * 0.65) { appendValue(out, Dna5(B(rng))); appendValue(out, truth[i]); }
        else if (u < err) continue;
        else appendValue(out, truth[i]);
    }
    return out;
}

int main(int argc, char ** argv)
{
    if (argc < 5) { fprintf(stderr, "usage: msa_profile len n_reads err seed [repeats]\n"); return 2; }
    unsigned len = atoi(argv[1]), nreads = atoi(argv[2]);
    double err = atof(argv[3]);
    unsigned seed = atoi(argv[4]);
    int reps = argc > 5 ? atoi(argv[5]) : 1;
    std::mt19937 rng(seed);
    for (int rep = 0; rep < reps; ++rep)
    {
        Dna5String truth;
        std::uniform_int_distribution<int> B(0, 3);
        for (unsigned i = 0; i < len; ++i) appendValue(truth, Dna5(B(rng)));
        StringSet<Dna5String> seqs;
        for (unsigned r = 0; r < nreads; ++r) appendValue(seqs, noisy(truth, err, rng));

        // ---- the library call, as build_consensus makes it (basics.h:489-494)
        Align<Dna5String> align;
        resize(rows(align), length(seqs));
        for (unsigned i = 0; i < length(seqs); ++i) assignSource(row(align, i), seqs[i]);
        auto t0 = Clock::now();
        globalMsaAlignment(align, SimpleScore(5, -3, -1, -3));
        auto t1 = Clock::now();

        // ---- the same stages, one by one (graph_align_tcoffee_msa.h:323-548 with the options of :553-562)
        typedef StringSet<Dna5String, Dependent<> > TDepSet;
        TDepSet seqSet = stringSet(align);
        typedef Graph<Alignment<TDepSet, void, WithoutEdgeId> > TOutGraph;
        TOutGraph gAlign(seqSet);
        typedef Size<TDepSet>::Type TSize;
        typedef Graph<Alignment<TDepSet, TSize> > TGraph;
        SimpleScore sc(5, -3, -1, -3);
        auto s0 = Clock::now();
        String<TSize> pList;
        selectPairs(seqSet, pList);
        String<double> distanceMatrix;
        String<Fragment<> > matches;
        String<int> scores;
        appendSegmentMatches(seqSet, pList, sc, matches, scores, distanceMatrix, GlobalPairwiseLibrary());
        auto s1 = Clock::now();
        size_t n_global = length(matches);
        _appendSegmentMatches(seqSet, pList, sc, matches, scores, DefaultAlignment());
        auto s2 = Clock::now();
        size_t n_all = length(matches);
        TGraph g(seqSet);
        buildAlignmentGraph(matches, scores, g, sc, ReScore());   // MsaOptions::rescore defaults to true
        auto s3 = Clock::now();
        Graph<Tree<double> > guideTree;
        for (unsigned i = 0; i < length(distanceMatrix); ++i)
            distanceMatrix[i] = static_cast<int64_t>(distanceMatrix[i] * 1e10) / 1e10;
        njTree(distanceMatrix, guideTree);   // MsaOptions::build defaults to 0 (graph_align_tcoffee_msa.h:137)
        auto s4 = Clock::now();
        tripletLibraryExtension(g);
        auto s5 = Clock::now();
        progressiveAlignment(g, guideTree, gAlign);
        auto s6 = Clock::now();
        Align<Dna5String> align2;
        convertAlignment(gAlign, align2);
        auto s7 = Clock::now();
        bool same = length(row(align, 0)) == length(row(align2, 0));
        for (unsigned r = 0; same && r < length(seqs); ++r)
            for (unsigned i = 0; same && i < length(row(align, r)); ++i)
                same = isGap(row(align, r), i) == isGap(row(align2, r), i);
        size_t cells = 0;
        for (unsigned a = 0; a < length(seqs); ++a)
            for (unsigned b = a + 1; b < length(seqs); ++b) cells += (size_t)length(seqs[a]) * length(seqs[b]);
        printf("{\"len\": %u, \"reads\": %u, \"err\": %.3f, \"library_call_ms\": %.1f, \"pairs\": %u, "
               "\"pair_cells\": %zu, \"global_pairwise_ms\": %.1f, \"local_pairwise_ms\": %.1f, \"matches_global\": %zu, "
               "\"matches_all\": %zu, \"build_graph_ms\": %.1f, \"guide_tree_ms\": %.2f, \"triplet_ms\": %.1f, "
               "\"progressive_ms\": %.1f, \"convert_ms\": %.2f, \"staged_equals_library\": %s}\n",
               len, nreads, err, ms(t0, t1), (unsigned)length(pList) / 2, cells, ms(s0, s1), ms(s1, s2), n_global, n_all,
               ms(s2, s3), ms(s3, s4), ms(s4, s5), ms(s5, s6), ms(s6, s7), same ? "true" : "false");
        fflush(stdout);
    }
    return 0;
}
