"""Synthetic end-to-end inputs of SURVEY.md section 8(d) config C1 (the reference ships no test data,
tests/sviper_test.sh only prints the help page): one contig, homozygous DELs and INSs, 30x 2x150 short reads,
20x noisy long reads whose CIGARs hold the SV as one D / I element, and a candidate VCF with jittered POS / SVLEN
plus records that take the FILTER paths (SKIP, FAIL0, FAIL1, FAIL2, FAIL4). Written as FASTA + .fai, two coordinate-sorted
BAM + BAI files (tests/bamwriter.py) and a VCF.

Used by tests/test_e2e_vcf.py to run the UNMODIFIED reference binary and the GPU build on the same files.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import bamwriter  # noqa: E402
ALPHA = np.frombuffer(b"ACGT", np.uint8)
COMP = np.array([3, 2, 1, 0], np.uint8)


def s(a):
    return ALPHA[a].tobytes().decode()


def cigar_str(ops):
    out, last, n = [], None, 0
    for o in ops:
        if o == last:
            n += 1
        else:
            if last is not None:
                out.append("%d%s" % (n, last))
            last, n = o, 1
    if last is not None:
        out.append("%d%s" % (n, last))
    return "".join(out)


def haplotype(ref, svs, ins_seqs):
    """the sample's sequence with, for every base, the reference coordinate it maps to and whether it is an inserted base"""
    alt, amap, akind = [], [], []          # akind: 0 = reference base, 1 = inserted base
    p = 0
    for typ, pos, ln in sorted(svs, key=lambda x: x[1]):
        alt.append(ref[p:pos]); amap.append(np.arange(p, pos)); akind.append(np.zeros(pos - p, np.uint8))
        if typ == "DEL":
            p = pos + ln
        else:
            alt.append(ins_seqs[pos]); amap.append(np.full(ln, pos)); akind.append(np.ones(ln, np.uint8))
            p = pos
    alt.append(ref[p:]); amap.append(np.arange(p, len(ref))); akind.append(np.zeros(len(ref) - p, np.uint8))
    return np.concatenate(alt), np.concatenate(amap), np.concatenate(akind)


def make(outdir, seed=1, contig_len=60000, cov_short=30, cov_long=20, k=400, only_filter_paths=False, extras=True, calls=None):
    """extras: records beyond the plain DEL / INS calls: an insertion of 1.2 kb; a deletion only the long reads
    carry (the short reads polish it away: FAIL5 in refine_variant, sviper.h:474-479); calls on contigs the short-read /
    long-read BAM does not know (FAIL6 / FAIL7, sviper.h:189-211); a call whose best supporting long read is too short to crop
    (the silent `return false` of sviper.h:327: FILTER and QUAL stay as they came in). FAIL3 (sviper.h:337-348) cannot be
    reached: the crop loop either appends a sequence or returns at :327 on its first element."""
    rng = np.random.default_rng(seed)
    os.makedirs(outdir, exist_ok=True)
    ref = rng.integers(0, 4, contig_len).astype(np.uint8)
    # (type, 0-based ref position of the first affected base / insertion point, length)
    lens = [100, 300, 300, 100, 800, 1200 if extras else 800]
    svs = [("DEL" if i % 2 == 0 else "INS", 6000 + 7000 * i, lens[i % 6]) for i in range(max(0, (contig_len - 18000) // 7000))]
    if calls is not None:
        svs = list(calls)                                # (make_c2: a given list of calls instead of the C1 mix)
    svs.append(("DEL", contig_len - 8000, 200))          # supported by long reads, but (almost) no short reads there: FAIL4
    ins_seqs = {pos: rng.integers(0, 4, ln).astype(np.uint8) for typ, pos, ln in svs if typ == "INS"}
    # a 45 bp insertion that only the long-read haplotype carries: every short read across it votes "gap", the indel
    # rounds delete it again and refine_variant no longer finds the call in the final alignment
    phantom = ("INS", 9500, 45) if (extras and len(svs) > 2 and not only_filter_paths) else None
    if phantom:
        ins_seqs[phantom[1]] = rng.integers(0, 4, phantom[2]).astype(np.uint8)
    alt, amap, akind = haplotype(ref, svs, ins_seqs)                                        # what the short reads sample
    alt_l, amap_l, akind_l = haplotype(ref, svs + ([phantom] if phantom else []), ins_seqs)  # what the long reads sample

    # ---- FASTA + .fai -----------------------------------------------------------------------------------------
    fa = os.path.join(outdir, "ref.fa")
    small = {nm: rng.integers(0, 4, 3000).astype(np.uint8) for nm in ("chrA", "chrB")} if extras else {}
    with open(fa, "w") as f, open(fa + ".fai", "w") as fi:
        off = 0
        for nm, sq in [("chr1", ref)] + sorted(small.items()):
            hdr = ">%s\n" % nm
            f.write(hdr); off += len(hdr)
            seq = s(sq)
            fi.write("%s\t%d\t%d\t60\t61\n" % (nm, len(seq), off))
            for i in range(0, len(seq), 60):
                f.write(seq[i:i + 60] + "\n"); off += len(seq[i:i + 60]) + 1
    long_refs = [("chr1", contig_len)]
    short_refs = [("chr1", contig_len)] + ([("chrB", 3000)] if extras else [])

    # ---- long reads: 8-15 kb of the haplotype, 10% errors, true CIGAR against the reference ------------------
    # no long read reaches into the last 3.5 kb (FAIL1 record there)
    recs = []
    n_long = int(cov_long * len(alt_l) / 11500)
    crop_victim = svs[2] if (extras and len(svs) > 3 and not only_filter_paths) else None
    for i in range(n_long + (1 if crop_victim else 0)):
        ln = int(rng.integers(8000, 15001))
        st = int(rng.integers(0, max(1, len(alt_l) - 3500 - ln)))
        mapq = int(rng.integers(20, 61))
        if i == n_long:
            # the read with the best mapping quality at this call carries a further 450-base insertion 150 bp behind the
            # deletion: its crop is more than k bases longer than expected, and polish_variant returns at sviper.h:327
            # without touching the record
            a0 = int(np.searchsorted(amap_l, crop_victim[1] - 2000)); ln = 4500; st = a0; mapq = 70
            junk_at = crop_victim[1] + crop_victim[2] + 150
        ops, opref, bases = [], [], []          # opref: reference coordinate of an M / D op (None for I)
        prev_ref = None
        u_all = rng.random(ln); nb_all = rng.integers(0, 4, (ln, 2))
        for j, a in enumerate(range(st, min(len(alt_l), st + ln))):
            r = int(amap_l[a])
            if akind_l[a] == 0 and prev_ref is not None and r > prev_ref + 1:
                ops.extend("D" * (r - prev_ref - 1)); opref.extend(range(prev_ref + 1, r))     # the deletion as one element
            if i == n_long and crop_victim and akind_l[a] == 0 and r == junk_at:
                for x in rng.integers(0, 4, 450):
                    ops.append("I"); opref.append(None); bases.append(x)
            if akind_l[a] == 1:
                ops.append("I"); opref.append(None); bases.append(alt_l[a])
                continue
            u = u_all[j]
            if u < 0.03:
                ops.append("M"); opref.append(r); bases.append((alt_l[a] + 1 + nb_all[j, 0] % 3) % 4)
            elif u < 0.065:
                ops.append("M"); opref.append(r); bases.append(alt_l[a])
                ops.append("I"); opref.append(None); bases.append(nb_all[j, 1])
            elif u < 0.10:
                ops.append("D"); opref.append(r)
            else:
                ops.append("M"); opref.append(r); bases.append(alt_l[a])
            prev_ref = r
        while ops and ops[0] != "M":                                       # an alignment starts and ends on a match
            if ops[0] == "I":
                bases.pop(0)
            ops.pop(0); opref.pop(0)
        while ops and ops[-1] != "M":
            if ops[-1] == "I":
                bases.pop()
            ops.pop(); opref.pop()
        first_ref = opref[0]
        recs.append(("long%04d" % i, 16 if rng.random() < 0.5 else 0, 0, first_ref, mapq,
                     cigar_str(ops), -1, -1, 0, s(np.array(bases, np.uint8)), "*"))
    recs.sort(key=lambda x: x[3])
    bamwriter.write_bam(os.path.join(outdir, "long.bam"), long_refs, recs)

    # ---- short reads: fragments of the haplotype, both mates placed at their reference coordinate -----------
    recs = []
    n_pairs = int(cov_short * len(alt) / 300)
    lo_gap, hi_gap = contig_len - 9500, contig_len - 6500                                           # almost no short reads here: FAIL4
    for i in range(n_pairs):
        fl = int(np.clip(rng.normal(430, 40), 320, 700))
        st = int(rng.integers(0, len(alt) - fl))
        if lo_gap - 700 < amap[st] < hi_gap and rng.random() < 0.99:
            continue
        a = alt[st:st + 150].copy(); b = alt[st + fl - 150:st + fl].copy()
        for r in (a, b):
            m = rng.random(150) < 0.005
            r[m] = (r[m] + 1 + rng.integers(0, 3, int(m.sum()))) % 4
        pa, pb = int(amap[st]), int(amap[st + fl - 150])
        qa = "".join(chr(33 + int(x)) for x in rng.integers(20, 40, 150))
        qb = "".join(chr(33 + int(x)) for x in rng.integers(20, 40, 150))
        first = rng.random() < 0.5                                          # which mate is read 1
        fa_, fb_ = (99, 147) if first else (163, 83)
        name = "sr%06d" % i
        recs.append((name, fa_, 0, pa, 60, "150M", 0, pb, fl, s(a), qa))
        recs.append((name, fb_, 0, pb, 60, "150M", 0, pa, -fl, s(b), qb))
    recs.sort(key=lambda x: x[3])
    bamwriter.write_bam(os.path.join(outdir, "short.bam"), short_refs, recs)

    # ---- candidate VCF: jittered calls + the FILTER paths -------------------------------------------------------
    vcf = os.path.join(outdir, "candidates.vcf")
    with open(vcf, "w") as f:
        f.write("##fileformat=VCFv4.2\n##contig=<ID=chr1,length=%d>\n" % contig_len)
        extra_calls = []
        if phantom:
            extra_calls.append("chr1\t%d\tsvP\tN\t<INS>\t.\tPASS\tSVTYPE=INS;END=%d;SVLEN=%d\tGT\t0/1\n" %
                               (phantom[1] + 2, phantom[1] + 2, phantom[2]))                                # FAIL5 candidate
        if extras:
            extra_calls.append("chrA\t1000\tsvA\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=1200;SVLEN=-200\tGT\t0/1\n")   # FAIL6
            extra_calls.append("chrB\t1000\tsvB\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=1200;SVLEN=-200\tGT\t0/1\n")   # FAIL7
        f.write('##INFO=<ID=SVTYPE,Number=1,Type=String,Description="Type">\n')
        f.write('##INFO=<ID=END,Number=1,Type=Integer,Description="End">\n')
        f.write('##INFO=<ID=SVLEN,Number=1,Type=Integer,Description="Length">\n')
        f.write('##FILTER=<ID=q10,Description="Quality below 10">\n')
        f.write('##FORMAT=<ID=GT,Number=1,Type=String,Description="Genotype">\n')
        f.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tSAMPLE\n")
        n = 0
        for typ, pos, ln in (svs[-1:] if only_filter_paths else svs):
            jp = pos + int(rng.integers(-10, 11))
            jl = max(30, int(ln * rng.uniform(0.9, 1.1)))
            end = jp + jl if typ == "DEL" else jp
            info = "SVTYPE=%s;END=%d;SVLEN=%s%d" % (typ, end, "-" if typ == "DEL" else "", jl)
            f.write("chr1\t%d\tsv%d\tN\t<%s>\t.\tPASS\t%s\tGT\t1/1\n" % (jp, n, typ, info)); n += 1
        f.write("chr1\t%d\tsv%d\tN\t<DUP>\t.\tPASS\tSVTYPE=DUP;END=%d;SVLEN=500\tGT\t0/1\n" % (contig_len - 14000, n, contig_len - 13500)); n += 1   # SKIP
        f.write("chr1\t%d\tsv%d\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=%d;SVLEN=-200\tGT\t0/1\n" % (contig_len - 12000, n, contig_len - 11800)); n += 1   # FAIL2
        f.write("chr1\t%d\tsv%d\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=%d;SVLEN=-200\tGT\t0/1\n" % (contig_len - 2000, n, contig_len - 1800)); n += 1   # FAIL1
        f.write("chr2\t1000\tsv%d\tN\t<DEL>\t.\tPASS\tSVTYPE=DEL;END=1200;SVLEN=-200\tGT\t0/1\n" % n); n += 1         # FAIL0
        for line in extra_calls:
            f.write(line); n += 1
    return dict(ref=fa, vcf=vcf, long=os.path.join(outdir, "long.bam"), short=os.path.join(outdir, "short.bam"), k=k,
                n_records=n, n_sv=0 if only_filter_paths else len(svs) - 1)


def make_c2(outdir, n_records=100, seed=2):
    """BASELINE.json configs[1] / SURVEY 8d C2 as FILES: n_records DELs with lengths log-uniform in 50..1000 bp, 7 kb apart on
    one contig, 30x 2x150 bp short reads, 20x long reads at 10% error, -k 400 (plus the FILTER-path records of make())."""
    rng = np.random.default_rng(seed * 7919)
    calls = [("DEL", 6000 + 7000 * i, int(round(np.exp(rng.uniform(np.log(50), np.log(1000)))))) for i in range(n_records)]
    contig_len = 6000 + 7000 * n_records + 18000
    d = make(outdir, seed=seed, contig_len=contig_len, extras=False, calls=calls)
    d["contig_len"] = contig_len
    return d


if __name__ == "__main__":
    import sys
    print(make(sys.argv[1] if len(sys.argv) > 1 else "/tmp/svp_e2e"))
