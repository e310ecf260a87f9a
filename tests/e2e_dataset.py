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


def make(outdir, seed=1, contig_len=60000, cov_short=30, cov_long=20, k=400, only_filter_paths=False):
    rng = np.random.default_rng(seed)
    os.makedirs(outdir, exist_ok=True)
    ref = rng.integers(0, 4, contig_len).astype(np.uint8)
    # (type, 0-based ref position of the first affected base / insertion point, length)
    lens = [100, 300, 300, 100, 800, 800]
    svs = [("DEL" if i % 2 == 0 else "INS", 6000 + 7000 * i, lens[i % 6]) for i in range(max(0, (contig_len - 18000) // 7000))]
    svs.append(("DEL", contig_len - 8000, 200))          # supported by long reads, but (almost) no short reads there: FAIL4
    # ---- the sample's haplotype, with for every alt base the reference coordinate it maps to -----------------
    alt, amap, akind = [], [], []          # akind: 0 = reference base, 1 = inserted base
    p = 0
    for typ, pos, ln in svs:
        alt.append(ref[p:pos]); amap.append(np.arange(p, pos)); akind.append(np.zeros(pos - p, np.uint8))
        if typ == "DEL":
            p = pos + ln
        else:
            alt.append(rng.integers(0, 4, ln).astype(np.uint8)); amap.append(np.full(ln, pos)); akind.append(np.ones(ln, np.uint8))
            p = pos
    alt.append(ref[p:]); amap.append(np.arange(p, contig_len)); akind.append(np.zeros(contig_len - p, np.uint8))
    alt = np.concatenate(alt); amap = np.concatenate(amap); akind = np.concatenate(akind)

    # ---- FASTA + .fai -----------------------------------------------------------------------------------------
    fa = os.path.join(outdir, "ref.fa")
    with open(fa, "w") as f:
        f.write(">chr1\n")
        seq = s(ref)
        for i in range(0, len(seq), 60):
            f.write(seq[i:i + 60] + "\n")
    with open(fa + ".fai", "w") as f:
        f.write("chr1\t%d\t6\t60\t61\n" % contig_len)

    # ---- long reads: 8-15 kb of the haplotype, 10% errors, true CIGAR against the reference ------------------
    # no long read reaches into the last 3.5 kb (FAIL1 record there)
    recs = []
    n_long = int(cov_long * len(alt) / 11500)
    for i in range(n_long):
        ln = int(rng.integers(8000, 15001))
        st = int(rng.integers(0, max(1, len(alt) - 3500 - ln)))
        ops, opref, bases = [], [], []          # opref: reference coordinate of an M / D op (None for I)
        prev_ref = None
        u_all = rng.random(ln); nb_all = rng.integers(0, 4, (ln, 2))
        for j, a in enumerate(range(st, min(len(alt), st + ln))):
            r = int(amap[a])
            if akind[a] == 0 and prev_ref is not None and r > prev_ref + 1:
                ops.extend("D" * (r - prev_ref - 1)); opref.extend(range(prev_ref + 1, r))     # the deletion as one element
            if akind[a] == 1:
                ops.append("I"); opref.append(None); bases.append(alt[a])
                continue
            u = u_all[j]
            if u < 0.03:
                ops.append("M"); opref.append(r); bases.append((alt[a] + 1 + nb_all[j, 0] % 3) % 4)
            elif u < 0.065:
                ops.append("M"); opref.append(r); bases.append(alt[a])
                ops.append("I"); opref.append(None); bases.append(nb_all[j, 1])
            elif u < 0.10:
                ops.append("D"); opref.append(r)
            else:
                ops.append("M"); opref.append(r); bases.append(alt[a])
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
        recs.append(("long%04d" % i, 16 if rng.random() < 0.5 else 0, 0, first_ref, int(rng.integers(20, 61)),
                     cigar_str(ops), -1, -1, 0, s(np.array(bases, np.uint8)), "*"))
    recs.sort(key=lambda x: x[3])
    bamwriter.write_bam(os.path.join(outdir, "long.bam"), [("chr1", contig_len)], recs)

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
    bamwriter.write_bam(os.path.join(outdir, "short.bam"), [("chr1", contig_len)], recs)

    # ---- candidate VCF: jittered calls + the FILTER paths -------------------------------------------------------
    vcf = os.path.join(outdir, "candidates.vcf")
    with open(vcf, "w") as f:
        f.write("##fileformat=VCFv4.2\n##contig=<ID=chr1,length=%d>\n" % contig_len)
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
    return dict(ref=fa, vcf=vcf, long=os.path.join(outdir, "long.bam"), short=os.path.join(outdir, "short.bam"), k=k,
                n_records=n, n_sv=0 if only_filter_paths else len(svs) - 1)


if __name__ == "__main__":
    import sys
    print(make(sys.argv[1] if len(sys.argv) > 1 else "/tmp/svp_e2e"))
