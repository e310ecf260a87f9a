"""Synthetic hot-path workloads (there is no network and the reference ships no test data,
tests/sviper_test.sh:14-17). Shapes follow BASELINE.json configs / SURVEY.md section 8(d).

A "variant case" is exactly what the reference hands to the hot path at sviper.h:426-460:
read pairs as sequenced (basics.h:306-401 undoes the BAM strand flip), the flanked long-read
consensus, and the reference part for the final alignment. Bases are uint8 0..4 = ACGTN.
"""
import numpy as np

COMP = np.array([3, 2, 1, 0, 4], np.uint8)


def revcomp(a):
    return COMP[a[::-1]]


def mutate(rng, seq, err, frac_sub=0.4, frac_ins=0.3):
    """errors at rate err: substitutions / single-base insertions / deletions"""
    out = []
    u = rng.random(len(seq))
    kind = rng.random(len(seq))
    nb = rng.integers(0, 4, len(seq))
    for i in range(len(seq)):
        if u[i] >= err:
            out.append(seq[i])
        elif kind[i] < frac_sub:
            out.append((seq[i] + 1 + nb[i] % 3) % 4)
        elif kind[i] < frac_sub + frac_ins:
            out.append(seq[i]); out.append(nb[i])
        # else deletion
    return np.array(out, np.uint8)


def make_variant(seed, kind="DEL", svlen=300, k=400, cov=30, read_len=150, cons_err=0.05, flank=500,
                 frag_mean=430.0, frag_sd=40.0, sub_rate=0.005, margin=300, n_rate=0.0):
    rng = np.random.default_rng(seed)
    core = 2 * k
    if kind == "DEL":
        g = rng.integers(0, 4, margin + flank + k + svlen + k + flank + margin).astype(np.uint8)
        cut = margin + flank + k
        alt = np.concatenate([g[:cut], g[cut + svlen:]])
        ref_part = g[margin:len(g) - margin]
    else:
        g = rng.integers(0, 4, margin + flank + core + flank + margin).astype(np.uint8)
        cut = margin + flank + k
        alt = np.concatenate([g[:cut], rng.integers(0, 4, svlen).astype(np.uint8), g[cut:]])
        ref_part = g[margin:len(g) - margin]
        core += svlen
    truth = alt[margin:len(alt) - margin]                       # flank + core + flank
    mid = mutate(rng, truth[flank:len(truth) - flank], cons_err)
    cons = np.concatenate([truth[:flank], mid, truth[len(truth) - flank:]])
    # read pairs from fragments overlapping the core
    n_pairs = max(1, int(round(cov * core / (2.0 * read_len))))
    r1, q1, r2, q2 = [], [], [], []
    lo, hi = margin + flank - read_len // 2, margin + flank + core + read_len // 2
    while len(r1) < n_pairs:
        fl = int(np.clip(rng.normal(frag_mean, frag_sd), 2 * read_len + 20, 700))
        st = int(rng.integers(lo - fl, hi))
        if st < 0 or st + fl > len(alt) or st + fl < lo or st > hi:
            continue
        a = alt[st:st + read_len].copy()
        b = revcomp(alt[st + fl - read_len:st + fl])
        for r in (a, b):
            m = rng.random(read_len) < sub_rate
            r[m] = (r[m] + 1 + rng.integers(0, 3, int(m.sum()))) % 4
            if n_rate > 0:
                r[rng.random(read_len) < n_rate] = 4
        qa = rng.integers(20, 40, read_len).astype(np.uint8)
        qb = rng.integers(20, 40, read_len).astype(np.uint8)
        qa[a == 4] = 0; qb[b == 4] = 0
        if rng.random() < 0.5:
            a, b, qa, qb = b, a, qb, qa
        r1.append(a); q1.append(qa); r2.append(b); q2.append(qb)
    return dict(kind=kind, svlen=svlen, k=k, flank=flank, ref_part=ref_part, consensus=cons, truth=truth,
                reads1=r1, quals1=q1, reads2=r2, quals2=q2, ins_mean=280.054, ins_sd=145.162)


def make_align_tasks(seed, n_windows=64, reads_per_window=256, read_len=150,
                     win_lens=(800, 1000, 1200, 1400, 1600, 1800, 2000), sub_rate=0.01, indel_rate=0.001):
    """BASELINE.json config 5: reads_per_window reads of read_len bp sampled from each window,
    1% substitutions, 0.1% single-base indels, half reverse-complemented, Phred 30."""
    rng = np.random.default_rng(seed)
    wins, reads, win_of = [], [], []
    for w in range(n_windows):
        L = int(win_lens[rng.integers(0, len(win_lens))])
        win = rng.integers(0, 4, L).astype(np.uint8)
        wins.append(win)
        for _ in range(reads_per_window):
            st = int(rng.integers(0, L - read_len - 8))
            src = win[st:st + read_len + 8]
            rd = mutate(rng, src, sub_rate + indel_rate, frac_sub=sub_rate / (sub_rate + indel_rate),
                        frac_ins=0.5 * indel_rate / (sub_rate + indel_rate))[:read_len]
            if len(rd) < read_len:
                rd = np.concatenate([rd, rng.integers(0, 4, read_len - len(rd)).astype(np.uint8)])
            if rng.random() < 0.5:
                rd = revcomp(rd)
            reads.append(np.ascontiguousarray(rd)); win_of.append(w)
    return dict(windows=wins, reads=reads, win_of=np.array(win_of, np.int32),
                quals=[np.full(read_len, 30, np.uint8) for _ in reads])
