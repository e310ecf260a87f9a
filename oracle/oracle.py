"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper over oracle/liboracle.so (the CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module. The product (sviper_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)
f64p = C.POINTER(C.c_double)


class OrcConfig(C.Structure):
    _fields_ = [("flank", C.c_int), ("ins_mean", C.c_double), ("ins_sd", C.c_double),
                ("min_coverage", C.c_double), ("baseQ_mean", C.c_double), ("fix_indels", C.c_int),
                ("mappQ_mean", C.c_double), ("subst", C.c_uint), ("ins", C.c_uint), ("del_", C.c_uint),
                ("rounds", C.c_uint)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_align_semiglobal.restype = C.c_int
        L.orc_msa_sw_fill.restype = None
        L.orc_msa_sw_fill.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p, u8p, C.POINTER(C.c_int)]
        L.orc_msa_global_fill.restype = C.c_int
        L.orc_msa_global_fill.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.POINTER(C.c_int)]
        L.orc_align_convex.restype = C.c_double
        L.orc_align_convex.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                       C.c_char_p, i32p]
        L.orc_assign_quality.restype = C.c_double
        L.orc_baseq_mean.restype = C.c_double
        L.orc_polish_round.restype = C.c_void_p
        L.orc_round_profile.restype = f64p
        L.orc_round_cov.restype = u32p
        L.orc_round_isgap.restype = u8p
        L.orc_round_out.restype = u8p
        for f in ("orc_round_Lp", "orc_round_profile", "orc_round_cov", "orc_round_isgap", "orc_round_out",
                  "orc_round_out_len", "orc_round_free"):
            getattr(L, f).argtypes = [C.c_void_p]
        L.orc_round_proper.argtypes = [C.c_void_p, C.c_int]
        L.orc_round_aln.argtypes = [C.c_void_p, C.c_int, C.c_int, i32p, i32p, i32p, C.c_char_p]
        _LIB = L
    return _LIB


_ENC = np.full(256, 4, np.uint8)
for _i, _c in enumerate("ACGT"):
    _ENC[ord(_c)] = _i
    _ENC[ord(_c.lower())] = _i
DEC = np.frombuffer(b"ACGTN", np.uint8)


def enc(s):
    """ASCII ACGTN -> uint8 0..4"""
    return _ENC[np.frombuffer(s.encode(), np.uint8)].copy()


def dec(a):
    return DEC[np.asarray(a, np.uint8)].tobytes().decode()


def _p(a, t=u8p):
    return a.ctypes.data_as(t)


def norm_quals(bases, quals):
    """Dna5Q semantics: clamp to 0..62, N carries quality 0 (seqan/basic/alphabet_residue.h:331-365)."""
    q = np.clip(np.asarray(quals, np.int64), 0, 62).astype(np.uint8)
    q[np.asarray(bases) == 4] = 0
    return q


def align_semiglobal(ref, read, mm=1, mx=-4, ge=-1, go=-6):
    ref = np.ascontiguousarray(ref, np.uint8); read = np.ascontiguousarray(read, np.uint8)
    ops = C.create_string_buffer(len(ref) + len(read) + 1)
    nops = C.c_int(); endc = C.c_int()
    s = lib().orc_align_semiglobal(_p(ref), len(ref), _p(read), len(read), mm, mx, ge, go, ops, C.byref(nops),
                                   C.byref(endc))
    return s, ops.raw[:nops.value].decode(), endc.value


def map_read(ref, read, qual):
    ref = np.ascontiguousarray(ref, np.uint8); read = np.ascontiguousarray(read, np.uint8)
    qual = np.ascontiguousarray(qual, np.uint8)
    ops = C.create_string_buffer(len(ref) + len(read) + 1)
    v = [C.c_int() for _ in range(6)]
    lib().orc_map_read(_p(read), _p(qual), len(read), _p(ref), len(ref), *[C.byref(x) for x in v[:5]], ops,
                       C.byref(v[5]))
    return dict(score=v[0].value, score_rc=v[1].value, is_rc=v[2].value, lead=v[3].value, end=v[4].value,
                ops=ops.raw[:v[5].value].decode())


def rows_from_ops(ops, ref, read):
    """gapped rows as ASCII, like operator<< on seqan::Gaps"""
    r, q, i, j = [], [], 0, 0
    for o in ops:
        if o == "I":
            r.append("-")
        else:
            r.append("ACGTN"[ref[i]]); i += 1
        if o == "D":
            q.append("-")
        else:
            q.append("ACGTN"[read[j]]); j += 1
    return "".join(r), "".join(q)


def pack_reads(reads, quals):
    """list of uint8 arrays -> (bases concat, quals concat, int32 offsets[n+1])"""
    off = np.zeros(len(reads) + 1, np.int32)
    off[1:] = np.cumsum([len(r) for r in reads])
    b = np.concatenate([np.asarray(r, np.uint8) for r in reads] + [np.zeros(0, np.uint8)])
    q = np.concatenate([np.asarray(x, np.uint8) for x in quals] + [np.zeros(0, np.uint8)])
    return np.ascontiguousarray(b), np.ascontiguousarray(q), off


def baseq_mean(q1, off1, q2, off2):
    return lib().orc_baseq_mean(_p(q1), _p(off1, i32p), _p(q2), _p(off2, i32p), len(off1) - 1)


def make_config(flank, ins_mean, ins_sd, baseQ_mean, min_coverage=4.0):
    c = OrcConfig()
    c.flank = flank; c.ins_mean = ins_mean; c.ins_sd = ins_sd; c.min_coverage = min_coverage
    c.baseQ_mean = baseQ_mean; c.fix_indels = 0; c.mappQ_mean = 0.0
    return c


def polish_round(cfg, r1, q1, off1, r2, q2, off2, ref, want_alns=False):
    L = lib()
    ref = np.ascontiguousarray(ref, np.uint8)
    np_ = len(off1) - 1
    h = L.orc_polish_round(C.byref(cfg), np_, _p(r1), _p(q1), _p(off1, i32p), _p(r2), _p(q2), _p(off2, i32p),
                           _p(ref), len(ref))
    Lp = L.orc_round_Lp(h)
    out = dict(
        Lp=Lp,
        profile=np.ctypeslib.as_array(L.orc_round_profile(h), (Lp, 6)).copy(),
        cov=np.ctypeslib.as_array(L.orc_round_cov(h), (Lp,)).copy(),
        isgap=np.ctypeslib.as_array(L.orc_round_isgap(h), (Lp,)).copy(),
        out=np.ctypeslib.as_array(L.orc_round_out(h), (L.orc_round_out_len(h),)).copy(),
        proper=np.array([L.orc_round_proper(h, i) for i in range(np_)], np.uint8),
    )
    if want_alns:
        alns = []
        buf = C.create_string_buffer(len(ref) + 100000)
        for i in range(np_):
            pair = []
            for w in (0, 1):
                s, src, rc = C.c_int(), C.c_int(), C.c_int()
                n = L.orc_round_aln(h, i, w, C.byref(s), C.byref(src), C.byref(rc), buf)
                pair.append(dict(score=s.value, score_rc=src.value, is_rc=rc.value, ops=buf.raw[:n].decode()))
            alns.append(pair)
        out["alns"] = alns
    L.orc_round_free(h)
    return out


def polish_to_perfection(cfg, r1, q1, off1, r2, q2, off2, ref):
    L = lib()
    ref = np.ascontiguousarray(ref, np.uint8)
    cap = len(ref) * 2 + 4096
    out = np.zeros(cap, np.uint8); cov = np.zeros(cap, np.uint32); covlen = C.c_int(0)
    n = L.orc_polish_to_perfection(C.byref(cfg), len(off1) - 1, _p(r1), _p(q1), _p(off1, i32p), _p(r2), _p(q2),
                                   _p(off2, i32p), _p(ref), len(ref), _p(out), cap, _p(cov, u32p), cap,
                                   C.byref(covlen))
    if n < 0:
        raise RuntimeError("oracle: polished sequence exceeds buffer")
    return out[:n].copy(), cov[:covlen.value].copy()


def align_convex(ref_part, polished, MM=90000.0, MX=-80000.0, GE=-1.0, GO=-100000.0):
    a = np.ascontiguousarray(ref_part, np.uint8); b = np.ascontiguousarray(polished, np.uint8)
    ops = C.create_string_buffer(len(a) + len(b) + 1)
    nops = C.c_int()
    s = lib().orc_align_convex(_p(a), len(a), _p(b), len(b), MM, MX, GE, GO, ops, C.byref(nops))
    return s, ops.raw[:nops.value].decode()


MSA_SCORES = (5, -3, -1, -3)   # match, mismatch, gap extend, gap open: SimpleScore(5, -3, -1, -3), basics.h:494


def fnv1a(buf):
    """64-bit FNV-1a of a byte buffer (what ref_harness prints for a trace matrix)"""
    h = 1469598103934665603
    for b in bytes(buf):
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def msa_sw_fill(h, v, forbidden_cells=(), scores=MSA_SCORES):
    """_alignSmithWaterman (graph_align_tcoffee_library.h:252-390). forbidden_cells: (col, row), 1-based.
    Returns (trace bytes, score_max, row, col, initialDir)."""
    a = np.ascontiguousarray(h, np.uint8); b = np.ascontiguousarray(v, np.uint8)
    n, m = len(a), len(b)
    forb = None
    if len(forbidden_cells):
        forb = np.zeros(n * m, np.uint8)
        for col, row in forbidden_cells:
            forb[(col - 1) * m + (row - 1)] = 1
    trace = np.zeros(n * ((m + 1) // 2) + 1, np.uint8)
    out = (C.c_int * 4)()
    lib().orc_msa_sw_fill(_p(a), n, _p(b), m, *scores, _p(forb) if forb is not None else None, _p(trace), out)
    return trace[:n * ((m + 1) // 2)], out[0], out[1], out[2], out[3]


def msa_global_fill(h, v, scores=MSA_SCORES):
    """trace matrix of the global pairwise library's Gotoh alignment. Returns (trace bytes (n+1)*(m+1), score, (S, H, V))"""
    a = np.ascontiguousarray(h, np.uint8); b = np.ascontiguousarray(v, np.uint8)
    n, m = len(a), len(b)
    trace = np.zeros((n + 1) * (m + 1), np.uint8)
    end = (C.c_int * 3)()
    sc = lib().orc_msa_global_fill(_p(a), n, _p(b), m, *scores, _p(trace), end)
    return trace, sc, (end[0], end[1], end[2])


def cigar(ops):
    out, last, n = [], None, 0
    for o in ops:
        if o != last:
            if n:
                out.append((last, n))
            last, n = o, 0
        n += 1
    if n:
        out.append((last, n))
    return out


def cigar_str(ops):
    return "".join("%d%s" % (n, o) for o, n in cigar(ops))


def evaluate_final_mapping(sv_type, pos, end, svlen, info, cigar, polished_text, begin_pos, k):
    """assign_quality + refine_variant (sviper.h:468-489) through the C restatement (oracle_eval.c).
    cigar: [(op, count)], polished_text: str. Returns dict(has, pos, end, svlen, qual, filter, info)."""
    L = lib()
    L.orc_assign_quality.restype = C.c_double
    ops = "".join(o for o, _ in cigar).encode()
    cnt = (C.c_uint * max(1, len(cigar)))(*[c for _, c in cigar])
    p, e, sl, rp = C.c_int(pos), C.c_int(end), C.c_int(svlen), C.c_int(0)
    has = L.orc_refine_variant(C.c_int(begin_pos), ops, cnt, C.c_int(len(cigar)), C.c_int(sv_type),
                               C.byref(p), C.byref(e), C.byref(sl), C.byref(rp))
    out = dict(has=has, pos=p.value, end=e.value, svlen=sl.value, filter="PASS" if has == 1 else "FAIL5", info=info,
               qual=L.orc_assign_quality(C.c_int(len(cigar)), C.c_int(k)))
    if has == 1:
        cap = len(info) * 8 + len(polished_text) * 4 + 256
        buf = C.create_string_buffer(cap)
        ins = polished_text[rp.value:rp.value + sl.value].encode() if sv_type == 2 else None
        L.orc_refine_info(info.encode(), C.c_int(sv_type), sl, e, ins, buf, C.c_size_t(cap))
        out["info"] = buf.value.decode()
    return out
