"""Python host side over the C ABI (include/sviper_b200.h). SViper itself is C++; this mirror exists so
that tests and bench.py read like the reference's call sites:

    mapping / map_single_read     include/sviper/mapping.h:84-179      -> Context.align_batch
    polish_to_perfection          include/sviper/polishing.h:319-346   -> Context.polish_batch
    final globalAlignment+CIGAR   include/sviper/sviper.h:452-466      -> Context.final_align_batch
    assign_quality                include/sviper/evaluate_final_mapping.h:17-28 -> assign_quality

Every call goes to the CUDA library; if it is missing or no GPU is present this module raises --
there is no CPU implementation behind it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libsviper_b200.so")

u8p, i32p, u32p, f64p = (C.POINTER(t) for t in (C.c_uint8, C.c_int32, C.c_uint32, C.c_double))


class AlignIn(C.Structure):
    _fields_ = [("n_reads", C.c_int32), ("read_bases", u8p), ("read_off", i32p), ("n_windows", C.c_int32),
                ("win_bases", u8p), ("win_off", i32p), ("n_tasks", C.c_int32), ("task_read", i32p),
                ("task_win", i32p)]


class AlignOut(C.Structure):
    _fields_ = [("score", i32p), ("score_rc", i32p), ("is_rc", i32p), ("lead", i32p), ("endpos", i32p),
                ("ops", C.c_char_p), ("ops_cap", C.c_int32), ("ops_len", i32p)]


class WindowIn(C.Structure):
    _fields_ = [("consensus", u8p), ("consensus_len", C.c_int32), ("ref_flank_length", C.c_int32),
                ("n_pairs", C.c_int32), ("reads1_bases", u8p), ("reads1_quals", u8p), ("reads1_off", i32p),
                ("reads2_bases", u8p), ("reads2_quals", u8p), ("reads2_off", i32p), ("baseQ_mean", C.c_double),
                ("insert_mean", C.c_double), ("insert_sd", C.c_double), ("min_coverage", C.c_double),
                ("ref_part", u8p), ("ref_part_len", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("status", C.c_int32), ("polished", u8p), ("polished_cap", C.c_int32), ("polished_len", C.c_int32),
                ("rounds", C.c_uint32), ("substituted_bases", C.c_uint32), ("inserted_bases", C.c_uint32),
                ("deleted_bases", C.c_uint32), ("mappQ_mean", C.c_double), ("cov_profile", u32p),
                ("cov_cap", C.c_int32), ("cov_len", C.c_int32), ("cigar_op", C.c_char_p), ("cigar_cnt", u32p),
                ("cigar_cap", C.c_int32), ("n_cigar", C.c_int32), ("final_score", C.c_double),
                ("alignment_rounds", C.c_uint32), ("reserved_", C.c_uint32)]


class VariantPos(C.Structure):
    _fields_ = [("sv_type", C.c_int32), ("ref_pos", C.c_int32), ("ref_pos_end", C.c_int32), ("sv_length", C.c_int32)]


SV_DEL, SV_INS = 1, 2


class BamRecord(C.Structure):
    _fields_ = [("begin_pos", C.c_int32), ("rid", C.c_int32), ("flag", C.c_uint16), ("mapq", C.c_uint8),
                ("cigar_op", C.c_char_p), ("cigar_cnt", C.POINTER(C.c_uint32)), ("n_cigar", C.c_int32),
                ("seq", C.c_char_p), ("seq_len", C.c_int32)]


class MsaPair(C.Structure):
    _fields_ = [("h", u8p), ("n", C.c_int32), ("v", u8p), ("m", C.c_int32)]


class MsaFillResult(C.Structure):
    _fields_ = [("score", C.c_int32), ("row", C.c_int32), ("col", C.c_int32), ("dir", C.c_int32), ("S", C.c_int32),
                ("H", C.c_int32), ("V", C.c_int32)]


class FinalIn(C.Structure):
    _fields_ = [("ref_part", u8p), ("ref_part_len", C.c_int32), ("seq", u8p), ("seq_len", C.c_int32)]


EXPORTS = ["svp_create", "svp_destroy", "svp_last_error", "svp_version", "svp_align_batch", "svp_align_rerun",
           "svp_polish_batch", "svp_polish_rerun", "svp_final_align_batch", "svp_assign_quality", "svp_submit_polish", "svp_wait_any",
           "svp_debug_round", "svp_probe_dpx", "svp_stats", "svp_set_lanes", "svp_refine_variant", "svp_refine_info",
           "svp_format_quality", "svp_record_supports_variant", "svp_read_region_boundaries",
           "svp_alignment_length_in_ref", "svp_cut_down_high_coverage", "svp_merge_record_group", "svp_k1_cells",
           "svp_flush", "svp_k1_failed_certificates", "svp_msa_begin", "svp_msa_global_fill", "svp_msa_local_fill", "svp_final_align_static", "svp_probe_fp64", "svp_engine_config", "svp_engine_stats", "svp_window_capacity"]

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "csrc"), "libsviper_b200.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("sviper_b200: %s is missing -- run __graft_entry__.build(); there is no CPU fallback"
                               % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.svp_create.restype = C.c_void_p
        L.svp_create.argtypes = [C.c_int, C.c_void_p]
        L.svp_destroy.argtypes = [C.c_void_p]
        L.svp_last_error.restype = C.c_char_p
        L.svp_last_error.argtypes = [C.c_void_p]
        L.svp_version.restype = C.c_char_p
        L.svp_align_batch.argtypes = [C.c_void_p, C.POINTER(AlignIn), C.POINTER(AlignOut), C.c_int]
        L.svp_align_rerun.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float)]
        L.svp_polish_batch.argtypes = [C.c_void_p, C.c_int32, C.POINTER(WindowIn), C.POINTER(Result)]
        L.svp_polish_rerun.argtypes = [C.c_void_p]
        L.svp_final_align_batch.argtypes = [C.c_void_p, C.c_int32, C.POINTER(FinalIn), C.POINTER(Result)]
        L.svp_assign_quality.restype = C.c_double
        L.svp_assign_quality.argtypes = [C.c_int32, C.c_int32]
        L.svp_submit_polish.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(WindowIn), C.POINTER(Result)]
        L.svp_wait_any.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.POINTER(Result))]
        L.svp_flush.argtypes = [C.c_void_p]
        L.svp_msa_begin.argtypes = [C.c_void_p, C.c_int32, C.POINTER(MsaPair), C.c_int32, C.c_int32, C.c_int32, C.c_int32]
        L.svp_msa_global_fill.argtypes = [C.c_void_p, C.c_int32, i32p, C.POINTER(u8p), C.POINTER(MsaFillResult)]
        L.svp_msa_local_fill.argtypes = [C.c_void_p, C.c_int32, i32p, C.POINTER(i32p), i32p, C.POINTER(u8p),
                                         C.POINTER(MsaFillResult)]
        L.svp_final_align_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                             C.c_int32, C.c_void_p]
        L.svp_engine_config.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        L.svp_engine_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.svp_window_capacity.restype = C.c_int32
        L.svp_window_capacity.argtypes = [C.c_int32]
        L.svp_debug_round.argtypes = [C.c_void_p, C.POINTER(WindowIn), C.c_int, f64p, u32p, u8p, C.c_int32, i32p, u8p,
                                      u8p, C.c_int32, i32p, f64p, u32p]
        L.svp_probe_dpx.argtypes = [C.c_void_p, f64p, f64p, f64p]
        L.svp_probe_fp64.argtypes = [C.c_void_p, f64p]
        L.svp_stats.argtypes = [C.c_void_p, i32p, f64p, C.c_int]
        L.svp_set_lanes.argtypes = [C.c_void_p, C.c_int32]
        L.svp_refine_variant.argtypes = [C.c_int32, C.c_char_p, u32p, C.c_int32, C.POINTER(VariantPos), i32p]
        L.svp_refine_info.argtypes = [C.c_char_p, C.POINTER(VariantPos), u8p, C.c_int32, C.c_int32, C.c_char_p, C.c_int32]
        L.svp_format_quality.argtypes = [C.c_double, C.c_char_p, C.c_int32]
        L.svp_record_supports_variant.argtypes = [C.c_int32, C.c_char_p, u32p, C.c_int32, C.POINTER(VariantPos)]
        L.svp_read_region_boundaries.argtypes = [C.c_int32, C.c_char_p, u32p, C.c_int32, C.c_int32, C.c_int32, i32p, i32p]
        L.svp_alignment_length_in_ref.argtypes = [C.c_char_p, u32p, C.c_int32]
        L.svp_cut_down_high_coverage.argtypes = [C.c_int32, i32p, u8p, i32p, C.c_int32, i32p]
        L.svp_k1_cells.argtypes = [C.c_void_p, f64p, C.c_int]
        L.svp_k1_failed_certificates.argtypes = [C.c_void_p, f64p, C.c_int]
        L.svp_merge_record_group.argtypes = [C.c_int32, C.POINTER(BamRecord), i32p, i32p, C.c_char_p, u32p, C.c_int32, i32p,
                                             C.c_char_p, C.c_int32, i32p]
        _lib = L
    return _lib


def _p(a, t=u8p):
    return a.ctypes.data_as(t)


def pack(seqs):
    """list of uint8 arrays -> (concatenated, int32 offsets)"""
    off = np.zeros(len(seqs) + 1, np.int32)
    if len(seqs):
        off[1:] = np.cumsum([len(s) for s in seqs])
    cat = np.concatenate([np.asarray(s, np.uint8) for s in seqs] + [np.zeros(0, np.uint8)])
    return np.ascontiguousarray(cat), off


def baseq_mean(quals1, quals2):
    """compute_baseQ_stats (polishing.h:197-234): host-side scalar prep, done by the caller at sviper.h:426.
    Note the reference divides by n*len(reads1[0]) + n*len(reads2[0])."""
    if not len(quals1):
        return 0.0
    # integer sums are exact in double, so any summation order equals the reference's sequential loop
    avg = float(sum(int(np.sum(q, dtype=np.int64)) for q in quals1) + sum(int(np.sum(q, dtype=np.int64)) for q in quals2))
    den = len(quals1) * len(quals1[0]) + len(quals2) * len(quals2[0])
    return avg / den if den else float("nan")


def assign_quality(n_cigar, flanking_region):
    return lib().svp_assign_quality(n_cigar, flanking_region)


def evaluate_final_mapping(variant, cigar, polished, begin_pos, flanking_region):
    """The tail of polish_variant (sviper.h:468-489): assign_quality + refine_variant on the final record.
    variant: dict(sv_type=SV_DEL|SV_INS, ref_pos, ref_pos_end, sv_length, info, filter); cigar: [(op, count)] as
    returned by Context.polish_batch; polished: uint8 bases; begin_pos = final_record.beginPos (sviper.h:461).
    Returns a new dict with quality, quality_str, filter, ref_pos, ref_pos_end, sv_length, info."""
    L = lib()
    out = dict(variant)
    out["quality"] = L.svp_assign_quality(len(cigar), flanking_region)
    buf = C.create_string_buffer(64)
    L.svp_format_quality(out["quality"], buf, 64)
    out["quality_str"] = buf.value.decode()
    ops = bytes(ord(o) if isinstance(o, str) else int(o) for o, _ in cigar)
    cnt = np.array([c for _, c in cigar], np.uint32)
    vp = VariantPos(variant["sv_type"], variant["ref_pos"], variant["ref_pos_end"], variant["sv_length"])
    ins_begin = C.c_int32(0)
    rc = L.svp_refine_variant(begin_pos, ops, _p(cnt, u32p), len(cigar), C.byref(vp), C.byref(ins_begin))
    if rc < 0:
        raise ValueError("svp_refine_variant: error %d (unknown CIGAR operation or SV type)" % rc)
    if rc == 0:
        out["filter"] = "FAIL5"
        return out
    pol = np.ascontiguousarray(polished, np.uint8)
    info = variant["info"].encode()
    need = L.svp_refine_info(info, C.byref(vp), _p(pol), len(pol), ins_begin.value, None, 0)
    if need < 0:
        raise ValueError("svp_refine_info: error %d" % need)
    ibuf = C.create_string_buffer(need + 1)
    L.svp_refine_info(info, C.byref(vp), _p(pol), len(pol), ins_begin.value, ibuf, need + 1)
    out.update(ref_pos=vp.ref_pos, ref_pos_end=vp.ref_pos_end, sv_length=vp.sv_length, info=ibuf.value.decode())
    return out


def _cigar_arrays(cigar):
    ops = bytes(ord(o) if isinstance(o, str) else int(o) for o, _ in cigar)
    return ops, np.array([c for _, c in cigar], np.uint32)


def record_supports_variant(begin_pos, cigar, sv_type, ref_pos, sv_length):
    """record_supports_variant (variant.h:239-275) for one long-read record; cigar: [(op, count)]"""
    ops, cnt = _cigar_arrays(cigar)
    vp = VariantPos(sv_type, ref_pos, ref_pos + sv_length if sv_type == SV_DEL else ref_pos, sv_length)
    rc = lib().svp_record_supports_variant(begin_pos, ops, _p(cnt, u32p), len(cigar), C.byref(vp))
    if rc < 0:
        raise ValueError("UNKNOWN CIGAR CHARACTER.")
    return bool(rc)


def get_read_region_boundaries(begin_pos, cigar, ref_region_begin, ref_region_end):
    """get_read_region_boundaries (basics.h:189-241): (read_begin, read_end) of a reference region"""
    ops, cnt = _cigar_arrays(cigar)
    b, e = C.c_int32(0), C.c_int32(0)
    rc = lib().svp_read_region_boundaries(begin_pos, ops, _p(cnt, u32p), len(cigar), ref_region_begin, ref_region_end,
                                          C.byref(b), C.byref(e))
    if rc < 0:
        raise ValueError("svp_read_region_boundaries: unknown CIGAR operation, or the region ends before the record")
    return b.value, e.value


def cut_down_high_coverage(records, mean_coverage):
    """cut_down_high_coverage (basics.h:403-442); records: [dict(begin_pos, mapq, cigar)] -> the kept ones, in the
    reference's output order"""
    L = lib()
    n = len(records)
    bp = np.array([r["begin_pos"] for r in records], np.int32)
    mq = np.array([r["mapq"] for r in records], np.uint8)
    ln = np.zeros(n, np.int32)
    for i, r in enumerate(records):
        ops, cnt = _cigar_arrays(r["cigar"])
        ln[i] = L.svp_alignment_length_in_ref(ops, _p(cnt, u32p), len(r["cigar"]))
    keep = np.zeros(max(1, n), np.int32)
    k = L.svp_cut_down_high_coverage(n, _p(bp, i32p), _p(mq), _p(ln, i32p), mean_coverage, _p(keep, i32p))
    if k < 0:
        raise ValueError("svp_cut_down_high_coverage: error %d" % k)
    return [records[i] for i in keep[:k]]


def merge_alignments(records):
    """merge_alignments (merge_split_alignments.h:536-561); records: [dict(name, flag, rid, begin_pos, mapq, cigar, seq)]
    sorted by name -> one merged record per name (the reference's output order)"""
    L = lib()
    out, b = [], 0
    while b < len(records):
        e = b + 1
        while e < len(records) and records[e]["name"] == records[b]["name"]:
            e += 1
        grp = records[b:e]
        keep, arr = [], (BamRecord * len(grp))()
        for i, r in enumerate(grp):
            ops, cnt = _cigar_arrays(r["cigar"])
            sq = r["seq"].encode()
            keep.append((ops, cnt, sq))
            arr[i] = BamRecord(r["begin_pos"], r["rid"], r["flag"], r["mapq"], ops, _p(cnt, u32p), len(cnt), sq, len(sq))
        ccap = sum(len(r["cigar"]) + 2 for r in grp) + 16
        scap = sum(len(r["seq"]) for r in grp) + 16
        oop, ocnt, oseq = C.create_string_buffer(ccap), np.zeros(ccap, np.uint32), C.create_string_buffer(scap)
        chosen, begin, ncig, slen = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        rc = L.svp_merge_record_group(len(grp), arr, C.byref(chosen), C.byref(begin), oop, _p(ocnt, u32p), ccap,
                                      C.byref(ncig), oseq, scap, C.byref(slen))
        if rc != 0:
            raise ValueError("svp_merge_record_group: error %d (the reference indexes out of bounds on this group)" % rc)
        m = dict(grp[chosen.value])
        m.update(begin_pos=begin.value, cigar=[(chr(oop.raw[i]), int(ocnt[i])) for i in range(ncig.value)],
                 seq=oseq.raw[:slen.value].decode())
        out.append(m)
        b = e
    return out


STAT_NAMES = ["k1_score", "k1_trace", "k2_traceback", "k3_pileup_vote", "k5_final", "k5_traceback", "pack", "other",
              "k1_score_mixed"]


class Context:
    def __init__(self, device=0, stream=None):
        self.L = lib()
        self.h = self.L.svp_create(device, stream)
        if not self.h:
            raise RuntimeError("sviper_b200: cannot create a context on CUDA device %d (no GPU?) -- no CPU fallback"
                               % device)
        self._keep = None

    def close(self):
        if self.h:
            self.L.svp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError("%s failed (%d): %s" % (what, rc, self.L.svp_last_error(self.h).decode()))

    # ---- map_single_read ---------------------------------------------------------------------
    def align_batch(self, reads, windows, task_read, task_win, traceback=True, ops_cap=0):
        rb, roff = pack(reads)
        wb, woff = pack(windows)
        tr = np.ascontiguousarray(task_read, np.int32); tw = np.ascontiguousarray(task_win, np.int32)
        n = len(tr)
        ain = AlignIn(len(reads), _p(rb), _p(roff, i32p), len(windows), _p(wb), _p(woff, i32p), n, _p(tr, i32p),
                      _p(tw, i32p))
        o = {k: np.zeros(n, np.int32) for k in ("score", "score_rc", "is_rc", "lead", "endpos", "ops_len")}
        opsbuf = C.create_string_buffer(max(1, n * ops_cap)) if ops_cap else None
        aout = AlignOut(_p(o["score"], i32p), _p(o["score_rc"], i32p), _p(o["is_rc"], i32p), _p(o["lead"], i32p),
                        _p(o["endpos"], i32p), C.cast(opsbuf, C.c_char_p) if opsbuf else None, ops_cap,
                        _p(o["ops_len"], i32p))
        self._keep = (rb, roff, wb, woff, tr, tw)
        self._check(self.L.svp_align_batch(self.h, C.byref(ain), C.byref(aout), 1 if traceback else 0), "svp_align_batch")
        if opsbuf is not None:
            raw = opsbuf.raw
            o["ops"] = [raw[t * ops_cap:t * ops_cap + min(ops_cap, int(o["ops_len"][t]))].decode() for t in range(n)]
        return o

    def align_rerun(self, traceback=False):
        ms = C.c_float(0)
        self._check(self.L.svp_align_rerun(self.h, 1 if traceback else 0, C.byref(ms)), "svp_align_rerun")
        return ms.value

    # ---- polish_to_perfection (+ final alignment) ------------------------------------------------
    @staticmethod
    def make_window(v, keep, with_final=True):
        """v: dict like synth.make_variant / tests.refcases.small_variant"""
        r1, o1 = pack(v["reads1"]); q1, _ = pack(v["quals1"])
        r2, o2 = pack(v["reads2"]); q2, _ = pack(v["quals2"])
        cons = np.ascontiguousarray(v["consensus"], np.uint8)
        bq = v.get("baseQ_mean")
        if bq is None:
            bq = baseq_mean(v["quals1"], v["quals2"])
        refp = np.ascontiguousarray(v["ref_part"], np.uint8) if (with_final and v.get("ref_part") is not None) else None
        keep.append((r1, o1, q1, r2, o2, q2, cons, refp))
        return WindowIn(_p(cons), len(cons), int(v["flank"]), len(v["reads1"]), _p(r1), _p(q1), _p(o1, i32p), _p(r2),
                        _p(q2), _p(o2, i32p), float(bq), float(v["ins_mean"]), float(v["ins_sd"]),
                        float(v.get("min_coverage", 4.0)), _p(refp) if refp is not None else None,
                        len(refp) if refp is not None else 0)

    @staticmethod
    def make_result(cap, keep, cigar_cap=4096):
        pol = np.zeros(cap, np.uint8); cov = np.zeros(3 * cap + 1024, np.uint32)
        cop = C.create_string_buffer(cigar_cap); ccnt = np.zeros(cigar_cap, np.uint32)
        keep.append((pol, cov, cop, ccnt))
        r = Result()
        r.polished = _p(pol); r.polished_cap = cap; r.cov_profile = _p(cov, u32p); r.cov_cap = len(cov)
        r.cigar_op = C.cast(cop, C.c_char_p); r.cigar_cnt = _p(ccnt, u32p); r.cigar_cap = cigar_cap
        return r, pol, cov, cop, ccnt

    def prepare(self, variants, with_final=True):
        """Builds the svp_window_in / svp_result arrays for a batch once (host buffers, as the reference holds its
        reads in memory before calling polish_to_perfection). run_prepared() is then the bare C-ABI call."""
        keep = []
        n = len(variants)
        win = (WindowIn * n)(*[self.make_window(v, keep, with_final) for v in variants])
        parts = [self.make_result(self.L.svp_window_capacity(len(v["consensus"])), keep) for v in variants]
        res = (Result * n)(*[p[0] for p in parts])
        h2d = sum(len(k[0]) + len(k[2]) + len(k[3]) + len(k[5]) + len(k[6]) + (len(k[7]) if k[7] is not None else 0)
                  for k in keep if len(k) == 8)
        return dict(n=n, win=win, res=res, parts=parts, keep=keep, h2d_bytes=int(h2d))

    def run_prepared(self, pb):
        self._check(self.L.svp_polish_batch(self.h, pb["n"], pb["win"], pb["res"]), "svp_polish_batch")

    # ---- asynchronous scheduler (svp_submit_polish / svp_flush / svp_wait_any) -----------------------
    def submit_prepared(self, pb, tag_base=0, flush=True):
        """queues every variant of a prepared batch (tag = tag_base + index) and returns without waiting"""
        for i in range(pb["n"]):
            self._check(self.L.svp_submit_polish(self.h, tag_base + i, C.byref(pb["win"][i]), C.byref(pb["res"][i])),
                        "svp_submit_polish")
        if flush:
            self._check(self.L.svp_flush(self.h), "svp_flush")

    def wait_any(self):
        """blocks for the next finished variant; returns its tag, or None when nothing is outstanding"""
        tag = C.c_uint64(); r = C.POINTER(Result)()
        rc = self.L.svp_wait_any(self.h, C.byref(tag), C.byref(r))
        if rc == 1:
            return None
        self._check(rc, "svp_wait_any")
        return tag.value

    def wait_n(self, n):
        tags = []
        for _ in range(n):
            t = self.wait_any()
            if t is None:
                raise RuntimeError("svp_wait_any: nothing outstanding after %d of %d results" % (len(tags), n))
            tags.append(t)
        return tags

    def engine_config(self, batch_target=0, max_batch=0):
        self._check(self.L.svp_engine_config(self.h, batch_target, max_batch), "svp_engine_config")

    def engine_stats(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self.L.svp_engine_stats(self.h, C.byref(a), C.byref(b)), "svp_engine_stats")
        return dict(batches=a.value, variants=b.value)

    def collect(self, pb, allow_failed=False):
        out = []
        for i in range(pb["n"]):
            r = pb["res"][i]
            _, pol, cov, cop, ccnt = pb["parts"][i]
            if r.status != 0:
                if allow_failed:
                    out.append(dict(status=r.status))
                    continue
                raise RuntimeError("variant %d failed with status %d" % (i, r.status))
            d = dict(polished=pol[:r.polished_len].copy(), rounds=r.rounds, alignment_rounds=r.alignment_rounds, subst=r.substituted_bases,
                     ins=r.inserted_bases, dele=r.deleted_bases, mappQ_mean=r.mappQ_mean,
                     cov=cov[:r.cov_len].copy(), n_cigar=r.n_cigar, final_score=r.final_score)
            if r.n_cigar:
                d["cigar"] = [(cop.raw[k:k + 1].decode(), int(ccnt[k])) for k in range(r.n_cigar)]
            out.append(d)
        return out

    def polish_batch(self, variants, with_final=True):
        pb = self.prepare(variants, with_final)
        self.run_prepared(pb)
        return self.collect(pb)

    def polish_rerun(self):
        self._check(self.L.svp_polish_rerun(self.h), "svp_polish_rerun")

    def final_align_batch(self, pairs, cigar_cap=4096):
        keep = []
        n = len(pairs)
        fin = (FinalIn * n)()
        parts = []
        for i, (a, b) in enumerate(pairs):
            a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
            keep.append((a, b))
            fin[i] = FinalIn(_p(a), len(a), _p(b), len(b))
            parts.append(self.make_result(16, keep, cigar_cap))
        res = (Result * n)(*[p[0] for p in parts])
        self._check(self.L.svp_final_align_batch(self.h, n, fin, res), "svp_final_align_batch")
        out = []
        for i in range(n):
            _, _, _, cop, ccnt = parts[i]
            r = res[i]
            if r.status != 0:
                raise RuntimeError("final alignment %d failed with status %d" % (i, r.status))
            out.append(dict(score=r.final_score, n_cigar=r.n_cigar,
                            cigar=[(cop.raw[k:k + 1].decode(), int(ccnt[k])) for k in range(r.n_cigar)]))
        return out

    def debug_round(self, v, fix_indels):
        keep = []
        w = self.make_window(v, keep, with_final=False)
        cap = 4 * len(v["consensus"]) + 4096
        prof = np.zeros((cap, 6), np.float64); cov = np.zeros(cap, np.uint32); isgap = np.zeros(cap, np.uint8)
        proper = np.zeros(max(1, len(v["reads1"])), np.uint8); seq = np.zeros(cap, np.uint8)
        ncols, olen, mq = C.c_int32(), C.c_int32(), C.c_double()
        cnt = np.zeros(3, np.uint32)
        self._check(self.L.svp_debug_round(self.h, C.byref(w), fix_indels, _p(prof, f64p), _p(cov, u32p), _p(isgap),
                                           cap, C.byref(ncols), _p(proper), _p(seq), cap, C.byref(olen), C.byref(mq),
                                           _p(cnt, u32p)), "svp_debug_round")
        n = ncols.value
        return dict(profile=prof[:n].copy(), cov=cov[:n].copy(), isgap=isgap[:n].copy(),
                    proper=proper[:len(v["reads1"])].copy(), out=seq[:olen.value].copy(), mappQ_mean=mq.value,
                    subst=int(cnt[0]), ins=int(cnt[1]), dele=int(cnt[2]), baseQ_mean=w.baseQ_mean)

    # ---- pairwise fills of build_consensus (svp_msa_*) ---------------------------------------------
    def msa_begin(self, pairs, scores=(5, -3, -1, -3)):
        """pairs: [(h, v)] uint8 arrays (h = columns = str[0], v = rows = str[1])"""
        self._msa_keep = [(np.ascontiguousarray(h, np.uint8), np.ascontiguousarray(v, np.uint8)) for h, v in pairs]
        arr = (MsaPair * len(pairs))(*[MsaPair(_p(h), len(h), _p(v), len(v)) for h, v in self._msa_keep])
        self._check(self.L.svp_msa_begin(self.h, len(pairs), arr, *scores), "svp_msa_begin")

    def _msa_fill(self, local, idx, new_cells):
        n = len(idx)
        ia = np.ascontiguousarray(idx, np.int32)
        sizes = []
        for p in idx:
            h, v = self._msa_keep[p]
            sizes.append(len(h) * ((len(v) + 1) // 2) if local else (len(h) + 1) * (len(v) + 1))
        bufs = [np.zeros(max(1, sz), np.uint8) for sz in sizes]
        tp = (u8p * n)(*[_p(b) for b in bufs])
        res = (MsaFillResult * n)()
        if local:
            cells = [np.ascontiguousarray(np.array(c, np.int32).reshape(-1), np.int32) if len(c) else np.zeros(2, np.int32)
                     for c in (new_cells or [[]] * n)]
            cp = (i32p * n)(*[_p(c, i32p) for c in cells])
            nn = np.array([len(c) for c in (new_cells or [[]] * n)], np.int32)
            rc = self.L.svp_msa_local_fill(self.h, n, _p(ia, i32p), cp, _p(nn, i32p), tp, res)
        else:
            rc = self.L.svp_msa_global_fill(self.h, n, _p(ia, i32p), tp, res)
        self._check(rc, "svp_msa_local_fill" if local else "svp_msa_global_fill")
        return [(b[:sz], dict(score=r.score, row=r.row, col=r.col, dir=r.dir, S=r.S, H=r.H, V=r.V)) for b, sz, r in zip(bufs, sizes, res)]

    def msa_global_fill(self, idx):
        return self._msa_fill(False, idx, None)

    def msa_local_fill(self, idx, new_cells=None):
        """new_cells: per listed pair, [(col, row), ...] to forbid before this round"""
        return self._msa_fill(True, idx, new_cells)

    def probe_dpx(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self._check(self.L.svp_probe_dpx(self.h, C.byref(a), C.byref(b), C.byref(c)), "svp_probe_dpx")
        d = dict(viaddmax=a.value, vimax3=b.value, mix=c.value)
        e = C.c_double()
        if self.L.svp_probe_fp64(self.h, C.byref(e)) == 0:
            d["dadd"] = e.value
        return d

    def k1_cells(self, reset=0, split=False):
        """DP cells computed by the score kernel for polish batches since the last reset (svp_k1_cells);
        split: (regular rounds, mixed rounds)"""
        v = (C.c_double * 2)()
        self._check(self.L.svp_k1_cells(self.h, v, reset), "svp_k1_cells")
        return v[0] + v[1] if not split else (v[0], v[1])

    def k1_failed_certificates(self, reset=0):
        v = C.c_double()
        self._check(self.L.svp_k1_failed_certificates(self.h, C.byref(v), reset), "svp_k1_failed_certificates")
        return v.value

    def set_lanes(self, n):
        """number of concurrent parts a polish batch is split into on this GPU (svp_set_lanes)"""
        self._check(self.L.svp_set_lanes(self.h, int(n)), "svp_set_lanes")

    def stats(self, reset=0):
        n = C.c_int32(); ms = (C.c_double * 9)()
        self.L.svp_stats(self.h, C.byref(n), ms, reset)
        return n.value, dict(zip(STAT_NAMES, list(ms)))
