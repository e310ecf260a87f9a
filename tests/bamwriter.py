"""Minimal BAM (BGZF) + BAI writer, test infrastructure for tests/e2e_dataset.py: this image has no samtools and the
BAI builder of the reference's vendored SeqAn crashes (bam_index_bai.h:771 carries a "does not work yet" note), so the
synthetic end-to-end inputs are written straight from Python, following the SAM/BAM specification (sections 4.2 and 5.2).
Records must be given sorted by (ref id, position)."""
import struct
import zlib

CIGAR_OPS = {c: i for i, c in enumerate("MIDNSHP=X")}
SEQ_CODE = {c: i for i, c in enumerate("=ACMGRSVTWYHKDBN")}


def reg2bin(beg, end):
    end -= 1
    if beg >> 14 == end >> 14: return ((1 << 15) - 1) // 7 + (beg >> 14)
    if beg >> 17 == end >> 17: return ((1 << 12) - 1) // 7 + (beg >> 17)
    if beg >> 20 == end >> 20: return ((1 << 9) - 1) // 7 + (beg >> 20)
    if beg >> 23 == end >> 23: return ((1 << 6) - 1) // 7 + (beg >> 23)
    if beg >> 26 == end >> 26: return ((1 << 3) - 1) // 7 + (beg >> 26)
    return 0


def parse_cigar(c):
    out, n = [], 0
    for ch in c:
        if ch.isdigit():
            n = n * 10 + ord(ch) - 48
        else:
            out.append((ch, n)); n = 0
    return out


def encode_record(name, flag, rid, pos, mapq, cigar, rnext, pnext, tlen, seq, qual):
    """pos / pnext 0-based; cigar a string or '*'; qual a Phred+33 string or '*'. Returns (bytes, end position)"""
    ops = parse_cigar(cigar) if cigar != "*" else []
    ref_len = sum(n for o, n in ops if o in "MDN=X")
    end = pos + max(1, ref_len)
    l_seq = len(seq) if seq != "*" else 0
    packed = bytearray((l_seq + 1) // 2)
    for i in range(l_seq):
        packed[i >> 1] |= SEQ_CODE[seq[i]] << (4 if i % 2 == 0 else 0)
    q = bytes([0xFF] * l_seq) if qual == "*" else bytes(ord(c) - 33 for c in qual)
    nm = name.encode() + b"\0"
    body = struct.pack("<iiBBHHHIiii", rid, pos, len(nm), mapq, reg2bin(pos, end), len(ops), flag, l_seq, rnext, pnext,
                       tlen) + nm + b"".join(struct.pack("<I", n << 4 | CIGAR_OPS[o]) for o, n in ops) + bytes(packed) + q
    return struct.pack("<i", len(body)) + body, end


def bgzf_block(data):
    co = zlib.compressobj(6, zlib.DEFLATED, -15)
    comp = co.compress(data) + co.flush()
    bsize = len(comp) + 25
    return (b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", bsize) + comp +
            struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data)))


BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def write_bam(path, refs, records):
    """refs: [(name, length)]; records: tuples for encode_record, sorted by (rid, pos). Writes path and path + '.bai'"""
    text = "@HD\tVN:1.6\tSO:coordinate\n" + "".join("@SQ\tSN:%s\tLN:%d\n" % r for r in refs)
    head = b"BAM\1" + struct.pack("<i", len(text)) + text.encode() + struct.pack("<i", len(refs))
    for nm, ln in refs:
        head += struct.pack("<i", len(nm) + 1) + nm.encode() + b"\0" + struct.pack("<i", ln)
    blocks = [bytearray(head)]          # the header gets its own block
    where = []                          # per record: (block index, offset in block)
    cur = bytearray()
    meta = []
    for r in records:
        b, end = encode_record(*r)
        assert len(b) < 0xFF00
        if len(cur) + len(b) > 0xFF00:
            blocks.append(cur); cur = bytearray()
        where.append((len(blocks), len(cur)))
        cur += b
        meta.append((r[2], r[3], end))
    blocks.append(cur)
    coff, out = [], bytearray()
    for b in blocks:
        coff.append(len(out))
        out += bgzf_block(bytes(b))
    eof_off = len(out)
    out += BGZF_EOF
    with open(path, "wb") as f:
        f.write(out)
    # ---- BAI -------------------------------------------------------------------------------------------------
    voff = [coff[bi] << 16 | o for bi, o in where] + [eof_off << 16]
    # a record that ends its block ends at the start of the next one
    vend = []
    for i, (bi, o) in enumerate(where):
        if i + 1 < len(where) and where[i + 1][0] == bi:
            vend.append(voff[i + 1])
        else:
            vend.append((coff[bi + 1] if bi + 1 < len(coff) else eof_off) << 16)
    bai = bytearray(b"BAI\1" + struct.pack("<i", len(refs)))
    for rid in range(len(refs)):
        bins, lin = {}, {}
        for i, (r, pos, end) in enumerate(meta):
            if r != rid:
                continue
            b = reg2bin(pos, end)
            ch = bins.setdefault(b, [])
            if ch and ch[-1][1] == voff[i]:
                ch[-1][1] = vend[i]
            else:
                ch.append([voff[i], vend[i]])
            for w in range(pos >> 14, ((end - 1) >> 14) + 1):
                lin.setdefault(w, voff[i])
        bai += struct.pack("<i", len(bins))
        for b in sorted(bins):
            bai += struct.pack("<Ii", b, len(bins[b]))
            for beg, e in bins[b]:
                bai += struct.pack("<QQ", beg, e)
        n_intv = max(lin) + 1 if lin else 0
        bai += struct.pack("<i", n_intv)
        last = 0
        for w in range(n_intv):
            last = lin.get(w, last)
            bai += struct.pack("<Q", last)
    bai += struct.pack("<Q", 0)
    with open(path + ".bai", "wb") as f:
        f.write(bai)
