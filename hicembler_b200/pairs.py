"""Pair records: the decoded form of the SAM/BAM fields the hot path reads.

One 24-byte little-endian record per alignment record; the fields are the
BAM core fields of ``bin/snap/SNAPLib/Bam.h:91-105`` that ``bam2clusters.py:145-146`` consumes
(RNAME, POS, MAPQ, RNEXT, PNEXT, len(SEQ)) plus FLAG for the ``samtools view -F`` filter
(``bam2clusters.py:108-111``).  ``pos``/``next_pos`` are 0-based as in BAM; SAM text is 1-based.
Optional side arrays (``tlen``, ``ref_span``) carry TLEN and the CIGAR reference span for the
contig-level path (``bam2clusters.py:183-189`` and the samtools region queries of ``:179``).
"""
from __future__ import annotations

import gzip
import io
import os
import struct
import zlib

import numpy as np

PAIR_DTYPE = np.dtype([("refID", "<i4"), ("pos", "<i4"), ("next_refID", "<i4"),
                       ("next_pos", "<i4"), ("l_seq", "<i4"), ("flag", "<u2"),
                       ("mapq", "u1"), ("pad", "u1")])
assert PAIR_DTYPE.itemsize == 24
SKIP_FLAG = 3980  # bam2clusters.py:108


PAIR16_DTYPE = np.dtype([("mid", "<u4"), ("next_mid", "<u4"), ("ref_flag", "<u4"), ("nref_mapq", "<u4")])
PAIR16_MAX_CONTIGS = (1 << 20) - 1


def pack_pairs16(rec):
    """24-byte records -> the 16-byte wire form of include/hicembler_b200.h (hc_pair16_t): read midpoints
    (get_window, bam2clusters.py:123: pos + len(SEQ)//2, len("*") == 1), 20-bit reference ids, FLAG, MAPQ.
    ``rec``: structured PAIR_DTYPE array, or an int32 [n, 6] numpy array / torch tensor (any device)."""
    try:
        import torch
    except ImportError:  # pragma: no cover
        torch = None
    if torch is not None and isinstance(rec, torch.Tensor):
        res = torch.empty((rec.shape[0], 4), dtype=torch.int32, device=rec.device)
        step = 1 << 25                                     # bounded temporaries for 5e8-record inputs
        for s0 in range(0, rec.shape[0], step):
            r = rec[s0:s0 + step].to(torch.int64)
            ref, pos, nref, npos, lseq, fm = (r[:, k] for k in range(6))
            if int(torch.maximum(ref.max(), nref.max())) >= PAIR16_MAX_CONTIGS:
                raise ValueError("hc_pair16_t holds reference ids below 2**20 - 1")
            half = torch.where(lseq > 0, lseq, torch.ones_like(lseq)) // 2
            out = torch.empty((r.shape[0], 4), dtype=torch.int64, device=rec.device)
            out[:, 0] = (pos + half) & 0xffffffff
            out[:, 1] = (npos + half) & 0xffffffff
            out[:, 2] = (ref & 0xfffff) | ((fm & 0xfff) << 20)
            out[:, 3] = (nref & 0xfffff) | (((fm >> 16) & 0xff) << 20)
            # two's complement wrap into int32 storage (torch has no uint32 arithmetic)
            res[s0:s0 + step] = torch.where(out >= 2 ** 31, out - 2 ** 32, out).to(torch.int32)
        return res
    if rec.dtype != PAIR_DTYPE:
        rec = np.ascontiguousarray(rec, dtype=np.int32).view(PAIR_DTYPE).reshape(-1)
    if len(rec) and max(int(rec["refID"].max()), int(rec["next_refID"].max())) >= PAIR16_MAX_CONTIGS:
        raise ValueError("hc_pair16_t holds reference ids below 2**20 - 1")
    half = np.where(rec["l_seq"] > 0, rec["l_seq"], 1).astype(np.int64) // 2
    out = np.empty(len(rec), dtype=PAIR16_DTYPE)
    out["mid"] = ((rec["pos"].astype(np.int64) + half) & 0xffffffff).astype(np.uint32)
    out["next_mid"] = ((rec["next_pos"].astype(np.int64) + half) & 0xffffffff).astype(np.uint32)
    out["ref_flag"] = (rec["refID"].astype(np.int64) & 0xfffff) | ((rec["flag"].astype(np.int64) & 0xfff) << 20)
    out["nref_mapq"] = (rec["next_refID"].astype(np.int64) & 0xfffff) | (rec["mapq"].astype(np.int64) << 20)
    return out


def unpack_pairs16(packed):
    """numpy restatement of hc_unpack_pairs16 (tests): pos = midpoint, l_seq = 0."""
    rec = np.zeros(len(packed), dtype=PAIR_DTYPE)
    r1 = (packed["ref_flag"] & 0xfffff).astype(np.int64)
    r2 = (packed["nref_mapq"] & 0xfffff).astype(np.int64)
    rec["refID"] = np.where(r1 == 0xfffff, -1, r1)
    rec["next_refID"] = np.where(r2 == 0xfffff, -1, r2)
    rec["pos"] = packed["mid"].view(np.int32)
    rec["next_pos"] = packed["next_mid"].view(np.int32)
    rec["flag"] = packed["ref_flag"] >> 20
    rec["mapq"] = (packed["nref_mapq"] >> 20) & 0xff
    return rec


def records_from_columns(refID, pos, next_refID, next_pos, l_seq, flag, mapq):
    rec = np.zeros(len(refID), dtype=PAIR_DTYPE)
    rec["refID"], rec["pos"], rec["next_refID"], rec["next_pos"] = refID, pos, next_refID, next_pos
    rec["l_seq"], rec["flag"], rec["mapq"] = l_seq, flag, mapq
    return rec


def records_to_sam(rec, names, tlen=None, with_header=False, sizes=None):
    """SAM text lines for ``rec`` -- what ``samtools view`` (no filter) would print for a BAM
    holding these records.  SEQ is ``l_seq`` bases ('*' when l_seq == 0), CIGAR ``<l_seq>M``."""
    out = []
    if with_header:
        out.append("@HD\tVN:1.0\tSO:coordinate\n")
        for n, s in zip(names, sizes):
            out.append("@SQ\tSN:%s\tLN:%d\n" % (n, s))
    seq_cache = {}
    for k in range(len(rec)):
        r = rec[k]
        L = int(r["l_seq"])
        if L not in seq_cache:
            seq_cache[L] = ("A" * L) if L else "*"
        rnext = "=" if r["next_refID"] == r["refID"] else names[r["next_refID"]]
        t = int(tlen[k]) if tlen is not None else (
            int(r["next_pos"]) - int(r["pos"]) if r["next_refID"] == r["refID"] else 0)
        out.append("r%d\t%d\t%s\t%d\t%d\t%s\t%s\t%d\t%d\t%s\t*\n" % (
            k, r["flag"], names[r["refID"]], r["pos"] + 1, r["mapq"],
            ("%dM" % L) if L else "*", rnext, r["next_pos"] + 1, t, seq_cache[L]))
    return out


_CIG_REF = {"M": 1, "D": 1, "N": 1, "=": 1, "X": 1}


def sam_to_records(lines, name2id):
    """Parse SAM text into (records, tlen, ref_span).  Records whose RNAME/RNEXT is '*' get -1."""
    import re
    cig = re.compile(r"(\d+)([MIDNSHP=X])")
    rows, tl, span = [], [], []
    for line in lines:
        if line.startswith("@"):
            continue
        f = line.rstrip("\n").split("\t")
        ref = name2id.get(f[2], -1)
        nref = ref if f[6] == "=" else name2id.get(f[6], -1)
        seq = f[9]
        rows.append((ref, int(f[3]) - 1, nref, int(f[7]) - 1, 0 if seq == "*" else len(seq),
                     int(f[1]), int(f[4]), 0))
        tl.append(int(f[8]))
        span.append(sum(int(n) for n, op in cig.findall(f[5]) if op in _CIG_REF) or 1)
    rec = np.array(rows, dtype=PAIR_DTYPE) if rows else np.zeros(0, dtype=PAIR_DTYPE)
    return rec, np.array(tl, dtype=np.int32), np.array(span, dtype=np.int32)


# ---------------------------------------------------------------------------------------
# BGZF / BAM (host I/O; replaces the ``samtools view`` text pipe, bam2clusters.py:108-117)
# layouts: bin/snap/SNAPLib/Bam.h:43-69 (header, reference), :91-105 (alignment core),
# :322-350 (BGZF block header).
# ---------------------------------------------------------------------------------------
_BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def _bgzf_block(data: bytes) -> bytes:
    comp = zlib.compressobj(6, zlib.DEFLATED, -15)
    cdata = comp.compress(data) + comp.flush()
    bsize = len(cdata) + 25
    hdr = struct.pack("<BBBBIBBHBBHH", 0x1F, 0x8B, 8, 4, 0, 0, 0xFF, 6, 66, 67, 2, bsize)
    return hdr + cdata + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))


def write_bam(path, rec, names, sizes, tlen=None):
    """Write a minimal coordinate-ordered BAM holding ``rec`` (SEQ = 'A'*l_seq, CIGAR l_seqM)."""
    body = io.BytesIO()
    text = "@HD\tVN:1.0\tSO:coordinate\n" + "".join(
        "@SQ\tSN:%s\tLN:%d\n" % (n, s) for n, s in zip(names, sizes))
    body.write(b"BAM\x01" + struct.pack("<i", len(text)) + text.encode())
    body.write(struct.pack("<i", len(names)))
    for n, s in zip(names, sizes):
        nb = n.encode() + b"\0"
        body.write(struct.pack("<i", len(nb)) + nb + struct.pack("<i", s))
    for k in range(len(rec)):
        r = rec[k]
        L = int(r["l_seq"])
        name = b"r%d\0" % k
        n_cig = 1 if L else 0
        t = int(tlen[k]) if tlen is not None else 0
        end = int(r["pos"]) + max(L, 1)
        bin_ = _reg2bin(int(r["pos"]), end)
        core = struct.pack("<iiBBHHHiiii", r["refID"], r["pos"], len(name), r["mapq"], bin_,
                           n_cig, r["flag"], L, r["next_refID"], r["next_pos"], t)
        cigar = struct.pack("<I", (L << 4) | 0) if L else b""
        seq = bytes([0x11]) * ((L + 1) // 2)
        qual = b"\xff" * L
        blk = core + name + cigar + seq + qual
        body.write(struct.pack("<i", len(blk)) + blk)
    raw = body.getvalue()
    with open(path, "wb") as out:
        for s in range(0, len(raw), 0xFF00):
            out.write(_bgzf_block(raw[s:s + 0xFF00]))
        out.write(_BGZF_EOF)


def write_bam_fixed(path, rec, names, sizes, level=1):
    """Vectorised BAM writer for synthetic benchmarks: every record has a 2-byte read name, one CIGAR operation
    (l_seq M), l_seq bases and qualities -- all records must share one l_seq, so that a record is a fixed-size
    struct (192 bytes at l_seq = 100) built with numpy; BGZF blocks of 0xFF00 bytes, zlib ``level``."""
    rec = np.ascontiguousarray(rec)
    L = int(rec["l_seq"][0]) if len(rec) else 0
    if len(rec) and not np.all(rec["l_seq"] == L):
        raise ValueError("write_bam_fixed needs one common l_seq")
    nseq, ncig = (L + 1) // 2, (1 if L else 0)
    body = 32 + 2 + 4 * ncig + nseq + L
    dt = np.dtype([("block_size", "<i4"), ("refID", "<i4"), ("pos", "<i4"), ("l_read_name", "u1"), ("mapq", "u1"),
                   ("bin", "<u2"), ("n_cigar", "<u2"), ("flag", "<u2"), ("l_seq", "<i4"), ("next_refID", "<i4"),
                   ("next_pos", "<i4"), ("tlen", "<i4"), ("name", "S2"), ("cigar", "<u4", (ncig,)),
                   ("seq", "u1", (nseq,)), ("qual", "u1", (L,))])
    assert dt.itemsize == 4 + body
    out = np.zeros(len(rec), dtype=dt)
    out["block_size"] = body
    for k in ("refID", "pos", "mapq", "flag", "l_seq", "next_refID", "next_pos"):
        out[k] = rec[k]
    out["l_read_name"] = 2
    out["name"] = b"r"
    out["n_cigar"] = ncig
    if ncig:
        out["cigar"] = (L << 4)
    out["seq"] = 0x11
    out["qual"] = 0xFF
    out["bin"] = 4680                    # (unused by the decoder; samtools recomputes it when indexing)
    text = "@HD\tVN:1.0\tSO:coordinate\n" + "".join("@SQ\tSN:%s\tLN:%d\n" % (n, s) for n, s in zip(names, sizes))
    head = io.BytesIO()
    head.write(b"BAM\x01" + struct.pack("<i", len(text)) + text.encode())
    head.write(struct.pack("<i", len(names)))
    for n, s in zip(names, sizes):
        nb = n.encode() + b"\0"
        head.write(struct.pack("<i", len(nb)) + nb + struct.pack("<i", s))
    raw = memoryview(head.getvalue() + out.tobytes())

    def block(data):
        comp = zlib.compressobj(level, zlib.DEFLATED, -15)
        cdata = comp.compress(data) + comp.flush()
        hdr = struct.pack("<BBBBIBBHBBHH", 0x1F, 0x8B, 8, 4, 0, 0, 0xFF, 6, 66, 67, 2, len(cdata) + 25)
        return hdr + cdata + struct.pack("<II", zlib.crc32(data) & 0xFFFFFFFF, len(data))
    with open(path, "wb") as fh:
        for s in range(0, len(raw), 0xFF00):
            fh.write(block(raw[s:s + 0xFF00]))
        fh.write(_BGZF_EOF)
    return len(raw)


def _reg2bin(beg, end):
    end -= 1
    if beg >> 14 == end >> 14: return ((1 << 15) - 1) // 7 + (beg >> 14)
    if beg >> 17 == end >> 17: return ((1 << 12) - 1) // 7 + (beg >> 17)
    if beg >> 20 == end >> 20: return ((1 << 9) - 1) // 7 + (beg >> 20)
    if beg >> 23 == end >> 23: return ((1 << 6) - 1) // 7 + (beg >> 23)
    if beg >> 26 == end >> 26: return ((1 << 3) - 1) // 7 + (beg >> 26)
    return 0


def bgzf_decompress(path) -> bytes:
    """Inflate every BGZF block of ``path`` (Bam.h:322-350) and return the concatenation."""
    with open(path, "rb") as f:
        buf = f.read()
    out, p = [], 0
    while p < len(buf):
        if buf[p:p + 4] != b"\x1f\x8b\x08\x04":
            raise ValueError("%s: not a BGZF block at offset %d" % (path, p))
        xlen = struct.unpack_from("<H", buf, p + 10)[0]
        q, bsize = p + 12, None
        while q < p + 12 + xlen:
            si1, si2, slen = struct.unpack_from("<BBH", buf, q)
            if si1 == 66 and si2 == 67:
                bsize = struct.unpack_from("<H", buf, q + 4)[0] + 1
            q += 4 + slen
        if bsize is None:
            raise ValueError("%s: BGZF block without BC subfield" % path)
        cdata = buf[p + 12 + xlen:p + bsize - 8]
        data = zlib.decompress(cdata, -15)
        crc, isize = struct.unpack_from("<II", buf, p + bsize - 8)
        if len(data) != isize or (zlib.crc32(data) & 0xFFFFFFFF) != crc:
            raise ValueError("%s: BGZF CRC/size mismatch" % path)
        out.append(data)
        p += bsize
    return b"".join(out)


def read_bam(path):
    """Decode a BAM into (names, sizes, records, tlen, ref_span)."""
    raw = bgzf_decompress(path)
    if raw[:4] != b"BAM\x01":
        raise ValueError("%s: bad BAM magic" % path)
    l_text = struct.unpack_from("<i", raw, 4)[0]
    p = 8 + l_text
    n_ref = struct.unpack_from("<i", raw, p)[0]
    p += 4
    names, sizes = [], []
    for _ in range(n_ref):
        l_name = struct.unpack_from("<i", raw, p)[0]
        names.append(raw[p + 4:p + 4 + l_name - 1].decode())
        sizes.append(struct.unpack_from("<i", raw, p + 4 + l_name)[0])
        p += 8 + l_name
    rows, tl, span = [], [], []
    n = len(raw)
    while p < n:
        bs = struct.unpack_from("<i", raw, p)[0]
        (refID, pos, l_rn, mapq, _bin, n_cig, flag, l_seq, nref, npos, tlen) = \
            struct.unpack_from("<iiBBHHHiiii", raw, p + 4)
        cp = p + 36 + l_rn
        s = 0
        for k in range(n_cig):
            v = struct.unpack_from("<I", raw, cp + 4 * k)[0]
            if (v & 15) in (0, 2, 3, 7, 8):
                s += v >> 4
        rows.append((refID, pos, nref, npos, l_seq, flag, mapq, 0))
        tl.append(tlen)
        span.append(s or 1)
        p += 4 + bs
    rec = np.array(rows, dtype=PAIR_DTYPE) if rows else np.zeros(0, dtype=PAIR_DTYPE)
    return names, sizes, rec, np.array(tl, dtype=np.int32), np.array(span, dtype=np.int32)


def read_bam_header(path):
    """(names, sizes) of a BAM: inflates only as many leading BGZF blocks as the header needs."""
    raw = b""
    with open(path, "rb") as f:
        def more():
            nonlocal raw
            head = f.read(18)
            if len(head) < 18 or head[:4] != b"\x1f\x8b\x08\x04":
                raise ValueError("%s: truncated BAM header" % path)
            xlen = struct.unpack_from("<H", head, 10)[0]
            extra = head[12:] + f.read(max(0, xlen - 6))
            bsize, q = None, 0
            while q + 4 <= len(extra):
                si1, si2, slen = struct.unpack_from("<BBH", extra, q)
                if si1 == 66 and si2 == 67:
                    bsize = struct.unpack_from("<H", extra, q + 4)[0] + 1
                q += 4 + slen
            if bsize is None:
                raise ValueError("%s: BGZF block without BC subfield" % path)
            raw += zlib.decompress(f.read(bsize - 12 - xlen - 8), -15)
            f.read(8)

        def need(n):
            while len(raw) < n:
                more()
        need(12)
        if raw[:4] != b"BAM\x01":
            raise ValueError("%s: bad BAM magic" % path)
        l_text = struct.unpack_from("<i", raw, 4)[0]
        need(12 + l_text)
        n_ref = struct.unpack_from("<i", raw, 8 + l_text)[0]
        p = 12 + l_text
        names, sizes = [], []
        for _ in range(n_ref):
            need(p + 4)
            l_name = struct.unpack_from("<i", raw, p)[0]
            need(p + 8 + l_name)
            names.append(raw[p + 4:p + 4 + l_name - 1].decode())
            sizes.append(struct.unpack_from("<i", raw, p + 4 + l_name)[0])
            p += 8 + l_name
    return names, sizes


def read_bam_native(path, wire=24, threads=0, side_arrays=True, pinned=False, chunk=1 << 22):
    """Decode a BAM with the native multi-threaded decoder (csrc/bam_decode.cu: hc_bam_open / hc_bam_next).
    Returns (names, sizes, records, tlen, ref_span) like read_bam; records are PAIR_DTYPE (wire=24) or
    PAIR16_DTYPE (wire=16: read midpoints, see pack_pairs16).  ``pinned=True`` returns the records as a pinned
    torch int32 tensor [n, wire/4] ready for the H2D copy (pipeline.host_pairs_to_dendrogram)."""
    import ctypes as C
    from . import _lib
    L = _lib.lib()
    names, sizes = read_bam_header(path)
    h, n_ref, ubytes = C.c_void_p(), C.c_int32(), C.c_int64()
    _lib.check(L.hc_bam_open(os.fsencode(path), int(threads), C.byref(h), C.byref(n_ref), C.byref(ubytes)))
    dt = PAIR_DTYPE if wire == 24 else PAIR16_DTYPE
    parts, tls, sps = [], [], []
    try:
        if n_ref.value != len(names):
            raise ValueError("%s: reference count mismatch" % path)
        while True:
            rec = np.empty(chunk, dtype=dt)
            tl = np.empty(chunk, dtype=np.int32) if side_arrays else None
            sp = np.empty(chunk, dtype=np.int32) if side_arrays else None
            n = C.c_int64()
            _lib.check(L.hc_bam_next(h, int(wire), rec.ctypes.data_as(C.c_void_p), chunk,
                                     tl.ctypes.data_as(C.c_void_p) if side_arrays else None,
                                     sp.ctypes.data_as(C.c_void_p) if side_arrays else None, C.byref(n)))
            if n.value == 0:
                break
            parts.append(rec[:n.value])
            if side_arrays:
                tls.append(tl[:n.value])
                sps.append(sp[:n.value])
    finally:
        L.hc_bam_close(h)
    rec = np.concatenate(parts) if parts else np.zeros(0, dtype=dt)
    tlen = (np.concatenate(tls) if tls else np.zeros(0, np.int32)) if side_arrays else None
    span = (np.concatenate(sps) if sps else np.zeros(0, np.int32)) if side_arrays else None
    if pinned:
        import torch
        t = torch.empty((len(rec), wire // 4), dtype=torch.int32, pin_memory=torch.cuda.is_available())
        t.numpy().view(dt).reshape(-1)[:] = rec
        rec = t
    return names, sizes, rec, tlen, span


def load_alignments(src, name2id=None):
    """Resolve one entry of the reference's ``bam`` list to (records, tlen, ref_span, names).

    ``src`` may be a BAM path, a SAM text path (.sam / .sam.gz), a numpy PAIR_DTYPE array, or a
    dict with keys records[, tlen, ref_span, names]."""
    if isinstance(src, dict):
        return src["records"], src.get("tlen"), src.get("ref_span"), src.get("names")
    if isinstance(src, np.ndarray):
        return src, None, None, None
    if isinstance(src, str):
        if src.endswith((".sam", ".sam.gz")):
            op = gzip.open if src.endswith(".gz") else open
            with op(src, "rt") as f:
                lines = f.readlines()
            names = [l.split("\t")[1][3:] for l in lines if l.startswith("@SQ")]
            n2i = name2id or {n: i for i, n in enumerate(names)}
            rec, tl, sp = sam_to_records(lines, n2i)
            return rec, tl, sp, names or None
        names, sizes, rec, tl, sp = read_bam_native(src)   # native zlib workers; read_bam is the restatement
        return rec, tl, sp, names
    raise TypeError("unsupported alignment source: %r" % type(src))
