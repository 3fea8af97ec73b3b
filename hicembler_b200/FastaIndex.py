"""Minimal host-side stand-in for the reference's FastaIndex (bin/FastaIndex/FastaIndex.py:36-360):
contig sizes, iteration order and N50-style statistics -- what the window tables need.  Iteration is
FILE order (the reference iterates a Python-2 dict, whose order is arbitrary)."""
from __future__ import annotations

import os


class FastaIndex(object):
    def __init__(self, handle, verbose=0):
        self.fasta = handle if isinstance(handle, str) else handle.name
        self.faidx = self.fasta + ".fai"
        self.genomeSize = 0
        self.id2stats = {}
        if os.path.isfile(self.faidx) and (not os.path.isfile(self.fasta) or
                                           os.stat(self.fasta).st_mtime <= os.stat(self.faidx).st_mtime):
            self._load_fai()
        else:
            self._generate_index()

    def _load_fai(self):
        # FastaIndex.py:123-139: rid, then integer stats; stats[0] is the sequence length
        for l in open(self.faidx):
            ldata = l.rstrip("\n").split("\t")
            if len(ldata) < 2:
                continue
            stats = [int(x) for x in ldata[1:]]
            self.id2stats[ldata[0]] = stats
            self.genomeSize += stats[0]

    def _generate_index(self):
        name, size = None, 0
        with open(self.fasta) as f:
            for line in f:
                if line.startswith(">"):
                    if name is not None:
                        self.id2stats[name] = [size]
                        self.genomeSize += size
                    name, size = line[1:].split()[0], 0
                else:
                    size += len(line.strip())
        if name is not None:
            self.id2stats[name] = [size]
            self.genomeSize += size

    def __len__(self):
        return len(self.id2stats)

    def __iter__(self):
        for seqid in self.id2stats:
            yield seqid

    _RC = str.maketrans("ACGTacgt", "TGCAtgca")

    def _scan_offsets(self):
        """byte range of every record's sequence lines (one pass over the FASTA, cached)"""
        self._seq_range = {}
        name, begin, pos = None, 0, 0
        with open(self.fasta, "rb") as f:
            for line in f:
                if line.startswith(b">"):
                    if name is not None:
                        self._seq_range[name] = (begin, pos)
                    name, begin = line[1:].split()[0].decode(), pos + len(line)
                pos += len(line)
        if name is not None:
            self._seq_range[name] = (begin, pos)

    def get_sequence(self, contig, reverse=False):
        """Return sequence of given contig, reverse-complemented if ``reverse`` -- FastaIndex.py:204-220
        (A/C/G/T in both cases are complemented, every other symbol is kept)."""
        if contig not in self.id2stats:
            raise KeyError(contig)
        if not hasattr(self, "_seq_range"):
            self._scan_offsets()
        begin, end = self._seq_range[contig]
        with open(self.fasta, "rb") as f:
            f.seek(begin)
            seq = b"".join(f.read(end - begin).split()).decode()
        return seq.translate(self._RC)[::-1] if reverse else seq

    def get_N_and_L(self, genomeFrac=0.5, return_L=False, genomeSize=None):
        """FastaIndex.py:304-325."""
        if not genomeSize:
            genomeSize = self.genomeSize
        totsize, size, i = 0, 0, 0
        for i, x in enumerate(sorted(self.id2stats.values(), reverse=True), 1):
            size = x[0]
            totsize += size
            if totsize >= genomeFrac * genomeSize:
                break
        return (size, i) if return_L else size

    def N50(self):
        return self.get_N_and_L(0.5)

    def N90(self):
        return self.get_N_and_L(0.9)
