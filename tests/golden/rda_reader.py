"""Minimal reader for R's RDX2 (.rda, XDR serialisation v2) data files.

Only used by tests/golden/make_fixtures.py, in the build container where
/root/reference exists, to turn the reference's bundled data frames
(data/alarm.rda, learning.test.rda, insurance.rda, gaussian.test.rda) into the
small .npz fixtures committed next to this file. Written from the
serialisation format documented in "R Internals", section 1.8.
"""
import bz2
import gzip
import lzma
import struct

import numpy as np

NILVALUE, REFSXP, GLOBALENV, EMPTYENV, BASEENV = 254, 255, 253, 242, 241
MISSINGARG, BASENAMESPACE, NAMESPACESXP = 251, 247, 249


class _Reader:
    def __init__(self, buf):
        self.b = buf
        self.i = 0
        self.refs = []

    def int(self):
        v = struct.unpack_from(">i", self.b, self.i)[0]
        self.i += 4
        return v

    def bytes(self, n):
        v = self.b[self.i:self.i + n]
        self.i += n
        return v

    def item(self):
        flags = self.int()
        typ = flags & 0xFF
        has_attr = bool(flags & (1 << 9))
        has_tag = bool(flags & (1 << 10))
        if typ == NILVALUE:
            return None
        if typ in (GLOBALENV, EMPTYENV, BASEENV, MISSINGARG, BASENAMESPACE):
            return None
        if typ == REFSXP:
            idx = flags >> 8
            if idx == 0:
                idx = self.int()
            return self.refs[idx - 1]
        if typ == 1:  # SYMSXP
            name = self.item()
            self.refs.append(name)
            return name
        if typ == 2 or typ == 6:  # LISTSXP / LANGSXP -> python list of (tag, value)
            out = []
            while True:
                attr = self.item() if has_attr else None
                tag = self.item() if has_tag else None
                car = self.item()
                out.append((tag, car))
                nxt = self.int()
                ntyp = nxt & 0xFF
                if ntyp == NILVALUE:
                    break
                if ntyp not in (2, 6):
                    raise ValueError("unexpected pairlist tail type %d" % ntyp)
                has_attr = bool(nxt & (1 << 9))
                has_tag = bool(nxt & (1 << 10))
            return out
        if typ == 9:  # CHARSXP
            n = self.int()
            if n == -1:
                return None
            return self.bytes(n).decode("latin-1")
        if typ in (10, 13):  # LGLSXP / INTSXP
            n = self.int()
            val = np.frombuffer(self.bytes(4 * n), dtype=">i4").astype(np.int32)
        elif typ == 14:
            n = self.int()
            val = np.frombuffer(self.bytes(8 * n), dtype=">f8").astype(np.float64)
        elif typ == 16:
            n = self.int()
            val = [self.item() for _ in range(n)]
        elif typ == 19 or typ == 20:
            n = self.int()
            val = [self.item() for _ in range(n)]
        else:
            raise ValueError("unsupported SEXP type %d at offset %d" % (typ, self.i))
        attrs = dict(self.item()) if has_attr else {}
        return {"value": val, "attr": attrs} if attrs else val


def _decompress(raw):
    if raw[:2] == b"\x1f\x8b":
        return gzip.decompress(raw)
    if raw[:3] == b"BZh":
        return bz2.decompress(raw)
    if raw[:6] == b"\xfd7zXZ\x00":
        return lzma.decompress(raw)
    return raw


def read_rda(path):
    """Return {object name: parsed object} for an .rda file."""
    buf = _decompress(open(path, "rb").read())
    if buf[:5] != b"RDX2\n" or buf[5:7] != b"X\n":
        raise ValueError("not an RDX2/XDR file: %r" % buf[:7])
    r = _Reader(buf)
    r.i = 7
    r.int()
    r.int()
    r.int()  # serialisation version, writer version, min reader version
    top = r.item()
    return {tag: val for tag, val in top}


def data_frame_columns(obj):
    """Split a parsed data.frame into (names, columns); factor columns come back
    as (int32 codes 1-based with INT_MIN = NA, [level labels])."""
    names = obj["attr"]["names"]
    cols = []
    for c in obj["value"]:
        if isinstance(c, dict) and "levels" in c["attr"]:
            cols.append((c["value"], list(c["attr"]["levels"])))
        elif isinstance(c, dict):
            cols.append(c["value"])
        else:
            cols.append(c)
    return list(names), cols
