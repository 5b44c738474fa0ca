"""ORACLE (test infrastructure, NOT product code) -- iden3 binfile / zkey / wtns readers + writers.

Restates src/binfile_utils.cpp:32-80 (container), src/zkey_utils.cpp:42-76 (Groth16 header),
src/zkey_utils.cpp:123-163 (UltraGroth header), src/wtns_utils.cpp:13-26 (wtns header) and the
section numbering used by src/prover.cpp:162-178,237-260,283-292.
"""
import struct

from . import bn254 as bn


class BinFile:
    """Container: 4-byte magic, u32 version, u32 nSections, then (u32 id, u64 size, payload)*."""

    def __init__(self, data: bytes, ftype: str, max_version: int):
        if len(data) < 12:
            raise ValueError("File is too short.")
        magic = data[:4].decode('latin1')
        if magic != ftype:
            raise ValueError(f"Invalid file type. It should be {ftype} and it is {magic}")
        self.version, nsec = struct.unpack_from('<II', data, 4)
        if self.version > max_version:
            raise ValueError(f"Invalid version. It should be <={max_version} and it is {self.version}")
        if len(data) < 12 + nsec * 12:
            raise ValueError(f"File is too short to contain {nsec} sections.")
        self.data = data
        self.sections = {}
        pos = 12
        for i in range(nsec):
            sid, ssize = struct.unpack_from('<IQ', data, pos)
            pos += 12
            self.sections.setdefault(sid, []).append((pos, ssize))
            pos += ssize
            if pos > len(data):
                raise ValueError(f"Section #{i} is invalid.")

    def section(self, sid, idx=0) -> memoryview:
        if sid not in self.sections:
            raise KeyError(f"Section does not exist: {sid}")
        off, size = self.sections[sid][idx]
        return memoryview(self.data)[off:off + size]

    def section_size(self, sid, idx=0) -> int:
        return self.sections[sid][idx][1]


def write_binfile(ftype: str, version: int, sections) -> bytes:
    """sections: list of (id, bytes)."""
    out = bytearray()
    out += ftype.encode('latin1')
    out += struct.pack('<II', version, len(sections))
    for sid, payload in sections:
        out += struct.pack('<IQ', sid, len(payload))
        out += payload
    return bytes(out)


def _le(x, n=32):
    return int(x).to_bytes(n, 'little')


def _rd(buf, off, n=32):
    return int.from_bytes(buf[off:off + n], 'little')


# -- point (de)serialisation: zkey stores affine coordinates in Montgomery form; infinity = zeros


def g1_to_zkey(P) -> bytes:
    if P is None:
        return b'\0' * 64
    return _le(P[0] * bn.MONT_R % bn.Q) + _le(P[1] * bn.MONT_R % bn.Q)


def g2_to_zkey(P) -> bytes:
    if P is None:
        return b'\0' * 128
    (xa, xb), (ya, yb) = P
    return b''.join(_le(v * bn.MONT_R % bn.Q) for v in (xa, xb, ya, yb))


def g1_from_zkey(buf, off=0):
    x = _rd(buf, off)
    y = _rd(buf, off + 32)
    if x == 0 and y == 0:
        return None
    return (x * bn.FQ_RINV % bn.Q, y * bn.FQ_RINV % bn.Q)


def g2_from_zkey(buf, off=0):
    v = [_rd(buf, off + 32 * i) for i in range(4)]
    if not any(v):
        return None
    v = [x * bn.FQ_RINV % bn.Q for x in v]
    return ((v[0], v[1]), (v[2], v[3]))


class ZKey:
    """Parsed Groth16 (protocol 1) or UltraGroth (protocol 1337) proving key."""

    def __init__(self, data: bytes):
        f = BinFile(data, 'zkey', 1)
        self.bin = f
        self.protocol = struct.unpack_from('<I', f.section(1), 0)[0]
        if self.protocol not in (1, 1337):
            raise ValueError("unknown protocol")
        h = f.section(2)
        pos = 0
        self.n8q = struct.unpack_from('<I', h, pos)[0]; pos += 4
        self.q = _rd(h, pos, self.n8q); pos += self.n8q
        self.n8r = struct.unpack_from('<I', h, pos)[0]; pos += 4
        self.r = _rd(h, pos, self.n8r); pos += self.n8r
        self.nVars, self.nPublic, self.domainSize = struct.unpack_from('<III', h, pos); pos += 12
        if self.protocol == 1337:
            self.num_indexes_c1, self.num_indexes_c2, self.rand_indx = struct.unpack_from('<III', h, pos)
            pos += 12
        self.alpha1 = g1_from_zkey(h, pos); pos += 64
        self.beta1 = g1_from_zkey(h, pos); pos += 64
        self.beta2 = g2_from_zkey(h, pos); pos += 128
        self.gamma2 = g2_from_zkey(h, pos); pos += 128
        if self.protocol == 1:
            self.delta1 = g1_from_zkey(h, pos); pos += 64
            self.delta2 = g2_from_zkey(h, pos); pos += 128
        else:
            self.round_delta1 = g1_from_zkey(h, pos); pos += 64
            self.round_delta2 = g2_from_zkey(h, pos); pos += 128
            self.final_delta1 = g1_from_zkey(h, pos); pos += 64
            self.final_delta2 = g2_from_zkey(h, pos); pos += 128
        self.nCoefs = f.section_size(4) // (12 + self.n8r)

    # lazily decoded sections -------------------------------------------------------------
    def g1_section(self, sid):
        s = self.bin.section(sid)
        return [g1_from_zkey(s, 64 * i) for i in range(len(s) // 64)]

    def g2_section(self, sid):
        s = self.bin.section(sid)
        return [g2_from_zkey(s, 128 * i) for i in range(len(s) // 128)]

    def u32_section(self, sid):
        s = self.bin.section(sid)
        return list(struct.unpack_from(f'<{len(s) // 4}I', s, 0))

    def coefs(self):
        """[(m, c, s, coef_raw)] -- the stored value is value*R^2 mod r (src/groth16.cpp:38,85-89)."""
        s = self.bin.section(4)
        out = []
        pos = 4  # first u32 is the count (src/groth16.cpp:38)
        for _ in range(self.nCoefs):
            m, c, sig = struct.unpack_from('<III', s, pos)
            out.append((m, c, sig, _rd(s, pos + 12)))
            pos += 44
        return out

    def ic(self):
        return self.g1_section(3)


class Wtns:
    def __init__(self, data: bytes):
        f = BinFile(data, 'wtns', 2)
        self.bin = f
        h = f.section(1)
        self.n8 = struct.unpack_from('<I', h, 0)[0]
        self.prime = _rd(h, 4, self.n8)
        self.nVars = struct.unpack_from('<I', h, 4 + self.n8)[0]

    def signals(self):
        s = self.bin.section(2)
        return [_rd(s, 32 * i) for i in range(len(s) // 32)]

    def u32(self, sid):
        s = self.bin.section(sid)
        return list(struct.unpack_from(f'<{len(s) // 4}I', s, 0))


def write_wtns(signals, extra_sections=()) -> bytes:
    hdr = struct.pack('<I', 32) + _le(bn.R) + struct.pack('<I', len(signals))
    secs = [(1, hdr), (2, b''.join(_le(v) for v in signals))]
    for sid, arr in extra_sections:
        secs.append((sid, struct.pack(f'<{len(arr)}I', *arr)))
    return write_binfile('wtns', 2, secs)
