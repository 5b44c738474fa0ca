"""ORACLE (test infrastructure, NOT product code) -- BN254 big-int arithmetic.

Independent CPU restatement of the arithmetic that the reference prover gets from
iden3/ffiasm (pinned fe3772e8e62f2235d308d445d09379a0fca8f5a9, absent from /root/reference)
plus the generated field code that IS present in the reference:

  * Fr / Fq Montgomery semantics ............ build/fr_raw_generic.cpp:5-8,107-148 (R = 2^256,
    4x64-bit limbs, single conditional subtraction), build/fq_raw_generic.cpp:5-8
  * Fq2 = Fq[u]/(u^2+1), element {a,b}=a+b*u  src/groth16.cpp:209-211,264-267
  * G1: y^2 = x^3 + 3, G2: y^2 = x^3 + 3/(9+u); point at infinity is all-zero bytes in zkeys
  * FFT root convention (ffiasm fft.hpp, RECALLED): w_{2^28} = 5^((r-1)/2^28), natural order
    in/out, ifft scaled by 1/n.  Pinned indirectly: oracle prove -> verify on testdata/.
  * pairing check used by the verifier ....... src/groth16.cpp:314-364,522-690 (port of Go
    bn256: optimal ate, loop 6u+2, Frobenius constants at src/groth16.cpp:302-311)
  * keccak256 (Ethereum padding 0x01) ........ src/keccak256.cpp

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.
Everything here is plain Python integers; nothing is shared with the CUDA/C++ product code.
"""

Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583  # base field
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617  # scalar field
MONT_R = 1 << 256
BN_U = 4965661367192848881  # src/groth16.cpp:618

FR_R2 = pow(MONT_R, 2, R)
FQ_R2 = pow(MONT_R, 2, Q)
FR_RINV = pow(MONT_R, -1, R)
FQ_RINV = pow(MONT_R, -1, Q)

# constants exactly as the reference spells them (build/fr_raw_generic.cpp:5-7, fq_raw_generic.cpp:5-7)
REF_FR_Q_LIMBS = [0x43e1f593f0000001, 0x2833e84879b97091, 0xb85045b68181585d, 0x30644e72e131a029]
REF_FR_R2_LIMBS = [0x1bb8e645ae216da7, 0x53fe3ab1e35c59e3, 0x8c49833d53bb8085, 0x0216d0b17f4e44a5]
REF_FR_NP = 0xc2e1f593efffffff
REF_FQ_Q_LIMBS = [0x3c208c16d87cfd47, 0x97816a916871ca8d, 0xb85045b68181585d, 0x30644e72e131a029]
REF_FQ_R2_LIMBS = [0xf32cfc5b538afa89, 0xb5e71911d44501fb, 0x47ab1eff0a417ff6, 0x06d89f71cab8351f]
REF_FQ_NP = 0x87d20782e4866389


def limbs_to_int(limbs):
    return sum(int(l) << (64 * i) for i, l in enumerate(limbs))


def int_to_limbs(x, n=4):
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(n)]


# ----------------------------------------------------------------------------------------------
# Montgomery primitives with the reference's exact semantics (inputs may be >= modulus)
# ----------------------------------------------------------------------------------------------

def mont_mul(a, b, mod):
    """Fr_rawMMul / Fq_rawMMul (build/fr_raw_generic.cpp:107-148): 4-round CIOS, then ONE
    conditional subtraction.  For a*b < mod*2^256 the pre-subtraction value is < 2*mod, so the
    result equals this closed form; tests pin it against the compiled reference code."""
    np_ = (-pow(mod, -1, MONT_R)) % MONT_R
    t = a * b
    m = (t * np_) % MONT_R
    u = (t + m * mod) >> 256
    u &= (1 << 256) - 1  # reference copies 4 limbs (drops any 5th limb)
    if u >= mod:
        u -= mod
    return u


def raw_add(a, b, mod):
    """Fr_rawAdd (build/fr_raw_generic.cpp:11-19)."""
    s = a + b
    carry = s >> 256
    s &= (1 << 256) - 1
    if carry or s >= mod:
        s = (s - mod) & ((1 << 256) - 1)
    return s


def raw_sub(a, b, mod):
    """Fr_rawSub (build/fr_raw_generic.cpp:31-39)."""
    d = a - b
    if d < 0:
        d = (d + (1 << 256) + mod) & ((1 << 256) - 1)
    return d


def raw_neg(a, mod):
    """Fr_rawNeg (build/fr_raw_generic.cpp:68-80)."""
    return (mod - a) if a != 0 else 0


def to_mont(a, mod):
    return mont_mul(a, pow(MONT_R, 2, mod), mod)


def from_mont(a, mod):
    """Fr_rawFromMontgomery (build/fr_raw_generic.cpp:155-190) == MMul(a, 1)."""
    return mont_mul(a, 1, mod)


# ----------------------------------------------------------------------------------------------
# Fq2
# ----------------------------------------------------------------------------------------------

def f2_add(x, y):
    return ((x[0] + y[0]) % Q, (x[1] + y[1]) % Q)


def f2_sub(x, y):
    return ((x[0] - y[0]) % Q, (x[1] - y[1]) % Q)


def f2_neg(x):
    return ((-x[0]) % Q, (-x[1]) % Q)


def f2_mul(x, y):
    a, b = x
    c, d = y
    return ((a * c - b * d) % Q, (a * d + b * c) % Q)


def f2_sqr(x):
    a, b = x
    return ((a * a - b * b) % Q, (2 * a * b) % Q)


def f2_scale(x, k):
    return ((x[0] * k) % Q, (x[1] * k) % Q)


def f2_inv(x):
    a, b = x
    n = pow((a * a + b * b) % Q, -1, Q)
    return ((a * n) % Q, (-b * n) % Q)


def f2_conj(x):
    return (x[0], (-x[1]) % Q)


def f2_pow(x, e):
    r = (1, 0)
    while e:
        if e & 1:
            r = f2_mul(r, x)
        x = f2_sqr(x)
        e >>= 1
    return r


F2_ZERO = (0, 0)
F2_ONE = (1, 0)
XI = (9, 1)
G2_B = f2_mul((3, 0), f2_inv(XI))

# Frobenius constants; asserted equal to the literals at src/groth16.cpp:302-311 in tests
XI_TO_P_MINUS1_OVER3 = f2_pow(XI, (Q - 1) // 3)
XI_TO_P_MINUS1_OVER2 = f2_pow(XI, (Q - 1) // 2)
XI_TO_PSQ_MINUS1_OVER3 = f2_pow(XI, (Q * Q - 1) // 3)

REF_XI_TO_P_MINUS1_OVER3 = (
    21575463638280843010398324269430826099269044274347216827212613867836435027261,
    10307601595873709700152284273816112264069230130616436755625194854815875713954)
REF_XI_TO_P_MINUS1_OVER2 = (
    2821565182194536844548159561693502659359617185244120367078079554186484126554,
    3505843767911556378687030309984248845540243509899259641013678093033130930403)
REF_XI_TO_PSQ_MINUS1_OVER3 = 21888242871839275220042445260109153167277707414472061641714758635765020556616


# ----------------------------------------------------------------------------------------------
# Curve arithmetic, generic over a field-ops table.  Points are Jacobian (X, Y, Z); None = infinity
# ----------------------------------------------------------------------------------------------

class _F1Ops:
    zero = 0
    one = 1

    @staticmethod
    def add(a, b): return (a + b) % Q
    @staticmethod
    def sub(a, b): return (a - b) % Q
    @staticmethod
    def mul(a, b): return (a * b) % Q
    @staticmethod
    def sqr(a): return (a * a) % Q
    @staticmethod
    def neg(a): return (-a) % Q
    @staticmethod
    def inv(a): return pow(a, -1, Q)
    @staticmethod
    def small(a, k): return (a * k) % Q
    @staticmethod
    def is_zero(a): return a == 0


class _F2Ops:
    zero = F2_ZERO
    one = F2_ONE
    add = staticmethod(f2_add)
    sub = staticmethod(f2_sub)
    mul = staticmethod(f2_mul)
    sqr = staticmethod(f2_sqr)
    neg = staticmethod(f2_neg)
    inv = staticmethod(f2_inv)
    small = staticmethod(f2_scale)

    @staticmethod
    def is_zero(a): return a == (0, 0)


class Curve:
    """Short-Weierstrass a=0 curve over F (Jacobian coordinates)."""

    def __init__(self, F, b):
        self.F = F
        self.b = b

    def is_on_curve(self, P):
        if P is None:
            return True
        F = self.F
        x, y = P
        return F.sub(F.sqr(y), F.add(F.mul(F.sqr(x), x), self.b)) == F.zero

    def to_jac(self, P):
        return None if P is None else (P[0], P[1], self.F.one)

    def to_affine(self, P):
        if P is None:
            return None
        F = self.F
        X, Y, Z = P
        zi = F.inv(Z)
        zi2 = F.sqr(zi)
        return (F.mul(X, zi2), F.mul(Y, F.mul(zi2, zi)))

    def neg(self, P):
        return None if P is None else (P[0], self.F.neg(P[1]), P[2])

    def dbl(self, P):
        if P is None:
            return None
        F = self.F
        X, Y, Z = P
        if F.is_zero(Y):
            return None
        A = F.sqr(X)
        B = F.sqr(Y)
        C = F.sqr(B)
        D = F.small(F.sub(F.sub(F.sqr(F.add(X, B)), A), C), 2)
        E = F.small(A, 3)
        Fv = F.sqr(E)
        X3 = F.sub(Fv, F.small(D, 2))
        Y3 = F.sub(F.mul(E, F.sub(D, X3)), F.small(C, 8))
        Z3 = F.small(F.mul(Y, Z), 2)
        return (X3, Y3, Z3)

    def add(self, P, Qp):
        if P is None:
            return Qp
        if Qp is None:
            return P
        F = self.F
        X1, Y1, Z1 = P
        X2, Y2, Z2 = Qp
        Z1Z1 = F.sqr(Z1)
        Z2Z2 = F.sqr(Z2)
        U1 = F.mul(X1, Z2Z2)
        U2 = F.mul(X2, Z1Z1)
        S1 = F.mul(Y1, F.mul(Z2, Z2Z2))
        S2 = F.mul(Y2, F.mul(Z1, Z1Z1))
        if U1 == U2:
            if S1 == S2:
                return self.dbl(P)
            return None
        H = F.sub(U2, U1)
        Rr = F.sub(S2, S1)
        HH = F.sqr(H)
        HHH = F.mul(H, HH)
        V = F.mul(U1, HH)
        X3 = F.sub(F.sub(F.sqr(Rr), HHH), F.small(V, 2))
        Y3 = F.sub(F.mul(Rr, F.sub(V, X3)), F.mul(S1, HHH))
        Z3 = F.mul(F.mul(Z1, Z2), H)
        return (X3, Y3, Z3)

    def add_affine(self, P, A):
        """P (Jacobian) + A (affine tuple or None)."""
        return self.add(P, self.to_jac(A))

    def mul(self, A, k):
        """k * A for an affine point A (or None); returns Jacobian."""
        if A is None or k == 0:
            return None
        if k < 0:
            return self.neg(self.mul(A, -k))
        base = self.to_jac(A)
        acc = None
        for bit in bin(k)[2:]:
            acc = self.dbl(acc)
            if bit == '1':
                acc = self.add(acc, base)
        return acc

    def mul_affine(self, A, k):
        return self.to_affine(self.mul(A, k))

    def msm_naive(self, points, scalars):
        acc = None
        for P, k in zip(points, scalars):
            if P is None or k == 0:
                continue
            acc = self.add(acc, self.mul(P, k))
        return acc

    def msm(self, points, scalars, c=None):
        """Sum k_i * P_i (affine inputs, None = infinity).  Plain unsigned-window Pippenger; the
        result (once affine) is independent of the algorithm, which is what makes it an oracle for
        ffiasm's multiMulByScalarMSM (call sites src/groth16.cpp:55-64,154)."""
        n = len(points)
        if n == 0:
            return None
        if n < 8:
            return self.msm_naive(points, scalars)
        if c is None:
            c = max(2, min(16, n.bit_length() - 2))
        nwin = (256 + c - 1) // c
        mask = (1 << c) - 1
        jac = [self.to_jac(P) for P in points]
        total = None
        for w in reversed(range(nwin)):
            if total is not None:
                for _ in range(c):
                    total = self.dbl(total)
            buckets = {}
            sh = w * c
            for P, k in zip(jac, scalars):
                if P is None:
                    continue
                d = (k >> sh) & mask
                if d:
                    cur = buckets.get(d)
                    buckets[d] = P if cur is None else self.add(cur, P)
            if not buckets:
                continue
            running = None
            acc = None
            prev = None
            for d in sorted(buckets.keys(), reverse=True):
                if prev is not None and running is not None:
                    gap = prev - d
                    # acc += running * gap
                    if gap == 1:
                        acc = self.add(acc, running)
                    else:
                        acc = self.add(acc, self.mul_jac(running, gap))
                running = self.add(running, buckets[d])
                prev = d
            acc = self.add(acc, self.mul_jac(running, prev))
            total = self.add(total, acc)
        return total

    def mul_jac(self, P, k):
        if P is None or k == 0:
            return None
        acc = None
        for bit in bin(k)[2:]:
            acc = self.dbl(acc)
            if bit == '1':
                acc = self.add(acc, P)
        return acc


G1 = Curve(_F1Ops, 3)
G2 = Curve(_F2Ops, G2_B)

G1_GEN = (1, 2)
G2_GEN = (
    (10857046999023057135944570762232829481370756359578518086990519993285655852781,
     11559732032986387107991004021392285783925812861821192530917403151452391805634),
    (8495653923123431417604973247489272438418190587263600148770280649306958101930,
     4082367875863433681332203403145435568316851327593401208105741076214120093531),
)


# ----------------------------------------------------------------------------------------------
# Pairing (optimal ate).  Fq12 = Fq[w]/(w^12 - 18 w^6 + 82); w^6 = 9 + u.
# ----------------------------------------------------------------------------------------------

def _f12_mul(a, b):
    t = [0] * 23
    for i, ai in enumerate(a):
        if ai == 0:
            continue
        for j, bj in enumerate(b):
            if bj:
                t[i + j] += ai * bj
    # reduce: w^12 = 18 w^6 - 82
    for k in range(22, 11, -1):
        v = t[k]
        if v:
            t[k - 6] += 18 * v
            t[k - 12] -= 82 * v
    return [x % Q for x in t[:12]]


def _f12_pow(a, e):
    r = [1] + [0] * 11
    while e:
        if e & 1:
            r = _f12_mul(r, a)
        a = _f12_mul(a, a)
        e >>= 1
    return r


F12_ONE = [1] + [0] * 11


def _embed_f2(x):
    """a + b*u -> (a - 9b) + b*w^6."""
    return ((x[0] - 9 * x[1]) % Q, x[1])


def _line(m, x1, y1, P):
    """Value at P=(xp,yp) in Fq of the line through the twisted point (x1 w^2, y1 w^3) with slope
    (m w):  m*w*(xp - x1 w^2) - (yp - y1 w^3)."""
    xp, yp = P
    me = _embed_f2(m)
    mx = _embed_f2(f2_mul(m, x1))
    ye = _embed_f2(y1)
    out = [0] * 12
    out[0] = (-yp) % Q
    # m*w*xp  -> w^1, w^7
    out[1] = (me[0] * xp) % Q
    out[7] = (me[1] * xp) % Q
    # -(m*x1) w^3 + y1 w^3 -> w^3, w^9
    out[3] = (ye[0] - mx[0]) % Q
    out[9] = (ye[1] - mx[1]) % Q
    return out


def _vertical(x1, P):
    """xp - x1 w^2 (used when the two points are opposite)."""
    xe = _embed_f2(x1)
    out = [0] * 12
    out[0] = P[0] % Q
    out[2] = (-xe[0]) % Q
    out[8] = (-xe[1]) % Q
    return out


def _g2_aff_dbl(T):
    x, y = T
    m = f2_mul(f2_scale(f2_sqr(x), 3), f2_inv(f2_scale(y, 2)))
    x3 = f2_sub(f2_sqr(m), f2_scale(x, 2))
    y3 = f2_sub(f2_mul(m, f2_sub(x, x3)), y)
    return (x3, y3), m


def _g2_aff_add(T, S):
    """Returns (T+S, slope) ; T+S None when opposite."""
    x1, y1 = T
    x2, y2 = S
    if x1 == x2:
        if y1 == y2:
            return _g2_aff_dbl(T)
        return None, None
    m = f2_mul(f2_sub(y2, y1), f2_inv(f2_sub(x2, x1)))
    x3 = f2_sub(f2_sub(f2_sqr(m), x1), x2)
    y3 = f2_sub(f2_mul(m, f2_sub(x1, x3)), y1)
    return (x3, y3), m


ATE_LOOP_COUNT = 6 * BN_U + 2  # 29793968203157093288


def miller_loop(Qa, Pa):
    """Miller loop of the optimal ate pairing, affine G2 point Qa, affine G1 point Pa.
    Same loop count / Frobenius corrections as src/groth16.cpp:522-609."""
    if Qa is None or Pa is None:
        return list(F12_ONE)
    T = Qa
    f = list(F12_ONE)
    for i in range(ATE_LOOP_COUNT.bit_length() - 2, -1, -1):
        T2, m = _g2_aff_dbl(T)
        f = _f12_mul(_f12_mul(f, f), _line(m, T[0], T[1], Pa))
        T = T2
        if (ATE_LOOP_COUNT >> i) & 1:
            T2, m = _g2_aff_add(T, Qa)
            f = _f12_mul(f, _line(m, T[0], T[1], Pa))
            T = T2
    Q1 = (f2_mul(f2_conj(Qa[0]), XI_TO_P_MINUS1_OVER3), f2_mul(f2_conj(Qa[1]), XI_TO_P_MINUS1_OVER2))
    nQ2 = (f2_mul(Qa[0], XI_TO_PSQ_MINUS1_OVER3), Qa[1])
    T2, m = _g2_aff_add(T, Q1)
    f = _f12_mul(f, _line(m, T[0], T[1], Pa))
    T = T2
    T2, m = _g2_aff_add(T, nQ2)
    if m is None:
        f = _f12_mul(f, _vertical(T[0], Pa))
    else:
        f = _f12_mul(f, _line(m, T[0], T[1], Pa))
    return f


FINAL_EXP = (Q ** 12 - 1) // R


def final_exponentiation(f):
    return _f12_pow(f, FINAL_EXP)


def pairing(Qa, Pa):
    return final_exponentiation(miller_loop(Qa, Pa))


def pairing_check(g1s, g2s):
    """prod e(g1_i, g2_i) == 1  (src/groth16.cpp:672-690).  Points affine, None = infinity."""
    acc = list(F12_ONE)
    for P, Qp in zip(g1s, g2s):
        if P is None or Qp is None:
            continue
        acc = _f12_mul(acc, miller_loop(Qp, P))
    return final_exponentiation(acc) == F12_ONE


# ----------------------------------------------------------------------------------------------
# NTT over Fr (ffiasm FFT<Fr> conventions)
# ----------------------------------------------------------------------------------------------

FR_S = 28
FR_NQR = 5
FR_ROOT_2S = pow(FR_NQR, (R - 1) >> FR_S, R)  # primitive 2^28-th root of unity


def fr_root(k, i=1):
    """fft->root(k, i) = w_{2^k}^i (call sites src/groth16.cpp:116,126,136)."""
    w = pow(FR_ROOT_2S, 1 << (FR_S - k), R)
    return pow(w, i, R)


def ntt(a, inverse=False):
    """Natural-order in, natural-order out; inverse includes the 1/n factor.  Canonical ints."""
    n = len(a)
    k = n.bit_length() - 1
    assert 1 << k == n
    a = list(a)
    # bit reversal
    j = 0
    for i in range(1, n):
        bit = n >> 1
        while j & bit:
            j ^= bit
            bit >>= 1
        j |= bit
        if i < j:
            a[i], a[j] = a[j], a[i]
    w_n = fr_root(k)
    if inverse:
        w_n = pow(w_n, -1, R)
    length = 2
    while length <= n:
        wl = pow(w_n, n // length, R)
        half = length >> 1
        tw = [1] * half
        for t in range(1, half):
            tw[t] = tw[t - 1] * wl % R
        for s in range(0, n, length):
            for t in range(half):
                u = a[s + t]
                v = a[s + t + half] * tw[t] % R
                a[s + t] = (u + v) % R
                a[s + t + half] = (u - v) % R
        length <<= 1
    if inverse:
        ninv = pow(n, -1, R)
        a = [x * ninv % R for x in a]
    return a


# ----------------------------------------------------------------------------------------------
# Keccak-256 (src/keccak256.cpp: FIPS202_KECCAK_256, rate 1088, suffix 0x01)
# ----------------------------------------------------------------------------------------------

_KECCAK_RC = []
_KECCAK_ROT = [[0] * 5 for _ in range(5)]


def _keccak_init():
    lfsr = 1
    for _ in range(24):
        rc = 0
        for j in range(7):
            bit = lfsr & 1
            lfsr = ((lfsr << 1) ^ ((lfsr >> 7) * 0x71)) & 0xFF
            if bit:
                rc ^= 1 << ((1 << j) - 1)
        _KECCAK_RC.append(rc)
    x, y = 1, 0
    for t in range(24):
        _KECCAK_ROT[x][y] = ((t + 1) * (t + 2) // 2) % 64
        x, y = y, (2 * x + 3 * y) % 5


_keccak_init()
_M64 = (1 << 64) - 1


def _rol(v, n):
    n %= 64
    return ((v << n) | (v >> (64 - n))) & _M64 if n else v


def _keccak_f(A):
    for rnd in range(24):
        C = [A[x][0] ^ A[x][1] ^ A[x][2] ^ A[x][3] ^ A[x][4] for x in range(5)]
        D = [C[(x - 1) % 5] ^ _rol(C[(x + 1) % 5], 1) for x in range(5)]
        A = [[A[x][y] ^ D[x] for y in range(5)] for x in range(5)]
        B = [[0] * 5 for _ in range(5)]
        for x in range(5):
            for y in range(5):
                B[y][(2 * x + 3 * y) % 5] = _rol(A[x][y], _KECCAK_ROT[x][y])
        A = [[B[x][y] ^ ((~B[(x + 1) % 5][y]) & B[(x + 2) % 5][y]) for y in range(5)] for x in range(5)]
        A[0][0] ^= _KECCAK_RC[rnd]
    return A


def keccak256(data: bytes) -> bytes:
    rate = 136
    msg = bytearray(data)
    msg.append(0x01)
    while len(msg) % rate:
        msg.append(0)
    msg[-1] |= 0x80
    A = [[0] * 5 for _ in range(5)]
    for off in range(0, len(msg), rate):
        blk = msg[off:off + rate]
        for i in range(rate // 8):
            x, y = i % 5, i // 5
            A[x][y] ^= int.from_bytes(blk[8 * i:8 * i + 8], 'little')
        A = _keccak_f(A)
    out = b''
    for i in range(4):
        x, y = i % 5, i // 5
        out += A[x][y].to_bytes(8, 'little')
    return out
