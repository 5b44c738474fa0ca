"""ORACLE (test infrastructure, NOT product code) -- Groth16 / UltraGroth prove + verify.

Step-for-step restatement, on Python integers, of
  Groth16::Prover::prove ............. src/groth16.cpp:48-203
  Groth16::Verifier::verify .......... src/groth16.cpp:314-364
  UltraGroth::derive_challenge ....... src/ultra_groth.cpp:32-58
  UltraGroth::compute_lookup ......... src/ultra_groth.cpp:62-106
  UltraGroth::execute_round .......... src/ultra_groth.cpp:161-184
  UltraGroth::execute_final_round .... src/ultra_groth.cpp:187-399
  UltraGroth::Prover::prove .......... src/ultra_groth.cpp:401-462
  UltraGroth::Verifier::verify ....... src/ultra_groth.cpp:581-648
  Proof::toJson / BuildPublicString .. src/groth16.cpp:217-250, src/ultra_groth.cpp:476-513,
                                       src/prover.cpp:89-117
The blinding scalars (r, s[, r_k]) are explicit arguments: the reference draws 31 random bytes
each (src/groth16.cpp:165-166), i.e. any value < 2^248.
"""
import json

from . import bn254 as bn
from .formats import ZKey, Wtns

R = bn.R


def h_coefficients(zk: ZKey, w):
    """a5..a8 of SURVEY 8(a): SpMV, c=a*b, (ifft, coset shift, fft) x3, a*b-c.  Returns the list the
    reference feeds to the H MSM (canonical integers, src/groth16.cpp:142-154)."""
    n = zk.domainSize
    a = [0] * n
    b = [0] * n
    r2inv = pow(bn.MONT_R, -2, R)
    for (m, c, s, coef) in zk.coefs():
        val = coef * r2inv % R  # the plain coefficient
        if m == 0:
            a[c] = (a[c] + w[s] * val) % R
        else:
            b[c] = (b[c] + w[s] * val) % R
    cc = [x * y % R for x, y in zip(a, b)]
    k = n.bit_length() - 1
    w2n = bn.fr_root(k + 1)
    shift = [1] * n
    for i in range(1, n):
        shift[i] = shift[i - 1] * w2n % R

    def to_coset(v):
        v = bn.ntt(v, inverse=True)
        v = [x * sft % R for x, sft in zip(v, shift)]
        return bn.ntt(v)

    a = to_coset(a)
    b = to_coset(b)
    cc = to_coset(cc)
    return [(x * y - z) % R for x, y, z in zip(a, b, cc)]


def _assemble(zk, pi_a, pib1, pi_b, pi_c, pih, r, s, delta1, delta2, extra_sub=None):
    """src/groth16.cpp:168-200 (and the extra -r_k*round_delta1 of src/ultra_groth.cpp:386-388)."""
    g1, g2 = bn.G1, bn.G2
    pi_a = g1.add_affine(pi_a, zk.alpha1)
    pi_a = g1.add(pi_a, g1.mul(delta1, r))
    pi_b = g2.add_affine(pi_b, zk.beta2)
    pi_b = g2.add(pi_b, g2.mul(delta2, s))
    pib1 = g1.add_affine(pib1, zk.beta1)
    pib1 = g1.add(pib1, g1.mul(delta1, s))
    pi_c = g1.add(pi_c, pih)
    pi_c = g1.add(pi_c, g1.mul_jac(pi_a, s))
    pi_c = g1.add(pi_c, g1.mul_jac(pib1, r))
    rs = (r * s) % R
    pi_c = g1.add(pi_c, g1.neg(g1.mul(delta1, rs)))
    if extra_sub is not None:
        pi_c = g1.add(pi_c, g1.neg(extra_sub))
    return g1.to_affine(pi_a), g2.to_affine(pi_b), g1.to_affine(pi_c)


def groth16_prove(zkey_bytes, wtns_bytes, r, s, want_intermediates=False):
    zk = ZKey(zkey_bytes)
    assert zk.protocol == 1
    wt = Wtns(wtns_bytes)
    assert wt.nVars == zk.nVars and wt.prime == R
    w = wt.signals()
    g1, g2 = bn.G1, bn.G2
    A = zk.g1_section(5)
    B1 = zk.g1_section(6)
    B2 = zk.g2_section(7)
    C = zk.g1_section(8)
    H = zk.g1_section(9)
    pi_a = g1.msm(A, w)
    pib1 = g1.msm(B1, w)
    pi_b = g2.msm(B2, w)
    pi_c = g1.msm(C, w[zk.nPublic + 1:])
    h = h_coefficients(zk, w)
    pih = g1.msm(H, h)
    inter = None
    if want_intermediates:
        inter = dict(msm_a=g1.to_affine(pi_a), msm_b1=g1.to_affine(pib1), msm_b2=g2.to_affine(pi_b),
                     msm_c=g1.to_affine(pi_c), msm_h=g1.to_affine(pih), h=h)
    pA, pB, pC = _assemble(zk, pi_a, pib1, pi_b, pi_c, pih, r, s, zk.delta1, zk.delta2)
    proof = dict(pi_a=pA, pi_b=pB, pi_c=pC, protocol='groth16')
    public = [w[i] for i in range(1, zk.nPublic + 1)]
    if want_intermediates:
        return proof, public, inter
    return proof, public


def derive_challenge(commitment):
    """keccak256(BE32(x) || BE32(y)) read big-endian, reduced mod r (src/ultra_groth.cpp:32-58)."""
    x, y = commitment
    digest = bn.keccak256(x.to_bytes(32, 'big') + y.to_bytes(32, 'big'))
    return int.from_bytes(digest, 'big') % R


def compute_lookup(w, chunks, freqs, wtns_idx, push_idx, rand):
    """src/ultra_groth.cpp:62-106; mutates and returns the witness list."""
    inv2 = [pow((i + rand) % R, -1, R) for i in range(len(freqs))]
    # RawFr::field.mul(frequencies[i], inv) converts the uint32_t frequency to `int` (build/fr.hpp:251 -> RawFr::set(int),
    # build/fr.cpp:209-223): a frequency >= 2^31 is read as the NEGATIVE number f - 2^32 and q is added once
    prod = [(freqs[i] - (1 << 32) if freqs[i] >= (1 << 31) else freqs[i]) * inv2[i] % R for i in range(len(freqs))]
    inv1 = [inv2[ch] for ch in chunks]
    push = [rand] + inv1 + inv2 + prod
    for wi, pi in zip(wtns_idx, push_idx):
        w[wi] = push[pi]
    return w


def ultragroth_prove(zkey_bytes, wtns_bytes, r_k, r, s, want_intermediates=False):
    zk = ZKey(zkey_bytes)
    assert zk.protocol == 1337
    wt = Wtns(wtns_bytes)
    assert wt.nVars == zk.nVars and wt.prime == R
    w = wt.signals()
    g1, g2 = bn.G1, bn.G2
    round_idx = zk.u32_section(10)[:zk.num_indexes_c1]
    final_idx = zk.u32_section(11)[:zk.num_indexes_c2]
    C1 = zk.g1_section(8)
    C2 = zk.g1_section(9)
    # round 1 (src/ultra_groth.cpp:415-425,161-184)
    round_w = [w[i] for i in round_idx]
    msm_round = g1.msm(C1[:len(round_w)], round_w)
    commit = g1.add(msm_round, g1.mul(zk.final_delta1, r_k))
    commit = g1.to_affine(commit)
    rand = derive_challenge(commit)
    w = compute_lookup(w, wt.u32(3), wt.u32(4), wt.u32(5), wt.u32(6), rand)
    final_w = [w[i] for i in final_idx]
    A = zk.g1_section(5)
    B1 = zk.g1_section(6)
    B2 = zk.g2_section(7)
    H = zk.g1_section(12)
    pi_a = g1.msm(A, w)
    pib1 = g1.msm(B1, w)
    pi_b = g2.msm(B2, w)
    pi_c = g1.msm(C2[:len(final_w)], final_w)
    h = h_coefficients(zk, w)
    pih = g1.msm(H, h)
    inter = None
    if want_intermediates:
        inter = dict(msm_round=g1.to_affine(msm_round), msm_a=g1.to_affine(pi_a), msm_b1=g1.to_affine(pib1),
                     msm_b2=g2.to_affine(pi_b), msm_c=g1.to_affine(pi_c), msm_h=g1.to_affine(pih),
                     h=h, rand=rand, witness=list(w))
    pA, pB, pC = _assemble(zk, pi_a, pib1, pi_b, pi_c, pih, r, s, zk.final_delta1, zk.final_delta2,
                           extra_sub=g1.mul(zk.round_delta1, r_k))
    proof = dict(pi_a=pA, pi_b=pB, pi_f=pC, pi_r=commit, protocol='ultragroth')
    public = [w[i] for i in range(1, zk.nPublic + 1) if i != zk.rand_indx]
    if want_intermediates:
        return proof, public, inter
    return proof, public


# -- JSON exactly as nlohmann::json::dump() prints it (sorted keys, no whitespace) ------------


def _g1j(P):
    return [str(P[0]), str(P[1]), "1"]


def _g2j(P):
    return [[str(P[0][0]), str(P[0][1])], [str(P[1][0]), str(P[1][1])], ["1", "0"]]


def proof_to_json(proof) -> str:
    d = {}
    for k, v in proof.items():
        if k == 'protocol':
            d[k] = v
        elif k == 'pi_b':
            d[k] = _g2j(v)
        else:
            d[k] = _g1j(v)
    return json.dumps(d, sort_keys=True, separators=(',', ':'))


def public_to_json(public) -> str:
    if not public:
        return "null"  # nlohmann: dump() of a default-constructed json
    return json.dumps([str(v) for v in public], separators=(',', ':'))


def proof_from_json(s):
    d = json.loads(s)
    out = {}
    for k, v in d.items():
        if k == 'protocol':
            out[k] = v
        elif k == 'pi_b':
            out[k] = ((int(v[0][0]), int(v[0][1])), (int(v[1][0]), int(v[1][1])))
        else:
            out[k] = (int(v[0]), int(v[1]))
    return out


# -- verifiers ---------------------------------------------------------------------------------


def groth16_verify(vk, proof, public) -> bool:
    """vk: dict(alpha1, beta2, gamma2, delta2, IC=[...]) affine ints (src/groth16.cpp:314-364)."""
    g1 = bn.G1
    if len(public) + 1 != len(vk['IC']):
        raise ValueError("len(inputs)+1 != len(vk.IC)")
    for P in (proof['pi_a'], proof['pi_c']):
        if not g1.is_on_curve(P):
            return False
    if not bn.G2.is_on_curve(proof['pi_b']):
        return False
    vkx = g1.to_jac(vk['IC'][0])
    for x, P in zip(public, vk['IC'][1:]):
        vkx = g1.add(vkx, g1.mul(P, x % R))
    vkx = g1.to_affine(vkx)
    neg = lambda P: None if P is None else (P[0], (-P[1]) % bn.Q)
    return bn.pairing_check(
        [proof['pi_a'], neg(vk['alpha1']), neg(vkx), neg(proof['pi_c'])],
        [proof['pi_b'], vk['beta2'], vk['gamma2'], vk['delta2']])


def ultragroth_verify(vk, proof, public) -> bool:
    """vk: dict(alpha1, beta2, gamma2, final_delta2, round_delta2, IC, IC_rand)
    (src/ultra_groth.cpp:581-648)."""
    g1 = bn.G1
    if len(public) + 1 != len(vk['IC']):
        raise ValueError("len(inputs) != len(vk.IC)")
    vkx = None
    for x, P in zip(public, vk['IC'][1:]):
        vkx = g1.add(vkx, g1.mul(P, x % R))
    rand = derive_challenge(proof['pi_r'])
    vkx = g1.add_affine(vkx, vk['IC'][0])
    vkx = g1.add(vkx, g1.mul(vk['IC_rand'], rand))
    vkx = g1.to_affine(vkx)
    neg = lambda P: None if P is None else (P[0], (-P[1]) % bn.Q)
    return bn.pairing_check(
        [proof['pi_a'], neg(vk['alpha1']), neg(vkx), neg(proof['pi_f']), neg(proof['pi_r'])],
        [proof['pi_b'], vk['beta2'], vk['gamma2'], vk['final_delta2'], vk['round_delta2']])


def vk_from_snarkjs_json(text):
    """testdata/verification_key.json field names (src/groth16.cpp:278-296)."""
    d = json.loads(text)
    g1 = lambda v: (int(v[0]), int(v[1]))
    g2 = lambda v: ((int(v[0][0]), int(v[0][1])), (int(v[1][0]), int(v[1][1])))
    return dict(alpha1=g1(d['vk_alpha_1']), beta2=g2(d['vk_beta_2']), gamma2=g2(d['vk_gamma_2']),
                delta2=g2(d['vk_delta_2']), IC=[g1(p) for p in d['IC']])


def vk_from_ultragroth_json(text):
    """UltraGroth verification key field names (src/ultra_groth.cpp:542-563)."""
    d = json.loads(text)
    g1 = lambda v: None if int(v[2]) == 0 else (int(v[0]), int(v[1]))
    g2 = lambda v: ((int(v[0][0]), int(v[0][1])), (int(v[1][0]), int(v[1][1])))
    return dict(alpha1=g1(d['vk_alpha_1']), beta2=g2(d['vk_beta_2']), gamma2=g2(d['vk_gamma_2']),
                final_delta2=g2(d['vk_delta_c2_2']), round_delta2=g2(d['vk_delta_c1_2']),
                IC=[g1(p) for p in d['IC']], IC_rand=g1(d['IC_rand']))
