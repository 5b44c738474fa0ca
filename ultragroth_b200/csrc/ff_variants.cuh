// Measured-and-rejected variants of the Montgomery product (kept with their microbenchmarks and parity tests, used by NO prover kernel):
//   fp_mul_ks ..... one Karatsuba level + separate reduction: 103 instead of 120 wide multiply-adds, 66.4 vs 67.6 G products/s (round 1)
//   fp_mul_dfma ... the m*p half of the reduction on the FP64 pipe: bit-exact, 51.1 vs 67.6 G products/s (round 2; DESIGN.md section 4,
//                   profiles/r02_ncu_dfma_pipes.txt)
#pragma once
#include "ff.cuh"

namespace ug {

// FP64 primitives of the hybrid product (fp_mul_dfma): fused multiply-add rounding TOWARDS ZERO and bit casts
#if defined(__CUDA_ARCH__)
UG_D double fma_rz(double a, double b, double c) { return __fma_rz(a, b, c); }
UG_D double dsub(double a, double b) { return __dsub_rn(a, b); }
UG_D double make_double(uint32_t hi, uint32_t lo) { return __hiloint2double((int)hi, (int)lo); }
UG_D uint64_t double_bits(double d) { return (uint64_t)__double_as_longlong(d); }
#else
}  // namespace ug
#include <cfenv>
#include <cmath>
#include <cstring>
namespace ug {
inline double fma_rz(double a, double b, double c) {  // tests only: the host FPU in round-towards-zero mode for one operation
    const int old = fegetround();
    fesetround(FE_TOWARDZERO);
    volatile double va = a, vb = b, vc = c;
    volatile double r = std::fma(va, vb, vc);
    fesetround(old);
    return r;
}
inline double dsub(double a, double b) { volatile double va = a, vb = b; return va - vb; }
inline double make_double(uint32_t hi, uint32_t lo) { uint64_t u = ((uint64_t)hi << 32) | lo; double d; memcpy(&d, &u, 8); return d; }
inline uint64_t double_bits(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
#endif


// ---- product-scanning variant: one Karatsuba level + separate Montgomery reduction ------------------------------
// fp_mul spends 64 of its 136 wide multiply-adds on x*y.  Splitting both operands in 128-bit halves needs three 4x4-limb
// products (48) and ~100 additions, which run on the ALU pipe that the bucket kernels leave 75 % idle; the 8 reduction rows
// (64 + 8) are unchanged.  Same value as fp_mul for every input with y < p (checked against the reference's field code
// through tests/emul on the CPU and the GPU parity tests).  UG_FP_MUL_KARATSUBA selects it for Fq (curve arithmetic).

// r[0..7] = a[0..3] * b[0..3].  Even-start limb pairs accumulate in ev (positions 0..7), odd-start pairs in od (positions
// 1..7), so that every (mad.lo.cc, madc.hi.cc) pair is an aligned IMAD.WIDE; the carry out of a row lands on a limb that
// holds at most earlier carries, so a plain addc ends the chain.
UG_HD void mul4x4(const uint32_t* a, const uint32_t* b, uint32_t* r) {
    uint32_t ev[8], od[7];
    // row 0 (x = b0): a0,a2 -> ev pairs (0,1),(2,3); a1,a3 -> od pairs (1,2),(3,4)
    mul_wide(a[0], b[0], ev[0], ev[1]);
    mul_wide(a[2], b[0], ev[2], ev[3]);
    mul_wide(a[1], b[0], od[0], od[1]);
    mul_wide(a[3], b[0], od[2], od[3]);
    // row 1 (x = b1): a0,a2 -> od pairs (1,2),(3,4), carry -> position 5; a1,a3 -> ev pairs (2,3),(4,5), carry -> 6
    od[0] = mad_lo_cc(a[0], b[1], od[0]); od[1] = madc_hi_cc(a[0], b[1], od[1]);
    od[2] = madc_lo_cc(a[2], b[1], od[2]); od[3] = madc_hi_cc(a[2], b[1], od[3]);
    od[4] = addc(0, 0);
    ev[2] = mad_lo_cc(a[1], b[1], ev[2]); ev[3] = madc_hi_cc(a[1], b[1], ev[3]);
    ev[4] = madc_lo_cc(a[3], b[1], 0); ev[5] = madc_hi_cc(a[3], b[1], 0);
    ev[6] = addc(0, 0);
    // row 2 (x = b2): a0,a2 -> ev pairs (2,3),(4,5), carry -> 6; a1,a3 -> od pairs (3,4),(5,6), carry -> 7
    ev[2] = mad_lo_cc(a[0], b[2], ev[2]); ev[3] = madc_hi_cc(a[0], b[2], ev[3]);
    ev[4] = madc_lo_cc(a[2], b[2], ev[4]); ev[5] = madc_hi_cc(a[2], b[2], ev[5]);
    ev[6] = addc(ev[6], 0);
    od[2] = mad_lo_cc(a[1], b[2], od[2]); od[3] = madc_hi_cc(a[1], b[2], od[3]);
    od[4] = madc_lo_cc(a[3], b[2], od[4]); od[5] = madc_hi_cc(a[3], b[2], 0);
    od[6] = addc(0, 0);
    // row 3 (x = b3): a0,a2 -> od pairs (3,4),(5,6), carry -> 7; a1,a3 -> ev pairs (4,5),(6,7)
    od[2] = mad_lo_cc(a[0], b[3], od[2]); od[3] = madc_hi_cc(a[0], b[3], od[3]);
    od[4] = madc_lo_cc(a[2], b[3], od[4]); od[5] = madc_hi_cc(a[2], b[3], od[5]);
    od[6] = addc(od[6], 0);
    ev[4] = mad_lo_cc(a[1], b[3], ev[4]); ev[5] = madc_hi_cc(a[1], b[3], ev[5]);
    ev[6] = madc_lo_cc(a[3], b[3], ev[6]); ev[7] = madc_hi_cc(a[3], b[3], 0);
    // r = ev + (od << 32)
    r[0] = ev[0];
    r[1] = add_cc(ev[1], od[0]);
#pragma unroll
    for (int k = 2; k < 7; k++) r[k] = addc_cc(ev[k], od[k - 1]);
    r[7] = addc(ev[7], od[6]);
}

template <class P>
UG_HD Fp<P> fp_mul_ks(const Fp<P>& x, const Fp<P>& y) {
    uint32_t T[16];
    {
        uint32_t zm[9], sa[4], sb[4];
        mul4x4(x.v, y.v, T);            // z0 = x_lo * y_lo
        mul4x4(x.v + 4, y.v + 4, T + 8);  // z2 = x_hi * y_hi
        sa[0] = add_cc(x.v[0], x.v[4]); sa[1] = addc_cc(x.v[1], x.v[5]); sa[2] = addc_cc(x.v[2], x.v[6]); sa[3] = addc_cc(x.v[3], x.v[7]);
        const uint32_t ca = 0u - addc(0, 0);  // all ones when x_lo + x_hi overflowed 128 bits
        sb[0] = add_cc(y.v[0], y.v[4]); sb[1] = addc_cc(y.v[1], y.v[5]); sb[2] = addc_cc(y.v[2], y.v[6]); sb[3] = addc_cc(y.v[3], y.v[7]);
        const uint32_t cb = 0u - addc(0, 0);
        mul4x4(sa, sb, zm);
        // (sa + ca*2^128) * (sb + cb*2^128) = sa*sb + 2^128 * (ca*sb + cb*sa) + 2^256 * (ca & cb)
        zm[4] = add_cc(zm[4], sb[0] & ca); zm[5] = addc_cc(zm[5], sb[1] & ca); zm[6] = addc_cc(zm[6], sb[2] & ca); zm[7] = addc_cc(zm[7], sb[3] & ca);
        zm[8] = addc(ca & cb & 1u, 0);
        zm[4] = add_cc(zm[4], sa[0] & cb); zm[5] = addc_cc(zm[5], sa[1] & cb); zm[6] = addc_cc(zm[6], sa[2] & cb); zm[7] = addc_cc(zm[7], sa[3] & cb);
        zm[8] = addc(zm[8], 0);
        // z1 = zm - z0 - z2  (>= 0, < 2^258)
        zm[0] = sub_cc(zm[0], T[0]);
#pragma unroll
        for (int k = 1; k < 8; k++) zm[k] = subc_cc(zm[k], T[k]);
        zm[8] = subc(zm[8], 0);
        zm[0] = sub_cc(zm[0], T[8]);
#pragma unroll
        for (int k = 1; k < 8; k++) zm[k] = subc_cc(zm[k], T[8 + k]);
        zm[8] = subc(zm[8], 0);
        // T += z1 << 128
        T[4] = add_cc(T[4], zm[0]);
#pragma unroll
        for (int k = 1; k < 9; k++) T[4 + k] = addc_cc(T[4 + k], zm[k]);
        T[13] = addc_cc(T[13], 0); T[14] = addc_cc(T[14], 0); T[15] = addc(T[15], 0);
    }
    return mont_reduce_512<P>(T);
}

// ---- FP64-assisted Montgomery product (sm_100a: B200 has a full-rate FP64 pipe that every kernel here leaves idle) -----------
// fp_mul spends 64 of its 136 wide multiply-adds on m*p, a product by a CONSTANT.  Here that half moves to the FP64 pipe:
//   T  = x*y                      16 limbs, 64 IMAD.WIDE (product scanning, even / odd accumulators)
//   m  = T_lo * (-p^-1) mod 2^256 low half product, 28 IMAD.WIDE + 8 IMAD
//   h  = floor(m*p / 2^256)       on the FP64 pipe: m and 16*p in five 52-bit limbs held as doubles; every limb product
//                                 a*b = H*2^52 + L is split EXACTLY by two round-towards-zero FMAs
//                                     hi = fma_rz(a, b, 2^104)               = 2^104 + H*2^52     (the mantissa IS H)
//                                     lo = fma_rz(a, b, (2^104 + 2^52) - hi) = 2^52 + L           (the mantissa IS L)
//                                 and the raw bit patterns are summed per 52-bit column as 64-bit integers.  Only the columns
//                                 that reach bit 260 of m*16p are formed: 15 full products (i+j >= 4) and 4 high halves (i+j = 3);
//                                 what is dropped is < 2^216 and is accounted for exactly, because the low 256 bits of m*p are
//                                 KNOWN: they are -T_lo mod 2^256 (that is what m is chosen for)
//   r  = T_hi + h + (T_lo != 0), one conditional subtraction -- the same value as fp_mul(x, y) for x, y < p.
// Per product: 92 IMAD.WIDE + 8 IMAD instead of 136 IMAD.WIDE on the multiplier pipe, 54 operations on the FP64 pipe, the rest ALU.

// m[0..7] = low 256 bits of t[0..7] * P::np[0..7] (compile-time constants): 28 wide products and 8 low words
template <class P>
UG_HD void mullo8x8_np(const uint32_t* t, uint32_t* m) {
    uint32_t ev[8], od[8];  // od[k] sits at limb k+1; limbs >= 8 are never formed
#pragma unroll
    for (int k = 0; k < 8; k++) { ev[k] = 0; od[k] = 0; }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t x = t[i];
        // row i: products np[k] * x at limb position i + k for k <= 7 - i.  Two carry chains per row, one over the even k and one over
        // the odd k; each runs over CONSECUTIVE limbs up to limb 7 of its accumulator (even positions -> ev, odd -> od), so the
        // carry of every pair is consumed by the next pair and the one past limb 7 is dropped (arithmetic mod 2^256).
#pragma unroll
        for (int k0 = 0; k0 < 2; k0++) {
            if (i + k0 > 7) continue;
            const bool even = ((i + k0) & 1) == 0;
            uint32_t* A = even ? ev : od;
            bool carry = false;
#pragma unroll
            for (int k = k0; i + k < 8; k += 2) {
                const int pos = i + k;
                const int o = even ? pos : pos - 1;
                if (pos == 7) {  // only the low word of this product is below 2^256 (pos 7 is odd: od[6])
                    A[o] = carry ? madc_lo(P::np(k), x, A[o]) : A[o] + mul_lo(P::np(k), x);
                } else {
                    A[o] = carry ? madc_lo_cc(P::np(k), x, A[o]) : mad_lo_cc(P::np(k), x, A[o]);
                    // the high word ends the chain when the next pair would start at limb 8 or beyond
                    if (pos + 2 < 8) A[o + 1] = madc_hi_cc(P::np(k), x, A[o + 1]);
                    else A[o + 1] = madc_hi(P::np(k), x, A[o + 1]);
                    carry = true;
                }
            }
        }
    }
    m[0] = ev[0];
    m[1] = add_cc(ev[1], od[0]);
#pragma unroll
    for (int k = 2; k < 7; k++) m[k] = addc_cc(ev[k], od[k - 1]);
    m[7] = addc(ev[7], od[6]);
}

template <class P>
UG_HD Fp<P> fp_mul_dfma(const Fp<P>& x, const Fp<P>& y) {
    uint32_t T[16];
    mul8x8(x.v, y.v, T);
    uint32_t m[8];
    mullo8x8_np<P>(T, m);
    // m in five 52-bit limbs as doubles: (2^52 + limb) assembled from bits, then - 2^52 (exact)
    const double two52 = 4503599627370496.0;
    double md[5];
    md[0] = dsub(make_double(0x43300000u | (m[1] & 0xfffffu), m[0]), two52);
    md[1] = dsub(make_double(0x43300000u | (((m[3] << 12) | (m[2] >> 20)) & 0xfffffu), (m[2] << 12) | (m[1] >> 20)), two52);
    md[2] = dsub(make_double(0x43300000u | ((m[4] >> 8) & 0xfffffu), (m[4] << 24) | (m[3] >> 8)), two52);
    md[3] = dsub(make_double(0x43300000u | (((m[6] << 4) | (m[5] >> 28)) & 0xfffffu), (m[5] << 4) | (m[4] >> 28)), two52);
    md[4] = dsub(make_double(0x43300000u | (m[7] >> 16), (m[7] << 16) | (m[6] >> 16)), two52);
    // column sums of raw bit patterns: C[k] collects position 52*k of m * 16p, k = 4..9
    const double c_hi = 20282409603651670423947251286016.0;            // 2^104
    const double c_lo = 20282409603651670423947251286016.0 + two52;    // 2^104 + 2^52
    const uint64_t bits_hi = 0x4670000000000000ull;                    // bit pattern of 2^104
    const uint64_t bits_lo = 0x4330000000000000ull;                    // bit pattern of 2^52
    uint64_t C[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 5; i++) {
#pragma unroll
        for (int j = 0; j < 5; j++) {
            if (i + j < 3) continue;
            const double hi = fma_rz(md[i], P::mod16_52(j), c_hi);
            C[i + j + 1 - 4] += double_bits(hi);
            if (i + j >= 4) {
                const double lo = fma_rz(md[i], P::mod16_52(j), dsub(c_lo, hi));
                C[i + j - 4] += double_bits(lo);
            }
        }
    }
    // strip the constants: column k-4 received n_hi(k) high halves (products with i+j = k-1) and n_lo(k) low halves (i+j = k)
    C[0] -= 4 * bits_hi + 5 * bits_lo;  // k = 4
    C[1] -= 5 * bits_hi + 4 * bits_lo;
    C[2] -= 4 * bits_hi + 3 * bits_lo;
    C[3] -= 3 * bits_hi + 2 * bits_lo;
    C[4] -= 2 * bits_hi + 1 * bits_lo;
    C[5] -= 1 * bits_hi;
    // S = sum C[k] 2^(52 (k+4)) = m*16p - D with 0 <= D < 2^216.  m*16p = h*2^260 + 16*L, L = -T_lo mod 2^256 known.  floor(S / 2^260) is
    // h, or h - 1 when the dropped part borrowed through bit 260, which shows as: bit 259 of S set while bit 255 of L is clear.
    const uint64_t mask52 = 0xfffffffffffffull;
    uint32_t tl_nz = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) tl_nz |= T[k];
    // bit 255 of L = -T_lo mod 2^256: L = 2^256 - T_lo for T_lo != 0
    uint32_t low7 = 0;
#pragma unroll
    for (int k = 0; k < 7; k++) low7 |= T[k];
    const uint32_t neg_top = (low7 ? ~T[7] : 0u - T[7]);  // top limb of -T_lo mod 2^256
    const uint32_t L_top_bit = tl_nz ? (neg_top >> 31) : 0u;
    const uint32_t S_top_bit = (uint32_t)(C[0] >> 51) & 1u;
    const uint64_t fix = (uint64_t)(S_top_bit & (L_top_bit ^ 1u)) + (tl_nz ? 1u : 0u);  // borrow repair + the carry of T_lo + (m*p)_lo
    uint64_t c = (C[0] >> 52) + C[1] + fix;
    uint64_t w0 = c & mask52; c = (c >> 52) + C[2];
    uint64_t w1 = c & mask52; c = (c >> 52) + C[3];
    uint64_t w2 = c & mask52; c = (c >> 52) + C[4];
    uint64_t w3 = c & mask52; c = (c >> 52) + C[5];
    uint64_t w4 = c;
    uint32_t h[8];
    h[0] = (uint32_t)w0;
    h[1] = (uint32_t)(w0 >> 32) | (uint32_t)(w1 << 20);
    h[2] = (uint32_t)(w1 >> 12);
    h[3] = (uint32_t)(w1 >> 44) | (uint32_t)(w2 << 8);
    h[4] = (uint32_t)(w2 >> 24) | (uint32_t)(w3 << 28);
    h[5] = (uint32_t)(w3 >> 4);
    h[6] = (uint32_t)(w3 >> 36) | (uint32_t)(w4 << 16);
    h[7] = (uint32_t)(w4 >> 16);
    Fp<P> r;
    r.v[0] = add_cc(T[8], h[0]);
#pragma unroll
    for (int k = 1; k < 7; k++) r.v[k] = addc_cc(T[8 + k], h[k]);
    r.v[7] = addc(T[15], h[7]);
    fp_final_sub(r);
    return r;
}

}  // namespace ug
