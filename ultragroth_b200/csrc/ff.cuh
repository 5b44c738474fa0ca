// BN254 Fr / Fq Montgomery arithmetic on 8 x 32-bit limbs for sm_100a.
//
// Replaces the reference's generated field code on the prover path:
//   Fr_rawMMul / Fq_rawMMul ......... build/fr.asm:372-539, build/fr_raw_generic.cpp:107-148
//   Fr_rawAdd / Sub / Neg ........... build/fr.asm:1452,2061,2233, build/fr_raw_generic.cpp:11-39,68-80
//   Fr_rawTo/FromMontgomery ......... build/fr.asm:831,957, build/fr_raw_generic.cpp:150-190
// Semantics kept bit-exact: R = 2^256, little-endian limbs, ONE conditional subtraction at the end
// of a multiplication, fully reduced results whenever the multiplicand (2nd operand) is < modulus.
//
// The multiplication is a CIOS whose partial products are split into an "even" and an "odd"
// accumulator so that every (mad.lo.cc, madc.hi.cc) pair lands on an aligned 64-bit register pair;
// ptxas turns each pair into one IMAD.WIDE.U32(.X) carrying through a predicate, i.e. 136
// wide multiply-adds per modular multiplication (64 for x*y, 64 for m*p, 8 for m).
//
// Every carry primitive also has a host definition (a software carry flag) so that tests can run
// the very same limb algorithms on the CPU build box, which has no GPU; the product library only
// ever runs them on the device.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define UG_HD __host__ __device__ __forceinline__
#define UG_D __device__ __forceinline__
#else
#define UG_HD inline
#define UG_D inline
#endif

namespace ug {

// ------------------------------------------------------------------------------------------------
// carry-chain primitives
// ------------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
UG_D uint32_t add_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t addc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t addc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t sub_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t subc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t subc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t mul_lo(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
UG_D uint32_t mul_hi(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
// full 32x32 -> 64 product in ONE IMAD.WIDE (mul.lo + mul.hi would be an IMAD and an IMAD.HI)
UG_D void mul_wide(uint32_t a, uint32_t b, uint32_t& lo, uint32_t& hi) {
    asm volatile("{\n\t.reg .u64 t;\n\tmul.wide.u32 t, %2, %3;\n\tmov.b64 {%0, %1}, t;\n\t}" : "=r"(lo), "=r"(hi) : "r"(a), "r"(b));
}
UG_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
UG_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
UG_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
UG_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
UG_D uint32_t madc_lo(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
#else
// host emulation of the PTX condition-code register (tests only)
static thread_local uint32_t g_cc = 0;
inline uint32_t add_cc(uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a + b; g_cc = (uint32_t)(s >> 32); return (uint32_t)s; }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a + b + g_cc; g_cc = (uint32_t)(s >> 32); return (uint32_t)s; }
inline uint32_t addc(uint32_t a, uint32_t b) { return a + b + g_cc; }
inline uint32_t sub_cc(uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a - b; g_cc = (uint32_t)(s >> 63); return (uint32_t)s; }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { uint64_t s = (uint64_t)a - b - g_cc; g_cc = (uint32_t)(s >> 63); return (uint32_t)s; }
inline uint32_t subc(uint32_t a, uint32_t b) { return a - b - g_cc; }
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
inline void mul_wide(uint32_t a, uint32_t b, uint32_t& lo, uint32_t& hi) { uint64_t p = (uint64_t)a * b; lo = (uint32_t)p; hi = (uint32_t)(p >> 32); }
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return add_cc(a * b, c); }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { return addc_cc(a * b, c); }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { return addc_cc(mul_hi(a, b), c); }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return addc(mul_hi(a, b), c); }
inline uint32_t madc_lo(uint32_t a, uint32_t b, uint32_t c) { return addc(a * b, c); }
#endif

// ------------------------------------------------------------------------------------------------
// field parameters (values: build/fr_raw_generic.cpp:5-7, build/fq_raw_generic.cpp:5-7)
// ------------------------------------------------------------------------------------------------
struct FrParams {
    static constexpr uint32_t NP0 = 0xefffffffu;  // low word of Fr_np = 0xc2e1f593efffffff
    UG_HD static constexpr uint32_t np(int i) {  // -r^-1 mod 2^256
        constexpr uint32_t m[8] = {0xefffffffu, 0xc2e1f593u, 0x4c6911b3u, 0x6586864bu, 0x99062391u, 0xe39a9828u, 0x0d8341b2u, 0x73f82f1du};
        return m[i];
    }
    UG_HD static constexpr double mod16_52(int i) {  // 16*r in 52-bit limbs (exact integers as doubles)
        constexpr double m[5] = {4320251768471568.0, 2385637178164193.0, 2275464320335492.0, 45871597467265.0, 851317936231194.0};
        return m[i];
    }
    UG_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u,
                                   0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    UG_HD static constexpr uint32_t r2(int i) {  // Fr_rawR2
        constexpr uint32_t m[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u,
                                   0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return m[i];
    }
    UG_HD static constexpr uint32_t one(int i) {  // 2^256 mod r
        constexpr uint32_t m[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u,
                                   0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
};

struct FqParams {
    static constexpr uint32_t NP0 = 0xe4866389u;  // low word of Fq_np = 0x87d20782e4866389
    UG_HD static constexpr uint32_t np(int i) {  // -q^-1 mod 2^256
        constexpr uint32_t m[8] = {0xe4866389u, 0x87d20782u, 0x1eca6ac9u, 0x9ede7d65u, 0x1833da80u, 0xd8afcbd0u, 0x91888c6bu, 0xf57a22b7u};
        return m[i];
    }
    UG_HD static constexpr double mod16_52(int i) {  // 16*q in 52-bit limbs (exact integers as doubles)
        constexpr double m[5] = {2464475987825776.0, 396312915557408.0, 2275464437044905.0, 45871597467265.0, 851317936231194.0};
        return m[i];
    }
    UG_HD static constexpr uint32_t mod(int i) {
        constexpr uint32_t m[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u,
                                   0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return m[i];
    }
    UG_HD static constexpr uint32_t r2(int i) {  // Fq_rawR2
        constexpr uint32_t m[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u,
                                   0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return m[i];
    }
    UG_HD static constexpr uint32_t one(int i) {  // 2^256 mod q
        constexpr uint32_t m[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u,
                                   0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return m[i];
    }
};

// ------------------------------------------------------------------------------------------------
// field element
// ------------------------------------------------------------------------------------------------
template <class P>
struct alignas(16) Fp {
    uint32_t v[8];
    using Params = P;

    UG_HD static Fp zero() { Fp r; for (int i = 0; i < 8; i++) r.v[i] = 0; return r; }
    UG_HD static Fp one() { Fp r; for (int i = 0; i < 8; i++) r.v[i] = P::one(i); return r; }
    UG_HD static Fp rsquared() { Fp r; for (int i = 0; i < 8; i++) r.v[i] = P::r2(i); return r; }
    UG_HD static Fp modulus() { Fp r; for (int i = 0; i < 8; i++) r.v[i] = P::mod(i); return r; }

    UG_HD bool is_zero() const {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o |= v[i];
        return o == 0;
    }
    UG_HD bool operator==(const Fp& b) const {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o |= v[i] ^ b.v[i];
        return o == 0;
    }
    UG_HD bool operator!=(const Fp& b) const { return !(*this == b); }
};

// r = (a >= p) ? a - p : a      (a < 2p)
template <class P>
UG_HD void fp_final_sub(Fp<P>& a) {
    uint32_t t[8];
    t[0] = sub_cc(a.v[0], P::mod(0));
#pragma unroll
    for (int i = 1; i < 8; i++) t[i] = subc_cc(a.v[i], P::mod(i));
    uint32_t borrow = subc(0, 0);  // 0xffffffff if a < p
#pragma unroll
    for (int i = 0; i < 8; i++) a.v[i] = borrow ? a.v[i] : t[i];
}

// Fr_rawAdd: add, subtract p if (carry or >= p).  Inputs < 2^256, canonical when both < p.
template <class P>
UG_HD Fp<P> fp_add(const Fp<P>& a, const Fp<P>& b) {
    Fp<P> r;
    r.v[0] = add_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < 8; i++) r.v[i] = addc_cc(a.v[i], b.v[i]);
    uint32_t carry = addc(0, 0);
    uint32_t t[8];
    t[0] = sub_cc(r.v[0], P::mod(0));
#pragma unroll
    for (int i = 1; i < 8; i++) t[i] = subc_cc(r.v[i], P::mod(i));
    uint32_t borrow = subc(0, 0);
    bool keep = (borrow != 0) && (carry == 0);
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = keep ? r.v[i] : t[i];
    return r;
}

// Fr_rawSub: subtract, add p on borrow.
template <class P>
UG_HD Fp<P> fp_sub(const Fp<P>& a, const Fp<P>& b) {
    Fp<P> r;
    r.v[0] = sub_cc(a.v[0], b.v[0]);
#pragma unroll
    for (int i = 1; i < 8; i++) r.v[i] = subc_cc(a.v[i], b.v[i]);
    uint32_t borrow = subc(0, 0);  // all ones on borrow
    r.v[0] = add_cc(r.v[0], P::mod(0) & borrow);
#pragma unroll
    for (int i = 1; i < 7; i++) r.v[i] = addc_cc(r.v[i], P::mod(i) & borrow);
    r.v[7] = addc(r.v[7], P::mod(7) & borrow);
    return r;
}

// Fr_rawNeg: p - a, with -0 = 0.
template <class P>
UG_HD Fp<P> fp_neg(const Fp<P>& a) {
    Fp<P> r;
    uint32_t nz = a.is_zero() ? 0u : 0xffffffffu;
    r.v[0] = sub_cc(P::mod(0) & nz, a.v[0]);
#pragma unroll
    for (int i = 1; i < 7; i++) r.v[i] = subc_cc(P::mod(i) & nz, a.v[i]);
    r.v[7] = subc(P::mod(7) & nz, a.v[7]);
    return r;
}

template <class P>
UG_HD Fp<P> fp_dbl(const Fp<P>& a) { return fp_add(a, a); }

// ---- Montgomery multiplication ---------------------------------------------------------------
// acc0 is aligned at column 0, acc1 at column 1.  x is the scanned operand (any 256-bit value),
// y the multiplicand (< p).  After each row the running value T = acc0 + acc1*2^32 is < 2^32 * 2p.

template <class P>
UG_HD void mont_reduce_row(uint32_t* a0, uint32_t* a1) {
    uint32_t m = mul_lo(a0[0], P::NP0);
    a1[0] = mad_lo_cc(P::mod(1), m, a1[0]);
    a1[1] = madc_hi_cc(P::mod(1), m, a1[1]);
    a1[2] = madc_lo_cc(P::mod(3), m, a1[2]);
    a1[3] = madc_hi_cc(P::mod(3), m, a1[3]);
    a1[4] = madc_lo_cc(P::mod(5), m, a1[4]);
    a1[5] = madc_hi_cc(P::mod(5), m, a1[5]);
    a1[6] = madc_lo_cc(P::mod(7), m, a1[6]);
    a1[7] = madc_hi_cc(P::mod(7), m, a1[7]);  // carry out is impossible (value < 2^32 * 2p); .cc only so that ptxas fuses the pair
    a0[0] = mad_lo_cc(P::mod(0), m, a0[0]);
    a0[1] = madc_hi_cc(P::mod(0), m, a0[1]);
    a0[2] = madc_lo_cc(P::mod(2), m, a0[2]);
    a0[3] = madc_hi_cc(P::mod(2), m, a0[3]);
    a0[4] = madc_lo_cc(P::mod(4), m, a0[4]);
    a0[5] = madc_hi_cc(P::mod(4), m, a0[5]);
    a0[6] = madc_lo_cc(P::mod(6), m, a0[6]);
    a0[7] = madc_hi_cc(P::mod(6), m, a0[7]);
    a1[7] = addc(a1[7], 0);
}

// prev0 (column-0 aligned, prev0[0] == 0 after the previous reduction) and prev1 (column 1) hold
// T*2^32; the shift by one word swaps the roles: prev1 becomes the column-0 accumulator and
// prev0[2..7] the column-1 accumulator, rebuilt in place inside prev0.
template <class P>
UG_HD void mont_mul_row(uint32_t* prev0, uint32_t* prev1, const uint32_t* y, uint32_t x) {
    prev1[0] = add_cc(prev1[0], prev0[1]);
    prev0[0] = madc_lo_cc(y[1], x, prev0[2]);
    prev0[1] = madc_hi_cc(y[1], x, prev0[3]);
    prev0[2] = madc_lo_cc(y[3], x, prev0[4]);
    prev0[3] = madc_hi_cc(y[3], x, prev0[5]);
    prev0[4] = madc_lo_cc(y[5], x, prev0[6]);
    prev0[5] = madc_hi_cc(y[5], x, prev0[7]);
    prev0[6] = madc_lo_cc(y[7], x, 0);
    prev0[7] = madc_hi_cc(y[7], x, 0);
    prev1[0] = mad_lo_cc(y[0], x, prev1[0]);
    prev1[1] = madc_hi_cc(y[0], x, prev1[1]);
    prev1[2] = madc_lo_cc(y[2], x, prev1[2]);
    prev1[3] = madc_hi_cc(y[2], x, prev1[3]);
    prev1[4] = madc_lo_cc(y[4], x, prev1[4]);
    prev1[5] = madc_hi_cc(y[4], x, prev1[5]);
    prev1[6] = madc_lo_cc(y[6], x, prev1[6]);
    prev1[7] = madc_hi_cc(y[6], x, prev1[7]);
    prev0[7] = addc(prev0[7], 0);
    mont_reduce_row<P>(prev1, prev0);
}

// Fr_rawMMul(x, y): x*y*2^-256 mod p.
template <class P>
UG_HD Fp<P> fp_mul(const Fp<P>& x, const Fp<P>& y) {
    uint32_t e[8], o[8];
    // row 0
    {
        uint32_t x0 = x.v[0];
#pragma unroll
        for (int k = 0; k < 8; k += 2) {
            mul_wide(y.v[k], x0, e[k], e[k + 1]);
            mul_wide(y.v[k + 1], x0, o[k], o[k + 1]);
        }
        mont_reduce_row<P>(e, o);
    }
#pragma unroll
    for (int i = 1; i < 8; i += 2) {
        mont_mul_row<P>(e, o, y.v, x.v[i]);      // now o is column 0, e column 1
        if (i + 1 < 8) mont_mul_row<P>(o, e, y.v, x.v[i + 1]);  // back: e column 0, o column 1
    }
    // after 7 more rows (odd count) the column-0 accumulator is o (o[0]==0), column-1 is e
    Fp<P> r;
    r.v[0] = add_cc(e[0], o[1]);
#pragma unroll
    for (int k = 1; k < 7; k++) r.v[k] = addc_cc(e[k], o[k + 1]);
    r.v[7] = addc(e[7], 0);
    fp_final_sub(r);
    return r;
}

// ---- separate (product-scanning) forms: a full 512-bit product and a stand-alone Montgomery reduction ----------------------
// Used where reductions can be SHARED between products (fq2_mul_lazy, curve.cuh); the measured-and-rejected whole-product variants
// built from the same parts (Karatsuba fp_mul_ks, FP64-assisted fp_mul_dfma) live in ff_variants.cuh.

// one fused "shift by a limb, bring in the next limb of T, reduce" row of the separate reduction: prev0 is the old column-0
// accumulator (prev0[0] == 0), prev1 the old column-1 accumulator; afterwards prev1 is column 0 (prev1[0] == 0), prev0 column 1
template <class P>
UG_HD void mont_sos_row(uint32_t* prev0, uint32_t* prev1, uint32_t tnext) {
    prev1[0] = add_cc(prev1[0], prev0[1]);
    const uint32_t m = mul_lo(prev1[0], P::NP0);
    prev0[0] = madc_lo_cc(P::mod(1), m, prev0[2]);
    prev0[1] = madc_hi_cc(P::mod(1), m, prev0[3]);
    prev0[2] = madc_lo_cc(P::mod(3), m, prev0[4]);
    prev0[3] = madc_hi_cc(P::mod(3), m, prev0[5]);
    prev0[4] = madc_lo_cc(P::mod(5), m, prev0[6]);
    prev0[5] = madc_hi_cc(P::mod(5), m, prev0[7]);
    prev0[6] = madc_lo_cc(P::mod(7), m, tnext);
    prev0[7] = madc_hi_cc(P::mod(7), m, 0);
    prev1[0] = mad_lo_cc(P::mod(0), m, prev1[0]);
    prev1[1] = madc_hi_cc(P::mod(0), m, prev1[1]);
    prev1[2] = madc_lo_cc(P::mod(2), m, prev1[2]);
    prev1[3] = madc_hi_cc(P::mod(2), m, prev1[3]);
    prev1[4] = madc_lo_cc(P::mod(4), m, prev1[4]);
    prev1[5] = madc_hi_cc(P::mod(4), m, prev1[5]);
    prev1[6] = madc_lo_cc(P::mod(6), m, prev1[6]);
    prev1[7] = madc_hi_cc(P::mod(6), m, prev1[7]);
    prev0[7] = addc(prev0[7], 0);
}

// T (16 limbs, < p * 2^256) -> T * 2^-256 mod p, fully reduced: eight fused "shift, bring in the next limb, reduce" rows.
// The 9-limb window starts as T[0..7]; after every row it moves up one limb and the next limb of T enters at window position 7
// (T[15] is added with the final fold).  64 + 8 multiplier operations.
template <class P>
UG_HD Fp<P> mont_reduce_512(const uint32_t* T) {
    uint32_t e[8], o[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { e[k] = T[k]; o[k] = 0; }
    mont_reduce_row<P>(e, o);
    mont_sos_row<P>(e, o, T[8]);   // o column 0, e column 1
    mont_sos_row<P>(o, e, T[9]);   // e column 0, o column 1
    mont_sos_row<P>(e, o, T[10]);
    mont_sos_row<P>(o, e, T[11]);
    mont_sos_row<P>(e, o, T[12]);
    mont_sos_row<P>(o, e, T[13]);
    mont_sos_row<P>(e, o, T[14]);  // o column 0 (o[0] == 0), e column 1
    Fp<P> r;
    r.v[0] = add_cc(e[0], o[1]);
#pragma unroll
    for (int k = 1; k < 7; k++) r.v[k] = addc_cc(e[k], o[k + 1]);
    r.v[7] = addc(e[7], T[15]);
    fp_final_sub(r);
    return r;
}

// T[0..15] = a[0..7] * b[0..7]
UG_HD void mul8x8(const uint32_t* a, const uint32_t* b, uint32_t* T) {
    uint32_t ev[16], od[16];  // od[k] sits at limb k+1
#pragma unroll
    for (int k = 0; k < 16; k++) { ev[k] = 0; od[k] = 0; }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t x = b[i];
        // limb pairs starting at an even position go to ev, at an odd position to od: every (lo, hi) pair is one aligned IMAD.WIDE
        uint32_t* A0 = (i & 1) ? od : ev;  // receives a[even k] * x at position i + k
        uint32_t* A1 = (i & 1) ? ev : od;  // receives a[odd k]  * x at position i + k
        const int o0 = (i & 1) ? i - 1 : i;      // index inside A0 of position i      (od index = position - 1)
        const int o1 = (i & 1) ? i + 1 : i;      // index inside A1 of position i + 1
        A0[o0 + 0] = mad_lo_cc(a[0], x, A0[o0 + 0]);
        A0[o0 + 1] = madc_hi_cc(a[0], x, A0[o0 + 1]);
        A0[o0 + 2] = madc_lo_cc(a[2], x, A0[o0 + 2]);
        A0[o0 + 3] = madc_hi_cc(a[2], x, A0[o0 + 3]);
        A0[o0 + 4] = madc_lo_cc(a[4], x, A0[o0 + 4]);
        A0[o0 + 5] = madc_hi_cc(a[4], x, A0[o0 + 5]);
        A0[o0 + 6] = madc_lo_cc(a[6], x, A0[o0 + 6]);
        A0[o0 + 7] = madc_hi_cc(a[6], x, A0[o0 + 7]);
        if (o0 + 8 < 16) A0[o0 + 8] = addc(A0[o0 + 8], 0);
        A1[o1 + 0] = mad_lo_cc(a[1], x, A1[o1 + 0]);
        A1[o1 + 1] = madc_hi_cc(a[1], x, A1[o1 + 1]);
        A1[o1 + 2] = madc_lo_cc(a[3], x, A1[o1 + 2]);
        A1[o1 + 3] = madc_hi_cc(a[3], x, A1[o1 + 3]);
        A1[o1 + 4] = madc_lo_cc(a[5], x, A1[o1 + 4]);
        A1[o1 + 5] = madc_hi_cc(a[5], x, A1[o1 + 5]);
        A1[o1 + 6] = madc_lo_cc(a[7], x, A1[o1 + 6]);
        A1[o1 + 7] = madc_hi_cc(a[7], x, A1[o1 + 7]);
        if (o1 + 8 < 16) A1[o1 + 8] = addc(A1[o1 + 8], 0);
    }
    T[0] = ev[0];
    T[1] = add_cc(ev[1], od[0]);
#pragma unroll
    for (int k = 2; k < 15; k++) T[k] = addc_cc(ev[k], od[k - 1]);
    T[15] = addc(ev[15], od[14]);
}

template <class P>
UG_HD Fp<P> fp_sqr(const Fp<P>& x) { return fp_mul(x, x); }

template <class P>
UG_HD Fp<P> fp_to_mont(const Fp<P>& a) { return fp_mul(a, Fp<P>::rsquared()); }

// Fr_rawFromMontgomery == MMul(a, 1): eight reduction rows on the value itself.
template <class P>
UG_HD Fp<P> fp_from_mont(const Fp<P>& a) {
    Fp<P> one_raw = Fp<P>::zero();
    one_raw.v[0] = 1;
    return fp_mul(a, one_raw);
}

// a^e for a 256-bit little-endian exponent given as limbs (used for inversion: e = p - 2)
template <class P>
UG_HD Fp<P> fp_inv(const Fp<P>& a) {
    // exponent p-2, square-and-multiply from the top bit
    Fp<P> r = Fp<P>::one();
    for (int i = 255; i >= 0; i--) {
        r = fp_sqr(r);
        uint32_t w = P::mod(i >> 5);
        if ((i >> 5) == 0) w -= 2;  // low word of p is odd and > 2 for both fields
        if ((w >> (i & 31)) & 1) r = fp_mul(r, a);
    }
    return r;
}

using Fr = Fp<FrParams>;
using Fq = Fp<FqParams>;

}  // namespace ug
