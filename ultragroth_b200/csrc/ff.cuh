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
