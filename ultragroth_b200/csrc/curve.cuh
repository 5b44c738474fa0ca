// BN254 G1 / G2 group arithmetic in extended-Jacobian (XYZZ) coordinates for sm_100a.
//
// Replaces ffiasm's Curve<F> (depends/ffiasm, absent from the reference tree; the reference's own
// use of the x,y,zz,zzz representation is visible at src/groth16.cpp:379-410).  Memory layouts match
// the zkey sections the reference casts to G1PointAffine / G2PointAffine (src/groth16.cpp:33-43):
//   G1 affine = {Fq x, Fq y}            64 B, Montgomery form, all-zero = point at infinity
//   G2 affine = {Fq2 x{a,b}, Fq2 y{a,b}} 128 B
// x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2; infinity <=> ZZ == 0.  Formulas: EFD madd-2008-s / add-2008-s /
// dbl-2008-s-1 (a = 0).
#pragma once
#include "ff.cuh"

namespace ug {

// ---- Fq2 = Fq[u]/(u^2+1), element {a,b} = a + b*u (src/groth16.cpp:209-211) -----------------
struct alignas(16) Fq2 {
    Fq a, b;
    UG_HD static Fq2 zero() { Fq2 r; r.a = Fq::zero(); r.b = Fq::zero(); return r; }
    UG_HD static Fq2 one() { Fq2 r; r.a = Fq::one(); r.b = Fq::zero(); return r; }
    UG_HD bool is_zero() const { return a.is_zero() && b.is_zero(); }
    UG_HD bool operator==(const Fq2& o) const { return a == o.a && b == o.b; }
};

// uniform free-function interface over Fq and Fq2
UG_HD Fq fadd(const Fq& x, const Fq& y) { return fp_add(x, y); }
UG_HD Fq fsub(const Fq& x, const Fq& y) { return fp_sub(x, y); }
UG_HD Fq fmul(const Fq& x, const Fq& y) { return fp_mul(x, y); }
UG_HD Fq fsqr(const Fq& x) { return fp_sqr(x); }
UG_HD Fq fneg(const Fq& x) { return fp_neg(x); }
UG_HD Fq fdbl(const Fq& x) { return fp_add(x, x); }
UG_HD Fq finv(const Fq& x) { return fp_inv(x); }

UG_HD Fr fadd(const Fr& x, const Fr& y) { return fp_add(x, y); }
UG_HD Fr fsub(const Fr& x, const Fr& y) { return fp_sub(x, y); }
UG_HD Fr fmul(const Fr& x, const Fr& y) { return fp_mul(x, y); }
UG_HD Fr fsqr(const Fr& x) { return fp_sqr(x); }

UG_HD Fq2 fadd(const Fq2& x, const Fq2& y) { Fq2 r; r.a = fp_add(x.a, y.a); r.b = fp_add(x.b, y.b); return r; }
UG_HD Fq2 fsub(const Fq2& x, const Fq2& y) { Fq2 r; r.a = fp_sub(x.a, y.a); r.b = fp_sub(x.b, y.b); return r; }
UG_HD Fq2 fneg(const Fq2& x) { Fq2 r; r.a = fp_neg(x.a); r.b = fp_neg(x.b); return r; }
UG_HD Fq2 fdbl(const Fq2& x) { Fq2 r; r.a = fp_add(x.a, x.a); r.b = fp_add(x.b, x.b); return r; }
// On the device the two Fq2 products are real calls when UG_FQ2_CALL is set: fully inlined, k_accumulate<Fq2> is
// ~190 KB of SASS, more than the SM's instruction cache holds (ncu: `no_instruction` was its top stall).
#if defined(__CUDA_ARCH__) && defined(UG_FQ2_CALL)
#define UG_FQ2_FN static __device__ __noinline__
#else
#define UG_FQ2_FN UG_HD
#endif
UG_FQ2_FN Fq2 fmul(const Fq2& x, const Fq2& y) {
    Fq ac = fp_mul(x.a, y.a);
    Fq bd = fp_mul(x.b, y.b);
    Fq s = fp_mul(fp_add(x.a, x.b), fp_add(y.a, y.b));
    Fq2 r;
    r.a = fp_sub(ac, bd);
    r.b = fp_sub(fp_sub(s, ac), bd);
    return r;
}
UG_FQ2_FN Fq2 fsqr(const Fq2& x) {
    Fq ab = fp_mul(x.a, x.b);
    Fq2 r;
    r.a = fp_mul(fp_add(x.a, x.b), fp_sub(x.a, x.b));
    r.b = fp_add(ab, ab);
    return r;
}
UG_HD Fq2 finv(const Fq2& x) {
    Fq n = fp_inv(fp_add(fp_sqr(x.a), fp_sqr(x.b)));
    Fq2 r;
    r.a = fp_mul(x.a, n);
    r.b = fp_neg(fp_mul(x.b, n));
    return r;
}

// ---- a*b - c*d with ONE Montgomery reduction (Fq) ---------------------------------------------------------------------------------
// Both products as 512-bit integers, their difference (+ p * 2^256 when negative) reduced once: 2*64 + 72 = 200 multiplier
// operations instead of 2*136.  Same canonical value as fsub(fmul(a, b), fmul(c, d)).
// MEASURED AND OFF BY DEFAULT (UG_LAZY_MULSUB turns it on): in the G1 mixed addition (Y3 = R*(Q - X3) - Y1*PPP) it costs the
// 128-register kernel a spill and the separate-reduction form its extra additions: msm_a 9.70 vs 9.55 ms, proof 74.2 vs 73.5 ms
// (tools/r02_ab.sh).  The same idea pays in Fq2, where it removes a whole reduction per product (fq2_mul_lazy below).
UG_HD Fq fmul_sub(const Fq& a, const Fq& b, const Fq& c, const Fq& d) {
#if !defined(UG_LAZY_MULSUB)
    return fp_sub(fp_mul(a, b), fp_mul(c, d));
#else
    uint32_t T0[16], T1[16];
    mul8x8(a.v, b.v, T0);
    mul8x8(c.v, d.v, T1);
    T0[0] = sub_cc(T0[0], T1[0]);
#pragma unroll
    for (int k = 1; k < 16; k++) T0[k] = subc_cc(T0[k], T1[k]);
    const uint32_t borrow = subc(0, 0);
    T0[8] = add_cc(T0[8], FqParams::mod(0) & borrow);
#pragma unroll
    for (int k = 1; k < 7; k++) T0[8 + k] = addc_cc(T0[8 + k], FqParams::mod(k) & borrow);
    T0[15] = addc(T0[15], FqParams::mod(7) & borrow);
    return mont_reduce_512<FqParams>(T0);
#endif
}

// ---- Fq2 product with LAZY reduction ------------------------------------------------------------------------------------
// (a + b u)(c + d u) = (ac - bd) + ((a+b)(c+d) - ac - bd) u.  The three Karatsuba products are formed as full 512-bit integers
// (mul8x8, 64 multiplier operations each) and only the two RESULT coordinates are Montgomery-reduced (mont_reduce_512, 72 each):
// 336 multiplier operations instead of the 408 of three complete Fq products.  ac - bd may be negative: p * 2^256 is added back, which
// keeps the value inside the reduction's input range [0, p * 2^256) and does not change it mod p; (a+b)(c+d) - ac - bd = ad + bc is
// < 2 p^2 < p * 2^256.  a+b and c+d are taken WITHOUT reduction (< 2p < 2^255).  Same canonical values as fmul(Fq2, Fq2).
UG_HD void fq2_mul_lazy(const Fq& xa, const Fq& xb, const Fq& ya, const Fq& yb, Fq& ra, Fq& rb) {
    uint32_t T0[16], T1[16], T2[16], sx[8], sy[8];
    mul8x8(xa.v, ya.v, T0);
    mul8x8(xb.v, yb.v, T1);
    sx[0] = add_cc(xa.v[0], xb.v[0]);
#pragma unroll
    for (int k = 1; k < 7; k++) sx[k] = addc_cc(xa.v[k], xb.v[k]);
    sx[7] = addc(xa.v[7], xb.v[7]);
    sy[0] = add_cc(ya.v[0], yb.v[0]);
#pragma unroll
    for (int k = 1; k < 7; k++) sy[k] = addc_cc(ya.v[k], yb.v[k]);
    sy[7] = addc(ya.v[7], yb.v[7]);
    mul8x8(sx, sy, T2);
    // T2 -= T0 + T1   (>= 0)
    T2[0] = sub_cc(T2[0], T0[0]);
#pragma unroll
    for (int k = 1; k < 15; k++) T2[k] = subc_cc(T2[k], T0[k]);
    T2[15] = subc(T2[15], T0[15]);
    T2[0] = sub_cc(T2[0], T1[0]);
#pragma unroll
    for (int k = 1; k < 15; k++) T2[k] = subc_cc(T2[k], T1[k]);
    T2[15] = subc(T2[15], T1[15]);
    // T0 -= T1, + p * 2^256 on borrow
    T0[0] = sub_cc(T0[0], T1[0]);
#pragma unroll
    for (int k = 1; k < 16; k++) T0[k] = subc_cc(T0[k], T1[k]);
    const uint32_t borrow = subc(0, 0);  // all ones when ac < bd
    T0[8] = add_cc(T0[8], FqParams::mod(0) & borrow);
#pragma unroll
    for (int k = 1; k < 7; k++) T0[8 + k] = addc_cc(T0[8 + k], FqParams::mod(k) & borrow);
    T0[15] = addc(T0[15], FqParams::mod(7) & borrow);
    ra = mont_reduce_512<FqParams>(T0);
    rb = mont_reduce_512<FqParams>(T2);
}

#if defined(__CUDACC__)
// Fq2c: the same field and layout as Fq2, but its two products are REAL calls (arguments and result by value, which the
// device ABI keeps in registers: no stack traffic).  A kernel instantiated on Fq2c is ~10x smaller than on Fq2:
// k_running_sum<Fq2> is 500 KB of SASS, several times the SM's instruction cache.
struct alignas(16) Fq2c {
    Fq a, b;
    __device__ __forceinline__ static Fq2c zero() { Fq2c r; r.a = Fq::zero(); r.b = Fq::zero(); return r; }
    __device__ __forceinline__ static Fq2c one() { Fq2c r; r.a = Fq::one(); r.b = Fq::zero(); return r; }
    __device__ __forceinline__ bool is_zero() const { return a.is_zero() && b.is_zero(); }
    __device__ __forceinline__ bool operator==(const Fq2c& o) const { return a == o.a && b == o.b; }
};
static_assert(sizeof(Fq2c) == sizeof(Fq2), "Fq2c must alias Fq2");
__device__ __forceinline__ Fq2c fadd(const Fq2c& x, const Fq2c& y) { Fq2c r; r.a = fp_add(x.a, y.a); r.b = fp_add(x.b, y.b); return r; }
__device__ __forceinline__ Fq2c fsub(const Fq2c& x, const Fq2c& y) { Fq2c r; r.a = fp_sub(x.a, y.a); r.b = fp_sub(x.b, y.b); return r; }
__device__ __forceinline__ Fq2c fneg(const Fq2c& x) { Fq2c r; r.a = fp_neg(x.a); r.b = fp_neg(x.b); return r; }
__device__ __forceinline__ Fq2c fdbl(const Fq2c& x) { Fq2c r; r.a = fp_add(x.a, x.a); r.b = fp_add(x.b, x.b); return r; }
// UG_FQ2_NO_LAZY restores three complete Fq products (measured on B200: 21.5 vs 25.1 G Fq2 products/s, B2 MSM 25.9 vs 22.95 ms)
static __device__ __noinline__ Fq2c fq2c_mul_call(Fq2c x, Fq2c y) {
#if !defined(UG_FQ2_NO_LAZY)
    Fq2c rl;
    fq2_mul_lazy(x.a, x.b, y.a, y.b, rl.a, rl.b);
    return rl;
#endif
    Fq ac = fp_mul(x.a, y.a);
    Fq bd = fp_mul(x.b, y.b);
    Fq s = fp_mul(fp_add(x.a, x.b), fp_add(y.a, y.b));
    Fq2c r;
    r.a = fp_sub(ac, bd);
    r.b = fp_sub(fp_sub(s, ac), bd);
    return r;
}
static __device__ __noinline__ Fq2c fq2c_sqr_call(Fq2c x) {
    Fq ab = fp_mul(x.a, x.b);
    Fq2c r;
    r.a = fp_mul(fp_add(x.a, x.b), fp_sub(x.a, x.b));
    r.b = fp_add(ab, ab);
    return r;
}
__device__ __forceinline__ Fq2c finv(const Fq2c& x) {
    Fq n = fp_inv(fp_add(fp_sqr(x.a), fp_sqr(x.b)));
    Fq2c r;
    r.a = fp_mul(x.a, n);
    r.b = fp_neg(fp_mul(x.b, n));
    return r;
}
__device__ __forceinline__ Fq2c fmul(const Fq2c& x, const Fq2c& y) { return fq2c_mul_call(x, y); }
__device__ __forceinline__ Fq2c fsqr(const Fq2c& x) { return fq2c_sqr_call(x); }
#endif

UG_HD Fq2 fmul_sub(const Fq2& a, const Fq2& b, const Fq2& c, const Fq2& d) { return fsub(fmul(a, b), fmul(c, d)); }
#if defined(__CUDACC__)
__device__ __forceinline__ Fq2c fmul_sub(const Fq2c& a, const Fq2c& b, const Fq2c& c, const Fq2c& d) { return fsub(fmul(a, b), fmul(c, d)); }
#endif

// ---- points ---------------------------------------------------------------------------------
template <class F>
struct alignas(16) Affine {
    F x, y;
    UG_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
    UG_HD static Affine inf() { Affine p; p.x = F::zero(); p.y = F::zero(); return p; }
};

template <class F>
struct alignas(16) XYZZ {
    F X, Y, ZZ, ZZZ;
    UG_HD bool is_inf() const { return ZZ.is_zero(); }
    UG_HD static XYZZ inf() { XYZZ p; p.X = F::zero(); p.Y = F::zero(); p.ZZ = F::zero(); p.ZZZ = F::zero(); return p; }
    UG_HD static XYZZ from_affine(const Affine<F>& a) {
        if (a.is_inf()) return inf();
        XYZZ p; p.X = a.x; p.Y = a.y; p.ZZ = F::one(); p.ZZZ = F::one(); return p;
    }
};

using G1Affine = Affine<Fq>;
using G2Affine = Affine<Fq2>;
using G1XYZZ = XYZZ<Fq>;
using G2XYZZ = XYZZ<Fq2>;

// 2 * (affine point), result XYZZ (mdbl-2008-s-1)
template <class F>
UG_HD XYZZ<F> pt_dbl_affine(const Affine<F>& p) {
    if (p.is_inf() || p.y.is_zero()) return XYZZ<F>::inf();
    XYZZ<F> r;
    F U = fdbl(p.y);
    F V = fsqr(U);
    F W = fmul(U, V);
    F S = fmul(p.x, V);
    F xx = fsqr(p.x);
    F M = fadd(fdbl(xx), xx);
    r.X = fsub(fsqr(M), fdbl(S));
    r.Y = fsub(fmul(M, fsub(S, r.X)), fmul(W, p.y));
    r.ZZ = V;
    r.ZZZ = W;
    return r;
}

// 2 * P (dbl-2008-s-1)
template <class F>
UG_HD XYZZ<F> pt_dbl(const XYZZ<F>& p) {
    if (p.is_inf() || p.Y.is_zero()) return XYZZ<F>::inf();
    XYZZ<F> r;
    F U = fdbl(p.Y);
    F V = fsqr(U);
    F W = fmul(U, V);
    F S = fmul(p.X, V);
    F xx = fsqr(p.X);
    F M = fadd(fdbl(xx), xx);
    r.X = fsub(fsqr(M), fdbl(S));
    r.Y = fsub(fmul(M, fsub(S, r.X)), fmul(W, p.Y));
    r.ZZ = fmul(V, p.ZZ);
    r.ZZZ = fmul(W, p.ZZZ);
    return r;
}

// acc += (neg ? -q : q), q affine (madd-2008-s); all corner cases handled
template <class F>
UG_HD void pt_madd(XYZZ<F>& acc, const Affine<F>& q, bool neg) {
    if (q.is_inf()) return;
    F qy = neg ? fneg(q.y) : q.y;
    if (acc.is_inf()) {
        acc.X = q.x; acc.Y = qy; acc.ZZ = F::one(); acc.ZZZ = F::one();
        return;
    }
    F U2 = fmul(q.x, acc.ZZ);
    F S2 = fmul(qy, acc.ZZZ);
    F Pp = fsub(U2, acc.X);
    F Rr = fsub(S2, acc.Y);
    if (Pp.is_zero()) {
        if (Rr.is_zero()) {
            Affine<F> t; t.x = q.x; t.y = qy;
            acc = pt_dbl_affine(t);
        } else {
            acc = XYZZ<F>::inf();
        }
        return;
    }
    F PP = fsqr(Pp);
    F PPP = fmul(Pp, PP);
    F Qv = fmul(acc.X, PP);
    F X3 = fsub(fsub(fsqr(Rr), PPP), fdbl(Qv));
    acc.Y = fmul_sub(Rr, fsub(Qv, X3), acc.Y, PPP);
    acc.X = X3;
    acc.ZZ = fmul(acc.ZZ, PP);
    acc.ZZZ = fmul(acc.ZZZ, PPP);
}

// acc += q (add-2008-s); all corner cases handled
template <class F>
UG_HD void pt_add(XYZZ<F>& acc, const XYZZ<F>& q) {
    if (q.is_inf()) return;
    if (acc.is_inf()) { acc = q; return; }
    F U1 = fmul(acc.X, q.ZZ);
    F U2 = fmul(q.X, acc.ZZ);
    F S1 = fmul(acc.Y, q.ZZZ);
    F S2 = fmul(q.Y, acc.ZZZ);
    F Pp = fsub(U2, U1);
    F Rr = fsub(S2, S1);
    if (Pp.is_zero()) {
        if (Rr.is_zero()) acc = pt_dbl(acc);
        else acc = XYZZ<F>::inf();
        return;
    }
    F PP = fsqr(Pp);
    F PPP = fmul(Pp, PP);
    F Qv = fmul(U1, PP);
    F X3 = fsub(fsub(fsqr(Rr), PPP), fdbl(Qv));
    acc.Y = fmul_sub(Rr, fsub(Qv, X3), S1, PPP);
    acc.X = X3;
    acc.ZZ = fmul(fmul(acc.ZZ, q.ZZ), PP);
    acc.ZZZ = fmul(fmul(acc.ZZZ, q.ZZZ), PPP);
}

#if defined(__CUDACC__)
// ---- points addressed through an accessor (coordinate-at-a-time) -------------------------------------------------
// The G2 kernels hold 64 registers of accumulator per XYZZ point and ran at 255 registers / 2 warps per scheduler.
// Keeping the accumulator in a per-thread column of shared memory (word k of thread t at slab[k*BS + t]: 128-bit
// accesses, conflict-free) and touching one coordinate at a time frees those registers for a third/fourth warp.
// coord: 0 = X, 1 = Y, 2 = ZZ, 3 = ZZZ.
template <class F, int BS>
struct SmemPoint {
    uint4* p;  // &slab[threadIdx.x]
    static constexpr int NV = sizeof(F) / 16;
    __device__ __forceinline__ F ld(int coord) const {
        F r;
        uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
        for (int k = 0; k < NV; k++) d[k] = p[(coord * NV + k) * BS];
        return r;
    }
    __device__ __forceinline__ void st(int coord, const F& v) const {
        const uint4* s = reinterpret_cast<const uint4*>(&v);
#pragma unroll
        for (int k = 0; k < NV; k++) p[(coord * NV + k) * BS] = s[k];
    }
    __device__ __forceinline__ bool is_inf() const {
        uint32_t o = 0;
#pragma unroll
        for (int k = 0; k < NV; k++) { uint4 w = p[(2 * NV + k) * BS]; o |= w.x | w.y | w.z | w.w; }
        return o == 0;
    }
    __device__ __forceinline__ void set_inf() const {
#pragma unroll
        for (int k = 0; k < 4 * NV; k++) p[k * BS] = make_uint4(0, 0, 0, 0);
    }
    __device__ __forceinline__ XYZZ<F> get() const { XYZZ<F> r; r.X = ld(0); r.Y = ld(1); r.ZZ = ld(2); r.ZZZ = ld(3); return r; }
    __device__ __forceinline__ void put(const XYZZ<F>& v) const { st(0, v.X); st(1, v.Y); st(2, v.ZZ); st(3, v.ZZZ); }
};
template <class F>
struct GlobalPoint {  // read-only XYZZ in global memory
    const XYZZ<F>* q;
    __device__ __forceinline__ F ld(int coord) const {
        F r;
        const uint4* s = reinterpret_cast<const uint4*>(q) + coord * (sizeof(F) / 16);
        uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
        for (int k = 0; k < (int)(sizeof(F) / 16); k++) d[k] = s[k];
        return r;
    }
    __device__ __forceinline__ bool is_inf() const { return ld(2).is_zero(); }
    __device__ __forceinline__ XYZZ<F> get() const { XYZZ<F> r; r.X = ld(0); r.Y = ld(1); r.ZZ = ld(2); r.ZZZ = ld(3); return r; }
};

// rare paths out of line: they would otherwise set the register budget of the hot loop
template <class F>
__device__ __noinline__ void pt_dbl_slow(XYZZ<F>* io) { *io = pt_dbl(*io); }
template <class F>
__device__ __noinline__ void pt_dbl_affine_slow(const Affine<F>* in, XYZZ<F>* out) { *out = pt_dbl_affine(*in); }

// acc += (neg ? -q : q), q affine in registers, acc behind an accessor (madd-2008-s)
template <class F, class A>
__device__ __forceinline__ void pt_madd_at(const A& acc, const Affine<F>& q, bool neg) {
    if (q.is_inf()) return;
    F qy = neg ? fneg(q.y) : q.y;
    if (acc.is_inf()) {
        acc.st(0, q.x); acc.st(1, qy); acc.st(2, F::one()); acc.st(3, F::one());
        return;
    }
    F Pp = fsub(fmul(q.x, acc.ld(2)), acc.ld(0));
    F Rr = fsub(fmul(qy, acc.ld(3)), acc.ld(1));
    if (Pp.is_zero()) {
        if (Rr.is_zero()) {
            Affine<F> t; t.x = q.x; t.y = qy;
            XYZZ<F> d;
            pt_dbl_affine_slow(&t, &d);
            acc.put(d);
        } else {
            acc.set_inf();
        }
        return;
    }
    F PP = fsqr(Pp);
    acc.st(2, fmul(acc.ld(2), PP));
    F PPP = fmul(Pp, PP);
    acc.st(3, fmul(acc.ld(3), PPP));
    F Qv = fmul(acc.ld(0), PP);
    F X3 = fsub(fsub(fsqr(Rr), PPP), fdbl(Qv));
    acc.st(0, X3);
    acc.st(1, fsub(fmul(Rr, fsub(Qv, X3)), fmul(acc.ld(1), PPP)));
}

// acc += q, both behind accessors (add-2008-s)
template <class F, class A, class Q>
__device__ __forceinline__ void pt_add_at(const A& acc, const Q& q) {
    if (q.is_inf()) return;
    if (acc.is_inf()) { acc.st(0, q.ld(0)); acc.st(1, q.ld(1)); acc.st(2, q.ld(2)); acc.st(3, q.ld(3)); return; }
    F U1 = fmul(acc.ld(0), q.ld(2));
    F Pp = fsub(fmul(q.ld(0), acc.ld(2)), U1);
    F S1 = fmul(acc.ld(1), q.ld(3));
    F Rr = fsub(fmul(q.ld(1), acc.ld(3)), S1);
    if (Pp.is_zero()) {
        if (Rr.is_zero()) {
            XYZZ<F> d = acc.get();
            pt_dbl_slow(&d);
            acc.put(d);
        } else {
            acc.set_inf();
        }
        return;
    }
    F PP = fsqr(Pp);
    acc.st(2, fmul(fmul(acc.ld(2), q.ld(2)), PP));
    F PPP = fmul(Pp, PP);
    acc.st(3, fmul(fmul(acc.ld(3), q.ld(3)), PPP));
    F Qv = fmul(U1, PP);
    F X3 = fsub(fsub(fsqr(Rr), PPP), fdbl(Qv));
    acc.st(0, X3);
    acc.st(1, fsub(fmul(Rr, fsub(Qv, X3)), fmul(S1, PPP)));
}
#endif  // __CUDACC__

// ---- Jacobian doubling chain (window-table build only: 2^c * P needs c doublings and ONE conversion back to affine) ----
// x = X/Z^2, y = Y/Z^3.  dbl-2009-l costs 2M + 5S against the 6M + 3S of the XYZZ doubling: 7 instead of 9 products in Fq and, with
// Fq2 squarings at 2 base products, 16 instead of 24 in Fq2.  Infinity <=> Z == 0 (BN254 has no points of order 2 in the prime-order
// subgroups, so Y == 0 cannot occur for valid bases; it maps to infinity as in pt_dbl).
template <class F>
struct alignas(16) Jac {
    F X, Y, Z;
    UG_HD bool is_inf() const { return Z.is_zero(); }
};
template <class F>
UG_HD Jac<F> jac_dbl_affine(const Affine<F>& p) {  // mdbl-2007-bl (Z1 = 1): 1M + 5S
    Jac<F> r;
    if (p.is_inf() || p.y.is_zero()) { r.X = F::one(); r.Y = F::one(); r.Z = F::zero(); return r; }
    F XX = fsqr(p.x), YY = fsqr(p.y), YYYY = fsqr(YY);
    F S = fdbl(fsub(fsub(fsqr(fadd(p.x, YY)), XX), YYYY));
    F M = fadd(fdbl(XX), XX);
    r.X = fsub(fsqr(M), fdbl(S));
    F y8 = fdbl(fdbl(fdbl(YYYY)));
    r.Y = fsub(fmul(M, fsub(S, r.X)), y8);
    r.Z = fdbl(p.y);
    return r;
}
template <class F>
UG_HD Jac<F> jac_dbl(const Jac<F>& p) {  // dbl-2009-l: 2M + 5S
    if (p.is_inf() || p.Y.is_zero()) { Jac<F> r; r.X = F::one(); r.Y = F::one(); r.Z = F::zero(); return r; }
    F A = fsqr(p.X), B = fsqr(p.Y), C = fsqr(B);
    F D = fdbl(fsub(fsub(fsqr(fadd(p.X, B)), A), C));
    F E = fadd(fdbl(A), A);
    Jac<F> r;
    r.X = fsub(fsqr(E), fdbl(D));
    F c8 = fdbl(fdbl(fdbl(C)));
    r.Y = fsub(fmul(E, fsub(D, r.X)), c8);
    r.Z = fdbl(fmul(p.Y, p.Z));
    return r;
}

template <class F>
UG_HD XYZZ<F> pt_neg(const XYZZ<F>& p) { XYZZ<F> r = p; r.Y = fneg(p.Y); return r; }

// XYZZ -> affine (one field inversion); infinity -> all-zero
template <class F>
UG_HD Affine<F> pt_to_affine(const XYZZ<F>& p) {
    if (p.is_inf()) return Affine<F>::inf();
    F izzz = finv(p.ZZZ);
    F izz = fsqr(fmul(izzz, p.ZZ));  // (ZZ/ZZZ)^2 = ZZ^2/ZZ^3 = 1/ZZ   (ZZ^3 = ZZZ^2)
    Affine<F> a;
    a.x = fmul(p.X, izz);
    a.y = fmul(p.Y, izzz);
    return a;
}

}  // namespace ug
