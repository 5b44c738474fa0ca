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
    acc.Y = fsub(fmul(Rr, fsub(Qv, X3)), fmul(acc.Y, PPP));
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
    acc.Y = fsub(fmul(Rr, fsub(Qv, X3)), fmul(S1, PPP));
    acc.X = X3;
    acc.ZZ = fmul(fmul(acc.ZZ, q.ZZ), PP);
    acc.ZZZ = fmul(fmul(acc.ZZZ, q.ZZZ), PPP);
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
