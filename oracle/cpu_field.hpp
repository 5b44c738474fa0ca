// ORACLE -- test infrastructure, NOT product code (see cpu_prover.cpp).  BN254 fields and curve arithmetic shared by the oracle's
// C++ translation units (cpu_prover.cpp: the prover restatement; cpu_synth.cpp: the synthetic-setup generator):
//   field semantics .... build/fr_raw_generic.cpp:107-148 (4-round CIOS, one conditional subtraction), :11-39 (add / sub),
//                        constants :5-8; build/fq_raw_generic.cpp likewise
//   Fq2 ................ Fq[u]/(u^2+1), element {a,b} = a + b u (src/groth16.cpp:209-211)
//   curve .............. y^2 = x^3 + b, Jacobian coordinates (the engine iden3/ffiasm @ fe3772e8 is absent; published formulas)
#pragma once
#include <cstdint>
#include <cstring>

typedef unsigned __int128 u128;
typedef uint64_t u64;

namespace {

// ---------------------------------------------------------------------------------------------
// prime fields (build/fr_raw_generic.cpp, build/fq_raw_generic.cpp)
// ---------------------------------------------------------------------------------------------
struct FrP {
    static constexpr u64 q[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    static constexpr u64 r2[4] = {0x1bb8e645ae216da7ull, 0x53fe3ab1e35c59e3ull, 0x8c49833d53bb8085ull, 0x0216d0b17f4e44a5ull};
    static constexpr u64 np = 0xc2e1f593efffffffull;
};
struct FqP {
    static constexpr u64 q[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    static constexpr u64 r2[4] = {0xf32cfc5b538afa89ull, 0xb5e71911d44501fbull, 0x47ab1eff0a417ff6ull, 0x06d89f71cab8351full};
    static constexpr u64 np = 0x87d20782e4866389ull;
};

template <class P>
struct El {
    u64 v[4];
    bool isZero() const { return (v[0] | v[1] | v[2] | v[3]) == 0; }
    bool operator==(const El& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2] && v[3] == o.v[3]; }
};

template <class P>
inline bool geq_q(const u64* a) {
    for (int i = 3; i >= 0; i--) {
        if (a[i] != P::q[i]) return a[i] > P::q[i];
    }
    return true;
}
template <class P>
inline void sub_q(u64* a) {
    u64 borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a[i] - P::q[i] - borrow;
        a[i] = (u64)d;
        borrow = (u64)(d >> 64) & 1;
    }
}
// Fr_rawAdd (fr_raw_generic.cpp:11-29)
template <class P>
inline El<P> add(const El<P>& a, const El<P>& b) {
    El<P> r;
    u128 c = 0;
    for (int i = 0; i < 4; i++) { c += (u128)a.v[i] + b.v[i]; r.v[i] = (u64)c; c >>= 64; }
    if (c || geq_q<P>(r.v)) sub_q<P>(r.v);
    return r;
}
// Fr_rawSub (:31-39)
template <class P>
inline El<P> sub(const El<P>& a, const El<P>& b) {
    El<P> r;
    u64 borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a.v[i] - b.v[i] - borrow;
        r.v[i] = (u64)d;
        borrow = (u64)(d >> 64) & 1;
    }
    if (borrow) {
        u128 c = 0;
        for (int i = 0; i < 4; i++) { c += (u128)r.v[i] + P::q[i]; r.v[i] = (u64)c; c >>= 64; }
    }
    return r;
}
template <class P>
inline El<P> neg(const El<P>& a) {
    if (a.isZero()) return a;
    El<P> z = {{0, 0, 0, 0}};
    return sub(z, a);
}
// Fr_rawMMul (:107-148): interleaved multiply / reduce rounds, product accumulated in 4+2 words
template <class P>
inline El<P> mul(const El<P>& a, const El<P>& b) {
    u64 t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) { c += (u128)a.v[j] * b.v[i] + t[j]; t[j] = (u64)c; c >>= 64; }
        c += t[4];
        t[4] = (u64)c;
        t[5] = (u64)(c >> 64);
        u64 m = t[0] * P::np;
        c = ((u128)m * P::q[0] + t[0]) >> 64;
        for (int j = 1; j < 4; j++) { c += (u128)m * P::q[j] + t[j]; t[j - 1] = (u64)c; c >>= 64; }
        c += t[4];
        t[3] = (u64)c;
        t[4] = t[5] + (u64)(c >> 64);
    }
    El<P> r = {{t[0], t[1], t[2], t[3]}};
    if (t[4] || geq_q<P>(r.v)) sub_q<P>(r.v);
    return r;
}
template <class P>
inline El<P> sqr(const El<P>& a) { return mul(a, a); }
template <class P>
inline El<P> R2() { return El<P>{{P::r2[0], P::r2[1], P::r2[2], P::r2[3]}}; }
template <class P>
inline El<P> toMont(const El<P>& a) { return mul(a, R2<P>()); }
template <class P>
inline El<P> fromMont(const El<P>& a) { return mul(a, El<P>{{1, 0, 0, 0}}); }
template <class P>
inline const El<P>& one() {  // R mod q, computed once (Fr_rawR2 * 1 reduced)
    static const El<P> v = toMont(El<P>{{1, 0, 0, 0}});
    return v;
}
template <class P>
inline El<P> fromU64(u64 x) { return toMont(El<P>{{x, 0, 0, 0}}); }
template <class P>
El<P> powv(El<P> b, const u64* e, int limbs) {
    El<P> r = one<P>();
    for (int i = limbs * 64 - 1; i >= 0; i--) {
        r = sqr(r);
        if ((e[i >> 6] >> (i & 63)) & 1) r = mul(r, b);
    }
    return r;
}
template <class P>
El<P> inv(const El<P>& a) {  // RawFr::inv uses mpz_invert (build/fr.cpp:238-250); a^(q-2) gives the same residue
    u64 e[4] = {P::q[0] - 2, P::q[1], P::q[2], P::q[3]};
    return powv(a, e, 4);
}

typedef El<FrP> Fr;
typedef El<FqP> Fq;

// ---- Fq2 = Fq[u]/(u^2+1) (src/groth16.cpp:209-211: element {a,b} = a + b u) ----
struct Fq2 {
    Fq a, b;
    bool isZero() const { return a.isZero() && b.isZero(); }
    bool operator==(const Fq2& o) const { return a == o.a && b == o.b; }
};
inline Fq2 add(const Fq2& x, const Fq2& y) { return {add(x.a, y.a), add(x.b, y.b)}; }
inline Fq2 sub(const Fq2& x, const Fq2& y) { return {sub(x.a, y.a), sub(x.b, y.b)}; }
inline Fq2 neg(const Fq2& x) { return {neg(x.a), neg(x.b)}; }
inline Fq2 mul(const Fq2& x, const Fq2& y) {
    Fq aa = mul(x.a, y.a), bb = mul(x.b, y.b);
    Fq cross = sub(sub(mul(add(x.a, x.b), add(y.a, y.b)), aa), bb);
    return {sub(aa, bb), cross};
}
inline Fq2 sqr(const Fq2& x) { return mul(x, x); }
inline Fq2 inv(const Fq2& x) {
    Fq n = inv(add(sqr(x.a), sqr(x.b)));
    return {mul(x.a, n), neg(mul(x.b, n))};
}
template <class F> inline F fone();
template <> inline Fq fone<Fq>() { return one<FqP>(); }
template <> inline Fq2 fone<Fq2>() { return {one<FqP>(), Fq{{0, 0, 0, 0}}}; }
template <class F> inline F fzero();
template <> inline Fq fzero<Fq>() { return Fq{{0, 0, 0, 0}}; }
template <> inline Fq2 fzero<Fq2>() { return {Fq{{0, 0, 0, 0}}, Fq{{0, 0, 0, 0}}}; }

// ---------------------------------------------------------------------------------------------
// curve y^2 = x^3 + b in Jacobian coordinates (x = X/Z^2, y = Y/Z^3), a = 0
// ---------------------------------------------------------------------------------------------
template <class F>
struct Aff { F x, y; bool isInf() const { return x.isZero() && y.isZero(); } };
template <class F>
struct Jac { F X, Y, Z; bool isInf() const { return Z.isZero(); } };

template <class F>
Jac<F> jinf() { return {fone<F>(), fone<F>(), fzero<F>()}; }
template <class F>
Jac<F> dbl(const Jac<F>& p) {  // dbl-2009-l
    if (p.isInf()) return p;
    F A = sqr(p.X), B = sqr(p.Y), C = sqr(B);
    F t = sub(sub(sqr(add(p.X, B)), A), C);
    F D = add(t, t);
    F E = add(add(A, A), A);
    F Fv = sqr(E);
    Jac<F> r;
    r.X = sub(Fv, add(D, D));
    F c8 = add(C, C); c8 = add(c8, c8); c8 = add(c8, c8);
    r.Y = sub(mul(E, sub(D, r.X)), c8);
    F yz = mul(p.Y, p.Z);
    r.Z = add(yz, yz);
    return r;
}
template <class F>
Jac<F> addj(const Jac<F>& p, const Jac<F>& q) {  // add-2007-bl, with the doubling / cancellation cases
    if (p.isInf()) return q;
    if (q.isInf()) return p;
    F Z1Z1 = sqr(p.Z), Z2Z2 = sqr(q.Z);
    F U1 = mul(p.X, Z2Z2), U2 = mul(q.X, Z1Z1);
    F S1 = mul(mul(p.Y, q.Z), Z2Z2), S2 = mul(mul(q.Y, p.Z), Z1Z1);
    if (U1 == U2) {
        if (S1 == S2) return dbl(p);
        return jinf<F>();
    }
    F H = sub(U2, U1);
    F I = add(H, H); I = sqr(I);
    F J = mul(H, I);
    F rr = sub(S2, S1); rr = add(rr, rr);
    F V = mul(U1, I);
    Jac<F> r;
    r.X = sub(sub(sqr(rr), J), add(V, V));
    F s1j = mul(S1, J);
    r.Y = sub(mul(rr, sub(V, r.X)), add(s1j, s1j));
    r.Z = mul(sub(sub(sqr(add(p.Z, q.Z)), Z1Z1), Z2Z2), H);
    return r;
}
template <class F>
Jac<F> addm(const Jac<F>& p, const Aff<F>& q) {  // mixed addition (madd-2007-bl)
    if (q.isInf()) return p;
    if (p.isInf()) return {q.x, q.y, fone<F>()};
    F Z1Z1 = sqr(p.Z);
    F U2 = mul(q.x, Z1Z1), S2 = mul(mul(q.y, p.Z), Z1Z1);
    if (U2 == p.X) {
        if (S2 == p.Y) return dbl(p);
        return jinf<F>();
    }
    F H = sub(U2, p.X), HH = sqr(H);
    F I = add(HH, HH); I = add(I, I);
    F J = mul(H, I);
    F rr = sub(S2, p.Y); rr = add(rr, rr);
    F V = mul(p.X, I);
    Jac<F> r;
    r.X = sub(sub(sqr(rr), J), add(V, V));
    F yj = mul(p.Y, J);
    r.Y = sub(mul(rr, sub(V, r.X)), add(yj, yj));
    r.Z = sub(sub(sqr(add(p.Z, H)), Z1Z1), HH);
    return r;
}
template <class F>
Jac<F> negj(const Jac<F>& p) { return {p.X, neg(p.Y), p.Z}; }
template <class F>
Aff<F> toAff(const Jac<F>& p) {
    if (p.isInf()) return {fzero<F>(), fzero<F>()};
    F zi = inv(p.Z), zi2 = sqr(zi);
    return {mul(p.X, zi2), mul(p.Y, mul(zi2, zi))};
}
template <class F>
Jac<F> mulScalar(const Jac<F>& p, const u64* k) {  // E.g1.mulByScalar: double-and-add over 256 bits
    Jac<F> r = jinf<F>();
    for (int i = 255; i >= 0; i--) {
        r = dbl(r);
        if ((k[i >> 6] >> (i & 63)) & 1) r = addj(r, p);
    }
    return r;
}

// omega_{2^k} = 5^((r-1)/2^k)  (root convention of ffiasm's FFT<Fr>)
inline Fr rootOfUnity(int k) {
    static const u64 e[4] = {0x9b9709143e1f593full, 0x181585d2833e8487ull, 0x131a029b85045b68ull, 0x000000030644e72eull};  // (r-1)/2^28
    Fr w = powv(fromU64<FrP>(5), e, 4);
    for (int i = 28; i > k; i--) w = sqr(w);
    return w;
}

}  // namespace
