// Host-side BN254 Fq / Fr / Fq2 on 4 x 64-bit limbs (unsigned __int128 CIOS).
//
// Used only for the O(1) tail of a proof, exactly the part the reference also runs serially on the CPU:
// window combination of an MSM, blinding and proof assembly (src/groth16.cpp:158-200,
// src/ultra_groth.cpp:170-183,337-396), the challenge / lookup scalars (src/ultra_groth.cpp:32-58) and
// decimal printing.  Memory layout is identical to the device types (32-byte little-endian Montgomery
// residues), so XYZZ<HFq> can be memcpy'd from XYZZ<Fq> results and the curve.cuh formulas are shared.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>

namespace ug {

typedef unsigned __int128 u128;

struct HFqParams {
    static constexpr uint64_t mod[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    static constexpr uint64_t r2[4] = {0xf32cfc5b538afa89ull, 0xb5e71911d44501fbull, 0x47ab1eff0a417ff6ull, 0x06d89f71cab8351full};
    static constexpr uint64_t one[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull, 0x0e0a77c19a07df2full};
    static constexpr uint64_t np = 0x87d20782e4866389ull;
};
struct HFrParams {
    static constexpr uint64_t mod[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    static constexpr uint64_t r2[4] = {0x1bb8e645ae216da7ull, 0x53fe3ab1e35c59e3ull, 0x8c49833d53bb8085ull, 0x0216d0b17f4e44a5ull};
    static constexpr uint64_t one[4] = {0xac96341c4ffffffbull, 0x36fc76959f60cd29ull, 0x666ea36f7879462eull, 0x0e0a77c19a07df2full};
    static constexpr uint64_t np = 0xc2e1f593efffffffull;
};

template <class P>
struct alignas(16) HFp {
    uint64_t v[4];
    static HFp zero() { HFp r; r.v[0] = r.v[1] = r.v[2] = r.v[3] = 0; return r; }
    static HFp one() { HFp r; for (int i = 0; i < 4; i++) r.v[i] = P::one[i]; return r; }
    static HFp rsquared() { HFp r; for (int i = 0; i < 4; i++) r.v[i] = P::r2[i]; return r; }
    bool is_zero() const { return (v[0] | v[1] | v[2] | v[3]) == 0; }
    bool operator==(const HFp& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2] && v[3] == o.v[3]; }
    bool operator!=(const HFp& o) const { return !(*this == o); }
};

template <class P>
inline bool hfp_geq_mod(const uint64_t* a) {
    for (int i = 3; i >= 0; i--) {
        if (a[i] > P::mod[i]) return true;
        if (a[i] < P::mod[i]) return false;
    }
    return true;
}
template <class P>
inline void hfp_sub_mod(uint64_t* a) {
    u128 borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a[i] - P::mod[i] - (uint64_t)borrow;
        a[i] = (uint64_t)d;
        borrow = (d >> 64) & 1;
    }
}

template <class P>
inline HFp<P> fadd(const HFp<P>& a, const HFp<P>& b) {
    HFp<P> r;
    u128 c = 0;
    for (int i = 0; i < 4; i++) { c += (u128)a.v[i] + b.v[i]; r.v[i] = (uint64_t)c; c >>= 64; }
    if (c || hfp_geq_mod<P>(r.v)) hfp_sub_mod<P>(r.v);
    return r;
}
template <class P>
inline HFp<P> fsub(const HFp<P>& a, const HFp<P>& b) {
    HFp<P> r;
    uint64_t borrow = 0;
    for (int i = 0; i < 4; i++) {
        u128 d = (u128)a.v[i] - b.v[i] - borrow;
        r.v[i] = (uint64_t)d;
        borrow = (uint64_t)(d >> 64) & 1;
    }
    if (borrow) {
        u128 c = 0;
        for (int i = 0; i < 4; i++) { c += (u128)r.v[i] + P::mod[i]; r.v[i] = (uint64_t)c; c >>= 64; }
    }
    return r;
}
template <class P>
inline HFp<P> fneg(const HFp<P>& a) { return a.is_zero() ? a : fsub(HFp<P>::zero(), a); }
template <class P>
inline HFp<P> fdbl(const HFp<P>& a) { return fadd(a, a); }

// x*y*2^-256 mod p, x scanned (any 256-bit value), y < p
template <class P>
inline HFp<P> fmul(const HFp<P>& x, const HFp<P>& y) {
    uint64_t t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)x.v[i] * y.v[j] + t[j];
            t[j] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[4] = (uint64_t)c;
        t[5] = (uint64_t)(c >> 64);
        uint64_t m = t[0] * P::np;
        c = (u128)m * P::mod[0] + t[0];
        c >>= 64;
        for (int j = 1; j < 4; j++) {
            c += (u128)m * P::mod[j] + t[j];
            t[j - 1] = (uint64_t)c;
            c >>= 64;
        }
        c += t[4];
        t[3] = (uint64_t)c;
        t[4] = t[5] + (uint64_t)(c >> 64);
    }
    HFp<P> r;
    for (int i = 0; i < 4; i++) r.v[i] = t[i];
    if (t[4] || hfp_geq_mod<P>(r.v)) hfp_sub_mod<P>(r.v);
    return r;
}
template <class P>
inline HFp<P> fsqr(const HFp<P>& x) { return fmul(x, x); }
// a*b - c*d (the device code reduces the difference of the two products once; the host has no reason to)
template <class P>
inline HFp<P> fmul_sub(const HFp<P>& a, const HFp<P>& b, const HFp<P>& c, const HFp<P>& d) { return fsub(fmul(a, b), fmul(c, d)); }
template <class P>
inline HFp<P> to_mont(const HFp<P>& a) { return fmul(a, HFp<P>::rsquared()); }
template <class P>
inline HFp<P> from_mont(const HFp<P>& a) { HFp<P> o = HFp<P>::zero(); o.v[0] = 1; return fmul(a, o); }

template <class P>
inline HFp<P> finv(const HFp<P>& a) {  // a^(p-2)
    uint64_t e[4];
    for (int i = 0; i < 4; i++) e[i] = P::mod[i];
    e[0] -= 2;
    HFp<P> r = HFp<P>::one();
    for (int i = 255; i >= 0; i--) {
        r = fsqr(r);
        if ((e[i >> 6] >> (i & 63)) & 1) r = fmul(r, a);
    }
    return r;
}

using HFq = HFp<HFqParams>;
using HFr = HFp<HFrParams>;

struct alignas(16) HFq2 {
    HFq a, b;
    static HFq2 zero() { HFq2 r; r.a = HFq::zero(); r.b = HFq::zero(); return r; }
    static HFq2 one() { HFq2 r; r.a = HFq::one(); r.b = HFq::zero(); return r; }
    bool is_zero() const { return a.is_zero() && b.is_zero(); }
    bool operator==(const HFq2& o) const { return a == o.a && b == o.b; }
};
inline HFq2 fadd(const HFq2& x, const HFq2& y) { HFq2 r; r.a = fadd(x.a, y.a); r.b = fadd(x.b, y.b); return r; }
inline HFq2 fsub(const HFq2& x, const HFq2& y) { HFq2 r; r.a = fsub(x.a, y.a); r.b = fsub(x.b, y.b); return r; }
inline HFq2 fneg(const HFq2& x) { HFq2 r; r.a = fneg(x.a); r.b = fneg(x.b); return r; }
inline HFq2 fdbl(const HFq2& x) { return fadd(x, x); }
inline HFq2 fmul(const HFq2& x, const HFq2& y) {
    HFq ac = fmul(x.a, y.a), bd = fmul(x.b, y.b);
    HFq s = fmul(fadd(x.a, x.b), fadd(y.a, y.b));
    HFq2 r; r.a = fsub(ac, bd); r.b = fsub(fsub(s, ac), bd); return r;
}
inline HFq2 fmul_sub(const HFq2& a, const HFq2& b, const HFq2& c, const HFq2& d) { return fsub(fmul(a, b), fmul(c, d)); }
inline HFq2 fsqr(const HFq2& x) {
    HFq ab = fmul(x.a, x.b);
    HFq2 r; r.a = fmul(fadd(x.a, x.b), fsub(x.a, x.b)); r.b = fadd(ab, ab); return r;
}
inline HFq2 finv(const HFq2& x) {
    HFq n = finv(fadd(fsqr(x.a), fsqr(x.b)));
    HFq2 r; r.a = fmul(x.a, n); r.b = fneg(fmul(x.b, n)); return r;
}

// canonical (non-Montgomery) 256-bit little-endian integer -> decimal string
inline std::string u256_to_dec(const uint64_t* limbs) {
    uint64_t t[4] = {limbs[0], limbs[1], limbs[2], limbs[3]};
    if ((t[0] | t[1] | t[2] | t[3]) == 0) return "0";
    std::string out;
    while (t[0] | t[1] | t[2] | t[3]) {
        // divide by 10^19
        const uint64_t D = 10000000000000000000ull;
        u128 rem = 0;
        for (int i = 3; i >= 0; i--) {
            u128 cur = (rem << 64) | t[i];
            t[i] = (uint64_t)(cur / D);
            rem = cur % D;
        }
        uint64_t chunk = (uint64_t)rem;
        bool last = (t[0] | t[1] | t[2] | t[3]) == 0;
        for (int k = 0; k < 19; k++) {
            if (last && chunk == 0) break;
            out.push_back(char('0' + chunk % 10));
            chunk /= 10;
        }
    }
    return std::string(out.rbegin(), out.rend());
}
template <class P>
inline std::string to_dec(const HFp<P>& mont) {
    HFp<P> c = from_mont(mont);
    return u256_to_dec(c.v);
}

}  // namespace ug
