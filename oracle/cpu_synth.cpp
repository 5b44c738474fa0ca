// ORACLE / CPU BASELINE -- test infrastructure, NOT product code.  Nothing under ultragroth_b200/ links, includes or calls this.
//
// CPU generator of the synthetic benchmark circuits WITH a valid trusted setup: the same circuit families, seeds and file layouts
// as the product's GPU generator (ugb_synth_build, csrc/synth.cu), written against the oracle's own field and curve code.  Two jobs:
//   * bench.py --impl reference and the cpu_baseline leg get their zkey / witness WITHOUT loading libultragroth_b200.so, so the CPU
//     arm really is a CPU-only program on the same workload (same seed => byte-identical files, checked by tests/test_gpu_synth.py);
//   * the -m "not gpu" tests get UltraGroth fixtures of any size on a machine without a GPU.
// File layouts follow the reference's loaders: zkey sections src/zkey_utils.cpp:42-76,123-163 and src/prover.cpp:160-189,237-260,
// wtns / .uwtns sections src/wtns_utils.cpp:13-26 and src/prover.cpp:283-292; vkey JSON field names src/groth16.cpp:278-296 and
// src/ultra_groth.cpp:542-563.
//
// Circuits.  Groth16 family (protocol 1):  (x_i + a_i) * x_i = x_{i+1} - k_i,  i < m,  one public output x_m.
// UltraGroth family (protocol 1337): the same chain plus a LogUp range lookup over K chunk signals c_j in [0,T) with T frequencies:
//   inv1_j*(c_j + rand) = 1, inv2_i*(rand + i) = 1, prod_i = f_i*inv2_i, sum_j inv1_j = sum_i prod_i, rand = the public signal at
//   rand_indx that compute_lookup (src/ultra_groth.cpp:62-106) fills in.
#include <omp.h>

#include <algorithm>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "cpu_field.hpp"

namespace {

struct SplitMix {
    u64 s;
    u64 next() {
        u64 z = (s += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    }
    Fr field() {  // four words, top nibble cleared (< 2^252 < r), then to Montgomery form
        Fr v;
        for (int i = 0; i < 4; i++) v.v[i] = next();
        v.v[3] &= 0x0fffffffffffffffull;
        return toMont(v);
    }
};

const Fr FR_ZERO = {{0, 0, 0, 0}};
inline Fr frU(u64 x) { return fromU64<FrP>(x); }
inline Fr frOne() { return one<FrP>(); }
Fr frPow(Fr b, u64 e) {
    Fr r = frOne();
    while (e) {
        if (e & 1) r = mul(r, b);
        b = sqr(b);
        e >>= 1;
    }
    return r;
}

struct Term { uint32_t row, sig; Fr val; };

struct Circuit {
    bool ultra = false;
    uint32_t nVars = 0, nPublic = 0, nConstraints = 0, rand_indx = 0;
    std::vector<Term> A, B, C;
    std::vector<Fr> witness;  // Montgomery form; the lookup signals stay zero (compute_lookup writes them)
    std::vector<uint32_t> round_idx, final_idx, chunks, freqs, wtns_idx, push_idx;
};

// m chain constraints on rows first_row.., x_i at signal first_sig + i (i < m), x_m at out_sig
void chain(Circuit& c, SplitMix& rng, uint32_t m, uint32_t first_sig, uint32_t out_sig, uint32_t first_row) {
    Fr x = rng.field();
    const Fr one = frOne();
    for (uint32_t i = 0; i < m; i++) {
        const uint32_t si = first_sig + i, sn = (i + 1 < m) ? first_sig + i + 1 : out_sig, row = first_row + i;
        c.witness[si] = x;
        const Fr a = frU(rng.next() & 0xffffffffull), k = frU(rng.next() & 0xffffffffull);
        c.A.push_back({row, si, one});
        c.A.push_back({row, 0, a});
        c.B.push_back({row, si, one});
        c.C.push_back({row, sn, one});
        c.C.push_back({row, 0, neg(k)});
        x = add(mul(add(x, a), x), k);
    }
    c.witness[out_sig] = x;
}

Circuit groth16_circuit(int logn, u64 seed) {
    if (logn < 2 || logn > 26) throw std::invalid_argument("synthetic circuit: logn must be in [2, 26]");
    Circuit c;
    const uint32_t n = 1u << logn;
    c.nPublic = 1;
    c.nConstraints = n - 2;
    c.nVars = n;
    c.witness.assign(n, FR_ZERO);
    c.witness[0] = frOne();
    SplitMix rng{seed};
    chain(c, rng, c.nConstraints, 2, 1, 0);
    return c;
}

Circuit ultra_circuit(int logn, u64 seed, int lookup_bits, uint32_t K) {
    Circuit c;
    c.ultra = true;
    const uint32_t n = 1u << logn, T = 1u << lookup_bits;
    c.nPublic = 2;
    c.rand_indx = 2;
    const u64 lookup_rows = (u64)K + 2ull * T + 1;
    if (lookup_rows + 4 > n) throw std::invalid_argument("synthetic UltraGroth circuit: lookup block does not fit the domain");
    const uint32_t m = n - 3 - (uint32_t)lookup_rows;
    c.nConstraints = n - 3;
    // signal map: 0 one | 1 out | 2 rand | chunks | freqs | chain x | inv1 | inv2 | prod
    const uint32_t s_out = 1, s_rand = 2, s_chunk = 3, s_freq = s_chunk + K, s_x = s_freq + T, s_inv1 = s_x + m, s_inv2 = s_inv1 + K,
                   s_prod = s_inv2 + T;
    c.nVars = s_prod + T;
    c.witness.assign(c.nVars, FR_ZERO);
    c.witness[0] = frOne();
    SplitMix rng{seed};
    chain(c, rng, m, s_x, s_out, 0);
    c.chunks.resize(K);
    c.freqs.assign(T, 0);
    for (uint32_t j = 0; j < K; j++) {
        const uint32_t v = (uint32_t)(rng.next() % T);
        c.chunks[j] = v;
        c.freqs[v]++;
        c.witness[s_chunk + j] = frU(v);
    }
    for (uint32_t i = 0; i < T; i++) c.witness[s_freq + i] = frU(c.freqs[i]);
    const Fr one = frOne();
    uint32_t row = m;
    for (uint32_t j = 0; j < K; j++, row++) {  // inv1_j * (c_j + rand) = 1
        c.A.push_back({row, s_inv1 + j, one});
        c.B.push_back({row, s_chunk + j, one});
        c.B.push_back({row, s_rand, one});
        c.C.push_back({row, 0, one});
    }
    for (uint32_t i = 0; i < T; i++, row++) {  // inv2_i * (rand + i) = 1
        c.A.push_back({row, s_inv2 + i, one});
        c.B.push_back({row, s_rand, one});
        if (i) c.B.push_back({row, 0, frU(i)});
        c.C.push_back({row, 0, one});
    }
    for (uint32_t i = 0; i < T; i++, row++) {  // f_i * inv2_i = prod_i
        c.A.push_back({row, s_freq + i, one});
        c.B.push_back({row, s_inv2 + i, one});
        c.C.push_back({row, s_prod + i, one});
    }
    // (sum_j inv1_j - sum_i prod_i) * 1 = 0: one wide row
    for (uint32_t j = 0; j < K; j++) c.A.push_back({row, s_inv1 + j, one});
    const Fr minus_one = neg(one);
    for (uint32_t i = 0; i < T; i++) c.A.push_back({row, s_prod + i, minus_one});
    c.B.push_back({row, 0, one});
    for (uint32_t s = s_chunk; s < s_x; s++) c.round_idx.push_back(s);   // committed in round 1: chunks and frequencies
    for (uint32_t s = s_x; s < c.nVars; s++) c.final_idx.push_back(s);   // everything else is private in the final round
    // push_vector = [rand | inv1 (K) | inv2 (T) | prod (T)]
    c.wtns_idx.push_back(s_rand); c.push_idx.push_back(0);
    for (uint32_t j = 0; j < K; j++) { c.wtns_idx.push_back(s_inv1 + j); c.push_idx.push_back(1 + j); }
    for (uint32_t i = 0; i < T; i++) { c.wtns_idx.push_back(s_inv2 + i); c.push_idx.push_back(1 + K + i); }
    for (uint32_t i = 0; i < T; i++) { c.wtns_idx.push_back(s_prod + i); c.push_idx.push_back(1 + K + T + i); }
    return c;
}

// ---- fixed-base scalar multiplication on the host: 16-bit windows, table[w][d] = d * 2^(16 w) * G (affine) ----
template <class F>
void batch_to_affine(const Jac<F>* in, size_t n, Aff<F>* out) {
    std::vector<F> pre(n);
    F acc = fone<F>();
    for (size_t i = 0; i < n; i++) {
        pre[i] = acc;
        if (!in[i].isInf()) acc = mul(acc, in[i].Z);
    }
    F iv = inv(acc);
    for (size_t i = n; i-- > 0;) {
        if (in[i].isInf()) { out[i] = {fzero<F>(), fzero<F>()}; continue; }
        F zi = mul(iv, pre[i]);
        iv = mul(iv, in[i].Z);
        F zi2 = sqr(zi);
        out[i] = {mul(in[i].X, zi2), mul(in[i].Y, mul(zi2, zi))};
    }
}

template <class F>
struct FixedBase {
    static constexpr int WB = 16, NW = 16;
    std::vector<Aff<F>> tab;  // [NW][2^WB]
    FixedBase(const Aff<F>& gen, int threads) {
        tab.resize((size_t)NW << WB);
        std::vector<Jac<F>> bases(NW);
        Jac<F> b = {gen.x, gen.y, fone<F>()};
        for (int w = 0; w < NW; w++) {
            bases[w] = b;
            for (int k = 0; k < WB; k++) b = dbl(b);
        }
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
        for (int w = 0; w < NW; w++) {
            std::vector<Jac<F>> row((size_t)1 << WB);
            row[0] = jinf<F>();
            Aff<F> ba = toAff(bases[w]);
            for (size_t d = 1; d < row.size(); d++) row[d] = addm(row[d - 1], ba);
            batch_to_affine(row.data(), row.size(), &tab[(size_t)w << WB]);
        }
    }
    // scalars in Montgomery form -> affine points, zkey layout
    void mul_all(const std::vector<Fr>& scalars_mont, std::vector<uint8_t>& out, int threads) const {
        const size_t n = scalars_mont.size();
        out.assign(n * sizeof(Aff<F>), 0);
        Aff<F>* o = (Aff<F>*)out.data();
        const size_t CH = 512;
#pragma omp parallel for schedule(dynamic, 8) num_threads(threads)
        for (size_t c0 = 0; c0 < n; c0 += CH) {
            const size_t cnt = std::min(CH, n - c0);
            Jac<F> acc[CH];
            for (size_t i = 0; i < cnt; i++) {
                Fr raw = fromMont(scalars_mont[c0 + i]);
                Jac<F> a = jinf<F>();
                for (int w = 0; w < NW; w++) {
                    const uint32_t d = (uint32_t)(raw.v[w >> 2] >> (16 * (w & 3))) & 0xffffu;
                    if (d) a = addm(a, tab[((size_t)w << WB) + d]);
                }
                acc[i] = a;
            }
            batch_to_affine(acc, cnt, o + c0);
        }
    }
};

Aff<Fq> g1_generator() { return {toMont(Fq{{1, 0, 0, 0}}), toMont(Fq{{2, 0, 0, 0}})}; }
Aff<Fq2> g2_generator() {  // the standard BN254 G2 generator (EIP-197)
    Aff<Fq2> g;
    g.x.a = toMont(Fq{{0x46debd5cd992f6edull, 0x674322d4f75edaddull, 0x426a00665e5c4479ull, 0x1800deef121f1e76ull}});
    g.x.b = toMont(Fq{{0x97e485b7aef312c2ull, 0xf1aa493335a9e712ull, 0x7260bfb731fb5d25ull, 0x198e9393920d483aull}});
    g.y.a = toMont(Fq{{0x4ce6cc0166fa7daaull, 0xe3d1e7690c43d37bull, 0x4aab71808dcb408full, 0x12c85ea5db8c6debull}});
    g.y.b = toMont(Fq{{0x55acdadcd122975bull, 0xbc4b313370b38ef3ull, 0xec9e99ad690c3395ull, 0x090689d0585ff075ull}});
    return g;
}

// ---- serialisation ----
void put32(std::vector<uint8_t>& v, uint32_t x) { uint8_t b[4]; memcpy(b, &x, 4); v.insert(v.end(), b, b + 4); }
void put64(std::vector<uint8_t>& v, u64 x) { uint8_t b[8]; memcpy(b, &x, 8); v.insert(v.end(), b, b + 8); }
void putb(std::vector<uint8_t>& v, const void* p, size_t n) { v.insert(v.end(), (const uint8_t*)p, (const uint8_t*)p + n); }
struct Container {
    std::vector<uint8_t> out;
    Container(const char* magic, uint32_t version, uint32_t nsections) { putb(out, magic, 4); put32(out, version); put32(out, nsections); }
    void add(uint32_t id, const std::vector<uint8_t>& body) { put32(out, id); put64(out, body.size()); putb(out, body.data(), body.size()); }
};

std::string dec256(const u64* limbs) {
    u64 v[4] = {limbs[0], limbs[1], limbs[2], limbs[3]};
    std::string s;
    while (v[0] | v[1] | v[2] | v[3]) {
        u64 rem = 0;
        for (int i = 3; i >= 0; i--) {
            u128 cur = ((u128)rem << 64) | v[i];
            v[i] = (u64)(cur / 10);
            rem = (u64)(cur % 10);
        }
        s.push_back((char)('0' + rem));
    }
    if (s.empty()) s = "0";
    std::reverse(s.begin(), s.end());
    return s;
}
std::string fq_dec(const Fq& m) { Fq r = fromMont(m); return dec256(r.v); }
std::string g1_dec(const uint8_t* p) {
    Aff<Fq> a;
    memcpy(&a, p, 64);
    if (a.isInf()) return "[\"0\",\"1\",\"0\"]";
    return "[\"" + fq_dec(a.x) + "\",\"" + fq_dec(a.y) + "\",\"1\"]";
}
std::string g2_dec(const uint8_t* p) {
    Aff<Fq2> a;
    memcpy(&a, p, 128);
    return "[[\"" + fq_dec(a.x.a) + "\",\"" + fq_dec(a.x.b) + "\"],[\"" + fq_dec(a.y.a) + "\",\"" + fq_dec(a.y.b) + "\"],[\"1\",\"0\"]]";
}

// scaled Lagrange-type basis: out[i] = k * g_i / (tau - g_i) with g_i = g0 * w^i, one shared inversion
void basis_at_tau(const Fr& tau, const Fr& g0, const Fr& w, const Fr& k, uint32_t n, std::vector<Fr>& out) {
    std::vector<Fr> g(n), den(n), pre(n);
    Fr cur = g0, acc = frOne();
    for (uint32_t i = 0; i < n; i++) {
        g[i] = cur;
        den[i] = sub(tau, cur);
        pre[i] = acc;
        acc = mul(acc, den[i]);
        cur = mul(cur, w);
    }
    Fr iv = inv(acc);
    out.resize(n);
    for (uint32_t i = n; i-- > 0;) {
        Fr di = mul(iv, pre[i]);
        iv = mul(iv, den[i]);
        out[i] = mul(mul(k, g[i]), di);
    }
}

void synth(int ultra, int logn, u64 seed, int lookup_bits, uint32_t n_chunks, int threads, std::vector<uint8_t>& zkey, std::vector<uint8_t>& wtns,
           std::string& vk_json) {
    Circuit c = ultra ? ultra_circuit(logn, seed, lookup_bits, n_chunks) : groth16_circuit(logn, seed);
    const uint32_t n = 1u << logn, nV = c.nVars, nP = c.nPublic, m = c.nConstraints;
    SplitMix tox{seed ^ 0x70c51cull};
    const Fr tau = tox.field(), alpha = tox.field(), beta = tox.field(), gamma = tox.field(), delta_f = tox.field(), delta_r = tox.field();

    // L_c(tau) = (tau^n - 1)/n * w^c / (tau - w^c)
    std::vector<Fr> L;
    basis_at_tau(tau, frOne(), rootOfUnity(logn), mul(sub(frPow(tau, n), frOne()), inv(frU(n))), n, L);
    std::vector<Fr> At(nV, FR_ZERO), Bt(nV, FR_ZERO), Ct(nV, FR_ZERO);
    for (const Term& t : c.A) At[t.sig] = add(At[t.sig], mul(t.val, L[t.row]));
    for (const Term& t : c.B) Bt[t.sig] = add(Bt[t.sig], mul(t.val, L[t.row]));
    for (const Term& t : c.C) Ct[t.sig] = add(Ct[t.sig], mul(t.val, L[t.row]));
    for (uint32_t i = 0; i <= nP; i++) At[i] = add(At[i], L[m + i]);  // snarkjs' public-input rows
    const Fr inv_gamma = inv(gamma), inv_df = inv(delta_f), inv_dr = inv(delta_r);
    std::vector<Fr> Kq(nV);
    for (uint32_t s = 0; s < nV; s++) Kq[s] = add(add(mul(beta, At[s]), mul(alpha, Bt[s])), Ct[s]);
    std::vector<Fr> ic(nP + 1), cq, c1q;
    for (uint32_t s = 0; s <= nP; s++) ic[s] = mul(Kq[s], inv_gamma);
    if (!c.ultra) {
        cq.resize(nV - nP - 1);
        for (uint32_t s = nP + 1; s < nV; s++) cq[s - nP - 1] = mul(Kq[s], inv_df);
    } else {
        c1q.resize(c.round_idx.size());
        for (size_t i = 0; i < c.round_idx.size(); i++) c1q[i] = mul(Kq[c.round_idx[i]], inv_dr);
        cq.resize(c.final_idx.size());
        for (size_t i = 0; i < c.final_idx.size(); i++) cq[i] = mul(Kq[c.final_idx[i]], inv_df);
    }
    // H_i = (tau^2n - 1) * g_i / (2 n delta_f (tau - g_i)),  g_i = omega_2n^(2i+1)
    std::vector<Fr> hq;
    {
        const Fr w2 = rootOfUnity(logn + 1);
        basis_at_tau(tau, w2, sqr(w2), mul(sub(frPow(tau, 2ull * n), frOne()), inv(mul(frU(2ull * n), delta_f))), n, hq);
    }

    FixedBase<Fq> fb1(g1_generator(), threads);
    FixedBase<Fq2> fb2(g2_generator(), threads);
    std::vector<uint8_t> h1, h2, pA, pB1, pB2, pC, pH, pIC, pC1;
    fb1.mul_all({alpha, beta, delta_f, delta_r}, h1, threads);
    fb2.mul_all({beta, gamma, delta_f, delta_r}, h2, threads);
    fb1.mul_all(At, pA, threads);
    fb1.mul_all(Bt, pB1, threads);
    fb2.mul_all(Bt, pB2, threads);
    fb1.mul_all(cq, pC, threads);
    fb1.mul_all(hq, pH, threads);
    fb1.mul_all(ic, pIC, threads);
    if (c.ultra) fb1.mul_all(c1q, pC1, threads);

    // ---- zkey ----
    Container zk("zkey", 1, c.ultra ? 13 : 10);
    { std::vector<uint8_t> s; put32(s, c.ultra ? 1337 : 1); zk.add(1, s); }
    {
        std::vector<uint8_t> s;
        put32(s, 32); putb(s, FqP::q, 32);
        put32(s, 32); putb(s, FrP::q, 32);
        put32(s, nV); put32(s, nP); put32(s, n);
        if (c.ultra) { put32(s, (uint32_t)c.round_idx.size()); put32(s, (uint32_t)c.final_idx.size()); put32(s, c.rand_indx); }
        putb(s, &h1[0], 64);     // alpha1
        putb(s, &h1[64], 64);    // beta1
        putb(s, &h2[0], 128);    // beta2
        putb(s, &h2[128], 128);  // gamma2
        if (c.ultra) { putb(s, &h1[192], 64); putb(s, &h2[384], 128); }  // round delta1 / delta2
        putb(s, &h1[128], 64);   // (final) delta1
        putb(s, &h2[256], 128);  // (final) delta2
        zk.add(2, s);
    }
    zk.add(3, pIC);
    {
        std::vector<uint8_t> s;
        const uint32_t ncoef = (uint32_t)(c.A.size() + c.B.size() + nP + 1);
        s.reserve(4 + (size_t)ncoef * 44);
        put32(s, ncoef);
        auto rec = [&](uint32_t mtx, uint32_t row, uint32_t sig, const Fr& val_mont) {
            put32(s, mtx); put32(s, row); put32(s, sig);
            Fr v = toMont(val_mont);  // the zkey stores value * R^2
            putb(s, &v, 32);
        };
        for (const Term& t : c.A) rec(0, t.row, t.sig, t.val);
        for (const Term& t : c.B) rec(1, t.row, t.sig, t.val);
        for (uint32_t i = 0; i <= nP; i++) rec(0, m + i, i, frOne());
        zk.add(4, s);
    }
    zk.add(5, pA);
    zk.add(6, pB1);
    zk.add(7, pB2);
    std::vector<uint8_t> contrib(64, 0);
    put32(contrib, 0);
    if (!c.ultra) {
        zk.add(8, pC);
        zk.add(9, pH);
        zk.add(10, contrib);
    } else {
        zk.add(8, pC1);
        zk.add(9, pC);
        std::vector<uint8_t> s;
        for (uint32_t v : c.round_idx) put32(s, v);
        zk.add(10, s);
        s.clear();
        for (uint32_t v : c.final_idx) put32(s, v);
        zk.add(11, s);
        zk.add(12, pH);
        zk.add(13, contrib);
    }
    zkey.swap(zk.out);

    // ---- wtns / .uwtns ----
    Container wt("wtns", 2, c.ultra ? 6 : 2);
    { std::vector<uint8_t> s; put32(s, 32); putb(s, FrP::q, 32); put32(s, nV); wt.add(1, s); }
    {
        std::vector<uint8_t> s((size_t)nV * 32);
        for (uint32_t i = 0; i < nV; i++) { Fr raw = fromMont(c.witness[i]); memcpy(&s[(size_t)i * 32], &raw, 32); }
        wt.add(2, s);
    }
    if (c.ultra) {
        auto u32sec = [&](uint32_t id, const std::vector<uint32_t>& v) {
            std::vector<uint8_t> s(v.size() * 4);
            if (!v.empty()) memcpy(s.data(), v.data(), s.size());
            wt.add(id, s);
        };
        u32sec(3, c.chunks);
        u32sec(4, c.freqs);
        u32sec(5, c.wtns_idx);
        u32sec(6, c.push_idx);
    }
    wtns.swap(wt.out);

    // ---- verification key ----
    std::string vk = "{\"protocol\":\"";
    vk += c.ultra ? "ultragroth" : "groth16";
    vk += "\",\"curve\":\"bn128\",\"nPublic\":" + std::to_string(c.ultra ? nP - 1 : nP);
    vk += ",\"vk_alpha_1\":" + g1_dec(&h1[0]);
    vk += ",\"vk_beta_2\":" + g2_dec(&h2[0]);
    vk += ",\"vk_gamma_2\":" + g2_dec(&h2[128]);
    if (!c.ultra) {
        vk += ",\"vk_delta_2\":" + g2_dec(&h2[256]);
    } else {
        vk += ",\"vk_delta_c2_2\":" + g2_dec(&h2[256]);
        vk += ",\"vk_delta_c1_2\":" + g2_dec(&h2[384]);
        vk += ",\"IC_rand\":" + g1_dec(&pIC[(size_t)c.rand_indx * 64]);
        vk += ",\"randIdx\":" + std::to_string(c.rand_indx);
    }
    vk += ",\"IC\":[";
    bool first = true;
    for (uint32_t s = 0; s <= nP; s++) {
        if (c.ultra && s == c.rand_indx) continue;
        if (!first) vk += ",";
        vk += g1_dec(&pIC[(size_t)s * 64]);
        first = false;
    }
    vk += "]}";
    vk_json.swap(vk);
}

}  // namespace

extern "C" {

// buffers are malloc'ed here; release with cpu_free
int cpu_synth_build(int ultra, int logn, u64 seed, int lookup_bits, uint32_t n_chunks, int threads, uint8_t** zkey, u64* zkey_size,
                    uint8_t** wtns, u64* wtns_size, char** vk_json) {
    try {
        if (threads <= 0) threads = omp_get_max_threads();
        std::vector<uint8_t> zk, wt;
        std::string vk;
        synth(ultra, logn, seed, lookup_bits, n_chunks, threads, zk, wt, vk);
        *zkey = (uint8_t*)malloc(zk.size());
        *wtns = (uint8_t*)malloc(wt.size());
        *vk_json = (char*)malloc(vk.size() + 1);
        if (!*zkey || !*wtns || !*vk_json) return 2;
        memcpy(*zkey, zk.data(), zk.size());
        memcpy(*wtns, wt.data(), wt.size());
        memcpy(*vk_json, vk.c_str(), vk.size() + 1);
        *zkey_size = zk.size();
        *wtns_size = wt.size();
        return 0;
    } catch (...) {
        return 1;
    }
}
void cpu_free(void* p) { free(p); }

}  // extern "C"
