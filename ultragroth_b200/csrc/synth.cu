// Synthetic circuits + trusted setup: builds a VALID zkey / witness /
// verification key for the circuit families the benchmarks are quoted on, so that proofs at 2^20..2^24
// constraints verify under the reference's verifier equation.  The reference has no setup tool (zkeys
// come from snarkjs); this is harness tooling that lives in the library because the ~5n fixed-base
// scalar multiplications need the GPU.
//
// Groth16 family (protocol 1):  (x_i + a_i) * x_i = x_{i+1} - k_i,  i < m,  one public output x_m.
// UltraGroth family (protocol 1337, layout of src/zkey_utils.cpp:123-163 and src/prover.cpp:237-292):
//   the same chain plus a LogUp range lookup: K chunk signals c_j in [0,T), T frequencies f_i,
//   inv1_j*(c_j + rand) = 1, inv2_i*(rand + i) = 1, prod_i = f_i*inv2_i, sum_j inv1_j = sum_i prod_i,
//   with rand the public signal at rand_indx that compute_lookup (src/ultra_groth.cpp:62-106) fills in.
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "curve.cuh"
#include "device.hpp"
#include "host_ff.hpp"
#include "memops.cuh"
#include "msm.hpp"
#include "ntt.hpp"

namespace ug {

// ---- fixed-base scalar multiplication: out[i] = scalar[i] * G ------------------------------------
// table[w*256 + d] = d * 2^(8w) * G (affine, d = 0 is infinity); 32 mixed additions per scalar.
template <class F>
__global__ void __launch_bounds__(128) k_fixed_base_mul(const Affine<F>* __restrict__ table, const uint8_t* __restrict__ scalars,
                                                        uint64_t n, Affine<F>* __restrict__ out) {
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* sp = reinterpret_cast<const uint4*>(scalars + i * 32);
    uint4 lo = __ldg(sp), hi = __ldg(sp + 1);
    uint32_t s[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = 0; w < 32; w++) {
        uint32_t d = (s[w >> 2] >> (8 * (w & 3))) & 0xffu;
        if (d) {
            Affine<F> p = ld_nc(&table[w * 256 + d]);
            pt_madd(acc, p, false);
        }
    }
    st_g(&out[i], pt_to_affine(acc));
}

template <class HF>
static std::vector<Affine<HF>> build_table(const Affine<HF>& gen) {
    std::vector<Affine<HF>> tab(32 * 256);
    XYZZ<HF> base = XYZZ<HF>::from_affine(gen);
    std::vector<XYZZ<HF>> row(256);
    for (int w = 0; w < 32; w++) {
        row[0] = XYZZ<HF>::inf();
        for (int d = 1; d < 256; d++) { row[d] = row[d - 1]; pt_add(row[d], base); }
        for (int d = 0; d < 256; d++) tab[w * 256 + d] = pt_to_affine(row[d]);
        for (int k = 0; k < 8; k++) base = pt_dbl(base);
    }
    return tab;
}

static HG1Affine g1_generator() {
    HG1Affine g;
    g.x = to_mont(HFq{{1, 0, 0, 0}});
    g.y = to_mont(HFq{{2, 0, 0, 0}});
    return g;
}
static HG2Affine g2_generator() {
    // the standard BN254 G2 generator (EIP-197), coordinates as little-endian 64-bit limbs
    HG2Affine g;
    g.x.a = to_mont(HFq{{0x46debd5cd992f6edull, 0x674322d4f75edaddull, 0x426a00665e5c4479ull, 0x1800deef121f1e76ull}});
    g.x.b = to_mont(HFq{{0x97e485b7aef312c2ull, 0xf1aa493335a9e712ull, 0x7260bfb731fb5d25ull, 0x198e9393920d483aull}});
    g.y.a = to_mont(HFq{{0x4ce6cc0166fa7daaull, 0xe3d1e7690c43d37bull, 0x4aab71808dcb408full, 0x12c85ea5db8c6debull}});
    g.y.b = to_mont(HFq{{0x55acdadcd122975bull, 0xbc4b313370b38ef3ull, 0xec9e99ad690c3395ull, 0x090689d0585ff075ull}});
    return g;
}

class FixedBase {
public:
    explicit FixedBase(DeviceCtx& ctx) : ctx_(ctx) {
        ctx.activate();
        auto t1 = build_table<HFq>(g1_generator());
        auto t2 = build_table<HFq2>(g2_generator());
        UG_CUDA_CHECK(cudaMalloc(&tab1_, t1.size() * sizeof(HG1Affine)));
        UG_CUDA_CHECK(cudaMalloc(&tab2_, t2.size() * sizeof(HG2Affine)));
        UG_CUDA_CHECK(cudaMemcpy(tab1_, t1.data(), t1.size() * sizeof(HG1Affine), cudaMemcpyHostToDevice));
        UG_CUDA_CHECK(cudaMemcpy(tab2_, t2.data(), t2.size() * sizeof(HG2Affine), cudaMemcpyHostToDevice));
    }
    ~FixedBase() { cudaFree(tab1_); cudaFree(tab2_); }
    // scalars: Montgomery-form HFr values; out: n affine points (64 / 128 bytes each) on the host
    void mul(int group, const HFr* scalars_mont, uint64_t n, uint8_t* out) {
        if (n == 0) return;
        ctx_.activate();
        std::vector<HFr> raw(n);
        for (uint64_t i = 0; i < n; i++) raw[i] = from_mont(scalars_mont[i]);
        const size_t psz = group == 1 ? 64 : 128;
        const uint64_t chunk = 1u << 22;
        void* d_s = ctx_.workspace("synth.scalars", (size_t)std::min(n, chunk) * 32);
        void* d_o = ctx_.workspace("synth.points", (size_t)std::min(n, chunk) * psz);
        for (uint64_t off = 0; off < n; off += chunk) {
            uint64_t cnt = std::min(chunk, n - off);
            UG_CUDA_CHECK(cudaMemcpyAsync(d_s, raw.data() + off, cnt * 32, cudaMemcpyHostToDevice, ctx_.stream));
            uint32_t blocks = (uint32_t)((cnt + 127) / 128);
            if (group == 1) k_fixed_base_mul<Fq><<<blocks, 128, 0, ctx_.stream>>>((const G1Affine*)tab1_, (const uint8_t*)d_s, cnt, (G1Affine*)d_o);
            else k_fixed_base_mul<Fq2><<<blocks, 128, 0, ctx_.stream>>>((const G2Affine*)tab2_, (const uint8_t*)d_s, cnt, (G2Affine*)d_o);
            UG_LAUNCH_CHECK(ctx_);
            UG_CUDA_CHECK(cudaMemcpyAsync(out + off * psz, d_o, cnt * psz, cudaMemcpyDeviceToHost, ctx_.stream));
            UG_CUDA_CHECK(cudaStreamSynchronize(ctx_.stream));
        }
    }

private:
    DeviceCtx& ctx_;
    void* tab1_ = nullptr;
    void* tab2_ = nullptr;
};

// ---- host-side circuit / setup ---------------------------------------------------------------------
struct Rng {  // splitmix64
    uint64_t s;
    uint64_t next() {
        uint64_t z = (s += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    }
    HFr field() {  // uniform-ish element of Fr, Montgomery form
        HFr v;
        for (int i = 0; i < 4; i++) v.v[i] = next();
        v.v[3] &= 0x0fffffffffffffffull;  // < 2^252 < r
        return to_mont(v);
    }
};

struct Term { uint32_t row, sig; HFr val; };  // val in Montgomery form

struct Circuit {
    uint32_t nVars = 0, nPublic = 0, nConstraints = 0;
    std::vector<Term> A, B, C;
    std::vector<HFr> witness;  // Montgomery form; rand-dependent signals left at zero
    // UltraGroth bookkeeping
    bool ultra = false;
    uint32_t rand_indx = 0;
    std::vector<uint32_t> round_idx, final_idx, chunks, freqs, wtns_idx, push_idx;
};

static void chain_body(Circuit& c, Rng& rng, uint32_t m, uint32_t first_sig, uint32_t out_sig, uint32_t first_row) {
    // x_0 = witness[first_sig], x_i at first_sig + i (i < m), x_m at out_sig
    HFr x = rng.field();
    auto sig = [&](uint32_t i) { return i < m ? first_sig + i : out_sig; };
    const HFr one = HFr::one();
    for (uint32_t i = 0; i < m; i++) {
        c.witness[sig(i)] = x;
        HFr a = hfr_from_u64(rng.next() & 0xffffffffull), k = hfr_from_u64(rng.next() & 0xffffffffull);
        uint32_t row = first_row + i;
        c.A.push_back({row, sig(i), one});
        c.A.push_back({row, 0, a});
        c.B.push_back({row, sig(i), one});
        c.C.push_back({row, sig(i + 1), one});
        c.C.push_back({row, 0, fneg(k)});
        x = fadd(fmul(fadd(x, a), x), k);
    }
    c.witness[out_sig] = x;
}

static Circuit make_groth16_circuit(int logn, uint64_t seed) {
    if (logn < 2 || logn > 26) throw std::invalid_argument("synthetic circuit: logn must be in [2, 26]");
    Circuit c;
    const uint32_t n = 1u << logn;
    c.nPublic = 1;
    c.nConstraints = n - 2;
    c.nVars = n;
    c.witness.assign(c.nVars, HFr::zero());
    c.witness[0] = HFr::one();
    Rng rng{seed};
    chain_body(c, rng, c.nConstraints, 2, 1, 0);
    return c;
}

static Circuit make_ultra_circuit(int logn, uint64_t seed, int lookup_bits, uint32_t K) {
    Circuit c;
    c.ultra = true;
    const uint32_t n = 1u << logn;
    const uint32_t T = 1u << lookup_bits;
    c.nPublic = 2;
    c.rand_indx = 2;
    const uint64_t lookup_rows = (uint64_t)K + 2ull * T + 1;
    if (lookup_rows + 3 + 1 > n) throw std::invalid_argument("synthetic UltraGroth circuit: lookup block does not fit the domain");
    const uint32_t m = n - 3 - (uint32_t)lookup_rows;  // chain constraints
    c.nConstraints = n - 3;
    // signal map
    const uint32_t s_out = 1, s_rand = 2;
    const uint32_t s_chunk = 3, s_freq = s_chunk + K, s_x = s_freq + T, s_inv1 = s_x + m, s_inv2 = s_inv1 + K, s_prod = s_inv2 + T;
    c.nVars = s_prod + T;
    c.witness.assign(c.nVars, HFr::zero());
    c.witness[0] = HFr::one();
    Rng rng{seed};
    chain_body(c, rng, m, s_x, s_out, 0);
    c.chunks.resize(K);
    c.freqs.assign(T, 0);
    for (uint32_t j = 0; j < K; j++) {
        uint32_t v = (uint32_t)(rng.next() % T);
        c.chunks[j] = v;
        c.freqs[v]++;
        c.witness[s_chunk + j] = hfr_from_u64(v);
    }
    for (uint32_t i = 0; i < T; i++) c.witness[s_freq + i] = hfr_from_u64(c.freqs[i]);
    const HFr one = HFr::one();
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
        if (i) c.B.push_back({row, 0, hfr_from_u64(i)});
        c.C.push_back({row, 0, one});
    }
    for (uint32_t i = 0; i < T; i++, row++) {  // f_i * inv2_i = prod_i
        c.A.push_back({row, s_freq + i, one});
        c.B.push_back({row, s_inv2 + i, one});
        c.C.push_back({row, s_prod + i, one});
    }
    // (sum_j inv1_j - sum_i prod_i) * 1 = 0 : one wide row, every term is written by compute_lookup
    for (uint32_t j = 0; j < K; j++) c.A.push_back({row, s_inv1 + j, one});
    const HFr minus_one = fneg(one);
    for (uint32_t i = 0; i < T; i++) c.A.push_back({row, s_prod + i, minus_one});
    c.B.push_back({row, 0, one});
    row++;
    // commitment rounds: chunks and frequencies first, everything else private in the final round
    for (uint32_t s = s_chunk; s < s_x; s++) c.round_idx.push_back(s);
    for (uint32_t s = s_x; s < c.nVars; s++) c.final_idx.push_back(s);
    // push_vector = [rand | inv1 (K) | inv2 (T) | prod (T)]
    c.wtns_idx.push_back(s_rand); c.push_idx.push_back(0);
    for (uint32_t j = 0; j < K; j++) { c.wtns_idx.push_back(s_inv1 + j); c.push_idx.push_back(1 + j); }
    for (uint32_t i = 0; i < T; i++) { c.wtns_idx.push_back(s_inv2 + i); c.push_idx.push_back(1 + K + i); }
    for (uint32_t i = 0; i < T; i++) { c.wtns_idx.push_back(s_prod + i); c.push_idx.push_back(1 + K + T + i); }
    return c;
}

static void put_u32(std::vector<uint8_t>& v, uint32_t x) { uint8_t b[4]; memcpy(b, &x, 4); v.insert(v.end(), b, b + 4); }
static void put_u64(std::vector<uint8_t>& v, uint64_t x) { uint8_t b[8]; memcpy(b, &x, 8); v.insert(v.end(), b, b + 8); }
static void put_bytes(std::vector<uint8_t>& v, const void* p, size_t n) { v.insert(v.end(), (const uint8_t*)p, (const uint8_t*)p + n); }

struct SectionWriter {
    std::vector<uint8_t> out;
    uint32_t count = 0;
    SectionWriter(const char* magic, uint32_t version, uint32_t nsections) {
        put_bytes(out, magic, 4);
        put_u32(out, version);
        put_u32(out, nsections);
    }
    void add(uint32_t id, const std::vector<uint8_t>& body) {
        put_u32(out, id);
        put_u64(out, body.size());
        put_bytes(out, body.data(), body.size());
        count++;
    }
};

static std::string g1_dec(const uint8_t* p) {
    HG1Affine a;
    memcpy(&a, p, 64);
    if (a.is_inf()) return "[\"0\",\"1\",\"0\"]";
    return "[\"" + to_dec(a.x) + "\",\"" + to_dec(a.y) + "\",\"1\"]";
}
static std::string g2_dec(const uint8_t* p) {
    HG2Affine a;
    memcpy(&a, p, 128);
    return "[[\"" + to_dec(a.x.a) + "\",\"" + to_dec(a.x.b) + "\"],[\"" + to_dec(a.y.a) + "\",\"" + to_dec(a.y.b) + "\"],[\"1\",\"0\"]]";
}

void synth_build(DeviceCtx& ctx, int ultra, int logn, uint64_t seed, int lookup_bits, uint32_t n_chunks, std::vector<uint8_t>& zkey,
                 std::vector<uint8_t>& wtns, std::string& vk_json) {
    Circuit c = ultra ? make_ultra_circuit(logn, seed, lookup_bits, n_chunks) : make_groth16_circuit(logn, seed);
    const uint32_t n = 1u << logn;
    const uint32_t nV = c.nVars, nP = c.nPublic, m = c.nConstraints;
    Rng tox{seed ^ 0x70c51cull};
    const HFr tau = tox.field(), alpha = tox.field(), beta = tox.field(), gamma = tox.field(), delta_f = tox.field(), delta_r = tox.field();

    // Lagrange basis at tau over the size-n domain: L_c = (tau^n - 1)/n * w^c / (tau - w^c), batch inverted
    std::vector<HFr> L(n);
    {
        const HFr w = fr_root_of_unity(logn);
        std::vector<HFr> wc(n), den(n), pre(n);
        HFr cur = HFr::one(), acc = HFr::one();
        for (uint32_t i = 0; i < n; i++) {
            wc[i] = cur;
            den[i] = fsub(tau, cur);
            pre[i] = acc;
            acc = fmul(acc, den[i]);
            cur = fmul(cur, w);
        }
        HFr inv = finv(acc);
        HFr tn = hfr_pow(tau, n);
        HFr k = fmul(fsub(tn, HFr::one()), finv(hfr_from_u64(n)));
        for (uint32_t i = n; i-- > 0;) {
            HFr di = fmul(inv, pre[i]);
            inv = fmul(inv, den[i]);
            L[i] = fmul(fmul(k, wc[i]), di);
        }
    }
    // per-signal polynomial evaluations at tau (snarkjs adds the public-input rows to A)
    std::vector<HFr> At(nV, HFr::zero()), Bt(nV, HFr::zero()), Ct(nV, HFr::zero());
    for (const Term& t : c.A) At[t.sig] = fadd(At[t.sig], fmul(t.val, L[t.row]));
    for (const Term& t : c.B) Bt[t.sig] = fadd(Bt[t.sig], fmul(t.val, L[t.row]));
    for (const Term& t : c.C) Ct[t.sig] = fadd(Ct[t.sig], fmul(t.val, L[t.row]));
    for (uint32_t i = 0; i <= nP; i++) At[i] = fadd(At[i], L[m + i]);
    // query scalars
    const HFr inv_gamma = finv(gamma), inv_df = finv(delta_f), inv_dr = finv(delta_r);
    std::vector<HFr> K(nV);
    for (uint32_t s = 0; s < nV; s++) K[s] = fadd(fadd(fmul(beta, At[s]), fmul(alpha, Bt[s])), Ct[s]);
    std::vector<HFr> ic(nP + 1), cq, c1q;
    for (uint32_t s = 0; s <= nP; s++) ic[s] = fmul(K[s], inv_gamma);
    if (!c.ultra) {
        cq.resize(nV - nP - 1);
        for (uint32_t s = nP + 1; s < nV; s++) cq[s - nP - 1] = fmul(K[s], inv_df);
    } else {
        c1q.resize(c.round_idx.size());
        for (size_t i = 0; i < c.round_idx.size(); i++) c1q[i] = fmul(K[c.round_idx[i]], inv_dr);
        cq.resize(c.final_idx.size());
        for (size_t i = 0; i < c.final_idx.size(); i++) cq[i] = fmul(K[c.final_idx[i]], inv_df);
    }
    // H_i = (tau^2n - 1) * g_i / (2 n delta_f (tau - g_i)),  g_i = omega_2n^(2i+1)   (odd-coset Lagrange basis times Z/delta)
    std::vector<HFr> hq(n);
    {
        const HFr w2 = fr_root_of_unity(logn + 1), w = fsqr(w2);
        std::vector<HFr> g(n), den(n), pre(n);
        HFr cur = w2, acc = HFr::one();
        for (uint32_t i = 0; i < n; i++) {
            g[i] = cur;
            den[i] = fsub(tau, cur);
            pre[i] = acc;
            acc = fmul(acc, den[i]);
            cur = fmul(cur, w);
        }
        HFr inv = finv(acc);
        HFr t2n = hfr_pow(tau, 2ull * n);
        HFr k = fmul(fsub(t2n, HFr::one()), finv(fmul(hfr_from_u64(2ull * n), delta_f)));
        for (uint32_t i = n; i-- > 0;) {
            HFr di = fmul(inv, pre[i]);
            inv = fmul(inv, den[i]);
            hq[i] = fmul(fmul(k, g[i]), di);
        }
    }

    FixedBase fb(ctx);
    auto g1pts = [&](const std::vector<HFr>& s) { std::vector<uint8_t> o(s.size() * 64); fb.mul(1, s.data(), s.size(), o.data()); return o; };
    auto g2pts = [&](const std::vector<HFr>& s) { std::vector<uint8_t> o(s.size() * 128); fb.mul(2, s.data(), s.size(), o.data()); return o; };
    std::vector<HFr> hdr1 = {alpha, beta, delta_f, delta_r};
    std::vector<HFr> hdr2 = {beta, gamma, delta_f, delta_r};
    std::vector<uint8_t> h1 = g1pts(hdr1), h2 = g2pts(hdr2);
    std::vector<uint8_t> pA = g1pts(At), pB1 = g1pts(Bt), pB2 = g2pts(Bt), pC = g1pts(cq), pH = g1pts(hq), pIC = g1pts(ic), pC1;
    if (c.ultra) pC1 = g1pts(c1q);

    // ---- zkey ----
    std::vector<uint8_t> q_le(32), r_le(32);
    memcpy(q_le.data(), HFqParams::mod, 32);
    memcpy(r_le.data(), HFrParams::mod, 32);
    SectionWriter zk("zkey", 1, c.ultra ? 13 : 10);
    { std::vector<uint8_t> s; put_u32(s, c.ultra ? 1337 : 1); zk.add(1, s); }
    {
        std::vector<uint8_t> s;
        put_u32(s, 32); put_bytes(s, q_le.data(), 32);
        put_u32(s, 32); put_bytes(s, r_le.data(), 32);
        put_u32(s, nV); put_u32(s, nP); put_u32(s, n);
        if (c.ultra) { put_u32(s, (uint32_t)c.round_idx.size()); put_u32(s, (uint32_t)c.final_idx.size()); put_u32(s, c.rand_indx); }
        put_bytes(s, &h1[0], 64);        // alpha1
        put_bytes(s, &h1[64], 64);       // beta1
        put_bytes(s, &h2[0], 128);       // beta2
        put_bytes(s, &h2[128], 128);     // gamma2
        if (c.ultra) { put_bytes(s, &h1[192], 64); put_bytes(s, &h2[384], 128); }  // round delta1 / delta2
        put_bytes(s, &h1[128], 64);      // (final) delta1
        put_bytes(s, &h2[256], 128);     // (final) delta2
        zk.add(2, s);
    }
    zk.add(3, pIC);
    {
        std::vector<uint8_t> s;
        uint32_t ncoef = (uint32_t)(c.A.size() + c.B.size() + nP + 1);
        s.reserve(4 + (size_t)ncoef * 44);
        put_u32(s, ncoef);
        auto rec = [&](uint32_t mtx, uint32_t row, uint32_t sig, const HFr& val_mont) {
            put_u32(s, mtx); put_u32(s, row); put_u32(s, sig);
            HFr v = to_mont(val_mont);  // value * R^2
            put_bytes(s, &v, 32);
        };
        for (const Term& t : c.A) rec(0, t.row, t.sig, t.val);
        for (const Term& t : c.B) rec(1, t.row, t.sig, t.val);
        for (uint32_t i = 0; i <= nP; i++) rec(0, m + i, i, HFr::one());
        zk.add(4, s);
    }
    zk.add(5, pA);
    zk.add(6, pB1);
    zk.add(7, pB2);
    if (!c.ultra) {
        zk.add(8, pC);
        zk.add(9, pH);
        std::vector<uint8_t> s(64, 0);
        put_u32(s, 0);
        zk.add(10, s);
    } else {
        zk.add(8, pC1);
        zk.add(9, pC);
        std::vector<uint8_t> s;
        for (uint32_t v : c.round_idx) put_u32(s, v);
        zk.add(10, s);
        s.clear();
        for (uint32_t v : c.final_idx) put_u32(s, v);
        zk.add(11, s);
        zk.add(12, pH);
        s.assign(64, 0);
        put_u32(s, 0);
        zk.add(13, s);
    }
    zkey.swap(zk.out);

    // ---- wtns (version 2; UltraGroth adds the four u32 lookup sections of src/prover.cpp:287-292) ----
    SectionWriter wt("wtns", 2, c.ultra ? 6 : 2);
    { std::vector<uint8_t> s; put_u32(s, 32); put_bytes(s, r_le.data(), 32); put_u32(s, nV); wt.add(1, s); }
    {
        std::vector<uint8_t> s((size_t)nV * 32);
        for (uint32_t i = 0; i < nV; i++) { HFr raw = from_mont(c.witness[i]); memcpy(&s[(size_t)i * 32], &raw, 32); }
        wt.add(2, s);
    }
    if (c.ultra) {
        auto u32sec = [&](uint32_t id, const std::vector<uint32_t>& v) { std::vector<uint8_t> s(v.size() * 4); if (!v.empty()) memcpy(s.data(), v.data(), s.size()); wt.add(id, s); };
        u32sec(3, c.chunks);
        u32sec(4, c.freqs);
        u32sec(5, c.wtns_idx);
        u32sec(6, c.push_idx);
    }
    wtns.swap(wt.out);

    // ---- verification key (field names of src/groth16.cpp:278-296 / src/ultra_groth.cpp:542-563) ----
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

void fixed_base_mul_host(DeviceCtx& ctx, int group, const void* scalars_raw, uint64_t n, void* out) {
    std::vector<HFr> m(n);
    for (uint64_t i = 0; i < n; i++) { HFr r; memcpy(&r, (const uint8_t*)scalars_raw + i * 32, 32); m[i] = to_mont(r); }
    FixedBase fb(ctx);
    fb.mul(group, m.data(), n, (uint8_t*)out);
}

}  // namespace ug
