// ORACLE / CPU BASELINE -- test infrastructure, NOT product code.  Nothing under ultragroth_b200/ links,
// includes or calls this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it (as the checker and as the timed CPU arm).
//
// A C++17 / OpenMP restatement of the reference's CPU prover, step for step:
//   Groth16::Prover::prove ............. src/groth16.cpp:48-203
//   UltraGroth::Prover::prove .......... src/ultra_groth.cpp:401-462 (execute_round :161-184,
//                                        derive_challenge :32-58, compute_lookup :62-106,
//                                        execute_final_round :187-399)
//   field semantics .................... build/fr_raw_generic.cpp:107-148 (4-round CIOS, one conditional
//                                        subtraction), :11-39 (add / sub), constants :5-8; fq likewise
//   file formats ....................... src/binfile_utils.cpp:32-80, src/zkey_utils.cpp:42-76,123-163,
//                                        src/wtns_utils.cpp:13-26, src/prover.cpp:237-260,283-292
// The arithmetic engine the reference calls for MSM / FFT / curve ops (iden3/ffiasm @ fe3772e8, absent from
// /root/reference) is restated from its published algorithm: Jacobian coordinates, windowed Pippenger with
// threads over windows, iterative radix-2 FFT with omega_n = 5^((r-1)/n).  This is "kind: port": generic
// C++ limbs, not ffiasm's x86-64 assembly, so it is slower than real rapidsnark.
//
// Parity: pinned by tests/test_oracle.py against the Python big-int oracle on the reference's
// testdata/ fixture (proof points and H coefficients bit-exact, proof verifies, tamper fails).
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "cpu_field.hpp"

namespace {

// ---------------------------------------------------------------------------------------------
// MSM: windowed Pippenger, threads over windows then over point slices (ffiasm's structure, restated)
// scalars: raw 32-byte little-endian integers at `stride` bytes apart
// ---------------------------------------------------------------------------------------------
inline uint32_t window_bits(const uint8_t* s, int pos, int c) {
    u64 acc = 0;
    int byte = pos >> 3, off = pos & 7;
    for (int k = 0; k < 8 && byte + k < 32; k++) acc |= (u64)s[byte + k] << (8 * k);
    return (uint32_t)((acc >> off) & ((1ull << c) - 1));
}

template <class F>
Jac<F> msm(const Aff<F>* bases, const uint8_t* scalars, size_t stride, size_t n, int threads) {
    if (n == 0) return jinf<F>();
    int c = n < 32 ? 3 : n < 1024 ? 7 : n < (1u << 15) ? 10 : n < (1u << 19) ? 13 : 16;
    const int W = (256 + c - 1) / c;
    // work units = (window, slice); slices keep all threads busy when W < threads
    int slices = std::max(1, (2 * threads + W - 1) / W);
    if (n < 4096) slices = 1;
    std::vector<Jac<F>> partial((size_t)W * slices);
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int unit = 0; unit < W * slices; unit++) {
        int w = unit / slices, sl = unit % slices;
        size_t lo = n * (size_t)sl / slices, hi = n * (size_t)(sl + 1) / slices;
        std::vector<Jac<F>> buckets((size_t)1 << c, jinf<F>());
        for (size_t i = lo; i < hi; i++) {
            uint32_t d = window_bits(scalars + i * stride, w * c, c);
            if (d) buckets[d] = addm(buckets[d], bases[i]);
        }
        Jac<F> run = jinf<F>(), sum = jinf<F>();
        for (uint32_t d = (1u << c) - 1; d >= 1; d--) {
            run = addj(run, buckets[d]);
            sum = addj(sum, run);
        }
        partial[unit] = sum;
    }
    Jac<F> acc = jinf<F>();
    for (int w = W - 1; w >= 0; w--) {
        for (int k = 0; k < c; k++) acc = dbl(acc);
        for (int sl = 0; sl < slices; sl++) acc = addj(acc, partial[(size_t)w * slices + sl]);
    }
    return acc;
}

// ---------------------------------------------------------------------------------------------
// FFT<Fr>: omega_{2^k} = 5^((r-1)/2^k); fft / ifft natural order, ifft scales by 1/n
// ---------------------------------------------------------------------------------------------
struct FFT {
    int logn;
    size_t n;
    std::vector<Fr> roots;  // omega_n^i, i < n/2 ... built once like FFT<Fr>'s table (src/groth16.hpp:108-110 builds 2*domainSize)
    explicit FFT(int k) : logn(k), n((size_t)1 << k) {
        roots.resize(std::max<size_t>(1, n / 2));
        Fr w = rootOfUnity(k);
        roots[0] = one<FrP>();
        for (size_t i = 1; i < roots.size(); i++) roots[i] = mul(roots[i - 1], w);
    }
    void transform(Fr* a, bool inverse, int threads) const {
        // bit reversal
#pragma omp parallel for schedule(static) num_threads(threads)
        for (size_t i = 0; i < n; i++) {
            size_t r = 0;
            for (int b = 0; b < logn; b++) r |= ((i >> b) & 1) << (logn - 1 - b);
            if (i < r) std::swap(a[i], a[r]);
        }
        for (int s = 0; s < logn; s++) {
            size_t half = (size_t)1 << s, step = n >> (s + 1);
#pragma omp parallel for schedule(static) num_threads(threads)
            for (size_t k = 0; k < n / 2; k++) {
                size_t j = k & (half - 1), blk = k >> s;
                size_t i0 = (blk << (s + 1)) | j, i1 = i0 + half;
                size_t ti = j * step;
                Fr w = (inverse && ti) ? roots[0] : roots[ti];
                if (inverse && ti) w = neg(roots[n / 2 - ti]);  // omega^-t = -omega^(n/2 - t)
                Fr v = mul(a[i1], w);
                Fr u = a[i0];
                a[i0] = add(u, v);
                a[i1] = sub(u, v);
            }
        }
        if (inverse) {
            Fr ninv = inv(fromU64<FrP>(n));
#pragma omp parallel for schedule(static) num_threads(threads)
            for (size_t i = 0; i < n; i++) a[i] = mul(a[i], ninv);
        }
    }
};

// ---------------------------------------------------------------------------------------------
// file formats
// ---------------------------------------------------------------------------------------------
struct Sections {
    std::map<uint32_t, std::pair<const uint8_t*, u64>> s;
    Sections(const uint8_t* data, u64 size, const char* magic) {
        if (size < 12 || memcmp(data, magic, 4) != 0) throw std::runtime_error("bad file");
        uint32_t ns;
        memcpy(&ns, data + 8, 4);
        u64 pos = 12;
        for (uint32_t i = 0; i < ns; i++) {
            if (pos + 12 > size) throw std::runtime_error("truncated section table");
            uint32_t id; u64 sz;
            memcpy(&id, data + pos, 4);
            memcpy(&sz, data + pos + 4, 8);
            pos += 12;
            if (sz > size - pos) throw std::runtime_error("truncated section");
            if (!s.count(id)) s[id] = {data + pos, sz};
            pos += sz;
        }
    }
    const uint8_t* data(uint32_t id) const { return s.at(id).first; }
    u64 size(uint32_t id) const { return s.at(id).second; }
};
inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

struct Zkey {
    bool ultra;
    uint32_t nVars, nPublic, domainSize, nCoefs, nC1 = 0, nC2 = 0, randIdx = 0;
    const uint8_t *alpha1, *beta1, *beta2, *gamma2, *rdelta1 = nullptr, *rdelta2 = nullptr, *delta1, *delta2;
    const uint8_t *coefs, *A, *B1, *B2, *C, *C1 = nullptr, *H;
    const uint8_t *idx1 = nullptr, *idx2 = nullptr;
    Zkey(const Sections& f) {
        uint32_t proto = rd32(f.data(1));
        if (proto != 1 && proto != 1337) throw std::runtime_error("unknown zkey protocol");
        ultra = proto == 1337;
        const uint8_t* p = f.data(2);
        p += 4 + 32 + 4 + 32;  // n8q, q, n8r, r
        nVars = rd32(p); nPublic = rd32(p + 4); domainSize = rd32(p + 8); p += 12;
        if (ultra) { nC1 = rd32(p); nC2 = rd32(p + 4); randIdx = rd32(p + 8); p += 12; }
        alpha1 = p; p += 64; beta1 = p; p += 64; beta2 = p; p += 128; gamma2 = p; p += 128;
        if (ultra) { rdelta1 = p; p += 64; rdelta2 = p; p += 128; }
        delta1 = p; p += 64; delta2 = p;
        nCoefs = (uint32_t)(f.size(4) / 44);
        coefs = f.data(4) + 4;
        A = f.data(5); B1 = f.data(6); B2 = f.data(7);
        if (!ultra) { C = f.data(8); H = f.data(9); }
        else { C1 = f.data(8); C = f.data(9); idx1 = f.data(10); idx2 = f.data(11); H = f.data(12); }
    }
};

void keccak_f(u64 st[25]) {
    static const u64 RC[24] = {0x1ull, 0x8082ull, 0x800000000000808aull, 0x8000000080008000ull, 0x808bull, 0x80000001ull, 0x8000000080008081ull,
                               0x8000000000008009ull, 0x8aull, 0x88ull, 0x80008009ull, 0x8000000aull, 0x8000808bull, 0x800000000000008bull,
                               0x8000000000008089ull, 0x8000000000008003ull, 0x8000000000008002ull, 0x8000000000000080ull, 0x800aull,
                               0x800000008000000aull, 0x8000000080008081ull, 0x8000000000008080ull, 0x80000001ull, 0x8000000080008008ull};
    for (int rnd = 0; rnd < 24; rnd++) {
        u64 C[5], D[5], B[25];
        for (int x = 0; x < 5; x++) C[x] = st[x] ^ st[x + 5] ^ st[x + 10] ^ st[x + 15] ^ st[x + 20];
        for (int x = 0; x < 5; x++) D[x] = C[(x + 4) % 5] ^ ((C[(x + 1) % 5] << 1) | (C[(x + 1) % 5] >> 63));
        for (int i = 0; i < 25; i++) st[i] ^= D[i % 5];
        // rho + pi: B[y][2x+3y] = rot(A[x][y], r[x][y]) walked along the (x,y) -> (y, 2x+3y) orbit
        int x = 1, y = 0;
        B[0] = st[0];
        for (int t = 0; t < 24; t++) {
            int rot = ((t + 1) * (t + 2) / 2) % 64;
            u64 v = st[x + 5 * y];
            v = rot ? ((v << rot) | (v >> (64 - rot))) : v;
            int nx = y, ny = (2 * x + 3 * y) % 5;
            B[nx + 5 * ny] = v;
            x = nx; y = ny;
        }
        for (int yy = 0; yy < 5; yy++)
            for (int xx = 0; xx < 5; xx++) st[xx + 5 * yy] = B[xx + 5 * yy] ^ (~B[(xx + 1) % 5 + 5 * yy] & B[(xx + 2) % 5 + 5 * yy]);
        st[0] ^= RC[rnd];
    }
}
void keccak256(const uint8_t* in, size_t len, uint8_t* out) {  // FIPS202_KECCAK_256 of src/keccak256.cpp (suffix 0x01)
    u64 st[25] = {0};
    uint8_t blk[136];
    while (len >= 136) {
        for (int i = 0; i < 17; i++) { u64 w; memcpy(&w, in + 8 * i, 8); st[i] ^= w; }
        keccak_f(st);
        in += 136; len -= 136;
    }
    memset(blk, 0, 136);
    memcpy(blk, in, len);
    blk[len] ^= 0x01;
    blk[135] ^= 0x80;
    for (int i = 0; i < 17; i++) { u64 w; memcpy(&w, blk + 8 * i, 8); st[i] ^= w; }
    keccak_f(st);
    memcpy(out, st, 32);
}

// ---------------------------------------------------------------------------------------------
// the prover
// ---------------------------------------------------------------------------------------------
struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double lap() {
        auto t1 = std::chrono::steady_clock::now();
        double s = std::chrono::duration<double>(t1 - t0).count();
        t0 = t1;
        return s;
    }
};

template <class F>
Jac<F> affBytes(const uint8_t* p) {
    Aff<F> a;
    memcpy(&a, p, sizeof(a));
    if (a.isInf()) return jinf<F>();
    return {a.x, a.y, fone<F>()};
}

// a5..a8: the H-MSM scalars (src/groth16.cpp:66-148)
// t_ms: [0] += spmv, [1] += transforms + pointwise, [2] += root-table construction
void h_scalars(const Zkey& zk, const Fr* w, std::vector<Fr>& a, int threads, double* t_ms) {
    const size_t n = zk.domainSize;
    int logn = 0;
    while (((size_t)1 << logn) < n) logn++;
    Timer tm;
    std::vector<Fr> b(n), c(n);
    a.assign(n, Fr{{0, 0, 0, 0}});
    std::fill(b.begin(), b.end(), Fr{{0, 0, 0, 0}});
    // striped locks exactly like the reference (:77-98: 1024 mutexes indexed by c % 1024)
    const int NLOCKS = 1024;
    std::vector<omp_lock_t> locks(NLOCKS);
    for (auto& l : locks) omp_init_lock(&l);
#pragma omp parallel for num_threads(threads)
    for (uint32_t k = 0; k < zk.nCoefs; k++) {
        const uint8_t* rec = zk.coefs + (size_t)k * 44;
        uint32_t m = rd32(rec), cc = rd32(rec + 4), s = rd32(rec + 8);
        Fr coef;
        memcpy(&coef, rec + 12, 32);
        Fr prod = mul(w[s], coef);
        std::vector<Fr>& dst = m == 0 ? a : b;
        omp_set_lock(&locks[cc % NLOCKS]);
        dst[cc] = add(dst[cc], prod);
        omp_unset_lock(&locks[cc % NLOCKS]);
    }
    for (auto& l : locks) omp_destroy_lock(&l);
#pragma omp parallel for num_threads(threads)
    for (size_t i = 0; i < n; i++) c[i] = mul(a[i], b[i]);
    if (t_ms) t_ms[0] += tm.lap() * 1e3;
    // root tables: the reference builds them once in the Prover constructor (src/groth16.hpp:108-110), so their
    // construction time is reported separately and kept out of the prove time
    FFT fft(logn);
    Fr w2n = rootOfUnity(logn + 1);
    std::vector<Fr> shift(n);  // root(log2n+1, i) (:114-118)
    shift[0] = one<FrP>();
    for (size_t i = 1; i < n; i++) shift[i] = mul(shift[i - 1], w2n);
    if (t_ms) t_ms[2] += tm.lap() * 1e3;
    for (std::vector<Fr>* v : {&a, &b, &c}) {
        fft.transform(v->data(), true, threads);
#pragma omp parallel for num_threads(threads)
        for (size_t i = 0; i < n; i++) (*v)[i] = mul((*v)[i], shift[i]);
        fft.transform(v->data(), false, threads);
    }
#pragma omp parallel for num_threads(threads)
    for (size_t i = 0; i < n; i++) a[i] = fromMont(sub(mul(a[i], b[i]), c[i]));
    if (t_ms) t_ms[1] += tm.lap() * 1e3;
}

// derive_challenge (src/ultra_groth.cpp:32-58): keccak256(BE32(x) || BE32(y)) of the canonical affine coordinates, digest read
// big-endian into 4 limbs UNREDUCED, then one Montgomery multiplication by R^2 (RawFr::fromMpz, build/fr.cpp:286-290)
Fr deriveChallenge(const Aff<Fq>& R) {
    uint8_t buf[64], dig[32];
    Fq x = fromMont(R.x), y = fromMont(R.y);
    for (int i = 0; i < 32; i++) {
        buf[i] = (uint8_t)(x.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
        buf[32 + i] = (uint8_t)(y.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
    }
    keccak256(buf, 64, dig);
    Fr raw = {{0, 0, 0, 0}};
    for (int i = 0; i < 32; i++) raw.v[(31 - i) >> 3] |= (u64)dig[i] << (8 * ((31 - i) & 7));
    return toMont(raw);
}

// compute_lookup (src/ultra_groth.cpp:62-106): w[wtns_idx[k]] = push[push_idx[k]], push = [rand | inv2[chunks[*]] | inv2 | prod]
void computeLookup(Fr* w, const uint32_t* chunks, size_t K, const uint32_t* freqs, size_t T, const uint32_t* widx, const uint32_t* pidx, size_t NI,
                   const Fr& rand, int threads) {
    std::vector<Fr> push(1 + K + 2 * T);
    push[0] = fromMont(rand);
#pragma omp parallel for num_threads(threads)
    for (size_t i = 0; i < T; i++) {
        Fr iv = inv(add(fromU64<FrP>(i), rand));
        push[1 + K + i] = fromMont(iv);
        // uint32_t -> int conversion of the reference (build/fr.hpp:251, RawFr::set(int) build/fr.cpp:209-223): f >= 2^31 is negative
        const int32_t fs = (int32_t)freqs[i];
        Fr fm = fs < 0 ? neg(fromU64<FrP>((u64)(-(int64_t)fs))) : fromU64<FrP>((u64)fs);
        push[1 + K + T + i] = fromMont(mul(fm, iv));
    }
    for (size_t j = 0; j < K; j++) push[1 + j] = push[1 + K + chunks[j]];
    for (size_t k = 0; k < NI; k++) w[widx[k]] = push[pidx[k]];
}

struct ProofOut {
    Aff<Fq> A, C, R;
    Aff<Fq2> B;
};

void assemble(const Zkey& zk, Jac<Fq> pi_a, Jac<Fq> pib1, Jac<Fq2> pi_b, Jac<Fq> pi_c, const Jac<Fq>& pih, const u64* r, const u64* s,
              const u64* rk, ProofOut& out) {
    Jac<Fq> alpha1 = affBytes<Fq>(zk.alpha1), beta1 = affBytes<Fq>(zk.beta1), delta1 = affBytes<Fq>(zk.delta1);
    Jac<Fq2> beta2 = affBytes<Fq2>(zk.beta2), delta2 = affBytes<Fq2>(zk.delta2);
    pi_a = addj(addj(pi_a, alpha1), mulScalar(delta1, r));
    pi_b = addj(addj(pi_b, beta2), mulScalar(delta2, s));
    pib1 = addj(addj(pib1, beta1), mulScalar(delta1, s));
    pi_c = addj(pi_c, pih);
    pi_c = addj(pi_c, mulScalar(pi_a, s));
    pi_c = addj(pi_c, mulScalar(pib1, r));
    Fr fr_r, fr_s;
    memcpy(&fr_r, r, 32);
    memcpy(&fr_s, s, 32);
    Fr rs = toMont(mul(fr_r, fr_s));  // E.fr.mul then toMontgomery (:191-192) == r*s mod q as a raw integer
    pi_c = addj(pi_c, negj(mulScalar(delta1, rs.v)));
    if (rk) pi_c = addj(pi_c, negj(mulScalar(affBytes<Fq>(zk.rdelta1), rk)));
    out.A = toAff(pi_a);
    out.B = toAff(pi_b);
    out.C = toAff(pi_c);
}

}  // namespace

extern "C" {

// out_points: A (64) | B (128) | C (64) | R (64, UltraGroth round commitment) affine Montgomery, zkey layout
// out_h: optional domainSize*32 bytes (the H-MSM scalars); times_ms: optional [msm, spmv, fft, prove total (without root tables), root tables]
// out_msms: optional 448 bytes, the RAW MSM results before blinding/assembly as affine Montgomery points in zkey layout:
//   A (64) | B1 (64) | B2 (128) | C (64) | H (64) | round C1 (64; all-zero for Groth16)   (src/groth16.cpp:53-64,153-154; src/ultra_groth.cpp:168)
int cpu_prove2(const uint8_t* zkey, u64 zkey_size, const uint8_t* wtns, u64 wtns_size, const uint8_t* r32, const uint8_t* s32,
               const uint8_t* rk32, int threads, uint8_t* out_points, uint8_t* out_h, uint8_t* out_witness, double* times_ms,
               uint8_t* out_msms) {
    try {
        if (threads <= 0) threads = omp_get_max_threads();
        Timer total;
        Sections zf(zkey, zkey_size, "zkey"), wf(wtns, wtns_size, "wtns");
        Zkey zk(zf);
        if (rd32(wf.data(1) + 4 + 32) != zk.nVars) return 3;
        std::vector<Fr> w(zk.nVars);
        memcpy(w.data(), wf.data(2), (size_t)zk.nVars * 32);
        u64 r[4], s[4], rk[4] = {0, 0, 0, 0};
        memcpy(r, r32, 32);
        memcpy(s, s32, 32);
        if (rk32) memcpy(rk, rk32, 32);
        double t_msm = 0, t_h[3] = {0, 0, 0};
        Timer tm;
        ProofOut out;
        memset(&out, 0, sizeof(out));
        std::vector<Fr> cw;  // scalars of the C MSM
        const Aff<Fq>* Cbase = (const Aff<Fq>*)zk.C;
        size_t nC;
        if (zk.ultra) {
            // round 1 (src/ultra_groth.cpp:415-425, 161-184)
            std::vector<Fr> rw(zk.nC1);
            for (uint32_t i = 0; i < zk.nC1; i++) rw[i] = w[rd32(zk.idx1 + 4 * (size_t)i)];
            Jac<Fq> commit = msm<Fq>((const Aff<Fq>*)zk.C1, (const uint8_t*)rw.data(), 32, zk.nC1, threads);
            if (out_msms) { Aff<Fq> t = toAff(commit); memcpy(out_msms + 384, &t, 64); }
            commit = addj(commit, mulScalar(affBytes<Fq>(zk.delta1), rk));
            out.R = toAff(commit);
            Fr rand = deriveChallenge(out.R);
            // compute_lookup (:62-106)
            computeLookup(w.data(), (const uint32_t*)wf.data(3), wf.size(3) / 4, (const uint32_t*)wf.data(4), wf.size(4) / 4,
                          (const uint32_t*)wf.data(5), (const uint32_t*)wf.data(6), wf.size(5) / 4, rand, threads);
            cw.resize(zk.nC2);
            for (uint32_t i = 0; i < zk.nC2; i++) cw[i] = w[rd32(zk.idx2 + 4 * (size_t)i)];
            nC = zk.nC2;
        } else {
            nC = zk.nVars - zk.nPublic - 1;
            cw.assign(w.begin() + zk.nPublic + 1, w.end());
        }
        t_msm += tm.lap() * 1e3;
        const uint8_t* ws = (const uint8_t*)w.data();
        Jac<Fq> pi_a = msm<Fq>((const Aff<Fq>*)zk.A, ws, 32, zk.nVars, threads);
        Jac<Fq> pib1 = msm<Fq>((const Aff<Fq>*)zk.B1, ws, 32, zk.nVars, threads);
        Jac<Fq2> pi_b = msm<Fq2>((const Aff<Fq2>*)zk.B2, ws, 32, zk.nVars, threads);
        Jac<Fq> pi_c = msm<Fq>(Cbase, (const uint8_t*)cw.data(), 32, nC, threads);
        t_msm += tm.lap() * 1e3;
        std::vector<Fr> h;
        h_scalars(zk, w.data(), h, threads, t_h);
        tm.lap();
        Jac<Fq> pih = msm<Fq>((const Aff<Fq>*)zk.H, (const uint8_t*)h.data(), 32, zk.domainSize, threads);
        t_msm += tm.lap() * 1e3;
        if (out_msms) {
            Aff<Fq> t;
            Aff<Fq2> t2 = toAff(pi_b);
            t = toAff(pi_a); memcpy(out_msms, &t, 64);
            t = toAff(pib1); memcpy(out_msms + 64, &t, 64);
            memcpy(out_msms + 128, &t2, 128);
            t = toAff(pi_c); memcpy(out_msms + 256, &t, 64);
            t = toAff(pih); memcpy(out_msms + 320, &t, 64);
            if (!zk.ultra) memset(out_msms + 384, 0, 64);
        }
        assemble(zk, pi_a, pib1, pi_b, pi_c, pih, r, s, zk.ultra ? rk : nullptr, out);
        if (out_points) {
            memcpy(out_points, &out.A, 64);
            memcpy(out_points + 64, &out.B, 128);
            memcpy(out_points + 192, &out.C, 64);
            memcpy(out_points + 256, &out.R, 64);
        }
        if (out_h) memcpy(out_h, h.data(), h.size() * 32);
        if (out_witness) memcpy(out_witness, w.data(), w.size() * 32);
        if (times_ms) {
            times_ms[0] = t_msm;
            times_ms[1] = t_h[0];
            times_ms[2] = t_h[1];
            times_ms[3] = total.lap() * 1e3 - t_h[2];  // prove time without the one-off root tables
            times_ms[4] = t_h[2];
        }
        return 0;
    } catch (...) {
        return 1;
    }
}

int cpu_prove(const uint8_t* zkey, u64 zkey_size, const uint8_t* wtns, u64 wtns_size, const uint8_t* r32, const uint8_t* s32,
              const uint8_t* rk32, int threads, uint8_t* out_points, uint8_t* out_h, uint8_t* out_witness, double* times_ms) {
    return cpu_prove2(zkey, zkey_size, wtns, wtns_size, r32, s32, rk32, threads, out_points, out_h, out_witness, times_ms, nullptr);
}

// standalone kernels of the sweep (BASELINE config 5): G1 MSM and Fr NTT on the CPU
int cpu_msm_g1(const uint8_t* bases, const uint8_t* scalars, u64 n, int threads, uint8_t* out_affine) {
    if (threads <= 0) threads = omp_get_max_threads();
    Aff<Fq> a = toAff(msm<Fq>((const Aff<Fq>*)bases, scalars, 32, n, threads));
    memcpy(out_affine, &a, 64);
    return 0;
}
int cpu_msm_g2(const uint8_t* bases, const uint8_t* scalars, u64 n, int threads, uint8_t* out_affine) {
    if (threads <= 0) threads = omp_get_max_threads();
    Aff<Fq2> a = toAff(msm<Fq2>((const Aff<Fq2>*)bases, scalars, 32, n, threads));
    memcpy(out_affine, &a, 128);
    return 0;
}
int cpu_ntt(uint8_t* data, int logn, int inverse, int threads) {
    if (threads <= 0) threads = omp_get_max_threads();
    FFT f(logn);
    f.transform((Fr*)data, inverse != 0, threads);
    return 0;
}
int cpu_field_mul(int field, const uint8_t* a, const uint8_t* b, uint8_t* out, u64 n) {
    for (u64 i = 0; i < n; i++) {
        if (field == 0) { Fr x, y; memcpy(&x, a + 32 * i, 32); memcpy(&y, b + 32 * i, 32); Fr z = mul(x, y); memcpy(out + 32 * i, &z, 32); }
        else { Fq x, y; memcpy(&x, a + 32 * i, 32); memcpy(&y, b + 32 * i, 32); Fq z = mul(x, y); memcpy(out + 32 * i, &z, 32); }
    }
    return 0;
}
void cpu_keccak256(const uint8_t* in, u64 len, uint8_t* out) { keccak256(in, len, out); }
// the two UltraGroth host steps on their own (pinned against the reference's compiled RawFr / RawFq / keccak in tests/test_oracle.py)
void cpu_derive_challenge(const uint8_t* commitment_xy_mont, uint8_t* rand_mont_out) {
    Aff<Fq> R;
    memcpy(&R, commitment_xy_mont, 64);
    Fr r = deriveChallenge(R);
    memcpy(rand_mont_out, &r, 32);
}
void cpu_compute_lookup(uint8_t* signals, const uint32_t* chunks, u64 K, const uint32_t* freqs, u64 T, const uint32_t* widx, const uint32_t* pidx,
                        u64 NI, const uint8_t* rand_mont) {
    Fr r;
    memcpy(&r, rand_mont, 32);
    computeLookup((Fr*)signals, chunks, K, freqs, T, widx, pidx, NI, r, 1);
}
int cpu_max_threads() { return omp_get_max_threads(); }

}  // extern "C"
