// The reference's prover C API (include/prover.h == src/prover.h) on top of the device prover.
//
// Mirrors src/prover.cpp: the same wrapper objects (Groth16Prover :144-216, UltraGrothProver :219-309),
// validation order, error codes and messages, buffer-size rules and strncpy output convention.  The O(1)
// tail of a proof -- blinding, proof assembly, affine conversion, JSON (src/groth16.cpp:158-250,
// src/ultra_groth.cpp:161-184,337-399,476-513) -- runs on the host, as it does in the reference.
#include "../../include/prover.h"
#include "../../include/ultragroth_b200.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <future>
#include <memory>
#include <mutex>
#include <random>
#include <string>

#include "hostio.hpp"
#include "prover_dev.hpp"

using namespace ug;

namespace ug { void set_last_error(const std::string& msg); }

namespace {

class ShortBufferException : public std::invalid_argument {
public:
    explicit ShortBufferException(const std::string& msg) : std::invalid_argument(msg) {}
};
class InvalidWitnessLengthException : public std::invalid_argument {
public:
    explicit InvalidWitnessLengthException(const std::string& msg) : std::invalid_argument(msg) {}
};

// UGB_BATCH_TRACE=1: host timestamps (ms since the batch began) at the seams of the pipelined batch, to stderr
struct BatchTrace {
    bool on = getenv("UGB_BATCH_TRACE") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void at(uint32_t k, const char* what) {
        if (!on) return;
        fprintf(stderr, "[batch] proof %u %-18s %8.3f ms\n", k, what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
};

void CopyError(char* error_msg, unsigned long long maxsize, const char* str) {
    if (error_msg) strncpy(error_msg, str, maxsize);
}

unsigned long long ProofBufferMinSize() { return 810; }
unsigned long long ProofBufferMinSizeUltraGroth() { return 1400; }
unsigned long long PublicBufferMinSize(unsigned long long count) { return count * 82 + 4; }

void CheckAndUpdateBufferSizes(unsigned long long proofCalcSize, unsigned long long* proofSize, unsigned long long publicCalcSize,
                               unsigned long long* publicSize, const std::string& type) {
    if (*proofSize < proofCalcSize)
        throw ShortBufferException("Proof buffer is too short. " + type + " size: " + std::to_string(proofCalcSize) +
                                   ", actual size: " + std::to_string(*proofSize));
    if (*publicSize < publicCalcSize)
        throw ShortBufferException("Public buffer is too short. " + type + " size: " + std::to_string(publicCalcSize) +
                                   ", actual size: " + std::to_string(*publicSize));
}

// ---- host curve helpers (O(1) work per proof) ----------------------------------------------------
struct Scalar { uint64_t v[4]; };

template <class HF>
XYZZ<HF> pt_mul(const XYZZ<HF>& p, const Scalar& k) {  // E.g1.mulByScalar: plain double-and-add
    XYZZ<HF> acc = XYZZ<HF>::inf();
    for (int i = 255; i >= 0; i--) {
        acc = pt_dbl(acc);
        if ((k.v[i >> 6] >> (i & 63)) & 1) pt_add(acc, p);
    }
    return acc;
}
template <class HF>
XYZZ<HF> from_affine_bytes(const uint8_t* p) {
    Affine<HF> a;
    memcpy(&a, p, sizeof(a));
    return XYZZ<HF>::from_affine(a);
}

bool scalar_geq_r(const Scalar& s) {
    for (int i = 3; i >= 0; i--) {
        if (s.v[i] > HFrParams::mod[i]) return true;
        if (s.v[i] < HFrParams::mod[i]) return false;
    }
    return true;
}
Scalar scalar_mod_r(Scalar s) {
    while (scalar_geq_r(s)) {
        unsigned __int128 borrow = 0;
        for (int i = 0; i < 4; i++) {
            unsigned __int128 d = (unsigned __int128)s.v[i] - HFrParams::mod[i] - (uint64_t)borrow;
            s.v[i] = (uint64_t)d;
            borrow = (d >> 64) & 1;
        }
    }
    return s;
}
// E.fr.mul(rs, r, s); E.fr.toMontgomery(rs, rs)  (src/groth16.cpp:191-192)  ==  r*s mod r as a raw integer
Scalar scalar_mul_mod_r(const Scalar& a, const Scalar& b) {
    HFr x, y;
    memcpy(&x, &a, 32);
    memcpy(&y, &b, 32);
    HFr m = fmul(x, y);   // a*b*R^-1
    HFr t = to_mont(m);   // a*b
    Scalar out;
    memcpy(&out, &t, 32);
    return out;
}

// randombytes_buf(&r.v[0], sizeof(r) - 1): 31 random bytes, top byte zero (src/groth16.cpp:165-166)
Scalar random_blinding() {
    Scalar s;
    memset(&s, 0, sizeof(s));
    FILE* f = fopen("/dev/urandom", "rb");
    bool ok = false;
    if (f) {
        ok = fread(&s, 1, 31, f) == 31;
        fclose(f);
    }
    if (!ok) {
        std::random_device rd;
        uint8_t* p = (uint8_t*)&s;
        for (int i = 0; i < 31; i++) p[i] = (uint8_t)rd();
    }
    return s;
}

std::string g1_json(const HG1Affine& p) { return "[\"" + to_dec(p.x) + "\",\"" + to_dec(p.y) + "\",\"1\"]"; }
std::string g2_json(const HG2Affine& p) {
    return "[[\"" + to_dec(p.x.a) + "\",\"" + to_dec(p.x.b) + "\"],[\"" + to_dec(p.y.a) + "\",\"" + to_dec(p.y.b) + "\"],[\"1\",\"0\"]]";
}

// BuildPublicString (src/prover.cpp:89-117): nlohmann dump() of an array of decimal strings ("null" when empty)
std::string public_json(const uint8_t* signals /* from index 1 */, uint32_t nPublic, uint32_t skip_index) {
    std::string out;
    bool any = false;
    for (uint32_t i = 1; i <= nPublic; i++) {
        if (i == skip_index) continue;
        Scalar s;
        memcpy(&s, signals + (size_t)(i - 1) * 32, 32);
        s = scalar_mod_r(s);
        out += any ? ",\"" : "[\"";
        out += u256_to_dec(s.v);
        out += "\"";
        any = true;
    }
    return any ? out + "]" : "null";
}

struct Blinding {
    bool fixed = false;
    Scalar r, s, rk;
};

// src/groth16.cpp:168-200 and src/ultra_groth.cpp:353-396
void assemble(const ZkeyHeader& h, const ProofMsms& m, const Scalar& r, const Scalar& s, const Scalar* rk, HG1Affine& A, HG2Affine& B,
              HG1Affine& C) {
    HG1XYZZ alpha1 = from_affine_bytes<HFq>(h.alpha1), beta1 = from_affine_bytes<HFq>(h.beta1), delta1 = from_affine_bytes<HFq>(h.final_delta1);
    HG2XYZZ beta2 = from_affine_bytes<HFq2>(h.beta2), delta2 = from_affine_bytes<HFq2>(h.final_delta2);
    HG1XYZZ pi_a = m.a, pib1 = m.b1, pi_c = m.c;
    HG2XYZZ pi_b = m.b2;
    // the seven scalar multiplications are independent double-and-add chains (0.1 ms each in G1, 0.4 ms in G2): the G2 one and the
    // ones that only need header points run on helper threads while this thread does the two that depend on pi_a / pib1 -- the
    // host tail of a proof is serial time in the single-call and sharded paths (the pipelined batch hides it under the next proof)
    const Scalar rs = scalar_mul_mod_r(r, s);
    auto f_b2 = std::async(std::launch::async, [&]() { return pt_mul(delta2, s); });
    auto f_g1 = std::async(std::launch::async, [&]() {
        HG1XYZZ t = pt_neg(pt_mul(delta1, rs));
        if (rk) pt_add(t, pt_neg(pt_mul(from_affine_bytes<HFq>(h.round_delta1), *rk)));
        return t;
    });
    pt_add(pi_a, alpha1);
    pt_add(pi_a, pt_mul(delta1, r));
    pt_add(pib1, beta1);
    pt_add(pib1, pt_mul(delta1, s));
    pt_add(pi_c, m.h);
    pt_add(pi_c, pt_mul(pi_a, s));
    pt_add(pi_c, pt_mul(pib1, r));
    pt_add(pi_c, f_g1.get());
    pt_add(pi_b, beta2);
    pt_add(pi_b, f_b2.get());
    A = pt_to_affine(pi_a);
    B = pt_to_affine(pi_b);
    C = pt_to_affine(pi_c);
}

int pick_device() {
    const char* e = getenv("UGB_DEVICE");
    if (e) return atoi(e);
    const char* lr = getenv("LOCAL_RANK");
    if (lr && getenv("UGB_DEVICE_FROM_LOCAL_RANK")) return atoi(lr);
    return 0;
}

class ProverBase {
public:
    // from a zkey file image (src/prover.cpp:144-216, 219-309)
    ProverBase(const void* zkey_buffer, unsigned long long zkey_size, bool ultra, uint32_t shard_rank = 0, uint32_t shard_world = 1,
               bool tables = true) {
        BinFile zkey(zkey_buffer, zkey_size, "zkey", 1);
        hdr = load_zkey_header(zkey, ultra);
        if (!prime_is_bn254_r(hdr.rPrime, hdr.n8r)) throw std::invalid_argument("zkey curve not supported");
        // the header points (alpha, beta, delta...) are copied: the zkey buffer is only borrowed during create
        // for everything that goes to HBM, and the reference's file-path variant frees it right after create
        hdr_copy.assign(zkey.data(2), zkey.data(2) + zkey.size(2));
        rebase(zkey.data(2), hdr_copy.data());
        dev.reset(new ProverDevice(pick_device(), zkey_data_from_file(zkey, hdr), shard_rank, shard_world, tables));
    }
    // from the section pointers of the reference's C++ class API (ugb_prover_create_from_sections)
    ProverBase(const ugb_zkey_sections& z, bool ultra, bool tables = true) {
        hdr.ultra = ultra;
        hdr.n8q = hdr.n8r = 32;
        hdr.nVars = z.nVars; hdr.nPublic = z.nPublic; hdr.domainSize = z.domainSize; hdr.nCoefs = z.nCoefs;
        hdr.num_indexes_c1 = z.round_indexes_count; hdr.num_indexes_c2 = z.final_round_indexes_count; hdr.rand_indx = z.challenge_index;
        // header points copied into one private block: alpha1 | beta1 | beta2 | delta1 | delta2 | round_delta1
        if (!z.vk_alpha1 || !z.vk_beta1 || !z.vk_beta2 || !z.vk_delta1 || !z.vk_delta2 || (ultra && !z.round_delta1))
            throw std::invalid_argument("Null verification-key point");
        hdr_copy.resize(64 + 64 + 128 + 64 + 128 + 64, 0);
        uint8_t* p = hdr_copy.data();
        memcpy(p, z.vk_alpha1, 64); hdr.alpha1 = p; p += 64;
        memcpy(p, z.vk_beta1, 64); hdr.beta1 = p; p += 64;
        memcpy(p, z.vk_beta2, 128); hdr.beta2 = p; p += 128;
        memcpy(p, z.vk_delta1, 64); hdr.final_delta1 = p; p += 64;
        memcpy(p, z.vk_delta2, 128); hdr.final_delta2 = p; p += 128;
        if (ultra) { memcpy(p, z.round_delta1, 64); hdr.round_delta1 = p; }
        ZkeyData d;
        d.ultra = ultra;
        d.nVars = z.nVars; d.nPublic = z.nPublic; d.domainSize = z.domainSize; d.nCoefs = z.nCoefs;
        d.coefs = (const uint8_t*)z.coefs;
        d.A = (const uint8_t*)z.pointsA; d.B1 = (const uint8_t*)z.pointsB1; d.B2 = (const uint8_t*)z.pointsB2;
        d.C = (const uint8_t*)z.pointsC; d.H = (const uint8_t*)z.pointsH; d.C1 = (const uint8_t*)z.round_pointsC;
        d.idx_c1 = (const uint8_t*)z.round_indexes; d.idx_c2 = (const uint8_t*)z.final_round_indexes;
        d.n_c1 = z.round_indexes_count; d.n_c2 = z.final_round_indexes_count;
        if (!d.coefs && d.nCoefs) throw std::invalid_argument("Null coefficient pointer");
        if (!d.A || !d.B1 || !d.B2 || !d.C || !d.H || (ultra && (!d.C1 || (!d.idx_c1 && d.n_c1) || (!d.idx_c2 && d.n_c2))))
            throw std::invalid_argument("Null zkey section pointer");
        dev.reset(new ProverDevice(pick_device(), d, 0, 1, tables));
    }
    virtual ~ProverBase() {}

    ZkeyHeader hdr;
    std::vector<uint8_t> hdr_copy;
    std::unique_ptr<ProverDevice> dev;
    Blinding blinding;
    // One prover object serialises its callers: all mutable proof state (witness, a/b/c, workspaces, pinned partial-sum slots, the
    // stream) lives in `dev`.  The reference allocates per call and so tolerates concurrent prove() calls; this keeps that safe.
    std::mutex mu;
    // parity taps of the last proof (ugb_prover_last_msms / _last_proof): the raw MSM results before blinding and assembly
    ProofMsms last_msms;
    bool has_last = false;
    std::string last_proof, last_public;
    bool is_ultra() const { return hdr.ultra; }

protected:
    void rebase(const uint8_t* old, const uint8_t* now) {
        auto mv = [&](const uint8_t*& p) { if (p) p = now + (p - old); };
        mv(hdr.qPrime); mv(hdr.rPrime); mv(hdr.alpha1); mv(hdr.beta1); mv(hdr.beta2); mv(hdr.gamma2);
        mv(hdr.round_delta1); mv(hdr.round_delta2); mv(hdr.final_delta1); mv(hdr.final_delta2);
    }
};

// sum of per-shard partial MSM results (point-range sharding: every shard returns its slice's XYZZ sums)
static ProofMsms sum_partials(const ProofMsms* parts, uint32_t count) {
    if (count == 0) throw std::invalid_argument("no partial sums");
    ProofMsms m = parts[0];
    for (uint32_t i = 1; i < count; i++) {
        pt_add(m.a, parts[i].a); pt_add(m.b1, parts[i].b1); pt_add(m.c, parts[i].c); pt_add(m.h, parts[i].h);
        pt_add(m.b2, parts[i].b2);
    }
    return m;
}

class Groth16Prover : public ProverBase {
public:
    Groth16Prover(const void* zkey_buffer, unsigned long long zkey_size, uint32_t shard_rank = 0, uint32_t shard_world = 1, bool tables = true)
        : ProverBase(zkey_buffer, zkey_size, false, shard_rank, shard_world, tables) {}
    explicit Groth16Prover(const ugb_zkey_sections& z) : ProverBase(z, false) {}

    const uint8_t* check_witness(const void* wtns_buffer, unsigned long long wtns_size) {
        BinFile wtns(wtns_buffer, wtns_size, "wtns", 2);
        WtnsHeader wh = load_wtns_header(wtns);
        if (hdr.nVars != wh.nVars)
            throw InvalidWitnessLengthException("Invalid witness length. Circuit: " + std::to_string(hdr.nVars) +
                                                ", witness: " + std::to_string(wh.nVars));
        if (!prime_is_bn254_r(wh.prime, wh.n8)) throw std::invalid_argument("different wtns curve");
        if (wtns.size(2) < (uint64_t)hdr.nVars * 32) throw std::range_error("wtns section too short: signals");
        return wtns.data(2);
    }
    // device half: the five MSMs over this prover's point range (the whole range when not sharded)
    void msms(const void* wtns_buffer, unsigned long long wtns_size, ProofMsms& m) {
        dev->prove_groth16(check_witness(wtns_buffer, wtns_size), m);
    }
    // host half: blinding and assembly (src/groth16.cpp:158-203)
    void finish_points(const ProofMsms& m, HG1Affine& A, HG2Affine& B, HG1Affine& C) {
        Scalar r = blinding.fixed ? blinding.r : random_blinding();
        Scalar s = blinding.fixed ? blinding.s : random_blinding();
        assemble(hdr, m, r, s, nullptr, A, B, C);
        last_msms = m;
        has_last = true;
    }
    // ... and JSON (src/groth16.cpp:205-250, src/prover.cpp:89-117)
    void finish(const ProofMsms& m, const uint8_t* w, std::string& stringProof, std::string& stringPublic) {
        HG1Affine A, C;
        HG2Affine B;
        finish_points(m, A, B, C);
        stringProof = "{\"pi_a\":" + g1_json(A) + ",\"pi_b\":" + g2_json(B) + ",\"pi_c\":" + g1_json(C) + ",\"protocol\":\"groth16\"}";
        stringPublic = public_json(w + 32, hdr.nPublic, 0);
    }
    // Groth16::Prover<Engine>::prove(FrElement* wtns) (src/groth16.cpp:48-203): nVars raw 32-byte signals in, proof points out
    void prove_raw(const uint8_t* w, HG1Affine& A, HG2Affine& B, HG1Affine& C) {
        ProofMsms m;
        dev->prove_groth16(w, m);
        finish_points(m, A, B, C);
    }
    void prove(const void* wtns_buffer, unsigned long long wtns_size, std::string& stringProof, std::string& stringPublic) {
        const uint8_t* w = check_witness(wtns_buffer, wtns_size);
        ProofMsms m;
        dev->prove_groth16(w, m);
        finish(m, w, stringProof, stringPublic);
    }
    // `count` proofs, pipelined (N3): witness k+1 is uploaded on a copy stream while proof k computes, and the host tail of proof k
    // (partial-sum Horner, blinding, assembly, JSON) runs after proof k+1 has been enqueued.  Same results as `count` prove() calls.
    void prove_batch(uint32_t count, const void* const* wtns_buffers, const unsigned long long* wtns_sizes, std::vector<std::string>& proofs,
                     std::vector<std::string>& publics) {
        std::vector<const uint8_t*> w(count);
        for (uint32_t k = 0; k < count; k++) w[k] = check_witness(wtns_buffers[k], wtns_sizes[k]);
        proofs.resize(count);
        publics.resize(count);
        if (count == 0) return;
        dev->batch_begin();
        dev->batch_upload(w[0], 0);
        std::string prev_sfx;
        for (uint32_t k = 0; k < count; k++) {
            const int slot = k & 1;
            dev->batch_use(slot);
            if (k + 1 < count) dev->batch_upload(w[k + 1], slot ^ 1);
            dev->phase_witness_msms();
            dev->phase_h_transform(7);
            dev->phase_h_msm(true);
            dev->batch_release(slot);
            const std::string sfx = dev->result_suffix();
            if (k > 0) {
                ProofMsms m;
                dev->phase_finish(m, &prev_sfx);
                finish(m, w[k - 1], proofs[k - 1], publics[k - 1]);
            }
            prev_sfx = sfx;
        }
        ProofMsms m;
        dev->phase_finish(m, &prev_sfx);
        finish(m, w[count - 1], proofs[count - 1], publics[count - 1]);
        dev->batch_end();
    }
    // a whole proof from the witness already resident in HBM (bench.py's device-resident leg): same device work and host tail
    void prove_resident(std::string& stringProof, std::string& stringPublic) {
        ProofMsms m;
        dev->prove_resident(m);
        std::vector<uint8_t> pub((size_t)(hdr.nPublic + 1) * 32);
        dev->download_witness(pub.data(), 0, hdr.nPublic + 1);
        finish(m, pub.data(), stringProof, stringPublic);
    }
    unsigned long long proofBufferMinSize() const { return ProofBufferMinSize(); }
    unsigned long long publicBufferMinSize() const { return PublicBufferMinSize(hdr.nPublic); }
};

class UltraGrothProver : public ProverBase {
public:
    UltraGrothProver(const void* zkey_buffer, unsigned long long zkey_size, uint32_t shard_rank = 0, uint32_t shard_world = 1, bool tables = true)
        : ProverBase(zkey_buffer, zkey_size, true, shard_rank, shard_world, tables) {}
    explicit UltraGrothProver(const ugb_zkey_sections& z) : ProverBase(z, true) {}

    // LookupInfo (src/prover.cpp:287-292, src/ultra_groth.hpp:22-61): four u32 arrays, BORROWED from the caller's wtns buffer
    // (any alignment); they are copied to HBM as they are and range-checked there (k_lookup_push)
    struct Lookup { const void *chunks = nullptr, *freqs = nullptr, *widx = nullptr, *pidx = nullptr; uint32_t chunks_len = 0, freqs_len = 0, idx_len = 0; };

    const uint8_t* check_witness(const BinFile& wtns) {
        WtnsHeader wh = load_wtns_header(wtns);
        if (hdr.nVars != wh.nVars)
            throw InvalidWitnessLengthException("Invalid witness length. Circuit: " + std::to_string(hdr.nVars) +
                                                ", witness: " + std::to_string(wh.nVars));
        if (!prime_is_bn254_r(wh.prime, wh.n8)) throw std::invalid_argument("different wtns curve");
        if (wtns.size(2) < (uint64_t)hdr.nVars * 32) throw std::range_error("wtns section too short: signals");
        return wtns.data(2);
    }
    void load_lookup(const BinFile& wtns, Lookup& lk) {
        lk.chunks = wtns.data(3); lk.freqs = wtns.data(4); lk.widx = wtns.data(5); lk.pidx = wtns.data(6);
        if ((wtns.size(3) >> 2) > 0xffffffffull || (wtns.size(4) >> 2) > 0xffffffffull || (wtns.size(5) >> 2) > 0xffffffffull)
            throw std::range_error("wtns lookup section too long");
        lk.chunks_len = (uint32_t)(wtns.size(3) >> 2);
        lk.freqs_len = (uint32_t)(wtns.size(4) >> 2);
        lk.idx_len = (uint32_t)(wtns.size(5) >> 2);
        // the reference iterates wtns_indxs_len entries of both arrays
        if ((wtns.size(6) >> 2) < lk.idx_len) throw std::range_error("wtns push index section shorter than witness index section");
    }
    const uint8_t* check_witness(const BinFile& wtns, Lookup& lk) {
        const uint8_t* w = check_witness(wtns);
        load_lookup(wtns, lk);
        return w;
    }
    void stage(const Lookup& lk) { dev->ultra_lookup_stage(lk.chunks, lk.chunks_len, lk.freqs, lk.freqs_len, lk.widx, lk.pidx, lk.idx_len); }
    // round 1, device half: upload + MSM(round_pointsC, round_wtns) over this prover's point range
    void round_msm(const uint8_t* w, HG1XYZZ& partial) {
        dev->ultra_upload_witness(w);
        dev->ultra_round_msm(partial);
    }
    // commitment = (sum of the round MSM) + r_k * final_delta1 (src/ultra_groth.cpp:161-184), then the challenge,
    // the lookup fill and the final-round MSMs over this prover's point range
    // r_k * final_delta1: O(1) host work that does not depend on the round MSM -- computed while the GPU still runs it
    HG1XYZZ rk_delta(const Scalar& rk) { return pt_mul(from_affine_bytes<HFq>(hdr.final_delta1), rk); }
    void final_msms(HG1XYZZ commit, const Scalar& rk, const Lookup* lk, HG1Affine& commit_aff, ProofMsms& m, const HG1XYZZ* rkd = nullptr) {
        m.round_c = commit;  // the raw round MSM (parity tap)
        pt_add(commit, rkd ? *rkd : rk_delta(rk));
        commit_aff = pt_to_affine(commit);
        HFr rand_mont = derive_challenge(commit_aff);  // src/ultra_groth.cpp:32-58
        if (lk) stage(*lk);
        dev->ultra_lookup_apply(rand_mont);
        HG1XYZZ raw_round = m.round_c;
        dev->ultra_final_msms(m);
        m.round_c = raw_round;
    }
    void finish_points(const ProofMsms& m, const Scalar& rk, HG1Affine& A, HG2Affine& B, HG1Affine& C) {
        Scalar r = blinding.fixed ? blinding.r : random_blinding();
        Scalar s = blinding.fixed ? blinding.s : random_blinding();
        assemble(hdr, m, r, s, &rk, A, B, C);
        last_msms = m;
        has_last = true;
    }
    // the public signals as they stand in HBM (the challenge sits among them, filled in by compute_lookup)
    std::vector<uint8_t> public_signals() {
        std::vector<uint8_t> pub((size_t)hdr.nPublic * 32);
        if (hdr.nPublic) dev->download_witness(pub.data(), 1, hdr.nPublic);
        return pub;
    }
    void finish_with_public(const ProofMsms& m, const Scalar& rk, const HG1Affine& commit_aff, const std::vector<uint8_t>& pub,
                            std::string& stringProof, std::string& stringPublic) {
        HG1Affine A, C;
        HG2Affine B;
        finish_points(m, rk, A, B, C);
        stringProof = "{\"pi_a\":" + g1_json(A) + ",\"pi_b\":" + g2_json(B) + ",\"pi_f\":" + g1_json(C) + ",\"pi_r\":" + g1_json(commit_aff) +
                      ",\"protocol\":\"ultragroth\"}";
        stringPublic = public_json(pub.data(), hdr.nPublic, hdr.rand_indx);
    }
    void finish(const ProofMsms& m, const Scalar& rk, const HG1Affine& commit_aff, std::string& stringProof, std::string& stringPublic) {
        finish_with_public(m, rk, commit_aff, public_signals(), stringProof, stringPublic);
    }
    // UltraGroth::Prover::prove(FrElement* wtns, LookupInfo&) (src/ultra_groth.cpp:401-462): raw signals + lookup arrays in, points out.
    // wtns_out: the reference fills the lookup signals into the CALLER's witness (compute_lookup writes `signals`); when non-null the
    // same nVars*32 bytes are written back.
    void prove_raw(const uint8_t* w, const Lookup& lk, HG1Affine& A, HG2Affine& B, HG1Affine& C, HG1Affine& R, uint8_t* wtns_out) {
        // the witness copy is asynchronous; the lookup arrays follow it on the same stream
        // the head of the witness and the round commitment first; the lookup arrays follow the round kernels on the stream (they share
        // the host->device copy engine with the tail of the witness, which travels on the copy stream meanwhile)
        dev->ultra_upload_witness(w);
        dev->ultra_round_launch();
        stage(lk);
        Scalar rk = blinding.fixed ? blinding.rk : random_blinding();
        const HG1XYZZ rkd = rk_delta(rk);
        HG1XYZZ commit;
        dev->ultra_round_finish(commit);
        ProofMsms m;
        final_msms(commit, rk, nullptr, R, m, &rkd);
        finish_points(m, rk, A, B, C);
        if (wtns_out) dev->download_witness(wtns_out, 0, hdr.nVars);
    }
    void prove(const void* wtns_buffer, unsigned long long wtns_size, std::string& stringProof, std::string& stringPublic) {
        BinFile wtns(wtns_buffer, wtns_size, "wtns", 2);
        const uint8_t* w = check_witness(wtns);
        dev->ultra_upload_witness(w);
        dev->ultra_round_launch();
        Lookup lk;
        load_lookup(wtns, lk);
        stage(lk);
        Scalar rk = blinding.fixed ? blinding.rk : random_blinding();
        const HG1XYZZ rkd = rk_delta(rk);
        HG1XYZZ commit;
        dev->ultra_round_finish(commit);
        HG1Affine commit_aff;
        ProofMsms m;
        final_msms(commit, rk, nullptr, commit_aff, m, &rkd);
        finish(m, rk, commit_aff, stringProof, stringPublic);
    }
    // `count` proofs, pipelined (N3): the upload of witness k+1 runs on a copy stream under the kernels of proof k, and the host tail of
    // proof k is worked off while the GPU already runs the round commitment of proof k+1.  Same results as `count` prove() calls.
    void prove_batch(uint32_t count, const void* const* wtns_buffers, const unsigned long long* wtns_sizes, std::vector<std::string>& proofs,
                     std::vector<std::string>& publics) {
        std::vector<const uint8_t*> w(count);
        std::vector<Lookup> lk(count);
        for (uint32_t k = 0; k < count; k++) {
            BinFile wtns(wtns_buffers[k], wtns_sizes[k], "wtns", 2);
            w[k] = check_witness(wtns, lk[k]);
        }
        proofs.resize(count);
        publics.resize(count);
        if (count == 0) return;
        BatchTrace tr;
        dev->batch_begin();
        dev->batch_upload(w[0], 0);
        ProofMsms prev_m;
        Scalar prev_rk;
        HG1Affine prev_commit;
        std::vector<uint8_t> prev_pub;
        for (uint32_t k = 0; k < count; k++) {
            const int slot = k & 1;
            tr.at(k, "begin");
            dev->batch_use(slot);
            stage(lk[k]);
            dev->ultra_round_launch();
            // only now: the next witness shares the host->device copy engine with the lookup arrays above, which gate the round MSM
            if (k + 1 < count) dev->batch_upload(w[k + 1], slot ^ 1);
            tr.at(k, "round enqueued");
            if (k > 0) {
                // proof k-1: its MSM results were waited for at the end of the previous iteration (they must leave the pinned slots
                // before proof k's final round is enqueued); the host tail runs here, under proof k's round commitment
                finish_with_public(prev_m, prev_rk, prev_commit, prev_pub, proofs[k - 1], publics[k - 1]);
            }
            Scalar rk = blinding.fixed ? blinding.rk : random_blinding();
            const HG1XYZZ rkd = rk_delta(rk);
            tr.at(k, "host tail k-1 done");
            HG1XYZZ commit;
            dev->ultra_round_finish(commit);
            tr.at(k, "round result");
            HG1Affine commit_aff;
            ProofMsms m;
            m.round_c = commit;
            pt_add(commit, rkd);
            commit_aff = pt_to_affine(commit);
            dev->ultra_lookup_apply(derive_challenge(commit_aff));
            dev->phase_witness_msms();
            dev->phase_h_transform(7);
            dev->phase_h_msm(true);
            dev->batch_release(slot);
            tr.at(k, "final enqueued");
            const HG1XYZZ raw_round = m.round_c;
            dev->phase_finish(m);
            m.round_c = raw_round;
            tr.at(k, "msm results");
            prev_m = m; prev_rk = rk; prev_commit = commit_aff;
            prev_pub = public_signals();  // before the next proof switches witness buffers
            if (k + 1 == count) finish_with_public(m, rk, commit_aff, prev_pub, proofs[k], publics[k]);
        }
        tr.at(count, "end");
        dev->batch_end();
    }
    // a whole proof (both rounds, challenge, lookup fill, assembly) from the witness and lookup arrays left in HBM by the previous proof
    void prove_resident(std::string& stringProof, std::string& stringPublic) {
        dev->ultra_round_launch();
        Scalar rk = blinding.fixed ? blinding.rk : random_blinding();
        const HG1XYZZ rkd = rk_delta(rk);
        HG1XYZZ commit;
        dev->ultra_round_finish(commit);
        HG1Affine commit_aff;
        ProofMsms m;
        final_msms(commit, rk, nullptr, commit_aff, m, &rkd);
        finish(m, rk, commit_aff, stringProof, stringPublic);
    }
    HG1Affine shard_commit_aff;  // sharded v2: the blinded round commitment between ugb_shard_lookup and ugb_shard_finish
    unsigned long long proofBufferMinSize() const { return ProofBufferMinSizeUltraGroth(); }
    unsigned long long publicBufferMinSize() const { return PublicBufferMinSize((unsigned long long)hdr.nPublic - 1); }

    static HFr derive_challenge(const HG1Affine& p) {
        uint8_t buf[64], digest[32];
        HFq x = from_mont(p.x), y = from_mont(p.y);
        for (int i = 0; i < 32; i++) {
            buf[i] = (uint8_t)(x.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
            buf[32 + i] = (uint8_t)(y.v[(31 - i) >> 3] >> (8 * ((31 - i) & 7)));
        }
        keccak256(buf, 64, digest);
        HFr raw = HFr::zero();  // digest read big-endian, stored unreduced, then one MMul by R^2
        for (int i = 0; i < 32; i++) raw.v[(31 - i) >> 3] |= (uint64_t)digest[i] << (8 * ((31 - i) & 7));
        return to_mont(raw);
    }
};

std::vector<uint8_t> read_file(const char* path) {
    if (!path) throw std::invalid_argument("Null zkey file path");
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw std::system_error(errno, std::generic_category(), std::string("open ") + path);
    std::streamsize n = f.tellg();
    f.seekg(0);
    std::vector<uint8_t> buf((size_t)n);
    if (n && !f.read((char*)buf.data(), n)) throw std::system_error(errno, std::generic_category(), std::string("read ") + path);
    return buf;
}

#define API_TRY try {
#define API_CATCH_CREATE                                                                                  \
    }                                                                                                     \
    catch (std::exception & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_ERROR; } \
    catch (std::exception * e) { CopyError(error_msg, error_msg_maxsize, e->what()); delete e; return PROVER_ERROR; } \
    catch (...) { CopyError(error_msg, error_msg_maxsize, "unknown error"); return PROVER_ERROR; }         \
    return PROVER_OK;
#define API_CATCH_PROVE                                                                                                    \
    }                                                                                                                      \
    catch (InvalidWitnessLengthException & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_INVALID_WITNESS_LENGTH; } \
    catch (ShortBufferException & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_ERROR_SHORT_BUFFER; } \
    catch (std::exception & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_ERROR; }                   \
    catch (std::exception * e) { CopyError(error_msg, error_msg_maxsize, e->what()); delete e; return PROVER_ERROR; }        \
    catch (...) { CopyError(error_msg, error_msg_maxsize, "unknown error"); return PROVER_ERROR; }                           \
    return PROVER_OK;

template <class P>
int prover_prove_impl(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, char* proof_buffer,
                      unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                      unsigned long long error_msg_maxsize) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    if (wtns_buffer == NULL) throw std::invalid_argument("Null witness buffer");
    if (proof_buffer == NULL) throw std::invalid_argument("Null proof buffer");
    if (proof_size == NULL) throw std::invalid_argument("Null proof size");
    if (public_buffer == NULL) throw std::invalid_argument("Null public buffer");
    if (public_size == NULL) throw std::invalid_argument("Null public size");
    P* prover = static_cast<P*>(static_cast<ProverBase*>(prover_object));
    std::lock_guard<std::mutex> guard(prover->mu);
    CheckAndUpdateBufferSizes(prover->proofBufferMinSize(), proof_size, prover->publicBufferMinSize(), public_size, "Minimum");
    std::string stringProof, stringPublic;
    prover->prove(wtns_buffer, wtns_size, stringProof, stringPublic);
    CheckAndUpdateBufferSizes(stringProof.length(), proof_size, stringPublic.length(), public_size, "Required");
    std::strncpy(proof_buffer, stringProof.c_str(), *proof_size);
    std::strncpy(public_buffer, stringPublic.c_str(), *public_size);
    API_CATCH_PROVE
}

template <class P>
int prover_create_impl(void** prover_object, const void* zkey_buffer, unsigned long long zkey_size, char* error_msg,
                       unsigned long long error_msg_maxsize, bool tables = true) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    if (zkey_buffer == NULL) throw std::invalid_argument("Null zkey buffer");
    *prover_object = static_cast<ProverBase*>(new P(zkey_buffer, zkey_size, 0, 1, tables));
    API_CATCH_CREATE
}

int public_size_impl(const void* zkey_buffer, unsigned long long zkey_size, bool ultra, unsigned long long* public_size, char* error_msg,
                     unsigned long long error_msg_maxsize) {
    API_TRY
    BinFile zkey(zkey_buffer, zkey_size, "zkey", 1);
    ZkeyHeader h = load_zkey_header(zkey, ultra);
    *public_size = PublicBufferMinSize(h.nPublic);
    API_CATCH_CREATE
}

}  // namespace

extern "C" {

int groth16_public_size_for_zkey_buf(const void* zkey_buffer, unsigned long long zkey_size, unsigned long long* public_size, char* error_msg,
                                     unsigned long long error_msg_maxsize) {
    return public_size_impl(zkey_buffer, zkey_size, false, public_size, error_msg, error_msg_maxsize);
}
int ultra_groth_public_size_for_zkey_buf(const void* zkey_buffer, unsigned long long zkey_size, unsigned long long* public_size,
                                         char* error_msg, unsigned long long error_msg_maxsize) {
    return public_size_impl(zkey_buffer, zkey_size, true, public_size, error_msg, error_msg_maxsize);
}
int groth16_public_size_for_zkey_file(const char* zkey_fname, unsigned long long* public_size, char* error_msg,
                                      unsigned long long error_msg_maxsize) {
    API_TRY
    std::vector<uint8_t> buf = read_file(zkey_fname);
    BinFile zkey(buf.data(), buf.size(), "zkey", 1);
    *public_size = PublicBufferMinSize(load_zkey_header(zkey, false).nPublic);
    API_CATCH_CREATE
}
int ultra_groth_public_size_for_zkey_file(const char* zkey_fname, unsigned long long* public_size, char* error_msg,
                                          unsigned long long error_msg_maxsize) {
    API_TRY
    std::vector<uint8_t> buf = read_file(zkey_fname);
    BinFile zkey(buf.data(), buf.size(), "zkey", 1);
    *public_size = PublicBufferMinSize(load_zkey_header(zkey, true).nPublic);
    API_CATCH_CREATE
}

void groth16_proof_size(unsigned long long* proof_size) { *proof_size = ProofBufferMinSize(); }
void ultra_groth_proof_size(unsigned long long* proof_size) { *proof_size = ProofBufferMinSizeUltraGroth(); }

int groth16_prover_create(void** prover_object, const void* zkey_buffer, unsigned long long zkey_size, char* error_msg,
                          unsigned long long error_msg_maxsize) {
    return prover_create_impl<Groth16Prover>(prover_object, zkey_buffer, zkey_size, error_msg, error_msg_maxsize);
}
int ultra_groth_prover_create(void** prover_object, const void* zkey_buffer, unsigned long long zkey_size, char* error_msg,
                              unsigned long long error_msg_maxsize) {
    return prover_create_impl<UltraGrothProver>(prover_object, zkey_buffer, zkey_size, error_msg, error_msg_maxsize);
}
int groth16_prover_create_zkey_file(void** prover_object, const char* zkey_file_path, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    std::vector<uint8_t> buf = read_file(zkey_file_path);
    *prover_object = static_cast<ProverBase*>(new Groth16Prover(buf.data(), buf.size()));
    API_CATCH_CREATE
}
int ultra_groth_prover_create_zkey_file(void** prover_object, const char* zkey_file_path, char* error_msg,
                                        unsigned long long error_msg_maxsize) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    std::vector<uint8_t> buf = read_file(zkey_file_path);
    *prover_object = static_cast<ProverBase*>(new UltraGrothProver(buf.data(), buf.size()));
    API_CATCH_CREATE
}

int groth16_prover_prove(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, char* proof_buffer,
                         unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                         unsigned long long error_msg_maxsize) {
    return prover_prove_impl<Groth16Prover>(prover_object, wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size,
                                            error_msg, error_msg_maxsize);
}
int ultra_groth_prover_prove(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, char* proof_buffer,
                             unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                             unsigned long long error_msg_maxsize) {
    return prover_prove_impl<UltraGrothProver>(prover_object, wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size,
                                               error_msg, error_msg_maxsize);
}

void groth16_prover_destroy(void* prover_object) {
    if (prover_object != NULL) delete static_cast<ProverBase*>(prover_object);
}
void ultra_groth_prover_destroy(void* prover_object) {
    if (prover_object != NULL) delete static_cast<ProverBase*>(prover_object);
}

int groth16_prover(const void* zkey_buffer, unsigned long long zkey_size, const void* wtns_buffer, unsigned long long wtns_size,
                   char* proof_buffer, unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                   unsigned long long error_msg_maxsize) {
    void* prover = NULL;
    // one proof only: the window tables (1.3 s to build at 2^22) cannot pay for themselves, so the one-level bases are used
    int error = prover_create_impl<Groth16Prover>(&prover, zkey_buffer, zkey_size, error_msg, error_msg_maxsize, false);
    if (error != PROVER_OK) return error;
    error = groth16_prover_prove(prover, wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size, error_msg,
                                 error_msg_maxsize);
    groth16_prover_destroy(prover);
    return error;
}
int ultra_groth_prover(const void* zkey_buffer, unsigned long long zkey_size, const void* wtns_buffer, unsigned long long wtns_size,
                       char* proof_buffer, unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size,
                       char* error_msg, unsigned long long error_msg_maxsize) {
    void* prover = NULL;
    int error = prover_create_impl<UltraGrothProver>(&prover, zkey_buffer, zkey_size, error_msg, error_msg_maxsize, false);
    if (error != PROVER_OK) return error;
    error = ultra_groth_prover_prove(prover, wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size, error_msg,
                                     error_msg_maxsize);
    ultra_groth_prover_destroy(prover);
    return error;
}
int groth16_prover_zkey_file(const char* zkey_file_path, const void* wtns_buffer, unsigned long long wtns_size, char* proof_buffer,
                             unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                             unsigned long long error_msg_maxsize) {
    std::vector<uint8_t> zkey;
    API_TRY
    zkey = read_file(zkey_file_path);
    }
    catch (std::exception & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_ERROR; }
    catch (...) { CopyError(error_msg, error_msg_maxsize, "unknown error"); return PROVER_ERROR; }
    return groth16_prover(zkey.data(), zkey.size(), wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size, error_msg,
                     error_msg_maxsize);
}
int ultra_groth_prover_zkey_file(const char* zkey_file_path, const void* wtns_buffer, unsigned long long wtns_size, char* proof_buffer,
                                 unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                                 unsigned long long error_msg_maxsize) {
    std::vector<uint8_t> zkey;
    API_TRY
    zkey = read_file(zkey_file_path);
    }
    catch (std::exception & e) { CopyError(error_msg, error_msg_maxsize, e.what()); return PROVER_ERROR; }
    catch (...) { CopyError(error_msg, error_msg_maxsize, "unknown error"); return PROVER_ERROR; }
    return ultra_groth_prover(zkey.data(), zkey.size(), wtns_buffer, wtns_size, proof_buffer, proof_size, public_buffer, public_size, error_msg,
                     error_msg_maxsize);
}

// ---- extensions declared in include/ultragroth_b200.h ---------------------------------------------
int ugb_prover_set_blinding(void* prover_object, const void* r32, const void* s32, const void* rk32) {
    if (!prover_object) return UGB_ERROR;
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    if (!r32 || !s32) { p->blinding.fixed = false; return UGB_OK; }
    p->blinding.fixed = true;
    memcpy(&p->blinding.r, r32, 32);
    memcpy(&p->blinding.s, s32, 32);
    if (rk32) memcpy(&p->blinding.rk, rk32, 32); else memset(&p->blinding.rk, 0, 32);
    return UGB_OK;
}
#define EXT_TRY try {
#define EXT_CATCH                                                                 \
    }                                                                             \
    catch (const std::exception& e) { ug::set_last_error(e.what()); return UGB_ERROR; } \
    catch (...) { ug::set_last_error("unknown error"); return UGB_ERROR; }        \
    return UGB_OK;

int ugb_prover_last_h(void* prover_object, void* h_out, uint64_t bytes) {
    EXT_TRY
    if (!prover_object || !h_out) throw std::invalid_argument("Null argument");
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    std::vector<uint8_t> h;
    p->dev->download_h(h);
    memcpy(h_out, h.data(), std::min<uint64_t>(bytes, h.size()));
    EXT_CATCH
}
// 448 bytes: A (64) | B1 (64) | B2 (128) | C (64) | H (64) | round C1 (64, zero for Groth16), affine Montgomery, zkey layout
int ugb_prover_last_msms(void* prover_object, void* out448) {
    EXT_TRY
    if (!prover_object || !out448) throw std::invalid_argument("Null argument");
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    if (!p->has_last) throw std::runtime_error("no proof has been made with this prover yet");
    const ProofMsms& m = p->last_msms;
    uint8_t* o = (uint8_t*)out448;
    HG1Affine t;
    HG2Affine t2 = pt_to_affine(m.b2);
    t = pt_to_affine(m.a); memcpy(o, &t, 64);
    t = pt_to_affine(m.b1); memcpy(o + 64, &t, 64);
    memcpy(o + 128, &t2, 128);
    t = pt_to_affine(m.c); memcpy(o + 256, &t, 64);
    t = pt_to_affine(m.h); memcpy(o + 320, &t, 64);
    if (p->is_ultra()) { t = pt_to_affine(m.round_c); memcpy(o + 384, &t, 64); }
    else memset(o + 384, 0, 64);
    EXT_CATCH
}
int ugb_prover_last_witness(void* prover_object, void* out, uint64_t first, uint64_t count) {
    EXT_TRY
    if (!prover_object || !out) throw std::invalid_argument("Null argument");
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    if (first + count > p->hdr.nVars) throw std::range_error("witness range out of bounds");
    p->dev->download_witness((uint8_t*)out, first, count);
    EXT_CATCH
}
int ugb_prover_set_timing(void* prover_object, int enable) {
    if (!prover_object) return UGB_ERROR;
    static_cast<ProverBase*>(prover_object)->dev->timing = enable != 0;
    return UGB_OK;
}
int ugb_prover_last_times(void* prover_object, float* out, int count) {
    if (!prover_object) return UGB_ERROR;
    const StageTimes& t = static_cast<ProverBase*>(prover_object)->dev->last_times;
    const float v[] = {t.upload, t.round, t.lookup, t.sort_w, t.msm_a, t.msm_b1, t.msm_b2, t.msm_c, t.spmv, t.intt, t.ntt, t.hfinal, t.sort_h, t.msm_h, t.total};
    for (int i = 0; i < count && i < (int)(sizeof(v) / sizeof(v[0])); i++) out[i] = v[i];
    return UGB_OK;
}
int ugb_prover_create_times(void* prover_object, double* out4) {
    if (!prover_object || !out4) return UGB_ERROR;
    for (int i = 0; i < 4; i++) out4[i] = static_cast<ProverBase*>(prover_object)->dev->create_ms[i];
    return UGB_OK;
}
int ugb_prover_inf_points(void* prover_object, uint64_t* out3) {
    if (!prover_object || !out3) return UGB_ERROR;
    for (int i = 0; i < 3; i++) out3[i] = static_cast<ProverBase*>(prover_object)->dev->inf_points[i];
    return UGB_OK;
}
// One WHOLE proof from the witness (and, UltraGroth, lookup arrays) left in HBM by the previous *_prover_prove call: both rounds,
// challenge, lookup fill, every MSM, the H pipeline, blinding, assembly and JSON -- everything except the host->device input copy.
// The proof is kept and can be read back with ugb_prover_last_proof.
int ugb_prover_prove_resident(void* prover_object) {
    EXT_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    if (!p->has_last) throw std::runtime_error("ugb_prover_prove_resident needs a previous *_prover_prove call (no witness in HBM yet)");
    if (p->is_ultra()) static_cast<UltraGrothProver*>(p)->prove_resident(p->last_proof, p->last_public);
    else static_cast<Groth16Prover*>(p)->prove_resident(p->last_proof, p->last_public);
    EXT_CATCH
}
// proof / public JSON of the last ugb_prover_prove_resident call (NUL terminated, truncated to the capacities)
int ugb_prover_last_proof(void* prover_object, char* proof_buffer, uint64_t proof_cap, char* public_buffer, uint64_t public_cap) {
    EXT_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    ProverBase* p = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(p->mu);
    if (proof_buffer && proof_cap) { strncpy(proof_buffer, p->last_proof.c_str(), proof_cap - 1); proof_buffer[proof_cap - 1] = 0; }
    if (public_buffer && public_cap) { strncpy(public_buffer, p->last_public.c_str(), public_cap - 1); public_buffer[public_cap - 1] = 0; }
    EXT_CATCH
}
int ugb_prover_set_profile(void* prover_object, int enable) {
    if (!prover_object) return UGB_ERROR;
    DeviceCtx& c = static_cast<ProverBase*>(prover_object)->dev->ctx();
    c.profile = enable != 0;
    if (enable) { c.kernel_stats.clear(); c.counters.clear(); }
    return UGB_OK;
}
// "name\tlaunches\tms\n" per kernel, NUL terminated; returns the number of bytes needed
uint64_t ugb_prover_kernel_stats(void* prover_object, char* out, uint64_t cap) {
    if (!prover_object) return 0;
    DeviceCtx& c = static_cast<ProverBase*>(prover_object)->dev->ctx();
    std::string s;
    for (auto& kv : c.kernel_stats) s += kv.first + "\t" + std::to_string(kv.second.launches) + "\t" + std::to_string(kv.second.ms) + "\n";
    for (auto& kv : c.counters) s += kv.first + "\t" + std::to_string(kv.second) + "\t0\n";
    if (out && cap) { strncpy(out, s.c_str(), cap - 1); out[cap - 1] = 0; }
    return s.size() + 1;
}
uint64_t ugb_prover_kernel_launches(void* prover_object) {
    if (!prover_object) return 0;
    return static_cast<ProverBase*>(prover_object)->dev->ctx().kernel_launches;
}
void* ugb_prover_stream(void* prover_object) {
    if (!prover_object) return nullptr;
    return (void*)static_cast<ProverBase*>(prover_object)->dev->ctx().stream;
}


// ---- the reference's C++ class API, C-ABI side (include/groth16.hpp / include/ultra_groth.hpp wrap these) ----
// Groth16::makeProver<Engine>(nVars, nPublic, domainSize, nCoefs, vk pointers, coefs, pointsA..H) (src/groth16.cpp:17-45) and
// UltraGroth::makeProver (src/ultra_groth.cpp:108-159): the zkey sections as HOST pointers; everything is copied to HBM here.
int ugb_prover_create_from_sections(void** prover_object, int ultra, const ugb_zkey_sections* sections, char* error_msg,
                                    unsigned long long error_msg_maxsize) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    if (sections == NULL) throw std::invalid_argument("Null zkey sections");
    if (ultra) *prover_object = static_cast<ProverBase*>(new UltraGrothProver(*sections));
    else *prover_object = static_cast<ProverBase*>(new Groth16Prover(*sections));
    API_CATCH_CREATE
}

// Groth16::Prover<Engine>::prove(FrElement* wtns) (src/groth16.cpp:48-203): nVars raw (non-Montgomery) 32-byte signals in; the proof
// points out in the reference's G1PointAffine / G2PointAffine layout (Montgomery coordinates, zkey layout).
int ugb_groth16_prove_raw(void* prover_object, const void* wtns_signals, void* out_A64, void* out_B128, void* out_C64, char* error_msg,
                          unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_signals || !out_A64 || !out_B128 || !out_C64) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (b->is_ultra()) throw std::invalid_argument("not a groth16 prover");
    HG1Affine A, C;
    HG2Affine B;
    static_cast<Groth16Prover*>(b)->prove_raw((const uint8_t*)wtns_signals, A, B, C);
    memcpy(out_A64, &A, 64); memcpy(out_B128, &B, 128); memcpy(out_C64, &C, 64);
    API_CATCH_PROVE
}

// UltraGroth::Prover<Engine>::prove(FrElement* wtns, LookupInfo&) (src/ultra_groth.cpp:401-462).  wtns_signals is IN/OUT like the
// reference's: compute_lookup's signals (src/ultra_groth.cpp:62-106) are written back into it.
int ugb_ultra_groth_prove_raw(void* prover_object, void* wtns_signals, const uint32_t* chunks, uint32_t chunks_len, const uint32_t* frequencies,
                              uint32_t frequencies_len, const uint32_t* wtns_indxs, const uint32_t* push_indxs, uint32_t wtns_indxs_len,
                              void* out_A64, void* out_B128, void* out_final_commitment64, void* out_round_commitment64, char* error_msg,
                              unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_signals || !out_A64 || !out_B128 || !out_final_commitment64 || !out_round_commitment64)
        throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    UltraGrothProver::Lookup lk;
    lk.chunks = chunks; lk.freqs = frequencies; lk.widx = wtns_indxs; lk.pidx = push_indxs;
    lk.chunks_len = chunks_len; lk.freqs_len = frequencies_len; lk.idx_len = wtns_indxs_len;
    HG1Affine A, C, R;
    HG2Affine B;
    static_cast<UltraGrothProver*>(b)->prove_raw((const uint8_t*)wtns_signals, lk, A, B, C, R, (uint8_t*)wtns_signals);
    memcpy(out_A64, &A, 64); memcpy(out_B128, &B, 128); memcpy(out_final_commitment64, &C, 64); memcpy(out_round_commitment64, &R, 64);
    API_CATCH_PROVE
}

// ---- sharded proving, phase by phase (declared in include/ultragroth_b200.h) ----
int ugb_shard_layout_get(void* prover_object, ugb_shard_layout* out) {
    EXT_TRY
    if (!prover_object || !out) throw std::invalid_argument("Null argument");
    ProverDevice& d = *static_cast<ProverBase*>(prover_object)->dev;
    out->w = d.dev_w(); out->a = d.dev_a(); out->b = d.dev_b(); out->c = d.dev_c();
    out->nVars = d.nVars; out->domainSize = d.domainSize;
    out->w_chunk = d.w_chunk(); out->h_chunk = d.h_chunk();
    out->rank = d.shard_rank(); out->world = d.shard_world();
    out->device = d.ctx().device;
    out->stream = (void*)d.ctx().stream;
    EXT_CATCH
}
int ugb_shard_upload_slice(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, char* error_msg,
                           unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_buffer) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (b->is_ultra()) {
        UltraGrothProver* p = static_cast<UltraGrothProver*>(b);
        BinFile wtns(wtns_buffer, wtns_size, "wtns", 2);
        UltraGrothProver::Lookup lk;
        const uint8_t* w = p->check_witness(wtns, lk);
        b->dev->upload_witness_slice(w);
        p->stage(lk);
    } else {
        b->dev->upload_witness_slice(static_cast<Groth16Prover*>(b)->check_witness(wtns_buffer, wtns_size));
    }
    API_CATCH_PROVE
}
int ugb_shard_round_partial(void* prover_object, void* round_partial_out128, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !round_partial_out128) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    HG1XYZZ part;
    b->dev->ultra_round_msm(part);
    memcpy(round_partial_out128, &part, sizeof(part));
    API_CATCH_PROVE
}
int ugb_shard_lookup(void* prover_object, const void* round_partials, uint32_t count, const void* rk32, char* error_msg,
                     unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !round_partials || !rk32 || count == 0) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    UltraGrothProver* p = static_cast<UltraGrothProver*>(b);
    std::vector<HG1XYZZ> rp(count);
    memcpy(rp.data(), round_partials, (size_t)count * sizeof(HG1XYZZ));
    HG1XYZZ commit = rp[0];
    for (uint32_t i = 1; i < count; i++) pt_add(commit, rp[i]);
    Scalar rk;
    memcpy(&rk, rk32, 32);
    pt_add(commit, pt_mul(from_affine_bytes<HFq>(p->hdr.final_delta1), rk));  // src/ultra_groth.cpp:161-184
    p->shard_commit_aff = pt_to_affine(commit);
    b->dev->ultra_lookup_apply(UltraGrothProver::derive_challenge(p->shard_commit_aff));
    API_CATCH_PROVE
}
int ugb_shard_h_transform(void* prover_object, int vector_mask, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    b->dev->phase_h_transform(vector_mask & 7);
    API_CATCH_PROVE
}
int ugb_shard_witness_msms(void* prover_object, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    b->dev->phase_witness_msms();
    API_CATCH_PROVE
}
int ugb_shard_h_msm(void* prover_object, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    b->dev->phase_h_msm(false);
    API_CATCH_PROVE
}
int ugb_shard_finish(void* prover_object, void* partial_out, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !partial_out) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    ProofMsms m;
    b->dev->phase_finish(m);
    // UltraGroth: round_c carries the blinded commitment to the assembling rank (identical on every shard)
    if (b->is_ultra()) m.round_c = XYZZ<HFq>::from_affine(static_cast<UltraGrothProver*>(b)->shard_commit_aff);
    else m.round_c = HG1XYZZ::inf();
    memcpy(partial_out, &m, sizeof(m));
    API_CATCH_PROVE
}

// ---- N3: `count` proofs on one prover object, pipelined; same outputs as `count` *_prover_prove calls (include/ultragroth_b200.h) ----
int ugb_prover_prove_batch(void* prover_object, uint32_t count, const void* const* wtns_buffers, const unsigned long long* wtns_sizes,
                           char* const* proof_buffers, unsigned long long* proof_sizes, char* const* public_buffers, unsigned long long* public_sizes,
                           char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object) throw std::invalid_argument("Null prover object");
    if (count && (!wtns_buffers || !wtns_sizes || !proof_buffers || !proof_sizes || !public_buffers || !public_sizes)) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    const unsigned long long pmin = b->is_ultra() ? ProofBufferMinSizeUltraGroth() : ProofBufferMinSize();
    const unsigned long long umin = PublicBufferMinSize(b->is_ultra() ? (unsigned long long)b->hdr.nPublic - 1 : b->hdr.nPublic);
    for (uint32_t k = 0; k < count; k++) {
        if (!wtns_buffers[k]) throw std::invalid_argument("Null witness buffer");
        if (!proof_buffers[k]) throw std::invalid_argument("Null proof buffer");
        if (!public_buffers[k]) throw std::invalid_argument("Null public buffer");
        CheckAndUpdateBufferSizes(pmin, &proof_sizes[k], umin, &public_sizes[k], "Minimum");
    }
    std::vector<std::string> proofs, publics;
    if (b->is_ultra()) static_cast<UltraGrothProver*>(b)->prove_batch(count, wtns_buffers, wtns_sizes, proofs, publics);
    else static_cast<Groth16Prover*>(b)->prove_batch(count, wtns_buffers, wtns_sizes, proofs, publics);
    for (uint32_t k = 0; k < count; k++) {
        CheckAndUpdateBufferSizes(proofs[k].length(), &proof_sizes[k], publics[k].length(), &public_sizes[k], "Required");
        std::strncpy(proof_buffers[k], proofs[k].c_str(), proof_sizes[k]);
        std::strncpy(public_buffers[k], publics[k].c_str(), public_sizes[k]);
    }
    API_CATCH_PROVE
}

void ugb_prover_destroy(void* prover_object) {
    if (prover_object != NULL) delete static_cast<ProverBase*>(prover_object);
}

int ugb_proof_points_to_json(int ultra, const void* A64, const void* B128, const void* C64, const void* R64, char* out, uint64_t cap) {
    EXT_TRY
    if (!A64 || !B128 || !C64 || !out || (ultra && !R64)) throw std::invalid_argument("Null argument");
    HG1Affine A, C, R;
    HG2Affine B;
    memcpy(&A, A64, 64); memcpy(&B, B128, 128); memcpy(&C, C64, 64);
    std::string js;
    if (ultra) {
        memcpy(&R, R64, 64);
        js = "{\"pi_a\":" + g1_json(A) + ",\"pi_b\":" + g2_json(B) + ",\"pi_f\":" + g1_json(C) + ",\"pi_r\":" + g1_json(R) + ",\"protocol\":\"ultragroth\"}";
    } else {
        js = "{\"pi_a\":" + g1_json(A) + ",\"pi_b\":" + g2_json(B) + ",\"pi_c\":" + g1_json(C) + ",\"protocol\":\"groth16\"}";
    }
    if (js.size() + 1 > cap) throw std::length_error("proof JSON buffer too short");
    memcpy(out, js.c_str(), js.size() + 1);
    EXT_CATCH
}

// ---- point-range sharded proving (declared in include/ultragroth_b200.h) ---------
// One prover object per GPU/process holds slice `rank` of `world` of every base array.  The MSM partial sums are
// plain host structs (ugb_prover_partial_size() bytes) that the caller exchanges (NCCL / gloo all-gather, or a
// host gather) and hands back to *_assemble on the rank that emits the proof.
uint64_t ugb_prover_partial_size(void) { return sizeof(ProofMsms); }

int ugb_prover_create_sharded(void** prover_object, const void* zkey_buffer, unsigned long long zkey_size, int ultra, uint32_t rank,
                              uint32_t world, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (prover_object == NULL) throw std::invalid_argument("Null prover object");
    if (zkey_buffer == NULL) throw std::invalid_argument("Null zkey buffer");
    if (ultra) *prover_object = static_cast<ProverBase*>(new UltraGrothProver(zkey_buffer, zkey_size, rank, world));
    else *prover_object = static_cast<ProverBase*>(new Groth16Prover(zkey_buffer, zkey_size, rank, world));
    API_CATCH_CREATE
}

int ugb_groth16_prove_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, void* partial_out,
                              char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_buffer || !partial_out) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (b->is_ultra()) throw std::invalid_argument("not a groth16 prover");
    ProofMsms m;
    static_cast<Groth16Prover*>(b)->msms(wtns_buffer, wtns_size, m);
    memcpy(partial_out, &m, sizeof(m));
    API_CATCH_PROVE
}

int ugb_groth16_prove_assemble(void* prover_object, const void* partials, uint32_t count, const void* wtns_buffer,
                               unsigned long long wtns_size, char* proof_buffer, unsigned long long* proof_size, char* public_buffer,
                               unsigned long long* public_size, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !partials || !wtns_buffer || !proof_buffer || !proof_size || !public_buffer || !public_size)
        throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (b->is_ultra()) throw std::invalid_argument("not a groth16 prover");
    Groth16Prover* prover = static_cast<Groth16Prover*>(b);
    CheckAndUpdateBufferSizes(prover->proofBufferMinSize(), proof_size, prover->publicBufferMinSize(), public_size, "Minimum");
    std::vector<ProofMsms> parts(count);
    memcpy(parts.data(), partials, (size_t)count * sizeof(ProofMsms));
    ProofMsms m = sum_partials(parts.data(), count);
    std::string stringProof, stringPublic;
    prover->finish(m, prover->check_witness(wtns_buffer, wtns_size), stringProof, stringPublic);
    CheckAndUpdateBufferSizes(stringProof.length(), proof_size, stringPublic.length(), public_size, "Required");
    std::strncpy(proof_buffer, stringProof.c_str(), *proof_size);
    std::strncpy(public_buffer, stringPublic.c_str(), *public_size);
    API_CATCH_PROVE
}

// UltraGroth, phase 1: witness upload + this shard's part of the round commitment MSM (128-byte XYZZ)
int ugb_ultra_round_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, void* round_partial_out,
                            char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_buffer || !round_partial_out) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    UltraGrothProver* prover = static_cast<UltraGrothProver*>(b);
    BinFile wtns(wtns_buffer, wtns_size, "wtns", 2);
    UltraGrothProver::Lookup lk;
    const uint8_t* w = prover->check_witness(wtns, lk);
    HG1XYZZ part;
    prover->round_msm(w, part);
    memcpy(round_partial_out, &part, sizeof(part));
    API_CATCH_PROVE
}

// phase 2: every shard sums ALL round partials, adds r_k * delta (rk32: the SAME 32-byte scalar on every shard), derives
// the challenge, fills the lookup signals in its witness copy and runs the final-round MSMs over its point range.
int ugb_ultra_final_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, const void* round_partials,
                            uint32_t count, const void* rk32, void* partial_out, char* error_msg, unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !wtns_buffer || !round_partials || !rk32 || !partial_out || count == 0) throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    UltraGrothProver* prover = static_cast<UltraGrothProver*>(b);
    BinFile wtns(wtns_buffer, wtns_size, "wtns", 2);
    UltraGrothProver::Lookup lk;
    prover->check_witness(wtns, lk);
    std::vector<HG1XYZZ> rp(count);
    memcpy(rp.data(), round_partials, (size_t)count * sizeof(HG1XYZZ));
    HG1XYZZ commit = rp[0];
    for (uint32_t i = 1; i < count; i++) pt_add(commit, rp[i]);
    Scalar rk;
    memcpy(&rk, rk32, 32);
    HG1Affine commit_aff;
    ProofMsms m;
    prover->final_msms(commit, rk, &lk, commit_aff, m);
    m.round_c = XYZZ<HFq>::from_affine(commit_aff);  // the blinded commitment, identical on every shard
    memcpy(partial_out, &m, sizeof(m));
    API_CATCH_PROVE
}

int ugb_ultra_prove_assemble(void* prover_object, const void* partials, uint32_t count, const void* rk32, char* proof_buffer,
                             unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                             unsigned long long error_msg_maxsize) {
    API_TRY
    if (!prover_object || !partials || !rk32 || !proof_buffer || !proof_size || !public_buffer || !public_size || count == 0)
        throw std::invalid_argument("Null argument");
    ProverBase* b = static_cast<ProverBase*>(prover_object);
    std::lock_guard<std::mutex> guard(b->mu);
    if (!b->is_ultra()) throw std::invalid_argument("not an ultragroth prover");
    UltraGrothProver* prover = static_cast<UltraGrothProver*>(b);
    CheckAndUpdateBufferSizes(prover->proofBufferMinSize(), proof_size, prover->publicBufferMinSize(), public_size, "Minimum");
    std::vector<ProofMsms> parts(count);
    memcpy(parts.data(), partials, (size_t)count * sizeof(ProofMsms));
    ProofMsms m = sum_partials(parts.data(), count);
    Scalar rk;
    memcpy(&rk, rk32, 32);
    HG1Affine commit_aff = pt_to_affine(parts[0].round_c);
    std::string stringProof, stringPublic;
    prover->finish(m, rk, commit_aff, stringProof, stringPublic);
    CheckAndUpdateBufferSizes(stringProof.length(), proof_size, stringPublic.length(), public_size, "Required");
    std::strncpy(proof_buffer, stringProof.c_str(), *proof_size);
    std::strncpy(public_buffer, stringPublic.c_str(), *public_size);
    API_CATCH_PROVE
}

}  // extern "C"
