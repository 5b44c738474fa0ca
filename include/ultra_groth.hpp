// UltraGroth::makeProver<Engine> / UltraGroth::Prover<Engine>::prove / UltraGroth::Proof<Engine> / UltraGroth::LookupInfo -- the
// reference's C++ class API for the two-round prover (src/ultra_groth.hpp:22-248, implementation src/ultra_groth.cpp:108-159,401-513),
// kept signature for signature and implemented on libultragroth_b200.so through the C ABI of include/ultragroth_b200.h.  This is what
// src/prover.cpp:237-309 binds.  See include/groth16.hpp for the differences (sections are copied to HBM; errors throw).
// execute_round / execute_final_round are not offered separately: the rounds share device state and run inside prove().
#ifndef UGB_ULTRA_GROTH_HPP
#define UGB_ULTRA_GROTH_HPP
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>

#include "ugb_engine.hpp"
#include "ultragroth_b200.h"

namespace UltraGroth {

// src/ultra_groth.hpp:22-61
struct LookupInfo {
    uint32_t* chunks;
    uint64_t chunks_len;
    uint32_t* frequencies;
    uint64_t frequencies_len;
    uint32_t* wtns_indxs;
    uint64_t wtns_indxs_len;
    uint32_t* push_indxs;
    uint64_t push_indxs_len;
    LookupInfo(uint32_t* chunks, uint64_t chunks_len, uint32_t* frequencies, uint64_t frequencies_len, uint32_t* wtns_indxs,
               uint64_t wtns_indxs_len, uint32_t* push_indxs, uint64_t push_indxs_len)
        : chunks(chunks), chunks_len(chunks_len), frequencies(frequencies), frequencies_len(frequencies_len), wtns_indxs(wtns_indxs),
          wtns_indxs_len(wtns_indxs_len), push_indxs(push_indxs), push_indxs_len(push_indxs_len) {}
};

template <typename Engine>
class Proof {
public:
    typename Engine::G1PointAffine A;
    typename Engine::G2PointAffine B;
    typename Engine::G1PointAffine final_commitment;
    typename Engine::G1PointAffine round_commitment;
    uint8_t* error = nullptr;
    uint32_t error_size = 0;
    explicit Proof(Engine&) {}
    std::string toJsonStr() {  // the toJson().dump() text of src/ultra_groth.cpp:476-513 (pi_a, pi_b, pi_f, pi_r, protocol)
        char buf[2560];
        if (ugb_proof_points_to_json(1, &A, &B, &final_commitment, &round_commitment, buf, sizeof(buf)) != UGB_OK) throw std::runtime_error(ugb_last_error());
        return std::string(buf);
    }
};

#pragma pack(push, 1)
template <typename Engine>
struct Coef {
    uint32_t m;
    uint32_t c;
    uint32_t s;
    typename Engine::FrElement coef;
};
#pragma pack(pop)

template <typename Engine>
class Prover {
    void* handle_ = nullptr;

public:
    Prover(Engine&, uint32_t nVars, uint32_t nPublic, uint32_t domainSize, uint64_t nCoefs, uint32_t* round_indexes, uint32_t round_indexes_count,
           uint32_t* final_round_indexes, uint32_t final_round_indexes_count, uint32_t challenge_index, typename Engine::G1PointAffine& alpha1,
           typename Engine::G1PointAffine& beta1, typename Engine::G2PointAffine& beta2, typename Engine::G1PointAffine& final_delta1,
           typename Engine::G2PointAffine& final_delta2, typename Engine::G1PointAffine& round_delta1, Coef<Engine>* coefs,
           typename Engine::G1PointAffine* pointsA, typename Engine::G1PointAffine* pointsB1, typename Engine::G2PointAffine* pointsB2,
           typename Engine::G1PointAffine* final_pointsC, typename Engine::G1PointAffine* round_pointsC, typename Engine::G1PointAffine* pointsH) {
        ugb_zkey_sections z{};
        z.nVars = nVars; z.nPublic = nPublic; z.domainSize = domainSize; z.nCoefs = nCoefs;
        z.vk_alpha1 = &alpha1; z.vk_beta1 = &beta1; z.vk_beta2 = &beta2; z.vk_delta1 = &final_delta1; z.vk_delta2 = &final_delta2;
        z.round_delta1 = &round_delta1;
        z.coefs = coefs; z.pointsA = pointsA; z.pointsB1 = pointsB1; z.pointsB2 = pointsB2; z.pointsC = final_pointsC; z.pointsH = pointsH;
        z.round_pointsC = round_pointsC;
        z.round_indexes = round_indexes; z.round_indexes_count = round_indexes_count;
        z.final_round_indexes = final_round_indexes; z.final_round_indexes_count = final_round_indexes_count;
        z.challenge_index = challenge_index;
        char err[512] = {0};
        if (ugb_prover_create_from_sections(&handle_, 1, &z, err, sizeof(err) - 1) != 0) throw std::runtime_error(err);
    }
    ~Prover() { if (handle_) ugb_prover_destroy(handle_); }
    Prover(const Prover&) = delete;
    Prover& operator=(const Prover&) = delete;

    // src/ultra_groth.cpp:401-462; wtns is in/out (compute_lookup writes the lookup signals into it)
    std::unique_ptr<Proof<Engine>> prove(typename Engine::FrElement* wtns, LookupInfo& lookupInfo) {
        std::unique_ptr<Proof<Engine>> p(new Proof<Engine>(Engine::instance()));
        char err[512] = {0};
        if (lookupInfo.chunks_len > 0xffffffffull || lookupInfo.frequencies_len > 0xffffffffull || lookupInfo.wtns_indxs_len > 0xffffffffull)
            throw std::runtime_error("lookup arrays too long");
        if (ugb_ultra_groth_prove_raw(handle_, wtns, lookupInfo.chunks, (uint32_t)lookupInfo.chunks_len, lookupInfo.frequencies,
                                      (uint32_t)lookupInfo.frequencies_len, lookupInfo.wtns_indxs, lookupInfo.push_indxs, (uint32_t)lookupInfo.wtns_indxs_len,
                                      &p->A, &p->B, &p->final_commitment, &p->round_commitment, err, sizeof(err) - 1) != 0)
            throw std::runtime_error(err);
        return p;
    }
    void* handle() const { return handle_; }
};

// src/ultra_groth.cpp:108-159, the overload src/prover.cpp:237-260 calls (22 arguments)
template <typename Engine>
std::unique_ptr<Prover<Engine>> makeProver(uint32_t nVars, uint32_t nPublic, uint32_t domainSize, uint64_t nCoefs, void* round_indexes,
                                           uint32_t round_indexes_count, void* final_round_indexes, uint32_t final_round_indexes_count,
                                           uint32_t challenge_index, void* vk_alpha1, void* vk_beta1, void* vk_beta2, void* final_delta1,
                                           void* final_delta2, void* round_delta1, void* coefs, void* pointsA, void* pointsB1, void* pointsB2,
                                           void* final_pointsC, void* round_pointsC, void* pointsH) {
    Prover<Engine>* p = new Prover<Engine>(
        Engine::instance(), nVars, nPublic, domainSize, nCoefs, (uint32_t*)round_indexes, round_indexes_count, (uint32_t*)final_round_indexes,
        final_round_indexes_count, challenge_index, *(typename Engine::G1PointAffine*)vk_alpha1, *(typename Engine::G1PointAffine*)vk_beta1,
        *(typename Engine::G2PointAffine*)vk_beta2, *(typename Engine::G1PointAffine*)final_delta1, *(typename Engine::G2PointAffine*)final_delta2,
        *(typename Engine::G1PointAffine*)round_delta1, (Coef<Engine>*)((uint64_t)coefs + 4), (typename Engine::G1PointAffine*)pointsA,
        (typename Engine::G1PointAffine*)pointsB1, (typename Engine::G2PointAffine*)pointsB2, (typename Engine::G1PointAffine*)final_pointsC,
        (typename Engine::G1PointAffine*)round_pointsC, (typename Engine::G1PointAffine*)pointsH);
    return std::unique_ptr<Prover<Engine>>(p);
}
// the 24-argument form DECLARED at src/ultra_groth.hpp:218-247 (two extra vk_delta pointers; the reference never defines it)
template <typename Engine>
std::unique_ptr<Prover<Engine>> makeProver(uint32_t nVars, uint32_t nPublic, uint32_t domainSize, uint64_t nCoefs, void* round_indexes,
                                           uint32_t round_indexes_count, void* final_round_indexes, uint32_t final_round_indexes_count,
                                           uint32_t challenge_index, void* vk_alpha1, void* vk_beta1, void* vk_beta2, void* /*vk_delta1*/,
                                           void* /*vk_delta2*/, void* final_delta1, void* final_delta2, void* round_delta1, void* coefs, void* pointsA,
                                           void* pointsB1, void* pointsB2, void* final_pointsC, void* round_pointsC, void* pointsH) {
    return makeProver<Engine>(nVars, nPublic, domainSize, nCoefs, round_indexes, round_indexes_count, final_round_indexes, final_round_indexes_count,
                              challenge_index, vk_alpha1, vk_beta1, vk_beta2, final_delta1, final_delta2, round_delta1, coefs, pointsA, pointsB1,
                              pointsB2, final_pointsC, round_pointsC, pointsH);
}

}  // namespace UltraGroth
#endif
