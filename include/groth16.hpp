// Groth16::makeProver<Engine> / Groth16::Prover<Engine>::prove / Groth16::Proof<Engine> -- the reference's C++ class API for the
// prover (src/groth16.hpp:15-136, implementation src/groth16.cpp:17-250), kept signature for signature and implemented on
// libultragroth_b200.so (sm_100a kernels) through the C ABI of include/ultragroth_b200.h.  This is what src/prover.cpp:160-189 and
// src/fullprover.cpp bind; a C++ embedder swaps `#include "groth16.hpp"` + AltBn128::Engine for this header + UgbAltBn128::Engine.
//
// Differences a maintainer should know: makeProver COPIES the sections to HBM (the reference keeps raw pointers into the mapped
// zkey), so the zkey buffer may be released after it returns; prove() throws std::runtime_error with the library's message on
// failure; Proof::toJsonStr() prints the canonical nlohmann dump (sorted keys, no spaces) that src/prover.cpp:504 produces through
// toJson().dump(), because that -- not the hand-written toJsonStr of src/groth16.cpp:205-214 -- is what the C API hands out.
#ifndef UGB_GROTH16_HPP
#define UGB_GROTH16_HPP
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>

#include "ugb_engine.hpp"
#include "ultragroth_b200.h"

namespace Groth16 {

template <typename Engine>
class Proof {
public:
    typename Engine::G1PointAffine A;
    typename Engine::G2PointAffine B;
    typename Engine::G1PointAffine C;
    explicit Proof(Engine&) {}
    std::string toJsonStr() {
        char buf[2048];
        if (ugb_proof_points_to_json(0, &A, &B, &C, nullptr, buf, sizeof(buf)) != UGB_OK) throw std::runtime_error(ugb_last_error());
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
    Prover(Engine&, uint32_t nVars, uint32_t nPublic, uint32_t domainSize, uint64_t nCoefs, typename Engine::G1PointAffine& vk_alpha1,
           typename Engine::G1PointAffine& vk_beta1, typename Engine::G2PointAffine& vk_beta2, typename Engine::G1PointAffine& vk_delta1,
           typename Engine::G2PointAffine& vk_delta2, Coef<Engine>* coefs, typename Engine::G1PointAffine* pointsA,
           typename Engine::G1PointAffine* pointsB1, typename Engine::G2PointAffine* pointsB2, typename Engine::G1PointAffine* pointsC,
           typename Engine::G1PointAffine* pointsH) {
        ugb_zkey_sections z{};
        z.nVars = nVars; z.nPublic = nPublic; z.domainSize = domainSize; z.nCoefs = nCoefs;
        z.vk_alpha1 = &vk_alpha1; z.vk_beta1 = &vk_beta1; z.vk_beta2 = &vk_beta2; z.vk_delta1 = &vk_delta1; z.vk_delta2 = &vk_delta2;
        z.coefs = coefs; z.pointsA = pointsA; z.pointsB1 = pointsB1; z.pointsB2 = pointsB2; z.pointsC = pointsC; z.pointsH = pointsH;
        char err[512] = {0};
        if (ugb_prover_create_from_sections(&handle_, 0, &z, err, sizeof(err) - 1) != 0) throw std::runtime_error(err);
    }
    ~Prover() { if (handle_) ugb_prover_destroy(handle_); }
    Prover(const Prover&) = delete;
    Prover& operator=(const Prover&) = delete;

    // src/groth16.cpp:48-203: wtns = nVars raw (non-Montgomery) signals
    std::unique_ptr<Proof<Engine>> prove(typename Engine::FrElement* wtns) {
        std::unique_ptr<Proof<Engine>> p(new Proof<Engine>(Engine::instance()));
        char err[512] = {0};
        if (ugb_groth16_prove_raw(handle_, wtns, &p->A, &p->B, &p->C, err, sizeof(err) - 1) != 0) throw std::runtime_error(err);
        return p;
    }
    void* handle() const { return handle_; }  // for the ugb_prover_* taps (blinding, last_msms ...)
};

// src/groth16.cpp:17-45
template <typename Engine>
std::unique_ptr<Prover<Engine>> makeProver(uint32_t nVars, uint32_t nPublic, uint32_t domainSize, uint64_t nCoefs, void* vk_alpha1, void* vk_beta1,
                                           void* vk_beta2, void* vk_delta1, void* vk_delta2, void* coefs, void* pointsA, void* pointsB1,
                                           void* pointsB2, void* pointsC, void* pointsH) {
    Prover<Engine>* p = new Prover<Engine>(
        Engine::instance(), nVars, nPublic, domainSize, nCoefs, *(typename Engine::G1PointAffine*)vk_alpha1, *(typename Engine::G1PointAffine*)vk_beta1,
        *(typename Engine::G2PointAffine*)vk_beta2, *(typename Engine::G1PointAffine*)vk_delta1, *(typename Engine::G2PointAffine*)vk_delta2,
        (Coef<Engine>*)((uint64_t)coefs + 4), (typename Engine::G1PointAffine*)pointsA, (typename Engine::G1PointAffine*)pointsB1,
        (typename Engine::G2PointAffine*)pointsB2, (typename Engine::G1PointAffine*)pointsC, (typename Engine::G1PointAffine*)pointsH);
    return std::unique_ptr<Prover<Engine>>(p);
}

}  // namespace Groth16
#endif
