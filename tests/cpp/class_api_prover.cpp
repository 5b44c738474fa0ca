// OURS (test program): a prover CLI written against boundary surface 2, the C++ class API -- the way src/prover.cpp:144-216 (Groth16Prover)
// and :219-309 (UltraGrothProver) use it -- with
//   * the REFERENCE's own loaders compiled in place (BinFileUtils::BinFile, ZKeyUtils::loadHeader / ultra_groth_loadHeader,
//     WtnsUtils::loadHeader: src/binfile_utils.cpp, src/zkey_utils.cpp, src/wtns_utils.cpp), and
//   * OUR include/groth16.hpp / include/ultra_groth.hpp (makeProver / Prover::prove / Proof) on UgbAltBn128::Engine in place of the
//     reference's headers on AltBn128::Engine.
// Built by oracle/ref_build.py into oracle/_ref/bin/class_api_prover; run by tests/test_gpu_dropin.py.
//   class_api_prover groth16|ultragroth <circuit.zkey> <witness.wtns> <proof.json> <public.json>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <vector>

#include "binfile_utils.hpp"
#include "groth16.hpp"
#include "ultra_groth.hpp"
#include "wtns_utils.hpp"
#include "zkey_utils.hpp"

using Engine = UgbAltBn128::Engine;

// BuildPublicString (src/prover.cpp:89-117) without nlohmann: decimal strings of the raw public signals
static std::string dec(const uint64_t* limbs) {
    uint64_t v[4] = {limbs[0], limbs[1], limbs[2], limbs[3]};
    std::string s;
    while (v[0] | v[1] | v[2] | v[3]) {
        unsigned __int128 rem = 0;
        for (int i = 3; i >= 0; i--) {
            unsigned __int128 cur = (rem << 64) | v[i];
            v[i] = (uint64_t)(cur / 10);
            rem = cur % 10;
        }
        s.insert(s.begin(), (char)('0' + (int)rem));
    }
    return s.empty() ? "0" : s;
}
static std::string public_json(const Engine::FrElement* w, uint32_t nPublic, uint32_t skip) {
    std::string out;
    bool any = false;
    for (uint32_t i = 1; i <= nPublic; i++) {
        if (i == skip) continue;
        out += any ? ",\"" : "[\"";
        out += dec(w[i].v) + "\"";
        any = true;
    }
    return any ? out + "]" : "null";
}

int main(int argc, char** argv) {
    if (argc != 6) {
        std::cerr << "Usage: class_api_prover groth16|ultragroth <circuit.zkey> <witness.wtns> <proof.json> <public.json>" << std::endl;
        return 1;
    }
    try {
        const bool ultra = std::string(argv[1]) == "ultragroth";
        BinFileUtils::BinFile zkey(std::string(argv[2]), "zkey", 1);
        BinFileUtils::BinFile wtns(std::string(argv[3]), "wtns", 2);
        auto wh = WtnsUtils::loadHeader(&wtns);
        std::vector<Engine::FrElement> w(wh->nVars);
        memcpy(w.data(), wtns.getSectionData(2), (size_t)wh->nVars * 32);  // prove() may write into it (UltraGroth)
        std::string proofJson, publicJson;
        if (!ultra) {
            auto h = ZKeyUtils::loadHeader(&zkey);
            if (h->nVars != wh->nVars) throw std::runtime_error("Invalid witness length");
            auto prover = Groth16::makeProver<Engine>(h->nVars, h->nPublic, h->domainSize, h->nCoefs, h->vk_alpha1, h->vk_beta1, h->vk_beta2,
                                                      h->vk_delta1, h->vk_delta2, zkey.getSectionData(4), zkey.getSectionData(5), zkey.getSectionData(6),
                                                      zkey.getSectionData(7), zkey.getSectionData(8), zkey.getSectionData(9));
            auto proof = prover->prove(w.data());
            proofJson = proof->toJsonStr();
            publicJson = public_json(w.data(), h->nPublic, 0);
        } else {
            auto h = ZKeyUtils::ultra_groth_loadHeader(&zkey);
            if (h->nVars != wh->nVars) throw std::runtime_error("Invalid witness length");
            auto prover = UltraGroth::makeProver<Engine>(h->nVars, h->nPublic, h->domainSize, h->nCoefs, zkey.getSectionData(10), h->num_indexes_c1,
                                                         zkey.getSectionData(11), h->num_indexes_c2, h->rand_indx, h->alpha1, h->beta1, h->beta2,
                                                         h->final_delta1, h->final_delta2, h->round_delta1, zkey.getSectionData(4),
                                                         zkey.getSectionData(5), zkey.getSectionData(6), zkey.getSectionData(7), zkey.getSectionData(9),
                                                         zkey.getSectionData(8), zkey.getSectionData(12));
            // LookupInfo from the .uwtns sections 3..6 (src/prover.cpp:283-292)
            auto u32 = [&](uint32_t id, std::vector<uint32_t>& v) {
                v.resize(wtns.getSectionSize(id) / 4);
                if (!v.empty()) memcpy(v.data(), wtns.getSectionData(id), v.size() * 4);
            };
            std::vector<uint32_t> chunks, freqs, widx, pidx;
            u32(3, chunks); u32(4, freqs); u32(5, widx); u32(6, pidx);
            UltraGroth::LookupInfo info(chunks.data(), chunks.size(), freqs.data(), freqs.size(), widx.data(), widx.size(), pidx.data(), pidx.size());
            auto proof = prover->prove(w.data(), info);
            proofJson = proof->toJsonStr();
            publicJson = public_json(w.data(), h->nPublic, h->rand_indx);  // w now holds the filled lookup signals, like the reference's
        }
        std::ofstream(argv[4]) << proofJson;
        std::ofstream(argv[5]) << publicJson;
    } catch (std::exception& e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
