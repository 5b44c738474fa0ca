// OURS (test infrastructure): extern "C" entry points over the REFERENCE's own host code, compiled where it lies under
// /root/reference and linked into oracle/_ref/libref_host.so by oracle/ref_build.py:
//   src/binfile_utils.cpp, src/fileloader.cpp ........ the iden3 binfile container reader
//   src/zkey_utils.cpp ............................... ZKeyUtils::loadHeader / ultra_groth_loadHeader (:42-76, :123-163)
//   src/wtns_utils.cpp ............................... WtnsUtils::loadHeader (:13-26)
//   build/fr.cpp, fr_generic.cpp, fr_raw_generic.cpp . RawFr (set(int), inv via mpz_invert, mul, fromMpz ...) and Fr_raw*
//   build/fq.cpp, fq_generic.cpp, fq_raw_generic.cpp . RawFq::toMpz
//   src/keccak256.cpp ................................ FIPS202_KECCAK_256
// Used to PIN the oracle (and through it the CUDA path) for UltraGroth, for which the reference holds no fixture: the synthetic
// protocol-1337 zkey / .uwtns files must parse under the reference's own loaders, and compute_lookup / derive_challenge must give
// the reference's digits.  Those two functions are `static` / templated inside src/ultra_groth.cpp, which cannot be compiled
// (it includes the absent ffiasm engine); their bodies are re-expressed here CALL FOR CALL on the reference's compiled RawFr /
// RawFq / keccak -- the same overloads, so the int conversions of RawFr::set(int) happen in the reference's code, not ours.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "binfile_utils.hpp"
#include "fq.hpp"
#include "fr.hpp"
#include "keccak256.h"
#include "wtns_utils.hpp"
#include "zkey_utils.hpp"

namespace {
std::string mpz_hex(mpz_t v) {
    char* s = mpz_get_str(0, 16, v);
    std::string r(s);
    free(s);
    return r;
}
int put(char* out, uint64_t cap, const std::string& s) {
    if (s.size() + 1 > cap) return 2;
    memcpy(out, s.c_str(), s.size() + 1);
    return 0;
}
}  // namespace

extern "C" {

// Parses a zkey image with the reference's BinFile + ZKeyUtils loaders and dumps what they saw as JSON text: every header field,
// the offsets of the header points inside the file, and the sizes of sections 3..13 (missing sections: -1).
int ref_zkey_header_dump(const void* data, uint64_t size, int ultra, char* out, uint64_t cap) {
    try {
        BinFileUtils::BinFile f(data, size, "zkey", 1);
        const uint8_t* base = (const uint8_t*)data;
        char buf[4096];
        std::string js;
        auto off = [&](void* p) { return (long long)((const uint8_t*)p - base); };
        if (!ultra) {
            auto h = ZKeyUtils::loadHeader(&f);
            snprintf(buf, sizeof(buf),
                     "{\"n8q\":%u,\"n8r\":%u,\"qPrime\":\"%s\",\"rPrime\":\"%s\",\"nVars\":%u,\"nPublic\":%u,\"domainSize\":%u,\"nCoefs\":%llu,"
                     "\"alpha1\":%lld,\"beta1\":%lld,\"beta2\":%lld,\"gamma2\":%lld,\"delta1\":%lld,\"delta2\":%lld",
                     h->n8q, h->n8r, mpz_hex(h->qPrime).c_str(), mpz_hex(h->rPrime).c_str(), h->nVars, h->nPublic, h->domainSize,
                     (unsigned long long)h->nCoefs, off(h->vk_alpha1), off(h->vk_beta1), off(h->vk_beta2), off(h->vk_gamma2), off(h->vk_delta1),
                     off(h->vk_delta2));
            js = buf;
        } else {
            auto h = ZKeyUtils::ultra_groth_loadHeader(&f);
            snprintf(buf, sizeof(buf),
                     "{\"n8q\":%u,\"n8r\":%u,\"qPrime\":\"%s\",\"rPrime\":\"%s\",\"nVars\":%u,\"nPublic\":%u,\"domainSize\":%u,\"nCoefs\":%llu,"
                     "\"num_indexes_c1\":%u,\"num_indexes_c2\":%u,\"rand_indx\":%u,"
                     "\"alpha1\":%lld,\"beta1\":%lld,\"beta2\":%lld,\"gamma2\":%lld,\"round_delta1\":%lld,\"round_delta2\":%lld,\"final_delta1\":%lld,"
                     "\"final_delta2\":%lld",
                     h->n8q, h->n8r, mpz_hex(h->qPrime).c_str(), mpz_hex(h->rPrime).c_str(), h->nVars, h->nPublic, h->domainSize,
                     (unsigned long long)h->nCoefs, h->num_indexes_c1, h->num_indexes_c2, h->rand_indx, off(h->alpha1), off(h->beta1), off(h->beta2),
                     off(h->gamma2), off(h->round_delta1), off(h->round_delta2), off(h->final_delta1), off(h->final_delta2));
            js = buf;
        }
        js += ",\"sections\":{";
        for (int id = 3; id <= 13; id++) {
            long long sz = -1, o = -1;
            try { sz = (long long)f.getSectionSize(id); o = off(f.getSectionData(id)); } catch (...) {}
            snprintf(buf, sizeof(buf), "%s\"%d\":[%lld,%lld]", id == 3 ? "" : ",", id, o, sz);
            js += buf;
        }
        js += "}}";
        return put(out, cap, js);
    } catch (std::exception& e) {
        put(out, cap, std::string("error: ") + e.what());
        return 1;
    }
}

int ref_wtns_header_dump(const void* data, uint64_t size, char* out, uint64_t cap) {
    try {
        BinFileUtils::BinFile f(data, size, "wtns", 2);
        auto h = WtnsUtils::loadHeader(&f);
        const uint8_t* base = (const uint8_t*)data;
        char buf[1024];
        snprintf(buf, sizeof(buf), "{\"n8\":%u,\"prime\":\"%s\",\"nVars\":%u,\"sections\":{", h->n8, mpz_hex(h->prime).c_str(), h->nVars);
        std::string js = buf;
        for (int id = 2; id <= 6; id++) {
            long long sz = -1, o = -1;
            try { sz = (long long)f.getSectionSize(id); o = (long long)((const uint8_t*)f.getSectionData(id) - base); } catch (...) {}
            snprintf(buf, sizeof(buf), "%s\"%d\":[%lld,%lld]", id == 2 ? "" : ",", id, o, sz);
            js += buf;
        }
        js += "}}";
        return put(out, cap, js);
    } catch (std::exception& e) {
        put(out, cap, std::string("error: ") + e.what());
        return 1;
    }
}

// derive_challenge (src/ultra_groth.cpp:32-58) on the reference's RawFq::toMpz, FIPS202_KECCAK_256 and RawFr::fromMpz.
// commitment_xy: the affine round commitment as stored (Montgomery x | y, 64 bytes); rand_out: the challenge, Montgomery form.
void ref_derive_challenge(const uint64_t* commitment_xy, uint64_t* rand_out) {
    RawFq::Element x, y;
    memcpy(x.v, commitment_xy, 32);
    memcpy(y.v, commitment_xy + 4, 32);
    uint8_t buffer[64], challenge[32];
    memset(buffer, 0, sizeof(buffer));
    mpz_t c;
    mpz_init(c);
    RawFq::field.toMpz(c, x);
    mpz_export(buffer + 0, NULL, 1, 8, 1, 0, c);
    RawFq::field.toMpz(c, y);
    mpz_export(buffer + 32, NULL, 1, 8, 1, 0, c);
    FIPS202_KECCAK_256(buffer, 64, challenge);
    mpz_t v;
    mpz_init(v);
    mpz_import(v, 32, 0, 1, -1, 0, challenge);
    RawFr::Element rand;
    RawFr::field.fromMpz(rand, v);
    memcpy(rand_out, rand.v, 32);
    mpz_clear(c);
    mpz_clear(v);
}

// compute_lookup (src/ultra_groth.cpp:62-106) on the reference's RawFr: signals is the witness (raw 32-byte values), in/out.
void ref_compute_lookup(uint64_t* signals, const uint32_t* chunks, uint32_t chunks_total, const uint32_t* frequencies, uint32_t lookup_size,
                        const uint32_t* wtns_indxs, const uint32_t* push_indxs, uint32_t wtns_indxs_len, const uint64_t* rand_mont) {
    RawFr::Element rand;
    memcpy(rand.v, rand_mont, 32);
    const uint64_t push_len = 1ull + chunks_total + 2ull * lookup_size;
    uint64_t* push = new uint64_t[push_len * 4];
    auto digits = [](const RawFr::Element& a, uint64_t* dst) {
        RawFr::Element t;
        Fr_rawFromMontgomery(t.v, a.v);
        memcpy(dst, t.v, 32);
    };
    digits(rand, push);
    uint64_t* inv2 = push + 4 * (1ull + chunks_total);
    uint64_t* prod = inv2 + 4ull * lookup_size;
    for (int i = 0; (uint32_t)i < lookup_size; i++) {
        RawFr::Element sum = RawFr::field.add(i, rand);  // add(int, Element): RawFr::set(int)
        RawFr::Element inv;
        RawFr::field.inv(inv, sum);                      // mpz_invert + MMul by R^3
        digits(inv, inv2 + 4ull * i);
        RawFr::Element p = RawFr::field.mul(frequencies[i], inv);  // uint32_t -> int -> RawFr::set(int)
        digits(p, prod + 4ull * i);
    }
    for (uint32_t j = 0; j < chunks_total; j++) memcpy(push + 4ull * (1 + j), inv2 + 4ull * chunks[j], 32);
    for (uint32_t k = 0; k < wtns_indxs_len; k++) memcpy(signals + 4ull * wtns_indxs[k], push + 4ull * push_indxs[k], 32);
    delete[] push;
}

// RawFr::set(int) and RawFr::inv for single values (Montgomery out)
void ref_fr_set_int(int value, uint64_t* out) {
    RawFr::Element r = RawFr::field.set(value);
    memcpy(out, r.v, 32);
}
void ref_fr_inv(const uint64_t* a_mont, uint64_t* out) {
    RawFr::Element a, r;
    memcpy(a.v, a_mont, 32);
    RawFr::field.inv(r, a);
    memcpy(out, r.v, 32);
}

}  // extern "C"
