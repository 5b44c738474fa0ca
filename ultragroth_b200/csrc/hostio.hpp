// Host-side file formats and helpers of the prover boundary (no CUDA here).
//
//   BinFile ........ iden3 binfile container, same checks and messages as src/binfile_utils.cpp:32-80
//   ZkeyHeader ..... Groth16 header (src/zkey_utils.cpp:42-76) and UltraGroth header (:123-163)
//   WtnsHeader ..... src/wtns_utils.cpp:13-26
//   keccak256 ...... Ethereum Keccak-256 (pad 0x01), what src/keccak256.cpp's FIPS202_KECCAK_256 computes
// Buffers are borrowed (pointers into the caller's zkey / wtns memory), exactly like the reference; the
// device prover copies what it needs to HBM at create time.
#pragma once
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace ug {

class BinFile {
public:
    struct Section { const uint8_t* start; uint64_t size; };

    BinFile(const void* data, uint64_t size, const std::string& type, uint32_t maxVersion) : addr_((const uint8_t*)data), size_(size) {
        const uint64_t headerSize = 12, minSectionSize = 12;
        if (size_ < headerSize) throw std::range_error("File is too short.");
        std::string t((const char*)addr_, 4);
        if (t != type) throw std::invalid_argument("Invalid file type. It should be " + type + " and it is " + t);
        uint64_t pos = 4;
        version_ = rd32(pos);
        pos += 4;
        if (version_ > maxVersion)
            throw std::invalid_argument("Invalid version. It should be <=" + std::to_string(maxVersion) + " and it is " + std::to_string(version_));
        uint32_t nSections = rd32(pos);
        pos += 4;
        if (size_ < headerSize + (uint64_t)nSections * minSectionSize)
            throw std::range_error("File is too short to contain " + std::to_string(nSections) + " sections.");
        for (uint32_t i = 0; i < nSections; i++) {
            if (pos + 12 > size_)
                throw std::range_error("Section #" + std::to_string(i) + " is invalid.. It ends at pos " + std::to_string(pos + 12) +
                                       " but should end before " + std::to_string(size_) + ".");
            uint32_t sType = rd32(pos);
            uint64_t sSize = rd64(pos + 4);
            pos += 12;
            sections_[sType].push_back(Section{addr_ + pos, sSize});
            if (sSize > size_ - pos)
                throw std::range_error("Section #" + std::to_string(i) + " is invalid.. It ends at pos " + std::to_string(pos + sSize) +
                                       " but should end before " + std::to_string(size_) + ".");
            pos += sSize;
        }
    }

    const Section& section(uint32_t id, uint32_t idx = 0) const {
        auto it = sections_.find(id);
        if (it == sections_.end()) throw std::range_error("Section does not exist: " + std::to_string(id));
        if (idx >= it->second.size())
            throw std::range_error("Section pos too big. There are " + std::to_string(it->second.size()) +
                                   " and it's trying to access section: " + std::to_string(idx));
        return it->second[idx];
    }
    const uint8_t* data(uint32_t id, uint32_t idx = 0) const { return section(id, idx).start; }
    uint64_t size(uint32_t id, uint32_t idx = 0) const { return section(id, idx).size; }
    bool has(uint32_t id) const { return sections_.count(id) != 0; }
    uint32_t version() const { return version_; }

private:
    uint32_t rd32(uint64_t pos) const { uint32_t v; memcpy(&v, addr_ + pos, 4); return v; }
    uint64_t rd64(uint64_t pos) const { uint64_t v; memcpy(&v, addr_ + pos, 8); return v; }
    const uint8_t* addr_;
    uint64_t size_;
    uint32_t version_ = 0;
    std::map<uint32_t, std::vector<Section>> sections_;
};

// bounds-checked cursor over one section
class SectionReader {
public:
    explicit SectionReader(const BinFile::Section& s) : p_(s.start), end_(s.start + s.size) {}
    uint32_t u32() { uint32_t v; memcpy(&v, take(4), 4); return v; }
    const uint8_t* take(uint64_t n) {
        if ((uint64_t)(end_ - p_) < n) throw std::range_error("Section is too short.");
        const uint8_t* r = p_;
        p_ += n;
        return r;
    }
private:
    const uint8_t* p_;
    const uint8_t* end_;
};

static const uint8_t BN254_R_LE[32] = {0x01, 0x00, 0x00, 0xf0, 0x93, 0xf5, 0xe1, 0x43, 0x91, 0x70, 0xb9, 0x79, 0x48, 0xe8, 0x33, 0x28,
                                       0x5d, 0x58, 0x81, 0x81, 0xb6, 0x45, 0x50, 0xb8, 0x29, 0xa0, 0x31, 0xe1, 0x72, 0x4e, 0x64, 0x30};

// PrimeIsValid (src/prover.cpp:73-86): little-endian prime of n8 bytes equals BN254 r
inline bool prime_is_bn254_r(const uint8_t* prime, uint32_t n8) {
    for (uint32_t i = 0; i < n8; i++) {
        uint8_t want = i < 32 ? BN254_R_LE[i] : 0;
        if (prime[i] != want) return false;
    }
    for (uint32_t i = n8; i < 32; i++) if (BN254_R_LE[i]) return false;
    return true;
}

struct ZkeyHeader {
    bool ultra = false;
    uint32_t n8q = 0, n8r = 0;
    const uint8_t* qPrime = nullptr;
    const uint8_t* rPrime = nullptr;
    uint32_t nVars = 0, nPublic = 0, domainSize = 0;
    uint64_t nCoefs = 0;
    // Groth16: alpha1, beta1, beta2, gamma2, delta1, delta2 (delta1/2 alias final_delta1/2)
    // UltraGroth adds round_delta1/2 and the index bookkeeping
    uint32_t num_indexes_c1 = 0, num_indexes_c2 = 0, rand_indx = 0;
    const uint8_t *alpha1 = nullptr, *beta1 = nullptr, *beta2 = nullptr, *gamma2 = nullptr;
    const uint8_t *round_delta1 = nullptr, *round_delta2 = nullptr, *final_delta1 = nullptr, *final_delta2 = nullptr;
};

inline ZkeyHeader load_zkey_header(const BinFile& f, bool ultra) {
    ZkeyHeader h;
    h.ultra = ultra;
    {
        SectionReader r(f.section(1));
        uint32_t protocol = r.u32();
        if (!ultra && protocol != 1) throw std::invalid_argument("zkey file is not groth16");
        if (ultra && protocol != 1337) throw std::invalid_argument("zkey file is not ultragroth");
    }
    SectionReader r(f.section(2));
    h.n8q = r.u32();
    h.qPrime = r.take(h.n8q);
    h.n8r = r.u32();
    h.rPrime = r.take(h.n8r);
    h.nVars = r.u32();
    h.nPublic = r.u32();
    h.domainSize = r.u32();
    if (ultra) {
        h.num_indexes_c1 = r.u32();
        h.num_indexes_c2 = r.u32();
        h.rand_indx = r.u32();
    }
    h.alpha1 = r.take((uint64_t)h.n8q * 2);
    h.beta1 = r.take((uint64_t)h.n8q * 2);
    h.beta2 = r.take((uint64_t)h.n8q * 4);
    h.gamma2 = r.take((uint64_t)h.n8q * 4);
    if (ultra) {
        h.round_delta1 = r.take((uint64_t)h.n8q * 2);
        h.round_delta2 = r.take((uint64_t)h.n8q * 4);
    }
    h.final_delta1 = r.take((uint64_t)h.n8q * 2);
    h.final_delta2 = r.take((uint64_t)h.n8q * 4);
    h.nCoefs = f.size(4) / (12 + (uint64_t)h.n8r);
    return h;
}

struct WtnsHeader {
    uint32_t n8 = 0;
    const uint8_t* prime = nullptr;
    uint32_t nVars = 0;
};

inline WtnsHeader load_wtns_header(const BinFile& f) {
    WtnsHeader h;
    SectionReader r(f.section(1));
    h.n8 = r.u32();
    h.prime = r.take(h.n8);
    h.nVars = r.u32();
    return h;
}

// ---- Keccak-256 (rate 1088, capacity 512, suffix 0x01) -------------------------------------------
inline void keccak_f1600(uint64_t st[25]) {
    static const uint64_t RC[24] = {
        0x0000000000000001ull, 0x0000000000008082ull, 0x800000000000808aull, 0x8000000080008000ull, 0x000000000000808bull, 0x0000000080000001ull,
        0x8000000080008081ull, 0x8000000000008009ull, 0x000000000000008aull, 0x0000000000000088ull, 0x0000000080008009ull, 0x000000008000000aull,
        0x000000008000808bull, 0x800000000000008bull, 0x8000000000008089ull, 0x8000000000008003ull, 0x8000000000008002ull, 0x8000000000000080ull,
        0x000000000000800aull, 0x800000008000000aull, 0x8000000080008081ull, 0x8000000000008080ull, 0x0000000080000001ull, 0x8000000080008008ull};
    static const int ROT[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
    static const int PIL[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
    for (int round = 0; round < 24; round++) {
        uint64_t bc[5];
        for (int i = 0; i < 5; i++) bc[i] = st[i] ^ st[i + 5] ^ st[i + 10] ^ st[i + 15] ^ st[i + 20];
        for (int i = 0; i < 5; i++) {
            uint64_t t = bc[(i + 4) % 5] ^ ((bc[(i + 1) % 5] << 1) | (bc[(i + 1) % 5] >> 63));
            for (int j = 0; j < 25; j += 5) st[j + i] ^= t;
        }
        uint64_t t = st[1];
        for (int i = 0; i < 24; i++) {
            int j = PIL[i];
            uint64_t b = st[j];
            st[j] = (t << ROT[i]) | (t >> (64 - ROT[i]));
            t = b;
        }
        for (int j = 0; j < 25; j += 5) {
            for (int i = 0; i < 5; i++) bc[i] = st[j + i];
            for (int i = 0; i < 5; i++) st[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
        }
        st[0] ^= RC[round];
    }
}

inline void keccak256(const uint8_t* in, size_t len, uint8_t out[32]) {
    uint64_t st[25];
    memset(st, 0, sizeof(st));
    const size_t rate = 136;
    uint8_t block[136];
    while (len >= rate) {
        for (size_t i = 0; i < rate / 8; i++) { uint64_t w; memcpy(&w, in + 8 * i, 8); st[i] ^= w; }
        keccak_f1600(st);
        in += rate;
        len -= rate;
    }
    memset(block, 0, rate);
    memcpy(block, in, len);
    block[len] ^= 0x01;
    block[rate - 1] ^= 0x80;
    for (size_t i = 0; i < rate / 8; i++) { uint64_t w; memcpy(&w, block + 8 * i, 8); st[i] ^= w; }
    keccak_f1600(st);
    memcpy(out, st, 32);
}

}  // namespace ug
