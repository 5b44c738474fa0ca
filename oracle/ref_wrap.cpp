// OURS (test infrastructure): extern "C" loops over the reference's own field functions, linked into
// oracle/_ref/libref_field.so together with the reference sources compiled where they lie.
#include <cstdint>
#include <cstring>
#include "fr_element.hpp"
#include "fq_element.hpp"
#include "keccak256.h"

void Fr_rawMMul(FrRawElement r, const FrRawElement a, const FrRawElement b);
void Fr_rawAdd(FrRawElement r, const FrRawElement a, const FrRawElement b);
void Fr_rawSub(FrRawElement r, const FrRawElement a, const FrRawElement b);
void Fr_rawNeg(FrRawElement r, const FrRawElement a);
void Fr_rawToMontgomery(FrRawElement r, const FrRawElement& a);
void Fr_rawFromMontgomery(FrRawElement r, const FrRawElement& a);
void Fq_rawMMul(FqRawElement r, const FqRawElement a, const FqRawElement b);
void Fq_rawAdd(FqRawElement r, const FqRawElement a, const FqRawElement b);
void Fq_rawSub(FqRawElement r, const FqRawElement a, const FqRawElement b);
void Fq_rawNeg(FqRawElement r, const FqRawElement a);
void Fq_rawToMontgomery(FqRawElement r, const FqRawElement& a);
void Fq_rawFromMontgomery(FqRawElement r, const FqRawElement& a);

extern "C" {
// op: 0 mul, 1 add, 2 sub, 3 neg, 4 toMontgomery, 5 fromMontgomery ; field: 0 Fr, 1 Fq
int ref_field_op(int field, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, long n) {
    for (long i = 0; i < n; i++) {
        uint64_t x[4], y[4], r[4];
        memcpy(x, a + 4 * i, 32);
        memcpy(y, b + 4 * i, 32);
        if (field == 0) {
            switch (op) {
                case 0: Fr_rawMMul(r, x, y); break; case 1: Fr_rawAdd(r, x, y); break; case 2: Fr_rawSub(r, x, y); break;
                case 3: Fr_rawNeg(r, x); break; case 4: Fr_rawToMontgomery(r, x); break; case 5: Fr_rawFromMontgomery(r, x); break;
                default: return 1;
            }
        } else {
            switch (op) {
                case 0: Fq_rawMMul(r, x, y); break; case 1: Fq_rawAdd(r, x, y); break; case 2: Fq_rawSub(r, x, y); break;
                case 3: Fq_rawNeg(r, x); break; case 4: Fq_rawToMontgomery(r, x); break; case 5: Fq_rawFromMontgomery(r, x); break;
                default: return 1;
            }
        }
        memcpy(out + 4 * i, r, 32);
    }
    return 0;
}
void ref_keccak256(const uint8_t* in, uint64_t len, uint8_t* out) { FIPS202_KECCAK_256(in, len, out); }
}
