// TEST INFRASTRUCTURE ONLY -- host build of the device limb algorithms (ff.cuh / curve.cuh) with a
// software carry flag, so that `pytest -m "not gpu"` can check the carry chains and the curve
// formulas on the GPU-less build box.  Never loaded by the ultragroth_b200 package; the product
// library runs these algorithms on the device only.
#include "../../ultragroth_b200/csrc/curve.cuh"
#include "../../ultragroth_b200/csrc/msm.cuh"
#include <cstring>
using namespace ug;

// op: 0 madd(acc, affine q, neg), 1 add(acc, xyzz q), 2 dbl(acc), 3 to_affine(acc) ; group 1/2
template <class F>
static void point_op(int op, const void* acc_in, const void* q_in, int neg, void* out) {
    XYZZ<F> acc; memcpy(&acc, acc_in, sizeof(acc));
    if (op == 0) { Affine<F> q; memcpy(&q, q_in, sizeof(q)); pt_madd(acc, q, neg != 0); memcpy(out, &acc, sizeof(acc)); }
    else if (op == 1) { XYZZ<F> q; memcpy(&q, q_in, sizeof(q)); pt_add(acc, q); memcpy(out, &acc, sizeof(acc)); }
    else if (op == 2) { acc = pt_dbl(acc); memcpy(out, &acc, sizeof(acc)); }
    else { Affine<F> a = pt_to_affine(acc); memcpy(out, &a, sizeof(a)); }
}

extern "C" {

// op: 0 mul, 1 add, 2 sub, 3 neg, 4 to_mont, 5 from_mont, 6 sqr, 7 inv ; field: 0 Fr, 1 Fq
int emul_field_op(int field, int op, const uint32_t* a, const uint32_t* b, uint32_t* out, long n) {
    for (long i = 0; i < n; i++) {
        if (field == 0) {
            Fr x, y, r; memcpy(x.v, a + 8 * i, 32); memcpy(y.v, b + 8 * i, 32);
            switch (op) {
                case 0: r = fp_mul(x, y); break; case 1: r = fp_add(x, y); break;
                case 2: r = fp_sub(x, y); break; case 3: r = fp_neg(x); break;
                case 4: r = fp_to_mont(x); break; case 5: r = fp_from_mont(x); break;
                case 6: r = fp_sqr(x); break; case 7: r = fp_inv(x); break; default: return 1;
            }
            memcpy(out + 8 * i, r.v, 32);
        } else {
            Fq x, y, r; memcpy(x.v, a + 8 * i, 32); memcpy(y.v, b + 8 * i, 32);
            switch (op) {
                case 0: r = fp_mul(x, y); break; case 1: r = fp_add(x, y); break;
                case 2: r = fp_sub(x, y); break; case 3: r = fp_neg(x); break;
                case 4: r = fp_to_mont(x); break; case 5: r = fp_from_mont(x); break;
                case 6: r = fp_sqr(x); break; case 7: r = fp_inv(x); break; default: return 1;
            }
            memcpy(out + 8 * i, r.v, 32);
        }
    }
    return 0;
}

int emul_point_op(int group, int op, const void* acc, const void* q, int neg, void* out) {
    if (group == 1) point_op<Fq>(op, acc, q, neg, out); else point_op<Fq2>(op, acc, q, neg, out);
    return 0;
}

// next_bucket (msm.cuh): the accumulation kernels' step to the next non-empty bucket.  starts: NB + 1 offsets.
uint32_t emul_next_bucket(const uint32_t* starts, uint32_t NB, uint32_t key, uint32_t pos, uint32_t* end_out) {
    uint32_t end = 0;
    next_bucket(starts, NB, key, pos, end);
    *end_out = end;
    return key;
}
// signed-digit recoding of one scalar (msm.cuh: scalar_reduce + scalar_digits): out[w] = signed digit of window w
void emul_scalar_digits(const uint32_t* scalar8, int c, int W, int32_t* out) {
    Scalar256 s;
    memcpy(s.v, scalar8, 32);
    scalar_reduce(s);
    for (int w = 0; w < W; w++) out[w] = 0;
    scalar_digits(s, c, W, [&](int w, uint32_t b, bool neg) { out[w] = neg ? -(int32_t)(b + 1) : (int32_t)(b + 1); });
}
}
