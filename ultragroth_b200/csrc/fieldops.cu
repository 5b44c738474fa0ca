// Element-wise Fr / Fq vector kernels (the parity kernels for build/fr.asm / fq.asm, and the
// pointwise loops of src/groth16.cpp:100-108).  HBM-bound: 96 B of traffic per element.
#include "fieldops.hpp"
#include "ff_variants.cuh"
#include "memops.cuh"
#include <algorithm>

namespace ug {

template <class F, int OP>
__global__ void __launch_bounds__(256) k_field_vec(const F* __restrict__ a, const F* __restrict__ b, F* __restrict__ out, uint64_t n) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        F x = ld_g(&a[i]);
        F r;
        if (OP == 0) r = fp_mul(x, ld_g(&b[i]));
        else if (OP == 1) r = fp_add(x, ld_g(&b[i]));
        else if (OP == 2) r = fp_sub(x, ld_g(&b[i]));
        else if (OP == 3) r = fp_neg(x);
        else if (OP == 4) r = fp_to_mont(x);
        else if (OP == 5) r = fp_from_mont(x);
        else if (OP == 6) r = fp_sqr(x);
        else r = fp_mul_dfma(x, ld_g(&b[i]));
        st_g(&out[i], r);
    }
}

template <class F>
static void launch_vec(DeviceCtx& ctx, int op, const void* a, const void* b, void* out, uint64_t n) {
    if (n == 0) return;
    ctx.activate();
    int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
    const F* A = (const F*)a; const F* Bp = (const F*)b; F* O = (F*)out;
    switch (op) {
        case 0: k_field_vec<F, 0><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 1: k_field_vec<F, 1><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 2: k_field_vec<F, 2><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 3: k_field_vec<F, 3><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 4: k_field_vec<F, 4><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 5: k_field_vec<F, 5><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 6: k_field_vec<F, 6><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        case 7: k_field_vec<F, 7><<<blocks, 256, 0, ctx.stream>>>(A, Bp, O, n); break;
        default: throw std::runtime_error("ultragroth_b200: unknown field op");
    }
    UG_LAUNCH_CHECK(ctx);
}

void field_vec_op(DeviceCtx& ctx, int field, int op, const void* a, const void* b, void* out, uint64_t n) {
    if (field == 0) launch_vec<Fr>(ctx, op, a, b, out, n);
    else if (field == 1) launch_vec<Fq>(ctx, op, a, b, out, n);
    else throw std::runtime_error("ultragroth_b200: unknown field");
}

}  // namespace ug
