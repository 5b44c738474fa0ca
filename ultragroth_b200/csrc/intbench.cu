// Integer-pipe microbenchmarks: the roofline denominators for the MSM / NTT kernels (SURVEY.md 8(d):
// "imad_peak must be microbenchmarked on the box").  Every thread runs 8 independent chains so the
// 4-cycle IMAD latency is covered without relying on occupancy.
#include "curve.cuh"
#include "ff_variants.cuh"
#include "device.hpp"

namespace ug {

constexpr int IB_ITERS = 4096;

__global__ void __launch_bounds__(256) k_bench_imad(uint32_t* out, uint32_t seed) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + threadIdx.x * 8 + i;
    uint32_t m = seed | 1u;
    for (int it = 0; it < IB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(a[(i + 1) & 7]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= a[i];
    if (s == 0x12345678u) out[0] = s;
}

__global__ void __launch_bounds__(256) k_bench_imad_wide(uint64_t* out, uint32_t seed) {
    uint64_t a[8];
    uint32_t x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = seed + threadIdx.x * 8 + i; x[i] = seed * (i + 3) + threadIdx.x; }
    for (int it = 0; it < IB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            // the multiplier is the low word of a neighbouring accumulator: nothing is loop invariant (a first version
            // with a constant multiplier was strength-reduced by ptxas to IADD3 pairs and overstated the rate 2x)
            uint32_t y = (uint32_t)a[(i + 1) & 7];
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(a[i]) : "r"(x[i]), "r"(y));
        }
    }
    uint64_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= a[i];
    if (s == 0x12345678u) out[0] = s;
}


// kind 3: the carry-linked row of the Montgomery product: mad.lo.cc / madc.hi.cc pairs, which ptxas fuses into
// one IMAD.WIDE.U32 (carry out) followed by IMAD.WIDE.U32.X (carry in + out).  4 wide multiply-adds per row,
// 4 independent rows per thread.
__global__ void __launch_bounds__(256) k_bench_imad_wide_x(uint32_t* out, uint32_t seed) {
    uint32_t a[4][8];
    uint32_t y[4];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) a[r][i] = seed + threadIdx.x * 8 + i + r;
#pragma unroll
    for (int i = 0; i < 4; i++) y[i] = seed * (i + 3) + threadIdx.x;
    uint32_t x = seed | 1u;
    for (int it = 0; it < IB_ITERS / 2; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            a[r][0] = mad_lo_cc(y[0], x, a[r][0]);
            a[r][1] = madc_hi_cc(y[0], x, a[r][1]);
            a[r][2] = madc_lo_cc(y[1], x, a[r][2]);
            a[r][3] = madc_hi_cc(y[1], x, a[r][3]);
            a[r][4] = madc_lo_cc(y[2], x, a[r][4]);
            a[r][5] = madc_hi_cc(y[2], x, a[r][5]);
            a[r][6] = madc_lo_cc(y[3], x, a[r][6]);
            a[r][7] = madc_hi(y[3], x, a[r][7]);
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) s ^= a[r][i];
    if (s == 0x12345678u) out[0] = s;
}

// kind 4: carry OUT only: (mad.lo.cc, madc.hi.cc) + addc into a separate word -> IMAD.WIDE.U32 writing a predicate,
// consumed by an IADD3.X on the ALU pipe.  8 independent accumulators per thread.
__global__ void __launch_bounds__(256) k_bench_imad_wide_cout(uint32_t* out, uint32_t seed) {
    uint32_t lo[8], hi[8], c[8], y[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { lo[i] = seed + threadIdx.x + i; hi[i] = seed * 3 + i; c[i] = 0; y[i] = seed * (i + 3) + threadIdx.x; }
    uint32_t x = seed | 1u;
    for (int it = 0; it < IB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            lo[i] = mad_lo_cc(y[i], x, lo[i]);
            hi[i] = madc_hi_cc(y[i], x, hi[i]);
            c[i] = addc(c[i], 0);
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= lo[i] ^ hi[i] ^ c[i];
    if (s == 0x12345678u) out[0] = s;
}

constexpr int MB_ITERS = 256;
__global__ void __launch_bounds__(256) k_bench_modmul(Fq* out, uint32_t seed) {
    Fq a[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        a[i] = Fq::one();
        a[i].v[0] += seed + threadIdx.x + i;
    }
    for (int it = 0; it < MB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = fp_mul(a[i], a[(i + 1) & 3]);
    }
    Fq s = fp_add(fp_add(a[0], a[1]), fp_add(a[2], a[3]));
    if (s.v[0] == 0x12345678u && s.v[7] == 1) out[0] = s;
}

// kind 5: the same chains on the Karatsuba + separate-reduction product (fp_mul_ks, ff_variants.cuh)
__global__ void __launch_bounds__(256) k_bench_modmul_ks(Fq* out, uint32_t seed) {
    Fq a[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        a[i] = Fq::one();
        a[i].v[0] += seed + threadIdx.x + i;
    }
    for (int it = 0; it < MB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = fp_mul_ks(a[i], a[(i + 1) & 3]);
    }
    Fq s = fp_add(fp_add(a[0], a[1]), fp_add(a[2], a[3]));
    if (s.v[0] == 0x12345678u && s.v[7] == 1) out[0] = s;
}

// kind 6: the FP64 pipe's own peak: 8 independent fma.rz.f64 chains per thread
__global__ void __launch_bounds__(256) k_bench_dfma(double* out, uint32_t seed) {
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = 1.0 + 1e-9 * (double)(seed + threadIdx.x * 8 + i);
    const double m = 1.0 + 1e-12 * (double)seed;
    for (int it = 0; it < IB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = __fma_rz(a[i], m, a[(i + 1) & 7] * 1e-300);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    if (s == 0.12345) out[0] = s;
}

// kind 7: kind 2's dependent chains on the FP64-assisted product (fp_mul_dfma, ff_variants.cuh)
__global__ void __launch_bounds__(256) k_bench_modmul_dfma(Fq* out, uint32_t seed) {
    Fq a[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        a[i] = Fq::one();
        a[i].v[0] += seed + threadIdx.x + i;
    }
    for (int it = 0; it < MB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) a[i] = fp_mul_dfma(a[i], a[(i + 1) & 3]);
    }
    Fq s = fp_add(fp_add(a[0], a[1]), fp_add(a[2], a[3]));
    if (s.v[0] == 0x12345678u && s.v[7] == 1) out[0] = s;
}

// kinds 8 / 9: dependent Fq2 product chains, three complete Fq products (Karatsuba) vs lazy reduction (fq2_mul_lazy, curve.cuh)
template <int LAZY>
__global__ void __launch_bounds__(256) k_bench_fq2mul(Fq2* out, uint32_t seed) {
    Fq2 a[2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
        a[i] = Fq2::one();
        a[i].a.v[0] += seed + threadIdx.x + i;
        a[i].b.v[1] += seed * 3 + i;
    }
    for (int it = 0; it < MB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 2; i++) {
            if (LAZY) { Fq2 r; fq2_mul_lazy(a[i].a, a[i].b, a[i ^ 1].a, a[i ^ 1].b, r.a, r.b); a[i] = r; }
            else a[i] = fmul(a[i], a[i ^ 1]);
        }
    }
    Fq2 s = fadd(a[0], a[1]);
    if (s.a.v[0] == 0x12345678u && s.b.v[7] == 1) out[0] = s;
}

double bench_int_pipe(DeviceCtx& ctx, int kind) {
    ctx.activate();
    void* d = ctx.workspace("intbench", 256);
    cudaEvent_t e0, e1;
    UG_CUDA_CHECK(cudaEventCreate(&e0));
    UG_CUDA_CHECK(cudaEventCreate(&e1));
    const int blocks = ctx.sm_count * 8;
    double ops = 0;
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        UG_CUDA_CHECK(cudaEventRecord(e0, ctx.stream));
        if (kind == 0) { k_bench_imad<<<blocks, 256, 0, ctx.stream>>>((uint32_t*)d, 7 + rep); ops = (double)blocks * 256 * 8 * IB_ITERS; }
        else if (kind == 1) { k_bench_imad_wide<<<blocks, 256, 0, ctx.stream>>>((uint64_t*)d, 7 + rep); ops = (double)blocks * 256 * 8 * IB_ITERS; }
        else if (kind == 3) { k_bench_imad_wide_x<<<blocks, 256, 0, ctx.stream>>>((uint32_t*)d, 7 + rep); ops = (double)blocks * 256 * 16 * (IB_ITERS / 2); }
        else if (kind == 4) { k_bench_imad_wide_cout<<<blocks, 256, 0, ctx.stream>>>((uint32_t*)d, 7 + rep); ops = (double)blocks * 256 * 8 * IB_ITERS; }
        else if (kind == 6) { k_bench_dfma<<<blocks, 256, 0, ctx.stream>>>((double*)d, 7 + rep); ops = (double)blocks * 256 * 8 * IB_ITERS; }
        else if (kind == 7) { k_bench_modmul_dfma<<<blocks, 256, 0, ctx.stream>>>((Fq*)d, 7 + rep); ops = (double)blocks * 256 * 4 * MB_ITERS; }
        else if (kind == 8) { k_bench_fq2mul<0><<<blocks, 256, 0, ctx.stream>>>((Fq2*)d, 7 + rep); ops = (double)blocks * 256 * 2 * MB_ITERS; }
        else if (kind == 9) { k_bench_fq2mul<1><<<blocks, 256, 0, ctx.stream>>>((Fq2*)d, 7 + rep); ops = (double)blocks * 256 * 2 * MB_ITERS; }
        else if (kind == 5) { k_bench_modmul_ks<<<blocks, 256, 0, ctx.stream>>>((Fq*)d, 7 + rep); ops = (double)blocks * 256 * 4 * MB_ITERS; }
        else { k_bench_modmul<<<blocks, 256, 0, ctx.stream>>>((Fq*)d, 7 + rep); ops = (double)blocks * 256 * 4 * MB_ITERS; }
        UG_LAUNCH_CHECK(ctx);
        UG_CUDA_CHECK(cudaEventRecord(e1, ctx.stream));
        UG_CUDA_CHECK(cudaEventSynchronize(e1));
        float ms;
        UG_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0) best = std::min(best, ms);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return ops / (best * 1e-3);
}

}  // namespace ug
