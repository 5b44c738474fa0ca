// A G1 MSM that allocates NOTHING: for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops,
// harnesses that forbid it).  No counterpart in the reference.
//
// Every other MSM entry point takes its bucket / partial-sum scratch from DeviceCtx::workspace (cudaMalloc on first use and on
// growth).  This variant keeps ALL of its scratch in one file-scope __device__ array of fixed size, so it is limited to
// MSM_NOALLOC_MAX_POINTS points and its bucket method is the plain one (8-bit unsigned windows, no digit sort):
//
//   k_na_buckets      grid (slices, 32 windows) x 256 threads.  Window w of a scalar is simply its byte w.  Thread b of a CTA owns bucket b
//                     of (window w, slice s): it scans the slice's byte digits in shared memory and adds the points whose digit is b
//                     (the scan is divergent and cheap, the mixed additions are warp-convergent).  Every thread stores its bucket,
//                     the point at infinity included, so no slot of the scratch has to be cleared beforehand.
//   k_na_window_sums  32 CTAs x 256 threads: bucket b summed over the slices, then sum_b b * B_b as sum_k 2^k (sum of the B_b with bit
//                     k of b set): eight tree sums in shared memory and a Horner step each.
//   k_na_horner       one thread: sum_w 2^(8 w) W_w (248 doublings, 31 additions), result XYZZ -> 128 bytes to the host.
//
// The scratch array is shared by every call on the device and keeps the last call's contents; nothing in it has to start at zero
// (each slot that is read was written earlier in the same call), but two calls that overlap on the GPU -- two contexts / streams on
// one device -- RACE on it: callers serialise.  Every GPU that loads this module holds the array until the process exits.
#include "curve.cuh"
#include "memops.cuh"
#include "msm.hpp"

#include <cstring>
#include <mutex>
#include <string>

namespace ug {

constexpr uint32_t NA_WINDOWS = 32;       // 8-bit windows over the 256 scalar bits (no reduction mod r needed: G1 has order r)
constexpr uint32_t NA_BUCKETS = 256;      // bucket b = digit value b; b = 0 stays empty
constexpr uint32_t NA_SLICE = 2048;       // points per CTA of k_na_buckets
constexpr uint32_t NA_MAX_SLICES = MSM_NOALLOC_MAX_POINTS / NA_SLICE;
constexpr size_t NA_PARTIAL_BYTES = (size_t)NA_WINDOWS * NA_MAX_SLICES * NA_BUCKETS * sizeof(G1XYZZ);  // 8 MiB
constexpr size_t NA_SUMS_BYTES = (size_t)(NA_WINDOWS + 1) * sizeof(G1XYZZ);                             // window sums + the result
constexpr size_t NA_SCRATCH_BYTES = NA_PARTIAL_BYTES + NA_SUMS_BYTES;
static_assert(MSM_NOALLOC_MAX_POINTS % NA_SLICE == 0, "whole slices");
static_assert(NA_SCRATCH_BYTES == MSM_NOALLOC_SCRATCH_BYTES, "msm.hpp states the footprint");

// In HOST code this name is only a stand-in for the symbol (not a device address): the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];

__global__ void __launch_bounds__(256) k_na_buckets(const G1Affine* __restrict__ bases, const uint8_t* __restrict__ scalars, uint32_t n,
                                                    G1XYZZ* __restrict__ partial) {
    __shared__ uint8_t dig[NA_SLICE];
    const uint32_t s = blockIdx.x, w = blockIdx.y, S = gridDim.x;
    const uint32_t lo = s * NA_SLICE;
    const uint32_t cnt = min(NA_SLICE, n - lo);
    for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) dig[i] = scalars[(size_t)(lo + i) * 32 + w];
    __syncthreads();
    const uint32_t b = threadIdx.x;
    G1XYZZ acc = G1XYZZ::inf();
    uint32_t i = b == 0 ? cnt : 0;  // digit 0 contributes nothing
    while (true) {
        while (i < cnt && dig[i] != b) i++;
        const bool have = i < cnt;
        if (!__any_sync(0xffffffffu, have)) break;
        if (have) {
            pt_madd(acc, ld_nc(&bases[lo + i]), false);
            i++;
        }
    }
    st_g(&partial[((size_t)w * S + s) * NA_BUCKETS + b], acc);
}

__global__ void __launch_bounds__(256) k_na_window_sums(const G1XYZZ* __restrict__ partial, uint32_t S, G1XYZZ* __restrict__ sums) {
    __shared__ uint4 tmp_raw[NA_BUCKETS * sizeof(G1XYZZ) / 16];  // 32 KB
    G1XYZZ* tmp = reinterpret_cast<G1XYZZ*>(tmp_raw);
    const uint32_t b = threadIdx.x, w = blockIdx.x;
    G1XYZZ mine = G1XYZZ::inf();
    for (uint32_t s = 0; s < S; s++) pt_add(mine, ld_g(&partial[((size_t)w * S + s) * NA_BUCKETS + b]));
    G1XYZZ wsum = G1XYZZ::inf();  // thread 0's
    for (int k = 7; k >= 0; k--) {
        tmp[b] = ((b >> k) & 1u) ? mine : G1XYZZ::inf();
        __syncthreads();
        for (uint32_t stride = NA_BUCKETS / 2; stride >= 1; stride >>= 1) {
            if (b < stride) {
                G1XYZZ t = tmp[b];
                pt_add(t, tmp[b + stride]);
                tmp[b] = t;
            }
            __syncthreads();
        }
        if (b == 0) {
            wsum = pt_dbl(wsum);
            pt_add(wsum, tmp[0]);
        }
        __syncthreads();
    }
    if (b == 0) st_g(&sums[w], wsum);
}

__global__ void k_na_horner(G1XYZZ* __restrict__ sums) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    G1XYZZ acc = ld_g(&sums[NA_WINDOWS - 1]);
    for (int w = (int)NA_WINDOWS - 2; w >= 0; w--) {
        for (int k = 0; k < 8; k++) acc = pt_dbl(acc);
        pt_add(acc, ld_g(&sums[w]));
    }
    st_g(&sums[NA_WINDOWS], acc);
}

namespace {
std::mutex g_na_mu;
constexpr int NA_MAX_DEVICES = 64;
unsigned char* g_na_addr[NA_MAX_DEVICES] = {};

// the scratch array's address ON THIS DEVICE: looked up once per device, with that device current
unsigned char* na_scratch(DeviceCtx& ctx) {
    ctx.activate();
    if (ctx.device < 0 || ctx.device >= NA_MAX_DEVICES) throw std::runtime_error("ultragroth_b200: device index out of range for the no-alloc MSM");
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_addr[ctx.device]) {
        void* p = nullptr;
        UG_CUDA_CHECK(cudaGetSymbolAddress(&p, g_na_scratch));
        if (!p) throw std::runtime_error("ultragroth_b200: cudaGetSymbolAddress gave no address for the no-alloc MSM scratch");
        g_na_addr[ctx.device] = (unsigned char*)p;
    }
    return g_na_addr[ctx.device];
}
}  // namespace

// sum_i scalars[i] * bases[i], i < n <= MSM_NOALLOC_MAX_POINTS; device pointers in, XYZZ result (128 B) to host memory.
// No cudaMalloc / cudaFree / cudaMallocHost / event creation on this path: three launches, one 128-byte copy, one stream sync.
void msm_g1_noalloc(DeviceCtx& ctx, const void* d_bases, const void* d_scalars, uint64_t n, void* h_out_xyzz) {
    if (n > MSM_NOALLOC_MAX_POINTS)
        throw std::invalid_argument("ultragroth_b200: ugb_msm_g1_noalloc handles at most " + std::to_string(MSM_NOALLOC_MAX_POINTS) + " points (fixed scratch of " +
                                    std::to_string(MSM_NOALLOC_SCRATCH_BYTES) + " bytes); got " + std::to_string(n) + " -- use ugb_msm_g1");
    if (!h_out_xyzz) throw std::invalid_argument("ultragroth_b200: null output pointer");
    if (n == 0) {
        memset(h_out_xyzz, 0, sizeof(G1XYZZ));  // the point at infinity
        return;
    }
    if (!d_bases || !d_scalars) throw std::invalid_argument("ultragroth_b200: null input pointer");
    unsigned char* scratch = na_scratch(ctx);
    G1XYZZ* partial = (G1XYZZ*)scratch;
    G1XYZZ* sums = (G1XYZZ*)(scratch + NA_PARTIAL_BYTES);
    const uint32_t S = (uint32_t)((n + NA_SLICE - 1) / NA_SLICE);
    k_na_buckets<<<dim3(S, NA_WINDOWS), 256, 0, ctx.stream>>>((const G1Affine*)d_bases, (const uint8_t*)d_scalars, (uint32_t)n, partial);
    UG_LAUNCH_CHECK(ctx);
    k_na_window_sums<<<NA_WINDOWS, 256, 0, ctx.stream>>>(partial, S, sums);
    UG_LAUNCH_CHECK(ctx);
    k_na_horner<<<1, 32, 0, ctx.stream>>>(sums);
    UG_LAUNCH_CHECK(ctx);
    UG_CUDA_CHECK(cudaMemcpyAsync(h_out_xyzz, sums + NA_WINDOWS, sizeof(G1XYZZ), cudaMemcpyDeviceToHost, ctx.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
}

}  // namespace ug
