// MSM driver: launches the kernels of msm.cuh and combines the per-bit sums on the host.
#include "msm.cuh"
#include "msm.hpp"

#include <algorithm>
#include <cstdlib>

namespace ug {

MsmConfig msm_pick_config(uint64_t n, int force_c) {
    int best_c = 0;
    if (force_c <= 0) {
        const char* env = getenv("UGB_MSM_C");
        if (env) force_c = atoi(env);
    }
    if (force_c > 0) {
        best_c = std::min(std::max(force_c, 2), 22);
    } else {
        // cost in modular multiplications: n*W mixed adds (10) + W*c*(B/2) bucket-bit adds (14)
        double best = 1e300;
        for (int c = 4; c <= 20; c++) {
            double W = (255 + c - 1) / c;
            double B = (double)(1u << (c - 1));
            double cost = (double)n * W * 10.0 + W * c * (B / 2) * 14.0 + W * B * 2.0 /* bucket memset + scan */;
            if (cost < best) { best = cost; best_c = c; }
        }
    }
    MsmConfig cfg;
    cfg.c = best_c;
    cfg.W = (255 + best_c - 1) / best_c;
    cfg.B = 1u << (best_c - 1);
    cfg.NB = (uint32_t)cfg.W * cfg.B;
    return cfg;
}

static void exclusive_scan_u32(DeviceCtx& ctx, const uint32_t* in, uint32_t n, uint32_t* out, uint32_t* out_total,
                               const std::string& slot) {
    uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    uint32_t* tiles = (uint32_t*)ctx.workspace(slot + ".scan_tiles", (size_t)ntiles * 4);
    UG_KERNEL(ctx, "k_scan_tile_sums", k_scan_tile_sums<<<ntiles, SCAN_BS, 0, ctx.stream>>>(in, n, tiles));
    UG_KERNEL(ctx, "k_scan_tiles", k_scan_tiles<<<1, SCAN_BS, 0, ctx.stream>>>(tiles, ntiles, out_total));
    UG_KERNEL(ctx, "k_scan_apply", k_scan_apply<<<ntiles, SCAN_BS, 0, ctx.stream>>>(in, n, tiles, out));
}

void msm_sort_digits(DeviceCtx& ctx, const uint8_t* d_scalars, const uint32_t* d_index, uint64_t n, int force_c,
                     const std::string& slot, MsmDigits& out) {
    ctx.activate();
    if (n >= (1ull << 31)) throw std::runtime_error("ultragroth_b200: MSM size must be < 2^31");
    MsmConfig cfg = msm_pick_config(n, force_c);
    if ((uint64_t)n * cfg.W >= (1ull << 32)) throw std::runtime_error("ultragroth_b200: MSM digit count overflows 32 bits");
    out.cfg = cfg;
    out.n = n;
    uint32_t* counts = (uint32_t*)ctx.workspace(slot + ".counts", (size_t)cfg.NB * 4);
    uint32_t* cursor = (uint32_t*)ctx.workspace(slot + ".cursor", (size_t)cfg.NB * 4);
    out.starts = (uint32_t*)ctx.workspace(slot + ".starts", ((size_t)cfg.NB + 1) * 4);
    out.entries = (uint32_t*)ctx.workspace(slot + ".entries", std::max<size_t>((size_t)n * cfg.W * 4, 16));
    UG_CUDA_CHECK(cudaMemsetAsync(counts, 0, (size_t)cfg.NB * 4, ctx.stream));
    if (n > 0) {
        int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_digits_count", k_digits_count<<<blocks, 256, 0, ctx.stream>>>(d_scalars, d_index, n, cfg.c, cfg.W, counts));
    }
    exclusive_scan_u32(ctx, counts, cfg.NB, out.starts, out.starts + cfg.NB, slot);
    UG_CUDA_CHECK(cudaMemcpyAsync(cursor, out.starts, (size_t)cfg.NB * 4, cudaMemcpyDeviceToDevice, ctx.stream));
    if (n > 0) {
        int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_digits_scatter", k_digits_scatter<<<blocks, 256, 0, ctx.stream>>>(d_scalars, d_index, n, cfg.c, cfg.W, cursor, out.entries));
    }
}

template <class F>
static void msm_launch_impl(DeviceCtx& ctx, const Affine<F>* d_bases, const MsmDigits& dg, const std::string& slot) {
    ctx.activate();
    const MsmConfig& cfg = dg.cfg;
    using P = XYZZ<F>;
    P* buckets = (P*)ctx.workspace("msm.buckets", (size_t)cfg.NB * sizeof(P));
    UG_CUDA_CHECK(cudaMemsetAsync(buckets, 0, (size_t)cfg.NB * sizeof(P), ctx.stream));

    // accumulate: T threads over the sorted entries
    int occ = 1;
    UG_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_accumulate<F>, 128, 0));
    occ = std::max(occ, 1);
    uint64_t max_entries = dg.n * (uint64_t)cfg.W;
    uint64_t want_threads = std::max<uint64_t>(128, (max_entries + 15) / 16);
    uint32_t blocks = (uint32_t)std::min<uint64_t>((uint64_t)ctx.sm_count * occ, (want_threads + 127) / 128);
    uint32_t T = blocks * 128;
    uint32_t K = 2 * T;
    // ping-pong partial lists
    uint32_t* keys_a = (uint32_t*)ctx.workspace("msm.pkeys_a", (size_t)K * 4);
    P* pts_a = (P*)ctx.workspace("msm.ppts_a", (size_t)K * sizeof(P));
    uint32_t K1 = 2 * ((K + SEG_BS - 1) / SEG_BS);
    uint32_t* keys_b = (uint32_t*)ctx.workspace("msm.pkeys_b", (size_t)K1 * 4);
    P* pts_b = (P*)ctx.workspace("msm.ppts_b", (size_t)K1 * sizeof(P));

    UG_KERNEL(ctx, (sizeof(F) == 32 ? "k_accumulate<G1>" : "k_accumulate<G2>"), k_accumulate<F><<<blocks, 128, 0, ctx.stream>>>(d_bases, dg.entries, dg.starts, cfg.NB, buckets, keys_a, pts_a));

    uint32_t* kin = keys_a; P* pin = pts_a; uint32_t* kout = keys_b; P* pout = pts_b;
    uint32_t Kcur = K;
    while (true) {
        uint32_t g = (Kcur + SEG_BS - 1) / SEG_BS;
        int final_pass = (g == 1);
        UG_KERNEL(ctx, (sizeof(F) == 32 ? "k_seg_reduce<G1>" : "k_seg_reduce<G2>"), k_seg_reduce<F><<<g, SEG_BS, 0, ctx.stream>>>(kin, pin, Kcur, cfg.NB, buckets, kout, pout, final_pass));
        if (final_pass) break;
        Kcur = 2 * g;
        std::swap(kin, kout);
        std::swap(pin, pout);
    }

    // bucket reduction to per-bit sums
    int S = std::max(1, std::min<int>(16, (int)(cfg.B / 2 / BITS_BS)));
    int nbits = cfg.W * cfg.c;
    P* slices = (P*)ctx.workspace("msm.slices", (size_t)nbits * S * sizeof(P));
    P* bits = (P*)ctx.workspace("msm.bits." + slot, (size_t)nbits * sizeof(P));
    dim3 grid(S, cfg.c, cfg.W);
    UG_KERNEL(ctx, (sizeof(F) == 32 ? "k_bucket_bits<G1>" : "k_bucket_bits<G2>"), k_bucket_bits<F><<<grid, BITS_BS, 0, ctx.stream>>>(buckets, cfg.c, S, slices));
    UG_KERNEL(ctx, (sizeof(F) == 32 ? "k_sum_slices<G1>" : "k_sum_slices<G2>"), k_sum_slices<F><<<(nbits + 63) / 64, 64, 0, ctx.stream>>>(slices, S, nbits, bits));
    // pinned result slot: nbits partial sums, then the number of digit entries this MSM processed (a counter
    // for the roofline arithmetic: every entry is one 4-byte index + one affine point gather + one mixed add)
    uint8_t* host = (uint8_t*)ctx.pinned("msm.bits." + slot, (size_t)nbits * sizeof(P) + 16);
    UG_CUDA_CHECK(cudaMemcpyAsync(host, bits, (size_t)nbits * sizeof(P), cudaMemcpyDeviceToHost, ctx.stream));
    UG_CUDA_CHECK(cudaMemcpyAsync(host + (size_t)nbits * sizeof(P), dg.starts + cfg.NB, 4, cudaMemcpyDeviceToHost, ctx.stream));
    ctx.record("msm." + slot);
}

void msm_g1_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot) {
    msm_launch_impl<Fq>(ctx, (const G1Affine*)d_bases, dg, slot);
}
void msm_g2_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot) {
    msm_launch_impl<Fq2>(ctx, (const G2Affine*)d_bases, dg, slot);
}
}  // namespace ug
