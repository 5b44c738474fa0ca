// MSM driver: launches the kernels of msm.cuh; the host combines the per-level / per-window partial sums (msm_host.cpp).
#include "msm.cuh"
#include "msm_variants.cuh"
#include "msm.hpp"

#include <algorithm>
#include <cstdlib>

namespace ug {

// default kernel variants, see "kernel variants" below
constexpr int ACC_VARIANT_G1 = 0, ACC_VARIANT_G2 = 12, RS_VARIANT_G1 = 0, RS_VARIANT_G2 = 12;

static int env_int_msm(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

MsmConfig msm_pick_config(uint64_t n, int force_c, bool pre) {
    int best_c = 0;
    if (force_c <= 0) force_c = env_int_msm(pre ? "UGB_MSM_PRE_C" : "UGB_MSM_C", 0);
    if (force_c > 0) {
        best_c = std::min(std::max(force_c, 2), 23);
    } else {
        // cost in modular multiplications: n*W mixed additions (10 each) + the bucket reduction.
        //   ad-hoc bases: W bucket sets, 2 additions (14 each) per bucket for the running sums + memset / histogram traffic (~2);
        //   resident tables: ONE bucket set, but its reduction tails (chunked running sums, bit sums, segmented merge) are latency-
        //   bound launches: calibrated on the witness-MSM phase of whole and of 1/2, 1/4, 1/8 sharded 2^22 provers they cost ~54
        //   products per bucket next to the accumulation of witness-like scalars (profiles/r02_window_calibration.txt; the old 30
        //   chose c = 22 where c = 20 proves 4 % faster).  A window that needs as many windows W as a smaller one only adds buckets
        //   (c = 18 against 17, 21 against 20: measured slower every time), so only the smallest c of each W is a candidate.
        const double per_bucket = pre ? 54.0 : 30.0;
        double best = 1e300;
        for (int c = 4; c <= (pre ? 22 : 20); c++) {
            const int Wi = (255 + c - 1) / c;
            if (pre && c > 4 && (255 + c - 2) / (c - 1) == Wi) continue;
            double W = Wi;
            double B = (double)(1u << (c - 1));
            double sets = pre ? 1.0 : W;
            if ((double)n * W >= 2147483648.0) continue;  // table index + sign bit must fit 32 bits
            double cost = (double)n * W * 10.0 + sets * B * per_bucket;
            if (cost < best) { best = cost; best_c = c; }
        }
        if (best_c == 0) best_c = 16;
    }
    MsmConfig cfg;
    cfg.c = best_c;
    cfg.W = (255 + best_c - 1) / best_c;
    cfg.B = 1u << (best_c - 1);
    cfg.pre = pre;
    cfg.NB = pre ? cfg.B : (uint32_t)cfg.W * cfg.B;
    return cfg;
}

MsmReducePlan msm_reduce_plan(const MsmConfig& cfg) {
    MsmReducePlan p;
    // about 64K threads over all windows, chunks of 4..256 buckets
    uint64_t total = (uint64_t)cfg.B * cfg.nwin();
    int mu = 2;
    while (mu < 8 && (total >> mu) > 65536) mu++;
    if (cfg.pre) {
        // one bucket set: a chunk of 2^mu buckets is a serial chain of 2 * 2^mu additions, the bit sums cost K * log2(K) / 2 additions
        // over K = B / 2^mu chunk sums; measured optimum (tools/r02_rs_mu.sh): mu = 2..3 at B = 2^16, 4 at 2^18, 5..6 at 2^21
        int lb = 0;
        while ((1ull << (lb + 1)) <= cfg.B) lb++;
        mu = std::min(8, std::max(2, (lb - 10) / 2));
    }
    const int forced = env_int_msm("UGB_RS_MU", 0);  // experiment knob: chunk = 2^mu buckets
    if (forced >= 1 && forced <= 10) mu = forced;
    p.mu = mu;
    p.K = (cfg.B + (1u << mu) - 1) >> mu;
    p.nbits = 0;
    while (p.K > 1 && ((p.K - 1) >> p.nbits) != 0) p.nbits++;
    return p;
}

static void exclusive_scan_u32(DeviceCtx& ctx, const uint32_t* in, uint32_t n, uint32_t* out, uint32_t* out_total,
                               const std::string& slot) {
    uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    uint32_t* tiles = (uint32_t*)ctx.workspace(slot + ".scan_tiles", (size_t)ntiles * 4);
    UG_KERNEL(ctx, "k_scan_tile_sums", k_scan_tile_sums<<<ntiles, SCAN_BS, 0, ctx.stream>>>(in, n, tiles));
    UG_KERNEL(ctx, "k_scan_tiles", k_scan_tiles<<<1, SCAN_BS, 0, ctx.stream>>>(tiles, ntiles, out_total));
    UG_KERNEL(ctx, "k_scan_apply", k_scan_apply<<<ntiles, SCAN_BS, 0, ctx.stream>>>(in, n, tiles, out));
}

void msm_inf_bitmap_async(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, uint32_t* d_bitmap, bool combine, uint32_t* d_count) {
    ctx.activate();
    if (n == 0) return;
    const int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
    if (group == 1) UG_KERNEL(ctx, "k_inf_bitmap", k_inf_bitmap<Fq><<<blocks, 256, 0, ctx.stream>>>((const G1Affine*)d_bases, (uint32_t)n, d_bitmap, combine ? 1 : 0, d_count));
    else UG_KERNEL(ctx, "k_inf_bitmap", k_inf_bitmap<Fq2c><<<blocks, 256, 0, ctx.stream>>>((const Affine<Fq2c>*)d_bases, (uint32_t)n, d_bitmap, combine ? 1 : 0, d_count));
}

uint64_t msm_bitmap_and(DeviceCtx& ctx, const uint32_t* d_x, const uint32_t* d_y, uint64_t n_points, uint32_t* d_out) {
    ctx.activate();
    const uint32_t words = (uint32_t)((n_points + 31) / 32);
    if (words == 0) return 0;
    uint32_t* d_count = (uint32_t*)ctx.workspace("msm.inf_count", 16);
    UG_CUDA_CHECK(cudaMemsetAsync(d_count, 0, 4, ctx.stream));
    const int blocks = (int)std::min<uint32_t>((words + 255) / 256, (uint32_t)ctx.sm_count * 4);
    UG_KERNEL(ctx, "k_bitmap_and", k_bitmap_and<<<blocks, 256, 0, ctx.stream>>>(d_x, d_y, words, d_out, d_count));
    uint32_t h = 0;
    UG_CUDA_CHECK(cudaMemcpyAsync(&h, d_count, 4, cudaMemcpyDeviceToHost, ctx.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
    return h;
}

uint64_t msm_inf_bitmap(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, uint32_t* d_bitmap, bool combine) {
    ctx.activate();
    if (n == 0) return 0;
    uint32_t* d_count = (uint32_t*)ctx.workspace("msm.inf_count", 16);
    UG_CUDA_CHECK(cudaMemsetAsync(d_count, 0, 4, ctx.stream));
    msm_inf_bitmap_async(ctx, group, d_bases, n, d_bitmap, combine, d_count);
    uint32_t h = 0;
    UG_CUDA_CHECK(cudaMemcpyAsync(&h, d_count, 4, cudaMemcpyDeviceToHost, ctx.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
    return h;
}

void msm_sort_digits(DeviceCtx& ctx, const uint8_t* d_scalars, const uint32_t* d_index, uint64_t n, int force_c,
                     const std::string& slot, MsmDigits& out, const MsmConfig* table_cfg, const uint32_t* d_skip) {
    ctx.activate();
    if (n >= (1ull << 31)) throw std::runtime_error("ultragroth_b200: MSM size must be < 2^31");
    MsmConfig cfg = table_cfg ? *table_cfg : msm_pick_config(n, force_c, false);
    if ((uint64_t)n * cfg.W >= (1ull << 31)) throw std::runtime_error("ultragroth_b200: MSM digit count overflows 31 bits");
    out.cfg = cfg;
    out.n = n;
    uint32_t* counts = (uint32_t*)ctx.workspace(slot + ".counts", (size_t)cfg.NB * 4);
    uint32_t* cursor = (uint32_t*)ctx.workspace(slot + ".cursor", (size_t)cfg.NB * 4);
    out.starts = (uint32_t*)ctx.workspace(slot + ".starts", ((size_t)cfg.NB + 1) * 4);
    out.entries = (uint32_t*)ctx.workspace(slot + ".entries", std::max<size_t>((size_t)n * cfg.W * 4, 16));
    const uint32_t key_stride = cfg.pre ? 0u : cfg.B;
    const uint32_t entry_stride = cfg.pre ? (uint32_t)n : 0u;
    UG_CUDA_CHECK(cudaMemsetAsync(counts, 0, (size_t)cfg.NB * 4, ctx.stream));
    if (n > 0) {
        int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_digits_count", k_digits_count<<<blocks, 256, 0, ctx.stream>>>(d_scalars, d_index, n, cfg.c, cfg.W, key_stride, counts, d_skip));
    }
    exclusive_scan_u32(ctx, counts, cfg.NB, out.starts, out.starts + cfg.NB, slot);
    UG_CUDA_CHECK(cudaMemcpyAsync(cursor, out.starts, (size_t)cfg.NB * 4, cudaMemcpyDeviceToDevice, ctx.stream));
    if (n > 0) {
        int blocks = (int)std::min<uint64_t>((n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_digits_scatter", k_digits_scatter<<<blocks, 256, 0, ctx.stream>>>(d_scalars, d_index, n, cfg.c, cfg.W, key_stride, entry_stride, cursor, out.entries, d_skip));
    }
}

// ---- kernel variants ------------------------------------------------------------------------------
// Selected by UGB_ACC_G1 / UGB_ACC_G2 / UGB_RS_G1 / UGB_RS_G2 (defaults above); only the variants that earned a place after
// the round-1 sweeps (tools/variants.sh) are compiled:
//    0   all-register kernel (k_accumulate / k_running_sum)                                           G1 default, G2
//   12   G2 only: accumulator(s) in shared memory, 2 CTAs/SM                                          G2 default
//   24   G1 only: batched-affine accumulation (k_accumulate_affine, 4 CTAs/SM)                        opt-in, slower
//   33   G2 only: batched-affine accumulation, 3 CTAs/SM                                              opt-in, slower
// G2 kernels are all instantiated on Fq2c (msm_g2_launch): Fq2's memory layout with CALLED products.
template <class F> struct AccSig { using Fn = void (*)(const Affine<F>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<F>*, uint32_t*, XYZZ<F>*, uint32_t, uint32_t); };
template <class F> struct RsSig { using Fn = void (*)(const XYZZ<F>*, uint32_t, uint32_t, uint32_t, uint32_t, XYZZ<F>*, XYZZ<F>*); };
template <class F> struct BaSig { using Fn = void (*)(const Affine<F>*, const uint32_t*, const uint32_t*, uint32_t, XYZZ<F>*, uint32_t*, XYZZ<F>*, uint32_t, uint32_t, uint32_t, uint4*, uint4*, uint32_t*, uint32_t*); };

template <class F>
static typename BaSig<F>::Fn select_affine(int variant) {
    if constexpr (sizeof(F) == 64) {
        if (variant == 33) return k_accumulate_affine<F, 3>;
    } else {
        if (variant == 24) return k_accumulate_affine<F, 4>;
    }
    throw std::runtime_error("ultragroth_b200: unknown affine accumulate variant");
}
template <class F>
static void select_accumulate(int variant, typename AccSig<F>::Fn& fn, size_t& smem) {
    smem = 0;
    if (variant == 0) { fn = k_accumulate<F>; return; }
    if constexpr (sizeof(F) == 64) {
        if (variant == 12) { fn = k_accumulate_s<F, 2>; smem = 128 * sizeof(XYZZ<F>); return; }
    }
    throw std::runtime_error("ultragroth_b200: unknown accumulate variant");
}
template <class F>
static void select_running_sum(int variant, typename RsSig<F>::Fn& fn, size_t& smem) {
    smem = 0;
    if (variant == 0) { fn = k_running_sum<F>; return; }
    if constexpr (sizeof(F) == 64) {
        if (variant == 12) { fn = k_running_sum_s<F, 2>; smem = 2 * 128 * sizeof(XYZZ<F>); return; }
    }
    throw std::runtime_error("ultragroth_b200: unknown running-sum variant");
}

template <class F>
// nb base arrays that pair with the SAME sorted digits (A, B1, C of a proof) go through together: one accumulate launch each
// into consecutive bucket sets, then ONE pass of the latency-bound tail (segmented merge, running sums, bit sums) over all of
// them -- the tail's serial chains are paid once instead of nb times and its grids are nb times fuller.
static void msm_launch_impl(DeviceCtx& ctx, const Affine<F>* const* d_bases_v, int nb, const MsmDigits* const* dgs, const std::string* slots) {
    ctx.activate();
    // dgs[m]: the sorted digits array m pairs with (all with the same window config; they differ only in which points were skipped)
    const MsmDigits& dg = *dgs[0];
    const MsmConfig& cfg = dg.cfg;
    for (int m = 1; m < nb; m++)
        if (dgs[m]->cfg.c != cfg.c || dgs[m]->cfg.NB != cfg.NB || dgs[m]->cfg.pre != cfg.pre) throw std::runtime_error("ultragroth_b200: MSM batch mixes window configurations");
    const bool g1 = sizeof(F) == 32;
    using P = XYZZ<F>;
    const uint32_t NBT = (uint32_t)nb * cfg.NB;  // buckets of the whole batch
    if ((uint64_t)nb * cfg.NB >= (1ull << 31)) throw std::runtime_error("ultragroth_b200: MSM batch has too many buckets");
    P* buckets = (P*)ctx.workspace("msm.buckets", (size_t)NBT * sizeof(P));
    UG_CUDA_CHECK(cudaMemsetAsync(buckets, 0, (size_t)NBT * sizeof(P), ctx.stream));

    // accumulate: T threads over the sorted entries.  Kernel variant: the XYZZ accumulator in shared memory at MINB CTAs
    // per SM (k_accumulate_s), or 0 = all-register k_accumulate.  UGB_ACC_G1 / UGB_ACC_G2 override the default.
    using AccFn = typename AccSig<F>::Fn;
    const int acc_variant = env_int_msm(g1 ? "UGB_ACC_G1" : "UGB_ACC_G2", g1 ? ACC_VARIANT_G1 : ACC_VARIANT_G2);
    const bool affine = acc_variant >= 20;
    AccFn acc_fn = k_accumulate<F>;
    typename BaSig<F>::Fn ba_fn = nullptr;
    size_t acc_smem = 0;
    int occ = 1;
    if (affine) {
        ba_fn = select_affine<F>(acc_variant);
        UG_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ba_fn, BA_BS, 0));
    } else {
        select_accumulate<F>(acc_variant, acc_fn, acc_smem);
        UG_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, acc_fn, 128, acc_smem));
    }
    occ = std::max(occ, 1);
    uint64_t max_entries = 0;
    for (int m = 0; m < nb; m++) max_entries = std::max<uint64_t>(max_entries, dgs[m]->n * (uint64_t)cfg.W);
    uint64_t want_threads = std::max<uint64_t>(128, (max_entries + 15) / 16);
    uint32_t blocks = (uint32_t)std::min<uint64_t>((uint64_t)ctx.sm_count * occ, (want_threads + 127) / 128);
    uint32_t T = blocks * 128;
    uint32_t K = 2 * T * (uint32_t)nb;
    // ping-pong partial lists
    uint32_t* keys_a = (uint32_t*)ctx.workspace("msm.pkeys_a", (size_t)K * 4);
    P* pts_a = (P*)ctx.workspace("msm.ppts_a", (size_t)K * sizeof(P));
    uint32_t K1 = 2 * ((K + SEG_BS - 1) / SEG_BS);
    uint32_t* keys_b = (uint32_t*)ctx.workspace("msm.pkeys_b", (size_t)K1 * 4);
    P* pts_b = (P*)ctx.workspace("msm.ppts_b", (size_t)K1 * sizeof(P));

    for (int m = 0; m < nb; m++) {
        P* bk = buckets + (size_t)m * cfg.NB;
        uint32_t* pk = keys_a + (size_t)m * 2 * T;
        P* pp = pts_a + (size_t)m * 2 * T;
        const uint32_t key_off = (uint32_t)m * cfg.NB;
        if (affine) {
            // per-thread scratch lists, sized for the largest possible entry count (the real count is only known on the device)
            const uint32_t Lmax = (uint32_t)((max_entries + T - 1) / T) + 1;
            uint4* list_mem = (uint4*)ctx.workspace("msm.ba_list", (size_t)T * Lmax * sizeof(Affine<F>));
            uint4* suf_mem = (uint4*)ctx.workspace("msm.ba_suf", (size_t)T * (Lmax / 2 + 1) * sizeof(F));
            uint32_t* keys_mem = (uint32_t*)ctx.workspace("msm.ba_keys", (size_t)T * Lmax * 4);
            uint32_t* pairs_mem = (uint32_t*)ctx.workspace("msm.ba_pairs", (size_t)T * (Lmax / 2 + 1) * 4);
            UG_KERNEL(ctx, (g1 ? "k_accumulate<G1>" : "k_accumulate<G2>"), ba_fn<<<blocks, BA_BS, 0, ctx.stream>>>(d_bases_v[m], dgs[m]->entries, dgs[m]->starts, cfg.NB, bk, pk, pp, key_off, NBT, Lmax, list_mem, suf_mem, keys_mem, pairs_mem));
        } else {
            UG_KERNEL(ctx, (g1 ? "k_accumulate<G1>" : "k_accumulate<G2>"), acc_fn<<<blocks, 128, acc_smem, ctx.stream>>>(d_bases_v[m], dgs[m]->entries, dgs[m]->starts, cfg.NB, bk, pk, pp, key_off, NBT));
        }
    }

    uint32_t* kin = keys_a; P* pin = pts_a; uint32_t* kout = keys_b; P* pout = pts_b;
    uint32_t Kcur = K;
    while (true) {
        uint32_t g = (Kcur + SEG_BS - 1) / SEG_BS;
        int final_pass = (g == 1);
        UG_KERNEL(ctx, (g1 ? "k_seg_reduce<G1>" : "k_seg_reduce<G2>"), k_seg_reduce<F><<<g, SEG_BS, 0, ctx.stream>>>(kin, pin, Kcur, NBT, buckets, kout, pout, final_pass));
        if (final_pass) break;
        Kcur = 2 * g;
        std::swap(kin, kout);
        std::swap(pin, pout);
    }

    // bucket reduction: one level of chunked running sums, then bit-decomposed plain sums of the chunk sums
    const MsmReducePlan plan = msm_reduce_plan(cfg);
    const int nwin1 = cfg.nwin(), L = plan.points();
    const int nwin = nwin1 * nb;  // bucket sets of the whole batch: window w of MSM m is set m*nwin1 + w
    const uint32_t K0 = plan.K;
    P* S = (P*)ctx.workspace("msm.Sa", (size_t)nwin * K0 * sizeof(P));
    P* R = (P*)ctx.workspace("msm.R", (size_t)nwin * K0 * sizeof(P));
    P* Rpart = (P*)ctx.workspace("msm.Rpart", (size_t)nwin * 64 * 33 * sizeof(P));
    P* sums = (P*)ctx.workspace("msm.sums", (size_t)L * nwin * sizeof(P));
    using RsFn = typename RsSig<F>::Fn;
    const int rs_variant = env_int_msm(g1 ? "UGB_RS_G1" : "UGB_RS_G2", g1 ? RS_VARIANT_G1 : RS_VARIANT_G2);
    RsFn rs_fn = k_running_sum<F>;
    size_t rs_smem = 0;
    select_running_sum<F>(rs_variant, rs_fn, rs_smem);
    if (rs_smem > 48 * 1024) UG_CUDA_CHECK(cudaFuncSetAttribute(rs_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem));
    {
        dim3 grid((K0 + 127) / 128, nwin);
        UG_KERNEL(ctx, (g1 ? "k_running_sum<G1>" : "k_running_sum<G2>"), rs_fn<<<grid, 128, rs_smem, ctx.stream>>>(buckets, cfg.B, cfg.B, 1u << plan.mu, K0, S, R));
        // sums[0][w] = sum_k R[k]
        const uint32_t G = std::min<uint32_t>(64, (K0 + SUM_BS - 1) / SUM_BS);
        if (G > 1) {
            UG_KERNEL(ctx, (g1 ? "k_sum_points<G1>" : "k_sum_points<G2>"), k_sum_points<F><<<dim3(G, nwin), SUM_BS, 0, ctx.stream>>>(R, K0, K0, Rpart));
            UG_KERNEL(ctx, (g1 ? "k_sum_points<G1>" : "k_sum_points<G2>"), k_sum_points<F><<<dim3(1, nwin), SUM_BS, 0, ctx.stream>>>(Rpart, G, G, sums));
        } else {
            UG_KERNEL(ctx, (g1 ? "k_sum_points<G1>" : "k_sum_points<G2>"), k_sum_points<F><<<dim3(1, nwin), SUM_BS, 0, ctx.stream>>>(R, K0, K0, sums));
        }
        // sums[1+b][w] = sum of S[k] over 1 <= k < K0 with bit b of k set
        if (plan.nbits > 0) {
            const uint32_t Gb = std::min<uint32_t>(32, (K0 + 1023) / 1024);
            if (Gb > 1) {
                UG_KERNEL(ctx, (g1 ? "k_bit_sums<G1>" : "k_bit_sums<G2>"), k_bit_sums<F><<<dim3(Gb, nwin, plan.nbits), SUM_BS, 0, ctx.stream>>>(S, K0, K0, Rpart));
                UG_KERNEL(ctx, (g1 ? "k_sum_points<G1>" : "k_sum_points<G2>"), k_sum_points<F><<<dim3(1, nwin * plan.nbits), SUM_BS, 0, ctx.stream>>>(Rpart, Gb, Gb, sums + nwin));
            } else {
                UG_KERNEL(ctx, (g1 ? "k_bit_sums<G1>" : "k_bit_sums<G2>"), k_bit_sums<F><<<dim3(1, nwin, plan.nbits), SUM_BS, 0, ctx.stream>>>(S, K0, K0, sums + nwin));
            }
        }
    }
    // pinned result slot per MSM: its L*nwin1 partial sums ([level][window]), then the number of digit entries it processed (a
    // counter for the roofline arithmetic: every entry is one 4-byte index + one affine point gather + one mixed add)
    const size_t bytes = (size_t)L * nwin1 * sizeof(P);
    if (bytes > (size_t)64 * 32 * sizeof(P)) throw std::runtime_error("ultragroth_b200: MSM reduction plan too deep");
    for (int m = 0; m < nb; m++) {
        uint8_t* host = (uint8_t*)ctx.pinned("msm.sums." + slots[m], (size_t)64 * 32 * sizeof(P) + 16);
        UG_CUDA_CHECK(cudaMemcpy2DAsync(host, (size_t)nwin1 * sizeof(P), sums + (size_t)m * nwin1, (size_t)nwin * sizeof(P),
                                        (size_t)nwin1 * sizeof(P), (size_t)L, cudaMemcpyDeviceToHost, ctx.stream));
        UG_CUDA_CHECK(cudaMemcpyAsync(host + bytes, dgs[m]->starts + cfg.NB, 4, cudaMemcpyDeviceToHost, ctx.stream));
    }
    for (int m = 0; m < nb; m++) ctx.record("msm." + slots[m]);
}

void msm_g1_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot) {
    const G1Affine* b = (const G1Affine*)d_bases;
    const MsmDigits* d = &dg;
    msm_launch_impl<Fq>(ctx, &b, 1, &d, &slot);
}
void msm_g1_launch_batch(DeviceCtx& ctx, const void* const* d_bases, int count, const MsmDigits* const* dgs, const std::string* slots) {
    if (count < 1 || count > 8) throw std::invalid_argument("ultragroth_b200: MSM batch of 1..8 base arrays");
    msm_launch_impl<Fq>(ctx, (const G1Affine* const*)d_bases, count, dgs, slots);
}
void msm_g1_launch_batch(DeviceCtx& ctx, const void* const* d_bases, int count, const MsmDigits& dg, const std::string* slots) {
    const MsmDigits* d[8] = {&dg, &dg, &dg, &dg, &dg, &dg, &dg, &dg};
    msm_g1_launch_batch(ctx, d_bases, count, d, slots);
}
void msm_g2_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot) {
    // every G2 kernel is instantiated on Fq2c (Fq2's layout, CALLED products): an inlined Fq2 kernel is 200-500 KB of SASS
    const Affine<Fq2c>* b = (const Affine<Fq2c>*)d_bases;
    const MsmDigits* d = &dg;
    msm_launch_impl<Fq2c>(ctx, &b, 1, &d, &slot);
}

// ---- resident window tables ----------------------------------------------------------------------
template <class F>
static void* build_table_impl(DeviceCtx& ctx, const Affine<F>* d_bases, uint64_t n, const MsmConfig& cfg) {
    ctx.activate();
    Affine<F>* table = nullptr;
    const size_t level_bytes = (size_t)n * sizeof(Affine<F>);
    UG_CUDA_CHECK(cudaMalloc((void**)&table, std::max<size_t>(level_bytes * cfg.W, 16)));
    if (n == 0) return table;
    UG_CUDA_CHECK(cudaMemcpyAsync(table, d_bases, level_bytes, cudaMemcpyDeviceToDevice, ctx.stream));
    Jac<F>* tmp = (Jac<F>*)ctx.workspace("msm.table_tmp", (size_t)n * sizeof(Jac<F>));
    const int blocks = (int)std::min<uint64_t>((n + 127) / 128, (uint64_t)ctx.sm_count * 8);
    const int blocks_aff = (int)std::min<uint64_t>((n + 128 * AFF_RUN - 1) / (128 * AFF_RUN), (uint64_t)ctx.sm_count * 8);
    for (int w = 1; w < cfg.W; w++) {
        UG_KERNEL(ctx, "k_table_shift", k_table_shift<F><<<blocks, 128, 0, ctx.stream>>>(table + (size_t)(w - 1) * n, (uint32_t)n, cfg.c, tmp));
        UG_KERNEL(ctx, "k_batch_to_affine", k_batch_to_affine<F><<<std::max(blocks_aff, 1), 128, 0, ctx.stream>>>(tmp, (uint32_t)n, table + (size_t)w * n));
    }
    // no synchronisation: the caller overlaps the next array's upload with these kernels (everything is in stream order)
    return table;
}

MsmTable msm_build_table(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, int force_c) {
    MsmTable t;
    t.n = n;
    t.cfg = msm_pick_config(n, force_c, true);
    if ((uint64_t)n * t.cfg.W >= (1ull << 31)) throw std::runtime_error("ultragroth_b200: MSM table too large for 31-bit indices");
    t.d_table = group == 1 ? build_table_impl<Fq>(ctx, (const G1Affine*)d_bases, n, t.cfg)
                           : build_table_impl<Fq2c>(ctx, (const Affine<Fq2c>*)d_bases, n, t.cfg);
    return t;
}
}  // namespace ug
