// C ABI of include/ultragroth_b200.h: thin, exception-free wrappers over the C++ host layer.
#include "../../include/ultragroth_b200.h"

#include <cstring>
#include <string>

#include "device.hpp"
#include "fieldops.hpp"
#include "msm.hpp"
#include "ntt.hpp"

#include <cstdlib>
#include <vector>
namespace ug {
double bench_int_pipe(DeviceCtx& ctx, int kind);
void synth_build(DeviceCtx& ctx, int ultra, int logn, uint64_t seed, int lookup_bits, uint32_t n_chunks, std::vector<uint8_t>& zkey,
                 std::vector<uint8_t>& wtns, std::string& vk_json);
void fixed_base_mul_host(DeviceCtx& ctx, int group, const void* scalars_raw, uint64_t n, void* out);
}

using namespace ug;

struct ugb_ctx {
    DeviceCtx dev;
    NttTables tabs;
    explicit ugb_ctx(int d) : dev(d), tabs(dev) {}
};

static thread_local std::string g_last_error;
namespace ug { void set_last_error(const std::string& msg) { g_last_error = msg; } }

#define UGB_TRY try {
#define UGB_CATCH                                                     \
    }                                                                 \
    catch (const std::exception& e) { g_last_error = e.what(); return UGB_ERROR; } \
    catch (...) { g_last_error = "unknown error"; return UGB_ERROR; } \
    return UGB_OK;

extern "C" {

const char* ugb_last_error(void) { return g_last_error.c_str(); }

int ugb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int ugb_ctx_create(int device, ugb_ctx** out) {
    UGB_TRY
    if (!out) throw std::invalid_argument("null ctx pointer");
    *out = new ugb_ctx(device);
    UGB_CATCH
}
void ugb_ctx_destroy(ugb_ctx* ctx) { delete ctx; }
int ugb_sync(ugb_ctx* ctx) { UGB_TRY ctx->dev.sync(); UGB_CATCH }
uint64_t ugb_kernel_launches(ugb_ctx* ctx) { return ctx->dev.kernel_launches; }
void* ugb_stream(ugb_ctx* ctx) { return (void*)ctx->dev.stream; }

int ugb_malloc(ugb_ctx* ctx, uint64_t bytes, void** dptr) {
    UGB_TRY
    ctx->dev.activate();
    UG_CUDA_CHECK(cudaMalloc(dptr, bytes ? bytes : 16));
    UGB_CATCH
}
int ugb_free(ugb_ctx* ctx, void* dptr) {
    UGB_TRY
    ctx->dev.activate();
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx->dev.stream));
    UG_CUDA_CHECK(cudaFree(dptr));
    UGB_CATCH
}
int ugb_memcpy_h2d(ugb_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    UGB_TRY
    ctx->dev.activate();
    UG_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->dev.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx->dev.stream));
    UGB_CATCH
}
int ugb_memcpy_d2h(ugb_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    UGB_TRY
    ctx->dev.activate();
    UG_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->dev.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx->dev.stream));
    UGB_CATCH
}

int ugb_field_vec_op(ugb_ctx* ctx, int field, int op, const void* d_a, const void* d_b, void* d_out, uint64_t n) {
    UGB_TRY
    field_vec_op(ctx->dev, field, op, d_a, d_b, d_out, n);
    UGB_CATCH
}

int ugb_msm_g1_xyzz(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
                    int window_bits, void* h_out) {
    UGB_TRY
    MsmDigits dg;
    msm_sort_digits(ctx->dev, (const uint8_t*)d_scalars, d_index, n, window_bits, "capi", dg);
    HG1XYZZ r;
    msm_accumulate_g1(ctx->dev, d_bases, dg, &r);
    memcpy(h_out, &r, sizeof(r));
    UGB_CATCH
}
int ugb_msm_g2_xyzz(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
                    int window_bits, void* h_out) {
    UGB_TRY
    MsmDigits dg;
    msm_sort_digits(ctx->dev, (const uint8_t*)d_scalars, d_index, n, window_bits, "capi", dg);
    HG2XYZZ r;
    msm_accumulate_g2(ctx->dev, d_bases, dg, &r);
    memcpy(h_out, &r, sizeof(r));
    UGB_CATCH
}
int ugb_msm_g1(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
               int window_bits, void* h_out) {
    HG1XYZZ r;
    int rc = ugb_msm_g1_xyzz(ctx, d_bases, d_scalars, d_index, n, window_bits, &r);
    if (rc) return rc;
    HG1Affine a = pt_to_affine(r);
    memcpy(h_out, &a, sizeof(a));
    return UGB_OK;
}
int ugb_msm_g2(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
               int window_bits, void* h_out) {
    HG2XYZZ r;
    int rc = ugb_msm_g2_xyzz(ctx, d_bases, d_scalars, d_index, n, window_bits, &r);
    if (rc) return rc;
    HG2Affine a = pt_to_affine(r);
    memcpy(h_out, &a, sizeof(a));
    return UGB_OK;
}
int ugb_msm_g1_noalloc(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, uint64_t n, void* h_out) {
    UGB_TRY
    if (!ctx) throw std::invalid_argument("null ctx");
    HG1XYZZ r;
    msm_g1_noalloc(ctx->dev, d_bases, d_scalars, n, &r);
    HG1Affine a = pt_to_affine(r);
    memcpy(h_out, &a, sizeof(a));
    UGB_CATCH
}
uint64_t ugb_msm_noalloc_max_points(void) { return MSM_NOALLOC_MAX_POINTS; }
int ugb_msm_plan(uint64_t n, int resident_table, int* c, int* windows, int* chunk_log2) {
    UGB_TRY
    const MsmConfig cfg = msm_pick_config(n, 0, resident_table != 0);
    const MsmReducePlan plan = msm_reduce_plan(cfg);
    if (c) *c = cfg.c;
    if (windows) *windows = cfg.W;
    if (chunk_log2) *chunk_log2 = plan.mu;
    UGB_CATCH
}
int ugb_g1_sum_xyzz_to_affine(const void* h_partials, uint64_t count, void* h_out) {
    UGB_TRY
    HG1XYZZ acc = HG1XYZZ::inf();
    for (uint64_t i = 0; i < count; i++) { HG1XYZZ p; memcpy(&p, (const char*)h_partials + i * sizeof(p), sizeof(p)); pt_add(acc, p); }
    HG1Affine a = pt_to_affine(acc);
    memcpy(h_out, &a, sizeof(a));
    UGB_CATCH
}
int ugb_g2_sum_xyzz_to_affine(const void* h_partials, uint64_t count, void* h_out) {
    UGB_TRY
    HG2XYZZ acc = HG2XYZZ::inf();
    for (uint64_t i = 0; i < count; i++) { HG2XYZZ p; memcpy(&p, (const char*)h_partials + i * sizeof(p), sizeof(p)); pt_add(acc, p); }
    HG2Affine a = pt_to_affine(acc);
    memcpy(h_out, &a, sizeof(a));
    UGB_CATCH
}

int ugb_ntt(ugb_ctx* ctx, void* d_data, int logn, int inverse) {
    UGB_TRY
    ntt_natural(ctx->dev, ctx->tabs, (Fr*)d_data, logn, inverse != 0);
    UGB_CATCH
}
int ugb_coset_transform(ugb_ctx* ctx, void* const* d_vecs, int count, int logn) {
    UGB_TRY
    if (count < 1 || count > 3) throw std::invalid_argument("count must be 1..3");
    Fr* v[3] = {nullptr, nullptr, nullptr};
    for (int i = 0; i < count; i++) v[i] = (Fr*)d_vecs[i];
    ntt_run(ctx->dev, ctx->tabs, v, count, logn, NTT_NATURAL_TO_BITREV, true, ctx->tabs.coset_inv_bitrev(logn), nullptr);
    ntt_run(ctx->dev, ctx->tabs, v, count, logn, NTT_BITREV_TO_NATURAL, false, nullptr, nullptr);
    UGB_CATCH
}
int ugb_bench_int_pipe(ugb_ctx* ctx, int kind, double* ops_per_s) {
    UGB_TRY
    *ops_per_s = bench_int_pipe(ctx->dev, kind);
    UGB_CATCH
}

int ugb_synth_build(ugb_ctx* ctx, int ultra, int logn, uint64_t seed, int lookup_bits, uint32_t n_chunks, void** zkey, uint64_t* zkey_size,
                    void** wtns, uint64_t* wtns_size, char** vk_json) {
    UGB_TRY
    std::vector<uint8_t> zk, wt;
    std::string vk;
    synth_build(ctx->dev, ultra, logn, seed, lookup_bits, n_chunks, zk, wt, vk);
    *zkey = malloc(zk.size());
    *wtns = malloc(wt.size());
    *vk_json = (char*)malloc(vk.size() + 1);
    if (!*zkey || !*wtns || !*vk_json) throw std::bad_alloc();
    memcpy(*zkey, zk.data(), zk.size());
    memcpy(*wtns, wt.data(), wt.size());
    memcpy(*vk_json, vk.c_str(), vk.size() + 1);
    *zkey_size = zk.size();
    *wtns_size = wt.size();
    UGB_CATCH
}
void ugb_host_free(void* p) { free(p); }
int ugb_fixed_base_mul(ugb_ctx* ctx, int group, const void* h_scalars, uint64_t n, void* h_out) {
    UGB_TRY
    if (group != 1 && group != 2) throw std::invalid_argument("group must be 1 or 2");
    fixed_base_mul_host(ctx->dev, group, h_scalars, n, h_out);
    UGB_CATCH
}

}  // extern "C"
