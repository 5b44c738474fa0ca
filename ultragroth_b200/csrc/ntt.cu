// NTT driver: twiddle tables resident in HBM and the pass schedule of ntt.cuh.
#include "ntt.cuh"
#include "ntt.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace ug {

HFr hfr_from_u64(uint64_t x) {
    HFr v = HFr::zero();
    v.v[0] = x;
    return to_mont(v);
}

HFr hfr_pow(HFr base, uint64_t e) {
    HFr r = HFr::one();
    while (e) {
        if (e & 1) r = fmul(r, base);
        base = fsqr(base);
        e >>= 1;
    }
    return r;
}

HFr fr_root_of_unity(int k) {
    if (k < 0 || k > 28) throw std::runtime_error("ultragroth_b200: BN254 Fr has 2-adicity 28");
    // (r-1)/2^28 = 0x30644e72e131a029b85045b68181585d2833e84879b97091_43e1f593f / 2^28 as limbs
    static const uint64_t e[4] = {0x9b9709143e1f593full, 0x181585d2833e8487ull, 0x131a029b85045b68ull, 0x000000030644e72eull};
    HFr g = hfr_from_u64(5);
    HFr r = HFr::one();
    for (int i = 255; i >= 0; i--) {
        r = fsqr(r);
        if ((e[i >> 6] >> (i & 63)) & 1) r = fmul(r, g);
    }
    for (int i = 28; i > k; i--) r = fsqr(r);
    return r;
}

NttTables::~NttTables() {
    cudaSetDevice(ctx_.device);
    for (int i = 0; i < 2; i++) if (S_[i]) cudaFree(S_[i]);
    for (auto& kv : W_) cudaFree(kv.second);
    for (auto& kv : coset_) cudaFree(kv.second);
}

const Fr* NttTables::stage_roots(bool inverse) {
    int k = inverse ? 1 : 0;
    if (S_[k]) return S_[k];
    ctx_.activate();
    const uint32_t N = 1u << NTT_MAX_STAGE_BITS;
    std::vector<HFr> tab(N);
    tab[0] = HFr::one();  // unused slot
    for (int u = 0; u < NTT_MAX_STAGE_BITS; u++) {
        HFr w = fr_root_of_unity(u + 1);
        if (inverse) w = finv(w);
        HFr cur = HFr::one();
        for (uint32_t j = 0; j < (1u << u); j++) {
            tab[(1u << u) + j] = cur;
            cur = fmul(cur, w);
        }
    }
    UG_CUDA_CHECK(cudaMalloc((void**)&S_[k], N * sizeof(Fr)));
    UG_CUDA_CHECK(cudaMemcpyAsync(S_[k], tab.data(), N * sizeof(Fr), cudaMemcpyHostToDevice, ctx_.stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx_.stream));
    return S_[k];
}

void NttTables::fill_powers(Fr* d_out, int logn, const HFr& base, const HFr& factor, bool bitrev) {
    ctx_.activate();
    const uint32_t n = 1u << logn;
    const int lo_bits = std::min(logn, 10);
    const uint32_t nlo = 1u << lo_bits, nhi = n >> lo_bits;
    std::vector<HFr> lo(nlo), hi(nhi);
    HFr cur = factor;
    for (uint32_t j = 0; j < nlo; j++) { lo[j] = cur; cur = fmul(cur, base); }
    HFr step = hfr_pow(base, nlo);
    cur = HFr::one();
    for (uint32_t i = 0; i < nhi; i++) { hi[i] = cur; cur = fmul(cur, step); }
    Fr* d_lo = (Fr*)ctx_.workspace("ntt.pow_lo", nlo * sizeof(Fr));
    Fr* d_hi = (Fr*)ctx_.workspace("ntt.pow_hi", (size_t)nhi * sizeof(Fr));
    UG_CUDA_CHECK(cudaMemcpyAsync(d_lo, lo.data(), nlo * sizeof(Fr), cudaMemcpyHostToDevice, ctx_.stream));
    UG_CUDA_CHECK(cudaMemcpyAsync(d_hi, hi.data(), (size_t)nhi * sizeof(Fr), cudaMemcpyHostToDevice, ctx_.stream));
    int blocks = (int)std::min<uint64_t>(((uint64_t)n + 255) / 256, (uint64_t)ctx_.sm_count * 8);
    UG_KERNEL(ctx_, "k_fill_powers", k_fill_powers<<<blocks, 256, 0, ctx_.stream>>>(d_out, n, logn, d_hi, d_lo, lo_bits, bitrev ? 1 : 0));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx_.stream));  // host vectors go out of scope
}

const Fr* NttTables::full_roots(int logn) {
    auto it = W_.find(logn);
    if (it != W_.end()) return it->second;
    ctx_.activate();
    Fr* d = nullptr;
    UG_CUDA_CHECK(cudaMalloc((void**)&d, ((size_t)1 << logn) * sizeof(Fr)));
    fill_powers(d, logn, fr_root_of_unity(logn), HFr::one(), false);
    W_[logn] = d;
    return d;
}

const Fr* NttTables::coset_inv_bitrev(int logn) {
    auto it = coset_.find(logn);
    if (it != coset_.end()) return it->second;
    ctx_.activate();
    Fr* d = nullptr;
    UG_CUDA_CHECK(cudaMalloc((void**)&d, ((size_t)1 << logn) * sizeof(Fr)));
    HFr ninv = finv(hfr_from_u64(1ull << logn));
    fill_powers(d, logn, fr_root_of_unity(logn + 1), ninv, true);
    coset_[logn] = d;
    return d;
}

// ---- pass schedule ------------------------------------------------------------------------------
struct PassShape { int lo, T, logC; };

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

// Splits log2(n) into passes of at most `maxT` stage bits over tiles of at most 2^logE elements.
static std::vector<PassShape> ntt_schedule(int logn) {
    const int logE = std::min(std::max(env_int("UGB_NTT_LOGE", 11), 6), 12);  // tile elements (2^11 = 64 KiB: 3 CTAs per SM overlap load/store with butterflies)
    std::vector<PassShape> passes;  // ordered from the low bits up
    if (logn <= logE) {
        passes.push_back({0, logn, 0});
        return passes;
    }
    int npass = (logn + logE - 1) / logE;
    if (npass < 2) npass = 2;
    // strided passes want >= 4 contiguous columns (128 B); if 2 passes cannot give that, use 3
    const int minLogC = std::min(std::max(env_int("UGB_NTT_MIN_LOGC", 1), 0), 4);
    // ... and no more than 2^maxLogC: once the pass count is fixed, smaller tiles (more CTAs per SM) hide barriers better
    const int maxLogC = std::min(std::max(env_int("UGB_NTT_MAX_LOGC", 1), minLogC), 11);
    while (true) {
        int base = logn / npass, rem = logn % npass;
        bool ok = true;
        std::vector<PassShape> out;
        int lo = 0;
        for (int i = 0; i < npass; i++) {
            int T = base + (i < rem ? 1 : 0);  // low passes take the larger share (they are contiguous)
            int logC = (i == 0) ? 0 : std::min(lo, logE - T);
            if (i > 0 && logC < std::min(minLogC, lo)) ok = false;
            logC = std::min(logC, maxLogC);
            out.push_back({lo, T, logC});
            lo += T;
        }
        if (ok || npass >= 4) { passes = out; break; }
        npass++;
    }
    return passes;
}

void ntt_run(DeviceCtx& ctx, NttTables& tabs, Fr* const* d_vecs, int count, int logn, NttOrder order, bool inverse,
             const Fr* d_scale, const HFr* cst) {
    if (count < 1 || count > 3) throw std::runtime_error("ultragroth_b200: ntt_run takes 1..3 vectors");
    if (logn < 0 || logn > 27) throw std::runtime_error("ultragroth_b200: NTT size out of range");
    ctx.activate();
    // the opt-in for > 48 KiB of dynamic shared memory is a per-DEVICE function attribute: remembered per context, not per process
    if (!ctx.flag("ntt.smem_optin")) {
        UG_CUDA_CHECK(cudaFuncSetAttribute(k_ntt_pass<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        UG_CUDA_CHECK(cudaFuncSetAttribute(k_ntt_pass<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        ctx.set_flag("ntt.smem_optin");
    }
    std::vector<PassShape> sched;
    if (logn == 0) sched.push_back({0, 0, 0});
    else sched = ntt_schedule(logn);
    const bool dif = (order == NTT_NATURAL_TO_BITREV);
    if (dif) std::reverse(sched.begin(), sched.end());  // DIF starts at the high bits
    const Fr* S = tabs.stage_roots(inverse);
    const Fr* W = sched.size() > 1 ? tabs.full_roots(logn) : nullptr;
    const int threads_env = env_int("UGB_NTT_THREADS", 0);  // 0: one thread per 8 tile elements, 128..256
    for (size_t k = 0; k < sched.size(); k++) {
        const PassShape& ps = sched[k];
        NttPass a;
        memset(&a, 0, sizeof(a));
        for (int i = 0; i < count; i++) a.data[i] = d_vecs[i];
        a.S = S;
        a.W = W;
        a.n = 1u << logn;
        a.logn = logn;
        a.lo = ps.lo;
        a.T = ps.T;
        a.logC = ps.logC;
        a.inverse = inverse ? 1 : 0;
        const bool last = (k + 1 == sched.size());
        // the fused scale rides on the contiguous (lo == 0) pass: last for DIF, first for DIT
        a.scale = (ps.lo == 0) ? d_scale : nullptr;
        a.use_const = (cst && last) ? 1 : 0;
        if (a.use_const) memcpy(&a.cst, cst, sizeof(Fr));
        const uint32_t E = 1u << (ps.T + ps.logC);
        const uint32_t tiles = (1u << logn) / E;
        const size_t smem = (size_t)E * 32;
        const int threads = threads_env > 0 ? std::min(std::max(threads_env, 32), 1024) : (int)std::min<uint32_t>(std::max<uint32_t>(E / 8, 128), 256);
        const int bt = (int)std::min<uint32_t>(threads, std::max<uint32_t>(32, E / 2));
        dim3 grid(tiles, count);
        const char* kname = dif ? (ps.lo ? "k_ntt_pass<dif,strided>" : "k_ntt_pass<dif,contig>")
                                : (ps.lo ? "k_ntt_pass<dit,strided>" : "k_ntt_pass<dit,contig>");
        // opt-in TMA variant of the contiguous pass (UGB_NTT_TMA=1): persistent CTAs, bulk copies in and out, mbarrier completion
        static const int use_tma = env_int("UGB_NTT_TMA", 0);
        if (use_tma && ps.lo == 0 && ps.logC == 0 && ps.T >= 6) {
            if (!ctx.flag("ntt.bulk_optin")) {
                UG_CUDA_CHECK(cudaFuncSetAttribute(k_ntt_pass_bulk<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                UG_CUDA_CHECK(cudaFuncSetAttribute(k_ntt_pass_bulk<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                ctx.set_flag("ntt.bulk_optin");
            }
            const uint32_t total = tiles * (uint32_t)count;
            const size_t bsmem = 2 * (size_t)E * 32 + 64;
            const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(4, (200 * 1024) / bsmem));
            const int bthreads = std::min(1024, std::max(128, env_int("UGB_NTT_TMA_THREADS", (int)std::min<uint32_t>(512, E / 4))));
            const uint32_t bgrid = std::min<uint32_t>(total, (uint32_t)(ctx.sm_count * per_sm));
            const char* bname = dif ? "k_ntt_pass_bulk<dif>" : "k_ntt_pass_bulk<dit>";
            if (dif) UG_KERNEL(ctx, bname, k_ntt_pass_bulk<true><<<bgrid, bthreads, bsmem, ctx.stream>>>(a, tiles, total));
            else UG_KERNEL(ctx, bname, k_ntt_pass_bulk<false><<<bgrid, bthreads, bsmem, ctx.stream>>>(a, tiles, total));
            continue;
        }
        if (dif) UG_KERNEL(ctx, kname, k_ntt_pass<true><<<grid, bt, smem, ctx.stream>>>(a));
        else UG_KERNEL(ctx, kname, k_ntt_pass<false><<<grid, bt, smem, ctx.stream>>>(a));
    }
}

void ntt_natural(DeviceCtx& ctx, NttTables& tabs, Fr* d_vec, int logn, bool inverse) {
    Fr* vecs[1] = {d_vec};
    HFr ninv;
    if (inverse) ninv = finv(hfr_from_u64(1ull << logn));
    ntt_run(ctx, tabs, vecs, 1, logn, NTT_NATURAL_TO_BITREV, inverse, nullptr, inverse ? &ninv : nullptr);
    if (logn >= 2) {
        const uint32_t n = 1u << logn;
        int blocks = (int)std::min<uint64_t>(((uint64_t)n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_bitrev_swap", k_bitrev_swap<<<blocks, 256, 0, ctx.stream>>>(d_vec, n, logn));
    }
}

}  // namespace ug
