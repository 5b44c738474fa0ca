// Device side of the Groth16 / UltraGroth prover: the data-parallel loops of
// Groth16::Prover::prove (src/groth16.cpp:48-154) and UltraGroth's rounds (src/ultra_groth.cpp:62-106,
// 161-335) as sm_100a kernels over HBM-resident zkey data.
#include "prover_dev.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <thread>

#include "curve.cuh"
#include "memops.cuh"

namespace ug {

// ---- a[c] = sum_{A row c} w[s]*coef, b[c] likewise, c[c] = a[c]*b[c]   (src/groth16.cpp:70-108) ----
// The reference scatters coefficient records under 1024 mutex stripes; here the records were
// bucketed by (matrix, row) at upload time, so one thread owns one constraint row: no atomics.
// coef holds value*R^2 (zkey convention), so fp_mul(raw witness, coef) is the Montgomery product.
__global__ void __launch_bounds__(256) k_spmv_abc(const uint32_t* __restrict__ row_ptr, const uint32_t* __restrict__ sig,
                                                  const Fr* __restrict__ coef, const Fr* __restrict__ w, uint32_t n,
                                                  Fr* __restrict__ a, Fr* __restrict__ b, Fr* __restrict__ c) {
    for (uint32_t row = blockIdx.x * blockDim.x + threadIdx.x; row < n; row += gridDim.x * blockDim.x) {
        Fr acc[2];
#pragma unroll
        for (int m = 0; m < 2; m++) {
            acc[m] = Fr::zero();
            uint32_t lo = row_ptr[m * n + row], hi = row_ptr[m * n + row + 1];
            for (uint32_t k = lo; k < hi; k++) {
                Fr wv = ld_nc(&w[sig[k]]);
                acc[m] = fp_add(acc[m], fp_mul(wv, ld_nc(&coef[k])));
            }
        }
        st_g(&a[row], acc[0]);
        st_g(&b[row], acc[1]);
        st_g(&c[row], fp_mul(acc[0], acc[1]));
    }
}

// Rows with more than SPMV_LONG_ROW terms (UltraGroth's "sum of all lookup terms" constraint has one term per
// looked-up signal) get a whole CTA each: strided partial sums, shared-memory tree, then c = a*b is refreshed.
constexpr uint32_t SPMV_LONG_ROW = 48;
constexpr int SPMV_LONG_BS = 512;
struct LongRow { uint32_t slot /* m*n + row */, lo, hi; };

// pass 1: one CTA per SEGMENT (<= SPMV_SEG terms) of a long row -> partial[seg]; pass 2 (k_spmv_long_finish): one CTA per
// long row sums its partials.  A 2^20-term lookup row is 128 CTAs instead of one.
constexpr uint32_t SPMV_SEG = 8192;
__device__ __forceinline__ Fr block_sum_fr(Fr acc, Fr* sh) {
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int d = SPMV_LONG_BS / 2; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = fp_add(sh[threadIdx.x], sh[threadIdx.x + d]);
        __syncthreads();
    }
    return sh[0];
}
__global__ void __launch_bounds__(SPMV_LONG_BS) k_spmv_long(const LongRow* __restrict__ segs, const uint32_t* __restrict__ sig,
                                                            const Fr* __restrict__ coef, const Fr* __restrict__ w,
                                                            Fr* __restrict__ partial) {
    __shared__ Fr sh[SPMV_LONG_BS];
    const LongRow r = segs[blockIdx.x];
    Fr acc = Fr::zero();
    for (uint32_t k = r.lo + threadIdx.x; k < r.hi; k += SPMV_LONG_BS) acc = fp_add(acc, fp_mul(ld_nc(&w[sig[k]]), ld_nc(&coef[k])));
    Fr tot = block_sum_fr(acc, sh);
    if (threadIdx.x == 0) st_g(&partial[blockIdx.x], tot);
}
// rows[i].lo..hi is the SEGMENT range of long row i
__global__ void __launch_bounds__(SPMV_LONG_BS) k_spmv_long_finish(const LongRow* __restrict__ rows, const Fr* __restrict__ partial,
                                                                   uint32_t n, Fr* __restrict__ a, Fr* __restrict__ b) {
    __shared__ Fr sh[SPMV_LONG_BS];
    const LongRow r = rows[blockIdx.x];
    Fr acc = Fr::zero();
    for (uint32_t k = r.lo + threadIdx.x; k < r.hi; k += SPMV_LONG_BS) acc = fp_add(acc, ld_g(&partial[k]));
    Fr tot = block_sum_fr(acc, sh);
    if (threadIdx.x == 0) {
        if (r.slot < n) st_g(&a[r.slot], tot);
        else st_g(&b[r.slot - n], tot);
    }
}
__global__ void k_spmv_fix_c(const LongRow* __restrict__ rows, uint32_t count, uint32_t n, const Fr* __restrict__ a,
                             const Fr* __restrict__ b, Fr* __restrict__ c) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint32_t row = rows[i].slot % n;
    st_g(&c[row], fp_mul(ld_g(&a[row]), ld_g(&b[row])));
}

// h[i] = fromMontgomery(a[i]*b[i] - c[i])   (src/groth16.cpp:142-148), written over a
__global__ void __launch_bounds__(256) k_h_final(Fr* __restrict__ a, const Fr* __restrict__ b, const Fr* __restrict__ c, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        Fr x = fp_mul(ld_g(&a[i]), ld_g(&b[i]));
        x = fp_sub(x, ld_g(&c[i]));
        st_g(&a[i], fp_from_mont(x));
    }
}

// ---- compute_lookup (src/ultra_groth.cpp:62-106) ----
// inv2[i] = fromMont((i + rand)^-1), prod[i] = fromMont(freq[i] * (i + rand)^-1): the reference runs one
// GMP mpz_invert per table entry; here every thread inverts a run of LOOKUP_RUN entries with
// Montgomery's trick (one Fermat inversion + 3 products per entry).
constexpr int LOOKUP_RUN = 8;

__global__ void __launch_bounds__(128) k_lookup_inverses(const uint32_t* __restrict__ freqs, uint32_t lookup_size, Fr rand_mont,
                                                         Fr* __restrict__ inv2, Fr* __restrict__ prod) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t base = t * LOOKUP_RUN;
    if (base >= lookup_size) return;
    const uint32_t cnt = min((uint32_t)LOOKUP_RUN, lookup_size - base);
    Fr val[LOOKUP_RUN], pre[LOOKUP_RUN];
    Fr acc = Fr::one();
    for (uint32_t k = 0; k < cnt; k++) {
        Fr iv = Fr::zero();
        iv.v[0] = base + k;
        val[k] = fp_add(fp_to_mont(iv), rand_mont);
        pre[k] = acc;
        // a zero (i + rand == 0) has no inverse; mpz_invert leaves 0 there, keep it out of the product
        if (!val[k].is_zero()) acc = fp_mul(acc, val[k]);
    }
    Fr inv = fp_inv(acc);
    for (int k = (int)cnt - 1; k >= 0; k--) {
        Fr r = Fr::zero();
        if (!val[k].is_zero()) {
            r = fp_mul(inv, pre[k]);
            inv = fp_mul(inv, val[k]);
        }
        // the reference converts the uint32_t frequency to `int` (build/fr.hpp:251 -> RawFr::set(int), build/fr.cpp:209-223):
        // a frequency >= 2^31 is the NEGATIVE number f - 2^32, i.e. -(2^32 - f) mod r
        const uint32_t fv = freqs[base + k];
        Fr f = Fr::zero();
        f.v[0] = (fv >> 31) ? 0u - fv : fv;
        f = fp_to_mont(f);
        if (fv >> 31) f = fp_neg(f);
        st_g(&inv2[base + k], fp_from_mont(r));
        st_g(&prod[base + k], fp_from_mont(fp_mul(f, r)));
    }
}

// signals[wtns_idx[k]] = push_vector[push_idx[k]], push_vector = [rand | inv2[chunks[*]] | inv2 | prod]
// The reference trusts the four index arrays (out-of-range entries are wild reads / writes there); here every index is checked where
// it is used and a violation sets *err (1: witness index, 2: push index, 4: chunk value) instead of touching memory; the host turns
// a non-zero flag into a std::range_error once the proof's results arrive.
__global__ void __launch_bounds__(256) k_lookup_push(const uint32_t* __restrict__ wtns_idx, const uint32_t* __restrict__ push_idx,
                                                     uint32_t count, const uint32_t* __restrict__ chunks, uint32_t chunks_total,
                                                     uint32_t lookup_size, uint32_t n_vars, Fr rand_raw, const Fr* __restrict__ inv2,
                                                     const Fr* __restrict__ prod, Fr* __restrict__ signals, uint32_t* __restrict__ err) {
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const uint32_t p = push_idx[k], wi = wtns_idx[k];
        if (wi >= n_vars) { atomicOr(err, 1u); continue; }
        Fr v;
        if (p == 0) v = rand_raw;
        else if (p <= chunks_total) {
            const uint32_t ch = chunks[p - 1];
            if (ch >= lookup_size) { atomicOr(err, 4u); continue; }
            v = ld_g(&inv2[ch]);
        } else if ((uint64_t)p <= (uint64_t)chunks_total + lookup_size) v = ld_g(&inv2[p - 1 - chunks_total]);
        else if ((uint64_t)p <= (uint64_t)chunks_total + 2ull * lookup_size) v = ld_g(&prod[p - 1 - chunks_total - lookup_size]);
        else { atomicOr(err, 2u); continue; }
        st_g(&signals[wi], v);
    }
}

// ---- upload helpers -------------------------------------------------------------------------------
static void* dev_alloc(size_t bytes) {
    void* p = nullptr;
    UG_CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
    return p;
}

static void* upload(DeviceCtx& ctx, const void* src, size_t bytes) {
    void* p = dev_alloc(bytes);
    if (bytes) UG_CUDA_CHECK(cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, ctx.stream));
    return p;
}

static int ilog2_exact(uint32_t n) {
    int k = 0;
    while ((1u << k) < n) k++;
    if ((1u << k) != n) throw std::invalid_argument("zkey domain size is not a power of two");
    return k;
}

ZkeyData zkey_data_from_file(const BinFile& zkey, const ZkeyHeader& h) {
    ZkeyData z;
    z.ultra = h.ultra;
    z.nVars = h.nVars; z.nPublic = h.nPublic; z.domainSize = h.domainSize;
    if (h.n8q != 32 || h.n8r != 32) throw std::invalid_argument("zkey field size not supported");
    auto sec = [&](uint32_t id, const uint8_t*& p, uint64_t& bytes) { p = zkey.data(id); bytes = zkey.size(id); };
    {
        const uint64_t sec_size = zkey.size(4);
        if (sec_size < 4) throw std::range_error("zkey section too short: coefs");
        z.coefs = zkey.data(4) + 4;
        z.coefs_bytes = sec_size - 4;
        z.nCoefs = std::min<uint64_t>(h.nCoefs, (sec_size - 4) / 44);
    }
    sec(5, z.A, z.A_bytes); sec(6, z.B1, z.B1_bytes); sec(7, z.B2, z.B2_bytes);
    if (!h.ultra) {
        sec(8, z.C, z.C_bytes); sec(9, z.H, z.H_bytes);
    } else {
        z.n_c1 = h.num_indexes_c1; z.n_c2 = h.num_indexes_c2;
        sec(8, z.C1, z.C1_bytes); sec(9, z.C, z.C_bytes); sec(10, z.idx_c1, z.idx_c1_bytes); sec(11, z.idx_c2, z.idx_c2_bytes); sec(12, z.H, z.H_bytes);
    }
    return z;
}

// out[idx[i] / world] = in[i] for the signals idx[i] this shard owns (idx[i] % world == rank): places the final-round C points at
// their witness slots (world = 1: every slot)
__global__ void __launch_bounds__(256) k_scatter_affine(const G1Affine* __restrict__ in, const uint32_t* __restrict__ idx, uint32_t count, uint32_t world,
                                                        uint32_t rank, G1Affine* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint32_t s = idx[i];
        if (s % world == rank) st_g(&out[s / world], ld_nc(&in[i]));
    }
}

// out[j] = first + j * stride: the signal of the j-th resident witness-paired point of a shard
__global__ void __launch_bounds__(256) k_iota_strided(uint32_t* __restrict__ out, uint32_t count, uint32_t first, uint32_t stride) {
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) out[j] = first + j * stride;
}

namespace {
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// Packed coefficient records -> CSR by (matrix, row), built by `nthreads` host threads.  Thread t owns the slots (m*n + c) of one
// contiguous range and scans ALL records for them, so the layout is deterministic (record order inside a row is kept) and no
// two threads ever write the same counter.
struct CoefCsr {
    std::vector<uint32_t> rp, sig;
    std::vector<uint8_t> val;
    std::vector<LongRow> longs, segs;
};
void build_csr(const ZkeyData& zk, CoefCsr& out) {
    const uint32_t n = zk.domainSize, nVars = zk.nVars;
    const uint64_t cnt = zk.nCoefs;
    const uint8_t* rec = zk.coefs;
    const size_t slots = (size_t)2 * n;
    unsigned hw = std::thread::hardware_concurrency();
    const int nthreads = (int)std::max<uint64_t>(1, std::min<uint64_t>({(uint64_t)(hw ? hw : 4), 32, cnt / 65536 + 1}));
    std::vector<uint32_t> len(slots, 0);
    std::vector<std::string> errs(nthreads);
    auto run = [&](auto fn) {
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; t++) th.emplace_back(fn, t);
        fn(0);
        for (auto& x : th) x.join();
        for (auto& e : errs) if (!e.empty()) throw std::range_error(e);
    };
    auto range = [&](int t, size_t& lo, size_t& hi) { lo = slots * t / nthreads; hi = slots * (t + 1) / nthreads; };
    run([&](int t) {
        size_t lo, hi;
        range(t, lo, hi);
        for (uint64_t k = 0; k < cnt; k++) {
            uint32_t m, c, s;
            memcpy(&m, rec + k * 44, 4); memcpy(&c, rec + k * 44 + 4, 4); memcpy(&s, rec + k * 44 + 8, 4);
            if (m > 1 || c >= n || s >= nVars) { if (t == 0) errs[0] = "zkey coefficient record out of range"; continue; }
            const size_t slot = (size_t)m * n + c;
            if (slot >= lo && slot < hi) len[slot]++;
        }
    });
    // short rows form the CSR proper; long rows keep an empty CSR row and live after `short_total`
    out.rp.assign(slots + 1, 0);
    for (size_t i = 0; i < slots; i++) out.rp[i + 1] = out.rp[i] + (len[i] > SPMV_LONG_ROW ? 0 : len[i]);
    uint32_t tail = out.rp[slots];
    std::vector<uint32_t> cur(out.rp.begin(), out.rp.end() - 1);
    for (size_t i = 0; i < slots; i++) {
        if (len[i] > SPMV_LONG_ROW) {
            const uint32_t seg0 = (uint32_t)out.segs.size();
            for (uint32_t o = 0; o < len[i]; o += SPMV_SEG) out.segs.push_back(LongRow{(uint32_t)i, tail + o, tail + std::min(len[i], o + SPMV_SEG)});
            out.longs.push_back(LongRow{(uint32_t)i, seg0, (uint32_t)out.segs.size()});
            cur[i] = tail;
            tail += len[i];
        }
    }
    out.sig.resize(cnt);
    out.val.resize(cnt * 32);
    run([&](int t) {
        size_t lo, hi;
        range(t, lo, hi);
        for (uint64_t k = 0; k < cnt; k++) {
            uint32_t m, c, s;
            memcpy(&m, rec + k * 44, 4); memcpy(&c, rec + k * 44 + 4, 4); memcpy(&s, rec + k * 44 + 8, 4);
            const size_t slot = (size_t)m * n + c;
            if (slot < lo || slot >= hi) continue;
            const uint32_t pos = cur[slot]++;
            out.sig[pos] = s;
            memcpy(&out.val[(size_t)pos * 32], rec + k * 44 + 12, 32);
        }
    });
}
}  // namespace

ProverDevice::ProverDevice(int device, const ZkeyData& zk, uint32_t shard_rank, uint32_t shard_world, bool tables)
    : nVars(zk.nVars), nPublic(zk.nPublic), domainSize(zk.domainSize), ultra_(zk.ultra) {
    if (shard_world == 0 || shard_rank >= shard_world) throw std::invalid_argument("bad shard rank / world");
    logn = ilog2_exact(domainSize);
    if (nPublic + 1 > nVars) throw std::invalid_argument("zkey header is inconsistent (nPublic >= nVars)");
    ctx_.reset(new DeviceCtx(device));
    // every device buffer below is a raw cudaMalloc pointer: a throw half way through must not leak them (a server that
    // retries *_prover_create on bad or oversized zkeys would exhaust HBM)
    try {
        construct(zk, shard_rank, shard_world, tables);
    } catch (...) {
        release();
        throw;
    }
}

void ProverDevice::construct(const ZkeyData& zk, uint32_t shard_rank, uint32_t shard_world, bool tables) {
    const double t_begin = now_ms();
    shard_rank_ = shard_rank;
    shard_world_ = shard_world;
    shard_range(nVars, shard_world, shard_rank, w_lo_, w_hi_);
    shard_range(domainSize, shard_world, shard_rank, h_lo_, h_hi_);
    own_cnt_ = nVars > shard_rank ? ((uint64_t)nVars - shard_rank + shard_world - 1) / shard_world : 0;
    const uint64_t wn = own_cnt_, hn = h_hi_ - h_lo_;
    tabs_.reset(new NttTables(*ctx_));
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    timing = getenv("UGB_TIMING") != nullptr;

    auto need = [&](uint64_t have, uint64_t bytes, const char* what) {
        if (have < bytes) throw std::range_error(std::string("zkey section too short: ") + what);
    };
    need(zk.coefs_bytes, zk.nCoefs * 44, "coefs");
    need(zk.A_bytes, (uint64_t)nVars * 64, "pointsA");
    need(zk.B1_bytes, (uint64_t)nVars * 64, "pointsB1");
    need(zk.B2_bytes, (uint64_t)nVars * 128, "pointsB2");
    need(zk.H_bytes, (uint64_t)domainSize * 64, "pointsH");
    // the coefficient CSR is host work: it runs on its own threads while the bases are uploaded and the window tables are built
    CoefCsr csr;
    std::string csr_err;
    double csr_ms = 0;
    std::thread csr_thread([&]() {
        const double t0 = now_ms();
        try { build_csr(zk, csr); } catch (std::exception& e) { csr_err = e.what(); if (csr_err.empty()) csr_err = "coefficient CSR build failed"; }
        csr_ms = now_ms() - t0;
    });
    struct Joiner { std::thread& t; ~Joiner() { if (t.joinable()) t.join(); } } joiner{csr_thread};

    // ---- bases: upload, zero-point maps, window tables -- pipelined ----
    // Uploads go on their own stream (the zkey buffer is pageable host memory: every copy blocks THIS thread while it is staged, but
    // not the GPU), and the table of an array is built on the compute stream as soon as its upload has landed: the GPU builds the table
    // of array i while the host pushes array i+1.  Nothing is synchronised or freed until the end (cudaFree would drain the device).
    cudaStream_t up = nullptr;
    UG_CUDA_CHECK(cudaStreamCreateWithFlags(&up, cudaStreamNonBlocking));
    std::vector<cudaEvent_t> up_events;
    std::vector<void*> garbage;  // level-0 base arrays superseded by their tables, scratch copies
    std::vector<std::vector<uint8_t>> staged;  // host gathers of a shard's interleaved points; declared before `cleanup` so that they outlive its stream sync
    struct Cleanup {
        cudaStream_t& up; std::vector<cudaEvent_t>& ev; std::vector<void*>& g; DeviceCtx& ctx;
        ~Cleanup() {
            cudaStreamSynchronize(up); cudaStreamSynchronize(ctx.stream);
            for (auto e : ev) cudaEventDestroy(e);
            for (void* p : g) cudaFree(p);
            cudaStreamDestroy(up);
        }
    } cleanup{up, up_events, garbage, ctx};
    auto upload_up = [&](const void* src, size_t bytes) -> void* {
        void* p = dev_alloc(bytes);
        if (bytes) UG_CUDA_CHECK(cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, up));
        return p;
    };
    // this shard's points of a witness-paired array: the whole array when unsharded, else points rank, rank + world, ... gathered on
    // the host (threads over ranges of j) into a buffer that lives until the end of construct
    auto upload_own = [&](const uint8_t* src, size_t elem) -> void* {
        if (shard_world == 1) return upload_up(src, (size_t)wn * elem);
        staged.emplace_back((size_t)wn * elem);
        uint8_t* dst = staged.back().data();
        const unsigned nt = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(8, wn >> 16));
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; t++)
            th.emplace_back([=]() {
                for (uint64_t j = wn * t / nt; j < wn * (t + 1) / nt; j++) memcpy(dst + j * elem, src + ((uint64_t)shard_rank + j * shard_world) * elem, elem);
            });
        for (auto& x : th) x.join();
        return upload_up(dst, (size_t)wn * elem);
    };
    auto landed = [&]() {  // the compute stream waits for everything pushed on the upload stream so far
        cudaEvent_t e;
        UG_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        up_events.push_back(e);
        UG_CUDA_CHECK(cudaEventRecord(e, up));
        UG_CUDA_CHECK(cudaStreamWaitEvent(ctx.stream, e, 0));
    };
    const char* pre_env = getenv("UGB_MSM_PRE");
    pre_ = tables && !(pre_env && atoi(pre_env) == 0);
    double t_tables = 0;
    auto tabulate = [&](void*& base, int group, uint64_t count, MsmConfig& cfg) {
        if (!pre_) return;
        const double t0 = now_ms();
        MsmTable t = msm_build_table(ctx, group, base, count);  // enqueued, not waited for
        garbage.push_back(base);
        base = t.d_table;
        cfg = t.cfg;
        t_tables += now_ms() - t0;
    };
    // zero-point maps: which witness-paired bases are the point at infinity (their polynomial vanishes at tau: signals that never
    // occur in that matrix).  A class with enough of them gets its own digit sort without those points: an infinity base costs its
    // lane a whole idle mixed addition in k_accumulate, a skipped digit costs nothing.  G1-only classes need >= 15 % to repay the
    // extra sort, the class holding the G2 array >= 4 %.  The counts are read back at the very end.
    const size_t bm_words = (size_t)((wn + 31) / 32) * 4 + 16;
    uint32_t* d_counts = (uint32_t*)dev_alloc(64);
    garbage.push_back(d_counts);
    UG_CUDA_CHECK(cudaMemsetAsync(d_counts, 0, 64, ctx.stream));
    uint32_t *bm_a = nullptr, *bm_b = nullptr, *bm_c = nullptr;
    auto bitmap = [&](uint32_t*& bm, int slot, int group, const void* bases, bool combine) {
        if (wn == 0) return;
        if (!bm) bm = (uint32_t*)dev_alloc(bm_words);
        if (combine) UG_CUDA_CHECK(cudaMemsetAsync(d_counts + slot, 0, 4, ctx.stream));
        msm_inf_bitmap_async(ctx, group, bases, wn, bm, combine, d_counts + slot);
    };
    MsmConfig tmp;
    if (pre_) ctx.workspace("msm.table_tmp", std::max<uint64_t>(wn, hn) * sizeof(Jac<Fq2c>) + 256);  // sized once: growing it later would drain the stream

    if (shard_world > 1 && wn) {
        own_idx_ = (uint32_t*)dev_alloc((size_t)wn * 4);
        const int blocks = (int)std::min<uint64_t>((wn + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_iota_strided", k_iota_strided<<<blocks, 256, 0, ctx.stream>>>(own_idx_, (uint32_t)wn, shard_rank, shard_world));
    }
    A_ = upload_own(zk.A, 64);
    landed();
    bitmap(bm_a, 0, 1, A_, false);
    tabulate(A_, 1, wn, cfg_w_);
    B1_ = upload_own(zk.B1, 64);
    landed();
    bitmap(bm_b, 1, 1, B1_, false);
    tabulate(B1_, 1, wn, tmp);
    B2_ = upload_own(zk.B2, 128);
    landed();
    bitmap(bm_b, 1, 2, B2_, true);
    tabulate(B2_, 2, wn, tmp);
    if (!ultra_) {
        const uint32_t nc = nVars - nPublic - 1;
        need(zk.C_bytes, (uint64_t)nc * 64, "pointsC");
        // C pairs with witness[nPublic+1 ..] (src/groth16.cpp:64): pad with points at infinity so that the
        // C MSM reuses the sorted digits of the full witness
        // slot i of the padded array = pointsC[i - (nPublic+1)] for i > nPublic, infinity below; keep this shard's slots
        C_ = dev_alloc((size_t)wn * 64);
        if (shard_world == 1) {
            UG_CUDA_CHECK(cudaMemsetAsync(C_, 0, std::max<size_t>((size_t)wn * 64, 16), up));
            const uint64_t first = (uint64_t)nPublic + 1;
            if (first < nVars)
                UG_CUDA_CHECK(cudaMemcpyAsync((uint8_t*)C_ + first * 64, zk.C, (size_t)(nVars - first) * 64, cudaMemcpyHostToDevice, up));
        } else if (wn) {
            staged.emplace_back((size_t)wn * 64);
            uint8_t* dst = staged.back().data();
            for (uint64_t j = 0; j < wn; j++) {
                const uint64_t sgn = (uint64_t)shard_rank + j * shard_world;
                if (sgn > nPublic) memcpy(dst + j * 64, zk.C + (sgn - nPublic - 1) * 64, 64);
                else memset(dst + j * 64, 0, 64);
            }
            UG_CUDA_CHECK(cudaMemcpyAsync(C_, dst, (size_t)wn * 64, cudaMemcpyHostToDevice, up));
        }
        c_shares_witness_sort_ = true;
        landed();
        bitmap(bm_c, 2, 1, C_, false);
        tabulate(C_, 1, wn, tmp);
    } else {
        n_c1_ = zk.n_c1;
        n_c2_ = zk.n_c2;
        need(zk.C1_bytes, (uint64_t)n_c1_ * 64, "round pointsC");
        need(zk.C_bytes, (uint64_t)n_c2_ * 64, "final pointsC");
        need(zk.idx_c1_bytes, (uint64_t)n_c1_ * 4, "round indexes");
        need(zk.idx_c2_bytes, (uint64_t)n_c2_ * 4, "final indexes");
        const uint8_t* i1 = zk.idx_c1;
        const uint8_t* i2 = zk.idx_c2;
        for (uint32_t i = 0; i < n_c1_; i++) {
            uint32_t v; memcpy(&v, i1 + (size_t)i * 4, 4);
            if (v >= nVars) throw std::range_error("round index out of range");
            round_hi_ = std::max<uint64_t>(round_hi_, (uint64_t)v + 1);
        }
        bool unique = true;
        {
            std::vector<bool> seen(nVars, false);
            for (uint32_t i = 0; i < n_c2_; i++) {
                uint32_t v; memcpy(&v, i2 + (size_t)i * 4, 4);
                if (v >= nVars) throw std::range_error("final round index out of range");
                if (seen[v]) unique = false;
                seen[v] = true;
            }
        }
        shard_range(n_c1_, shard_world, shard_rank, c1_lo_, c1_hi_);
        idx_c1_ = (uint32_t*)upload_up(zk.idx_c1, (size_t)n_c1_ * 4);
        idx_c2_ = (uint32_t*)upload_up(zk.idx_c2, (size_t)n_c2_ * 4);
        if (unique) {
            // final_wtns[i] = wtns[final_idx[i]] (src/ultra_groth.cpp:439-445): place C2[i] at slot final_idx[i]
            // of an nVars-long array of points at infinity, so the final C MSM shares the witness digits (scatter on the device)
            void* c2 = upload_up(zk.C, (size_t)n_c2_ * 64);
            garbage.push_back(c2);
            C_ = dev_alloc((size_t)wn * 64);
            landed();
            UG_CUDA_CHECK(cudaMemsetAsync(C_, 0, std::max<size_t>((size_t)wn * 64, 16), ctx.stream));
            if (n_c2_) {
                const int blocks = (int)std::min<uint64_t>(((uint64_t)n_c2_ + 255) / 256, (uint64_t)ctx.sm_count * 8);
                UG_KERNEL(ctx, "k_scatter_affine", k_scatter_affine<<<blocks, 256, 0, ctx.stream>>>((const G1Affine*)c2, idx_c2_, n_c2_, shard_world, shard_rank, (G1Affine*)C_));
            }
            c_shares_witness_sort_ = true;
            bitmap(bm_c, 2, 1, C_, false);
            tabulate(C_, 1, wn, tmp);
        } else {
            if (shard_world > 1) throw std::invalid_argument("sharded UltraGroth prover needs unique final-round indexes");
            C_ = upload_up(zk.C, (size_t)n_c2_ * 64);
            c_shares_witness_sort_ = false;
            landed();
            tabulate(C_, 1, n_c2_, cfg_c2_);
        }
        C1_ = upload_up(zk.C1 + c1_lo_ * 64, (size_t)(c1_hi_ - c1_lo_) * 64);
        landed();
        tabulate(C1_, 1, c1_hi_ - c1_lo_, cfg_c1_);
    }
    H_ = upload_up(zk.H + h_lo_ * 64, (size_t)hn * 64);
    landed();
    tabulate(H_, 1, hn, cfg_h_);
    create_ms[0] = now_ms() - t_begin - t_tables;  // host time spent pushing the bases (the tables build underneath)
    // coefficient records {u32 m, u32 c, u32 s, Fr coef} packed(1) (src/groth16.cpp:38, src/groth16.hpp:42-50) -> CSR by (m, c)
    csr_thread.join();
    if (!csr_err.empty()) throw std::range_error(csr_err);
    create_ms[1] = csr_ms;
    nCoefs_ = zk.nCoefs;
    n_long_ = (uint32_t)csr.longs.size();
    n_long_segs_ = (uint32_t)csr.segs.size();
    if (n_long_) {
        long_rows_ = upload_up(csr.longs.data(), csr.longs.size() * sizeof(LongRow));
        long_segs_ = upload_up(csr.segs.data(), csr.segs.size() * sizeof(LongRow));
        long_partial_ = dev_alloc((size_t)n_long_segs_ * 32);
    }
    row_ptr_ = (uint32_t*)upload_up(csr.rp.data(), csr.rp.size() * 4);
    coef_sig_ = (uint32_t*)upload_up(csr.sig.data(), csr.sig.size() * 4);
    coef_val_ = upload_up(csr.val.data(), csr.val.size());
    UG_CUDA_CHECK(cudaStreamSynchronize(up));  // the host vectors of the CSR go out of scope below

    // read the zero-point counts back and keep the maps that pay for a sort of their own
    {
        uint32_t h_counts[3] = {0, 0, 0};
        UG_CUDA_CHECK(cudaMemcpyAsync(h_counts, d_counts, 12, cudaMemcpyDeviceToHost, ctx.stream));
        UG_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
        const char* e = getenv("UGB_MSM_SKIP_INF");
        const bool enabled = !(e && atoi(e) == 0) && wn > 0;
        auto keep = [&](uint32_t*& skip, uint32_t* bm, uint64_t& out_count, uint32_t count, double thr) {
            out_count = count;
            if (bm && enabled && (double)count >= thr * (double)wn) skip = bm;
            else if (bm) garbage.push_back(bm);
        };
        keep(skip_a_, bm_a, inf_points[0], h_counts[0], 0.15);
        keep(skip_b_, bm_b, inf_points[1], h_counts[1], 0.04);
        keep(skip_c_, bm_c, inf_points[2], h_counts[2], 0.15);
        // two classes whose zero points nearly coincide (>= 90 % of the larger map survives the AND) share ONE sort on the common map:
        // the few points only one of them could have skipped cost less than a digit sort (in the UltraGroth benchmark circuit A has
        // no terms for the chunk signals and the final-round C has none for chunks + frequencies: the same set but for 2^16 signals)
        auto merge = [&](uint32_t*& x, uint32_t*& y, uint64_t cx, uint64_t cy) -> bool {
            if (!x || !y || x == y) return false;
            uint32_t* both = (uint32_t*)dev_alloc(bm_words);
            const uint64_t c = msm_bitmap_and(ctx, x, y, wn, both);
            if ((double)c >= 0.9 * (double)std::max(cx, cy)) {
                garbage.push_back(x);
                if (y != x) garbage.push_back(y);
                x = y = both;
                return true;
            }
            garbage.push_back(both);
            return false;
        };
        // at most one pair is merged (a third class joining would need its map re-pointed as well)
        if (!merge(skip_a_, skip_c_, inf_points[0], inf_points[2]))
            if (!merge(skip_a_, skip_b_, inf_points[0], inf_points[1])) merge(skip_b_, skip_c_, inf_points[1], inf_points[2]);
    }
    create_ms[2] = now_ms() - t_begin - create_ms[0];  // the rest of the wait until every table is built
    w_ = dev_alloc((size_t)w_chunk() * shard_world_ * 32);  // padded so that equal-sized slices can be all-gathered in place
    // padded like w_: slices of the transformed vectors travel between shards in equal chunks
    a_ = dev_alloc((size_t)h_chunk() * shard_world_ * 32);
    b_ = dev_alloc((size_t)h_chunk() * shard_world_ * 32);
    c_ = dev_alloc((size_t)h_chunk() * shard_world_ * 32);
    // twiddles resident before the first proof
    tabs_->stage_roots(false);
    tabs_->stage_roots(true);
    if (logn > 0) {
        tabs_->coset_inv_bitrev(logn);
        if (logn > 12) tabs_->full_roots(logn);
    }
    ev_.resize(16);
    for (auto& e : ev_) UG_CUDA_CHECK(cudaEventCreate(&e));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx.stream));
    create_ms[3] = now_ms() - t_begin;
}

void ProverDevice::release() {
    if (!ctx_) return;
    cudaSetDevice(ctx_->device);
    cudaStreamSynchronize(ctx_->stream);
    void** ptrs[] = {&long_rows_, &long_segs_, &long_partial_, &A_, &B1_, &B2_, &C_, &H_, &C1_, (void**)&idx_c1_, (void**)&idx_c2_, (void**)&row_ptr_,
                     (void**)&skip_a_, (void**)&skip_b_, (void**)&skip_c_, (void**)&own_idx_,
                     (void**)&coef_sig_, &coef_val_, &w_, &a_, &b_, &c_};
    if (skip_b_ == skip_a_) skip_b_ = nullptr;  // merged classes share one map
    if (skip_c_ == skip_a_ || skip_c_ == skip_b_) skip_c_ = nullptr;
    if (w_alt_) {  // w_ may be on either of the two batch buffers
        w_ = wbuf_[0];
        cudaFree(w_alt_);
        w_alt_ = nullptr;
        for (int i = 0; i < 2; i++) { cudaEventDestroy(ev_up_[i]); cudaEventDestroy(ev_free_[i]); }
    }
    if (copy_stream_) { cudaStreamSynchronize(copy_stream_); cudaStreamDestroy(copy_stream_); copy_stream_ = nullptr; }
    if (ev_rest_) { cudaEventDestroy(ev_rest_); ev_rest_ = nullptr; }
    for (void** p : ptrs) if (*p) { cudaFree(*p); *p = nullptr; }
    for (auto& e : ev_) if (e) cudaEventDestroy(e);
    ev_.clear();
    cudaGetLastError();
}

ProverDevice::~ProverDevice() { release(); }

void ProverDevice::download_h(std::vector<uint8_t>& out) {
    ctx_->activate();
    out.resize((size_t)domainSize * 32);
    UG_CUDA_CHECK(cudaMemcpyAsync(out.data(), a_, out.size(), cudaMemcpyDeviceToHost, ctx_->stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx_->stream));
}

void ProverDevice::download_witness(uint8_t* dst, uint64_t first, uint64_t count) {
    ctx_->activate();
    UG_CUDA_CHECK(cudaMemcpyAsync(dst, (uint8_t*)w_ + first * 32, count * 32, cudaMemcpyDeviceToHost, ctx_->stream));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx_->stream));
}

// Everything from the (device-resident) witness to the five MSM results, in four phases that the sharded prover (several GPUs, one
// point range each) interleaves with its exchanges; one GPU runs them back to back.  Stage-clock events: 0 start, 1 sorts, 2-5 the
// witness MSMs, 6 SpMV, 7 iNTT + coset, 8 NTT, 9 pointwise, 10 H sort, 11 H MSM.
void ProverDevice::mark_stage(int idx) {
    if (timing) UG_CUDA_CHECK(cudaEventRecord(ev_[idx], ctx_->stream));
}

// phase 1: digit sorts of the witness slice + the A, B1, B2, C bucket MSMs (enqueued, not waited for)
void ProverDevice::phase_witness_msms() {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    Phase& ph = ph_;
    ph = Phase();
    mark_stage(0);
    // the four witness MSMs read the same scalars (src/groth16.cpp:53-64): ONE digit sort serves every array that has no zero-point
    // map of its own; arrays with many points at infinity get a sort that skips them (skip_a_ / skip_b_ / skip_c_)
    // resident point j pairs with signal own_idx_[j] (null = j: unsharded) of the whole witness
    const uint8_t* wsl = (const uint8_t*)w_;
    const uint32_t* widx = own_idx_;
    const uint64_t wn = own_cnt_;
    const MsmConfig* wcfg = pre_ ? &cfg_w_ : nullptr;
    const bool c_in_w = c_shares_witness_sort_;
    if (!skip_a_ || !skip_b_ || (c_in_w && !skip_c_)) msm_sort_digits(ctx, wsl, widx, wn, 0, "w", ph.dw, wcfg);
    // one sort per DISTINCT map (classes merged at create share a pointer)
    if (skip_a_) msm_sort_digits(ctx, wsl, widx, wn, 0, "wa", ph.dwa, wcfg, skip_a_);
    if (skip_b_) {
        if (skip_b_ == skip_a_) ph.dwb = ph.dwa;
        else msm_sort_digits(ctx, wsl, widx, wn, 0, "wb", ph.dwb, wcfg, skip_b_);
    }
    if (c_in_w && skip_c_) {
        if (skip_c_ == skip_a_) ph.dwc = ph.dwa;
        else if (skip_c_ == skip_b_) ph.dwc = ph.dwb;
        else msm_sort_digits(ctx, wsl, widx, wn, 0, "wc", ph.dwc, wcfg, skip_c_);
    }
    ph.dA = skip_a_ ? &ph.dwa : &ph.dw;
    ph.dB = skip_b_ ? &ph.dwb : &ph.dw;
    ph.dC = c_in_w ? (skip_c_ ? &ph.dwc : &ph.dw) : &ph.dc;
    mark_stage(1);
    static const bool batch_abc = !(getenv("UGB_MSM_BATCH") && atoi(getenv("UGB_MSM_BATCH")) == 0);
    ph.batched = c_in_w && batch_abc;
    if (ph.batched) {
        // A, B1 and C go through together: one shared reduction tail (msm_g1_launch_batch); the stage clock is split evenly
        const void* bases[3] = {A_, B1_, C_};
        const MsmDigits* dgs[3] = {ph.dA, ph.dB, ph.dC};
        const std::string slots[3] = {"a" + sfx_, "b1" + sfx_, "c" + sfx_};
        msm_g1_launch_batch(ctx, bases, 3, dgs, slots);
        mark_stage(2);
        mark_stage(3);
        msm_g2_launch(ctx, B2_, *ph.dB, "b2" + sfx_);
        mark_stage(4);
    } else {
        msm_g1_launch(ctx, A_, *ph.dA, "a" + sfx_);
        mark_stage(2);
        msm_g1_launch(ctx, B1_, *ph.dB, "b1" + sfx_);
        mark_stage(3);
        msm_g2_launch(ctx, B2_, *ph.dB, "b2" + sfx_);
        mark_stage(4);
        if (c_in_w) {
            msm_g1_launch(ctx, C_, *ph.dC, "c" + sfx_);
        } else {
            msm_sort_digits(ctx, (const uint8_t*)w_, idx_c2_, n_c2_, 0, "cidx", ph.dc, pre_ ? &cfg_c2_ : nullptr);
            msm_g1_launch(ctx, C_, ph.dc, "c" + sfx_);
        }
    }
    mark_stage(5);
}

// phase 2: a, b, c = A w, B w, (A w) o (B w) (src/groth16.cpp:66-108), then ifft -> coset shift -> fft (:110-140) of the vectors
// selected by `mask` (bit 0: a, bit 1: b, bit 2: c).  A shard that owns no vector passes 0 and only keeps the stage clock in step.
void ProverDevice::phase_h_transform(int mask) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    const uint32_t n = domainSize;
    if (mask & 7) {
        int blocks = (int)std::min<uint64_t>(((uint64_t)n + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_spmv_abc", k_spmv_abc<<<blocks, 256, 0, ctx.stream>>>(row_ptr_, coef_sig_, (const Fr*)coef_val_, (const Fr*)w_, n, (Fr*)a_, (Fr*)b_, (Fr*)c_));
        if (n_long_) {
            UG_KERNEL(ctx, "k_spmv_long", k_spmv_long<<<n_long_segs_, SPMV_LONG_BS, 0, ctx.stream>>>((const LongRow*)long_segs_, coef_sig_, (const Fr*)coef_val_, (const Fr*)w_, (Fr*)long_partial_));
            UG_KERNEL(ctx, "k_spmv_long_finish", k_spmv_long_finish<<<n_long_, SPMV_LONG_BS, 0, ctx.stream>>>((const LongRow*)long_rows_, (const Fr*)long_partial_, n, (Fr*)a_, (Fr*)b_));
            UG_KERNEL(ctx, "k_spmv_fix_c", k_spmv_fix_c<<<(n_long_ + 127) / 128, 128, 0, ctx.stream>>>((const LongRow*)long_rows_, n_long_, n, (const Fr*)a_, (const Fr*)b_, (Fr*)c_));
        }
    }
    mark_stage(6);
    Fr* all[3] = {(Fr*)a_, (Fr*)b_, (Fr*)c_};
    Fr* vecs[3] = {nullptr, nullptr, nullptr};
    int count = 0;
    for (int v = 0; v < 3; v++) if (mask & (1 << v)) vecs[count++] = all[v];
    // ifft + coset shift (+ 1/n) fused, output bit-reversed; then fft from bit-reversed to natural
    if (count) ntt_run(ctx, *tabs_, vecs, count, logn, NTT_NATURAL_TO_BITREV, true, tabs_->coset_inv_bitrev(logn), nullptr);
    mark_stage(7);
    if (count) ntt_run(ctx, *tabs_, vecs, count, logn, NTT_BITREV_TO_NATURAL, false, nullptr, nullptr);
    mark_stage(8);
}

// phase 3: h = fromMontgomery(a*b - c) (:142-148) and the H MSM over this shard's range.  whole_h: form all domainSize scalars
// (one GPU; also what the ugb_prover_last_h tap reads) instead of only [h_lo_, h_hi_).
void ProverDevice::phase_h_msm(bool whole_h) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    const uint64_t lo = whole_h ? 0 : h_lo_, cnt = whole_h ? domainSize : h_hi_ - h_lo_;
    if (cnt) {
        int blocks = (int)std::min<uint64_t>((cnt + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_h_final", k_h_final<<<blocks, 256, 0, ctx.stream>>>((Fr*)a_ + lo, (const Fr*)b_ + lo, (const Fr*)c_ + lo, (uint32_t)cnt));
    }
    mark_stage(9);
    msm_sort_digits(ctx, (const uint8_t*)a_ + h_lo_ * 32, nullptr, h_hi_ - h_lo_, 0, "h", ph_.dh, pre_ ? &cfg_h_ : nullptr);
    mark_stage(10);
    msm_g1_launch(ctx, H_, ph_.dh, "h" + sfx_);
    mark_stage(11);
}

// phase 4: wait for the partial sums and finish them on the host
void ProverDevice::phase_finish(ProofMsms& out, const std::string* sfx) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    Phase& ph = ph_;
    const std::string& x = sfx ? *sfx : sfx_;  // a pipelined batch finishes proof k-1 after proof k was enqueued (other result slots)
    msm_g1_finish(ctx, *ph.dA, "a" + x, &out.a);
    msm_g1_finish(ctx, *ph.dB, "b1" + x, &out.b1);
    msm_g2_finish(ctx, *ph.dB, "b2" + x, &out.b2);
    msm_g1_finish(ctx, *ph.dC, "c" + x, &out.c);
    msm_g1_finish(ctx, ph.dh, "h" + x, &out.h);
    check_lookup_error();  // the flag travelled before the MSM results on the same stream
    ctx.prof_collect();
    if (timing) {
        UG_CUDA_CHECK(cudaEventSynchronize(ev_[11]));
        float* dst[] = {&last_times.sort_w, &last_times.msm_a, &last_times.msm_b1, &last_times.msm_b2, &last_times.msm_c, &last_times.spmv,
                        &last_times.intt, &last_times.ntt, &last_times.hfinal, &last_times.sort_h, &last_times.msm_h};
        for (int i = 0; i < 11; i++) UG_CUDA_CHECK(cudaEventElapsedTime(dst[i], ev_[i], ev_[i + 1]));
        if (ph.batched) {  // the batched A/B1/C launch was timed as one interval (recorded under msm_a)
            const float each = last_times.msm_a / 3.0f;
            last_times.msm_a = last_times.msm_b1 = last_times.msm_c = each;
        }
        UG_CUDA_CHECK(cudaEventElapsedTime(&last_times.total, ev_[0], ev_[11]));
    }
}

void ProverDevice::h_pipeline_and_msms(ProofMsms& out, bool ultra) {
    (void)ultra;
    phase_witness_msms();
    phase_h_transform(7);
    phase_h_msm(true);
    phase_finish(out);
}

// sharded v2, phase 0: only this shard's slice of the witness crosses PCIe; the caller then all-gathers the slices between the
// GPUs (NVLink) into every shard's full copy (w_ is padded to world * w_chunk signals for that)
void ProverDevice::upload_witness_slice(const uint8_t* witness) {
    ctx_->activate();
    if (w_hi_ > w_lo_)
        UG_CUDA_CHECK(cudaMemcpyAsync((uint8_t*)w_ + w_lo_ * 32, witness + w_lo_ * 32, (size_t)(w_hi_ - w_lo_) * 32, cudaMemcpyHostToDevice, ctx_->stream));
}

// ---- pipelined batches: two witness buffers and a copy stream, so that the upload of witness k+1 runs under the
// kernels of proof k; the caller (prover_api.cpp) also defers the host tail of proof k until proof k+1 has been enqueued ----
void ProverDevice::batch_begin() {
    ctx_->activate();
    if (!w_alt_) {
        w_alt_ = dev_alloc((size_t)w_chunk() * shard_world_ * 32);
        if (!copy_stream_) UG_CUDA_CHECK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) {
            UG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_up_[i], cudaEventDisableTiming));
            UG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_free_[i], cudaEventDisableTiming));
        }
    }
    wbuf_[0] = w_;
    wbuf_[1] = w_alt_;
    // both buffers are free once everything enqueued so far has run
    for (int i = 0; i < 2; i++) UG_CUDA_CHECK(cudaEventRecord(ev_free_[i], ctx_->stream));
}
// host -> device copy of a whole witness into buffer `slot` on the copy stream (waits until the proof that last used it is done)
void ProverDevice::batch_upload(const uint8_t* witness, int slot) {
    ctx_->activate();
    UG_CUDA_CHECK(cudaStreamWaitEvent(copy_stream_, ev_free_[slot], 0));
    UG_CUDA_CHECK(cudaMemcpyAsync(wbuf_[slot], witness, (size_t)nVars * 32, cudaMemcpyHostToDevice, copy_stream_));
    UG_CUDA_CHECK(cudaEventRecord(ev_up_[slot], copy_stream_));
}
// the proof about to be enqueued reads buffer `slot`
void ProverDevice::batch_use(int slot) {
    ctx_->activate();
    UG_CUDA_CHECK(cudaStreamWaitEvent(ctx_->stream, ev_up_[slot], 0));
    w_ = wbuf_[slot];
    sfx_ = slot ? "#1" : "";
}
// every kernel that reads buffer `slot` has been enqueued
void ProverDevice::batch_release(int slot) {
    UG_CUDA_CHECK(cudaEventRecord(ev_free_[slot], ctx_->stream));
}
void ProverDevice::batch_end() {
    ctx_->activate();
    UG_CUDA_CHECK(cudaStreamSynchronize(copy_stream_));
    UG_CUDA_CHECK(cudaStreamSynchronize(ctx_->stream));
    sfx_.clear();  // w_ stays on whichever buffer holds the last witness (prove_resident reads it)
}

void ProverDevice::prove_groth16(const uint8_t* witness, ProofMsms& out) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    UG_CUDA_CHECK(cudaMemcpyAsync(w_, witness, (size_t)nVars * 32, cudaMemcpyHostToDevice, ctx.stream));
    h_pipeline_and_msms(out, false);
}

void ProverDevice::prove_resident(ProofMsms& out) {
    ctx_->activate();
    if (ultra_) ultra_round_msm(out.round_c);  // same MSM shape as a real proof's first round
    h_pipeline_and_msms(out, ultra_);
}

void ProverDevice::ultra_upload_witness(const uint8_t* witness) {
    ctx_->activate();
    // the round commitment only reads the signals of the round index list (a prefix of the witness in practice): that part goes first,
    // on the proof's stream; the rest follows on the copy stream and is awaited by the first kernel that needs it (ultra_lookup_apply),
    // so it travels while the round MSM runs
    const uint64_t head = std::min<uint64_t>(round_hi_, nVars);
    if (head == 0 || head * 2 > nVars) {
        UG_CUDA_CHECK(cudaMemcpyAsync(w_, witness, (size_t)nVars * 32, cudaMemcpyHostToDevice, ctx_->stream));
        return;
    }
    if (!copy_stream_) UG_CUDA_CHECK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
    if (!ev_rest_) UG_CUDA_CHECK(cudaEventCreateWithFlags(&ev_rest_, cudaEventDisableTiming));
    // the copy stream must not overwrite signals that kernels of the PREVIOUS proof on this stream may still read
    UG_CUDA_CHECK(cudaEventRecord(ev_rest_, ctx_->stream));
    UG_CUDA_CHECK(cudaStreamWaitEvent(copy_stream_, ev_rest_, 0));
    UG_CUDA_CHECK(cudaMemcpyAsync(w_, witness, (size_t)head * 32, cudaMemcpyHostToDevice, ctx_->stream));
    UG_CUDA_CHECK(cudaMemcpyAsync((uint8_t*)w_ + head * 32, witness + head * 32, (size_t)(nVars - head) * 32, cudaMemcpyHostToDevice, copy_stream_));
    UG_CUDA_CHECK(cudaEventRecord(ev_rest_, copy_stream_));
    rest_pending_ = true;
}

void ProverDevice::ultra_round_launch() {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    // round_wtns[i] = wtns[round_indexes[i]] (src/ultra_groth.cpp:415-419) folded into the scalar fetch
    msm_sort_digits(ctx, (const uint8_t*)w_, idx_c1_ + c1_lo_, c1_hi_ - c1_lo_, 0, "round", d_round_, pre_ ? &cfg_c1_ : nullptr);
    msm_g1_launch(ctx, C1_, d_round_, "round");
}
void ProverDevice::ultra_round_finish(HG1XYZZ& out) {
    ctx_->activate();
    msm_g1_finish(*ctx_, d_round_, "round", &out);
}
void ProverDevice::ultra_round_msm(HG1XYZZ& out) {
    ultra_round_launch();
    ultra_round_finish(out);
}

void ProverDevice::ultra_lookup_stage(const void* chunks, uint32_t chunks_len, const void* freqs, uint32_t freqs_len,
                                      const void* wtns_idx, const void* push_idx, uint32_t idx_len) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    // straight copies on the proof's stream (the caller's arrays are borrowed until the proof's final synchronisation); all range
    // checks happen on the device, where the indices are used (k_lookup_push)
    uint32_t* d_chunks = (uint32_t*)ctx.workspace("lk.chunks", (size_t)chunks_len * 4 + 16);
    uint32_t* d_freqs = (uint32_t*)ctx.workspace("lk.freqs", (size_t)freqs_len * 4 + 16);
    uint32_t* d_widx = (uint32_t*)ctx.workspace("lk.widx", (size_t)idx_len * 4 + 16);
    uint32_t* d_pidx = (uint32_t*)ctx.workspace("lk.pidx", (size_t)idx_len * 4 + 16);
    if (chunks_len) UG_CUDA_CHECK(cudaMemcpyAsync(d_chunks, chunks, (size_t)chunks_len * 4, cudaMemcpyHostToDevice, ctx.stream));
    if (freqs_len) UG_CUDA_CHECK(cudaMemcpyAsync(d_freqs, freqs, (size_t)freqs_len * 4, cudaMemcpyHostToDevice, ctx.stream));
    if (idx_len) {
        UG_CUDA_CHECK(cudaMemcpyAsync(d_widx, wtns_idx, (size_t)idx_len * 4, cudaMemcpyHostToDevice, ctx.stream));
        UG_CUDA_CHECK(cudaMemcpyAsync(d_pidx, push_idx, (size_t)idx_len * 4, cudaMemcpyHostToDevice, ctx.stream));
    }
    lk_chunks_len_ = chunks_len; lk_freqs_len_ = freqs_len; lk_idx_len_ = idx_len;
}

void ProverDevice::ultra_lookup_apply(const HFr& rand_mont) {
    DeviceCtx& ctx = *ctx_;
    ctx.activate();
    if (rest_pending_) {  // the tail of the witness (ultra_upload_witness) must have landed before anything reads or fills it
        UG_CUDA_CHECK(cudaStreamWaitEvent(ctx.stream, ev_rest_, 0));
        rest_pending_ = false;
    }
    const uint32_t chunks_len = lk_chunks_len_, freqs_len = lk_freqs_len_, idx_len = lk_idx_len_;
    uint32_t* d_chunks = (uint32_t*)ctx.workspace("lk.chunks", (size_t)chunks_len * 4 + 16);
    uint32_t* d_freqs = (uint32_t*)ctx.workspace("lk.freqs", (size_t)freqs_len * 4 + 16);
    uint32_t* d_widx = (uint32_t*)ctx.workspace("lk.widx", (size_t)idx_len * 4 + 16);
    uint32_t* d_pidx = (uint32_t*)ctx.workspace("lk.pidx", (size_t)idx_len * 4 + 16);
    Fr* d_inv2 = (Fr*)ctx.workspace("lk.inv2", (size_t)freqs_len * 32 + 32);
    Fr* d_prod = (Fr*)ctx.workspace("lk.prod", (size_t)freqs_len * 32 + 32);
    uint32_t* d_err = (uint32_t*)ctx.workspace("lk.err", 16);
    uint32_t* h_err = (uint32_t*)ctx.pinned("lk.err", 16);
    UG_CUDA_CHECK(cudaMemsetAsync(d_err, 0, 4, ctx.stream));
    Fr rm, rr;
    memcpy(&rm, &rand_mont, 32);
    HFr raw = from_mont(rand_mont);
    memcpy(&rr, &raw, 32);
    if (freqs_len) {
        uint32_t threads = (freqs_len + LOOKUP_RUN - 1) / LOOKUP_RUN;
        UG_KERNEL(ctx, "k_lookup_inverses", k_lookup_inverses<<<(threads + 127) / 128, 128, 0, ctx.stream>>>(d_freqs, freqs_len, rm, d_inv2, d_prod));
    }
    if (idx_len) {
        int blocks = (int)std::min<uint64_t>(((uint64_t)idx_len + 255) / 256, (uint64_t)ctx.sm_count * 8);
        UG_KERNEL(ctx, "k_lookup_push", k_lookup_push<<<blocks, 256, 0, ctx.stream>>>(d_widx, d_pidx, idx_len, d_chunks, chunks_len, freqs_len, nVars, rr, d_inv2, d_prod, (Fr*)w_, d_err));
    }
    UG_CUDA_CHECK(cudaMemcpyAsync(h_err, d_err, 4, cudaMemcpyDeviceToHost, ctx.stream));
    lookup_err_pending_ = true;
}

void ProverDevice::check_lookup_error() {
    if (!lookup_err_pending_) return;
    lookup_err_pending_ = false;
    const uint32_t e = *(const uint32_t*)ctx_->pinned("lk.err", 16);
    if (e & 1u) throw std::range_error("lookup witness index out of range");
    if (e & 2u) throw std::range_error("lookup push index out of range");
    if (e & 4u) throw std::range_error("lookup chunk index out of range");
}

void ProverDevice::ultra_final_msms(ProofMsms& out) {
    ctx_->activate();
    h_pipeline_and_msms(out, true);
}

}  // namespace ug
