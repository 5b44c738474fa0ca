// Device-resident prover state: everything Groth16::Prover / UltraGroth::Prover keep as raw zkey
// pointers (src/groth16.hpp:74-114, src/ultra_groth.hpp:108-206) lives here in HBM instead, uploaded
// once at *_prover_create (north star item 4: bases and twiddles stay resident across proofs).
#pragma once
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "device.hpp"
#include "hostio.hpp"
#include "msm.hpp"
#include "ntt.hpp"

namespace ug {

// The MSM results of one proof, as XYZZ partial sums on the host (one per shard when several GPUs
// hold disjoint point ranges; the caller adds them up).
struct ProofMsms {
    HG1XYZZ a, b1, c, h, round_c;
    HG2XYZZ b2;
};

struct StageTimes {  // milliseconds of device time per stage of the last proof (CUDA events)
    float upload = 0, round = 0, lookup = 0, sort_w = 0, msm_a = 0, msm_b1 = 0, msm_b2 = 0, msm_c = 0, spmv = 0, intt = 0, ntt = 0,
          hfinal = 0, sort_h = 0, msm_h = 0, total = 0;
};

// What the device prover needs from a zkey: borrowed HOST pointers, valid only during ProverDevice construction.  Filled either
// from a parsed zkey file (zkey_data_from_file: the section casts of src/prover.cpp:160-189,237-260) or from the section
// pointers a C++ embedder hands to Groth16::makeProver / UltraGroth::makeProver (src/groth16.hpp:118-136, src/ultra_groth.hpp:218-247).
// *_bytes: how many bytes are readable behind the pointer (SIZE_MAX = trusted, as the reference's class API trusts its caller).
struct ZkeyData {
    bool ultra = false;
    uint32_t nVars = 0, nPublic = 0, domainSize = 0;
    uint64_t nCoefs = 0;
    const uint8_t* coefs = nullptr;  // nCoefs packed 44-byte records {u32 m, u32 c, u32 s, Fr coef} (after the u32 count of section 4)
    const uint8_t *A = nullptr, *B1 = nullptr, *B2 = nullptr, *C = nullptr /* Groth16 C / UltraGroth final C */, *H = nullptr, *C1 = nullptr /* round C */;
    const uint8_t *idx_c1 = nullptr, *idx_c2 = nullptr;  // u32 lists (UltraGroth)
    uint32_t n_c1 = 0, n_c2 = 0;
    uint64_t coefs_bytes = UINT64_MAX, A_bytes = UINT64_MAX, B1_bytes = UINT64_MAX, B2_bytes = UINT64_MAX, C_bytes = UINT64_MAX, H_bytes = UINT64_MAX,
             C1_bytes = UINT64_MAX, idx_c1_bytes = UINT64_MAX, idx_c2_bytes = UINT64_MAX;
};
ZkeyData zkey_data_from_file(const BinFile& zkey, const ZkeyHeader& hdr);

class ProverDevice {
public:
    // shard_rank / shard_world: point sharding: this device keeps only its share of every base array resident and
    // its MSMs return PARTIAL sums; the witness, the coefficients and the H pipeline are replicated on every shard.  H and the round
    // bases are cut into contiguous ranges shard_range(count, world, rank); the witness-paired bases (A, B1, B2, padded C) are dealt
    // out INTERLEAVED -- shard r owns signals r, r + world, r + 2 world, ... -- because signal order is circuit order: a contiguous
    // range hands one shard the constants and small values and another the full-width ones (7 vs 14 ms of bucket work on 8 GPUs).
    // tables = false: skip the window-table build (one-shot *_prover calls: the 2^(c*k) multiples only pay off across proofs)
    ProverDevice(int device, const ZkeyData& zk, uint32_t shard_rank = 0, uint32_t shard_world = 1, bool tables = true);
    // slice `rank` of [0, n): equal chunks of ceil(n / world) (the last ones may be short or empty), so that slices can be exchanged
    // with equal-sized all-gather / scatter collectives
    static uint64_t shard_chunk(uint64_t n, uint32_t world) { return (n + world - 1) / world; }
    static void shard_range(uint64_t n, uint32_t world, uint32_t rank, uint64_t& lo, uint64_t& hi) {
        const uint64_t chunk = shard_chunk(n, world);
        lo = std::min<uint64_t>(n, rank * chunk);
        hi = std::min<uint64_t>(n, lo + chunk);
    }
    uint64_t w_chunk() const { return shard_chunk(nVars, shard_world_); }
    uint64_t h_chunk() const { return shard_chunk(domainSize, shard_world_); }
    uint32_t shard_rank() const { return shard_rank_; }
    uint32_t shard_world() const { return shard_world_; }
    void* dev_w() const { return w_; }
    void* dev_a() const { return a_; }
    void* dev_b() const { return b_; }
    void* dev_c() const { return c_; }

    // ---- the proof in phases (the sharded prover interleaves its exchanges between them; see prover_dev.cu) ----
    void upload_witness_slice(const uint8_t* witness);
    void phase_witness_msms();
    void phase_h_transform(int mask);
    void phase_h_msm(bool whole_h);
    void phase_finish(ProofMsms& out, const std::string* sfx = nullptr);
    // pipelined batches (N3): see prover_dev.cu
    void batch_begin();
    void batch_upload(const uint8_t* witness, int slot);
    void batch_use(int slot);
    void batch_release(int slot);
    void batch_end();
    std::string result_suffix() const { return sfx_; }
    ~ProverDevice();
    double create_ms[4] = {0, 0, 0, 0};  // create-time breakdown: base uploads, coefficient CSR, window tables, total

    DeviceCtx& ctx() { return *ctx_; }
    NttTables& tables() { return *tabs_; }

    // ---- Groth16 (src/groth16.cpp:48-203 up to the blinding) ----
    // witness: nVars raw 32-byte scalars on the host
    void prove_groth16(const uint8_t* witness, ProofMsms& out);
    // the same device work on the witness already resident in HBM from the previous proof (no host copies
    // in: bench.py's device-resident throughput leg)
    void prove_resident(ProofMsms& out);

    // ---- UltraGroth (src/ultra_groth.cpp:401-462) ----
    void ultra_upload_witness(const uint8_t* witness);
    void ultra_round_msm(HG1XYZZ& out);                                   // execute_round's MSM (:168)
    void ultra_round_launch();                                            // ... enqueued only / host half
    void ultra_round_finish(HG1XYZZ& out);
    // compute_lookup (:62-106) in two halves: _stage copies the four u32 index arrays to HBM (asynchronous; the host arrays -- any
    // alignment -- must stay valid until the proof's final synchronisation), _apply runs the kernels once the challenge is known;
    // indices are range-checked on the device and a violation surfaces as std::range_error from ultra_final_msms
    void ultra_lookup_stage(const void* chunks, uint32_t chunks_len, const void* freqs, uint32_t freqs_len, const void* wtns_idx,
                            const void* push_idx, uint32_t idx_len);
    void ultra_lookup_apply(const HFr& rand_mont);
    void ultra_final_msms(ProofMsms& out);                                // execute_final_round's MSMs + H
    void download_witness(uint8_t* dst, uint64_t first, uint64_t count);  // signals after the lookup fill

    // debug / parity taps
    void download_h(std::vector<uint8_t>& out);  // the H-MSM scalars of the last proof (canonical integers)
    StageTimes last_times;
    bool timing = false;

    uint32_t nVars, nPublic, domainSize;
    int logn;

private:
    struct Phase {  // sorted digits of the proof in flight (they point into named workspaces)
        MsmDigits dw, dwa, dwb, dwc, dc, dh;
        const MsmDigits *dA = nullptr, *dB = nullptr, *dC = nullptr;
        bool batched = false;
    } ph_;
    uint32_t shard_rank_ = 0, shard_world_ = 1;
    MsmDigits d_round_;
    std::string sfx_;  // suffix of the pinned MSM result slots of the proof being enqueued ("" / "#1": two proofs can be in flight)
    void *w_alt_ = nullptr, *wbuf_[2] = {nullptr, nullptr};
    cudaStream_t copy_stream_ = nullptr;
    cudaEvent_t ev_rest_ = nullptr;   // tail of a split witness upload (ultra_upload_witness)
    bool rest_pending_ = false;
    uint64_t round_hi_ = 0;           // 1 + the largest signal index the round commitment reads
    cudaEvent_t ev_up_[2] = {nullptr, nullptr}, ev_free_[2] = {nullptr, nullptr};
    void mark_stage(int idx);
    void construct(const ZkeyData& zk, uint32_t shard_rank, uint32_t shard_world, bool tables);
    void release();
    void h_pipeline_and_msms(ProofMsms& out, bool ultra);
    void check_lookup_error();
    bool lookup_err_pending_ = false;
    std::unique_ptr<DeviceCtx> ctx_;
    std::unique_ptr<NttTables> tabs_;
    bool ultra_;
    uint32_t n_c1_ = 0, n_c2_ = 0;
    // [w_lo_, w_hi_): the contiguous witness chunk this shard uploads (upload_witness_slice); H holds points [h_lo_, h_hi_), C1
    // [c1_lo_, c1_hi_); the witness-paired arrays hold the own_cnt_ points of signals rank + j * world (own_idx_[j], null when world = 1)
    uint64_t w_lo_ = 0, w_hi_ = 0, h_lo_ = 0, h_hi_ = 0, c1_lo_ = 0, c1_hi_ = 0;
    uint64_t own_cnt_ = 0;
    uint32_t* own_idx_ = nullptr;
    // bases
    void *A_ = nullptr, *B1_ = nullptr, *B2_ = nullptr, *C_ = nullptr, *H_ = nullptr, *C1_ = nullptr;
    bool c_shares_witness_sort_ = false;
    // resident window tables (msm_build_table): every *_ base pointer above then holds W levels instead of one
    bool pre_ = false;
    MsmConfig cfg_w_, cfg_h_, cfg_c1_, cfg_c2_;
    uint32_t *idx_c1_ = nullptr, *idx_c2_ = nullptr;
    // bitmaps of the points at infinity among the witness-paired bases (null: too few to pay for a sort of their own): an array whose
    // polynomial is zero for many signals (B in most circuits) gets its own digit sort that skips them
    uint32_t *skip_a_ = nullptr, *skip_b_ = nullptr, *skip_c_ = nullptr;
public:
    uint64_t inf_points[3] = {0, 0, 0};  // points at infinity in A, in B1 and B2 together, in the padded C (this shard's range)
private:
    // coefficient matrix, CSR by (matrix, row)
    uint32_t* row_ptr_ = nullptr;  // [2n + 1]
    uint32_t* coef_sig_ = nullptr;
    void* coef_val_ = nullptr;
    uint64_t nCoefs_ = 0;
    void* long_rows_ = nullptr;  // rows too long for one thread (see k_spmv_long)
    uint32_t n_long_ = 0;
    void* long_segs_ = nullptr;     // the long rows cut into <= SPMV_SEG-term segments, one CTA each
    void* long_partial_ = nullptr;  // one Fr partial sum per segment
    uint32_t n_long_segs_ = 0;
    // staged lookup arrays (ultra_lookup_stage)
    uint32_t lk_chunks_len_ = 0, lk_freqs_len_ = 0, lk_idx_len_ = 0;
    // per-proof buffers
    void *w_ = nullptr, *a_ = nullptr, *b_ = nullptr, *c_ = nullptr;
    std::vector<cudaEvent_t> ev_;
};

}  // namespace ug
