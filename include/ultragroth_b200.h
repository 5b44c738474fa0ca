/* ultragroth_b200 -- internal CUDA C ABI.
 *
 * This is what a host prover calls INSTEAD of iden3/ffiasm: the reference's prover reaches its
 * arithmetic engine through
 *     E.g1/g2.multiMulByScalarMSM ...... src/groth16.cpp:55,58,61,64,154; src/ultra_groth.cpp:168,203-234,329
 *     FFT<Fr>::fft / ifft / root ....... src/groth16.cpp:110-140;          src/ultra_groth.cpp:288-309
 *     E.fr.mul/add/sub/fromMontgomery .. src/groth16.cpp:70-108,142-148;   build/fr.asm (Fr_raw*)
 * and each entry point below names the call it replaces.  Plain C types only: pointers, sizes, an int
 * status (0 = ok, nonzero = error, text via ugb_last_error()).  No exception crosses this boundary.
 * Device pointers are raw CUDA device addresses (cudaMalloc / torch .data_ptr()); all work is enqueued
 * on the context's own stream and every call that returns host data synchronises that stream.
 * There is NO CPU fallback: without a B200 (sm_100) device ugb_ctx_create fails.
 */
#ifndef ULTRAGROTH_B200_H
#define ULTRAGROTH_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ugb_ctx ugb_ctx;

#define UGB_OK 0
#define UGB_ERROR 1

/* field selectors / vector op codes for ugb_field_vec_op */
#define UGB_FIELD_FR 0
#define UGB_FIELD_FQ 1
#define UGB_OP_MUL 0        /* Fr_rawMMul        build/fr_raw_generic.cpp:107 */
#define UGB_OP_ADD 1        /* Fr_rawAdd         build/fr_raw_generic.cpp:11  */
#define UGB_OP_SUB 2        /* Fr_rawSub         build/fr_raw_generic.cpp:31  */
#define UGB_OP_NEG 3        /* Fr_rawNeg         build/fr_raw_generic.cpp:68  */
#define UGB_OP_TO_MONT 4    /* Fr_rawToMontgomery   build/fr_raw_generic.cpp:150 */
#define UGB_OP_FROM_MONT 5  /* Fr_rawFromMontgomery build/fr_raw_generic.cpp:155 */
#define UGB_OP_SQR 6        /* Fr_rawMSquare     build/fr_raw_generic.cpp:150 */
#define UGB_OP_MUL_DFMA 7   /* Fr_rawMMul through the FP64-assisted product (fp_mul_dfma, csrc/ff.cuh): same values */

const char* ugb_last_error(void);
int ugb_device_count(void);

int ugb_ctx_create(int device, ugb_ctx** out);
void ugb_ctx_destroy(ugb_ctx* ctx);
int ugb_sync(ugb_ctx* ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t ugb_kernel_launches(ugb_ctx* ctx);
/* raw cudaStream_t of the context, for CUDA-event timing on the launching stream */
void* ugb_stream(ugb_ctx* ctx);

int ugb_malloc(ugb_ctx* ctx, uint64_t bytes, void** dptr);
int ugb_free(ugb_ctx* ctx, void* dptr);
int ugb_memcpy_h2d(ugb_ctx* ctx, void* dst_dev, const void* src_host, uint64_t bytes);
int ugb_memcpy_d2h(ugb_ctx* ctx, void* dst_host, const void* src_dev, uint64_t bytes);

/* out[i] = a[i] (op) b[i] over n 32-byte elements, all device pointers (b ignored for unary ops).
 * Replaces the E.fr.* loops of src/groth16.cpp:100-108 and is the parity kernel for build/fr.asm. */
int ugb_field_vec_op(ugb_ctx* ctx, int field, int op, const void* d_a, const void* d_b, void* d_out, uint64_t n);

/* result = sum_i scalar[index ? index[i] : i] * base[i], i < n.
 * d_bases: n affine points in zkey layout (Montgomery; G1 64 B, G2 128 B; all-zero = infinity);
 * d_scalars: raw 32-byte little-endian integers; d_index: optional u32 gather list (may be NULL);
 * window_bits: 0 = choose automatically.  h_out: affine result in the same layout (host memory).
 * Replaces E.g1/g2.multiMulByScalarMSM (src/groth16.cpp:55-64,154). */
int ugb_msm_g1(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
               int window_bits, void* h_out);
int ugb_msm_g2(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
               int window_bits, void* h_out);
/* same, result left in extended-Jacobian XYZZ form (G1 128 B, G2 256 B): the partial sum one GPU
 * contributes when the point range is sharded across GPUs */
int ugb_msm_g1_xyzz(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
                    int window_bits, void* h_out_xyzz);
int ugb_msm_g2_xyzz(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_index, uint64_t n,
                    int window_bits, void* h_out_xyzz);
/* ugb_msm_g1 for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same inputs and result as ugb_msm_g1 with d_index = NULL (bit-identical affine bytes), for n <= UGB_MSM_NOALLOC_MAX_POINTS; a larger
 * n is refused (UGB_ERROR, ugb_last_error names the limit).  It calls no cudaMalloc / cudaFree / cudaMallocHost: all of its scratch
 * is one fixed __device__ array of UGB_MSM_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that runs this entry point holds
 * until the process exits.  The array is shared by ALL calls on a device: calls that overlap on the GPU (two contexts, two host
 * threads) race on it -- serialise them.  The call returns after its own stream has drained, like ugb_msm_g1. */
#define UGB_MSM_NOALLOC_MAX_POINTS 16384u
#define UGB_MSM_NOALLOC_SCRATCH_BYTES (8u * 1024u * 1024u + 4224u)
int ugb_msm_g1_noalloc(ugb_ctx* ctx, const void* d_bases, const void* d_scalars, uint64_t n, void* h_out);
uint64_t ugb_msm_noalloc_max_points(void);
/* host-side reduction of `count` XYZZ partial sums (from several GPUs / ranks) to one affine point */
int ugb_g1_sum_xyzz_to_affine(const void* h_partials, uint64_t count, void* h_out_affine);
int ugb_g2_sum_xyzz_to_affine(const void* h_partials, uint64_t count, void* h_out_affine);

/* In-place Fr transform of 2^logn Montgomery-form elements on the device, natural order in and out;
 * inverse != 0 includes the 1/n scale.  Replaces FFT<Fr>::fft / ifft (src/groth16.cpp:112,122,132 and
 * :120,130,140; src/ultra_groth.cpp:289-309). */
int ugb_ntt(ugb_ctx* ctx, void* d_data, int logn, int inverse);
/* The fused H-pipeline transform of src/groth16.cpp:110-140 on `count` (1..3) vectors:
 * ifft, multiply by omega_2n^i, fft -- evaluations on the odd coset, natural order in and out. */
int ugb_coset_transform(ugb_ctx* ctx, void* const* d_vecs, int count, int logn);

/* ---- extensions on the reference's prover objects (void* from *_prover_create, include/prover.h) ----
 * The reference draws r, s (and UltraGroth's r_k) from randombytes_buf (src/groth16.cpp:165-166,
 * src/ultra_groth.cpp:173,345-346) and has no way to pin them; parity tests need one.  r32/s32/rk32 are
 * 32-byte little-endian raw integers (< 2^248 to stay inside the reference's range); NULL r32 or s32
 * returns the prover to random blinding.
 * The hook exists for PARITY TESTS only: a proof made with reused blinding scalars is not zero-knowledge. */
int ugb_prover_set_blinding(void* prover_object, const void* r32, const void* s32, const void* rk32);
/* the H-MSM scalars (canonical 32-byte integers, domainSize of them) of the last proof */
int ugb_prover_last_h(void* prover_object, void* h_out, uint64_t bytes);
/* the RAW MSM results of the last proof, before blinding and assembly (src/groth16.cpp:53-64,153-154; src/ultra_groth.cpp:168):
 * 448 bytes = A (64) | B1 (64) | B2 (128) | C (64) | H (64) | round C1 (64; zero for Groth16), affine Montgomery, zkey layout */
int ugb_prover_last_msms(void* prover_object, void* out448);
/* `count` 32-byte signals of the witness as it stands in HBM after the last proof (UltraGroth: lookup signals filled in) */
int ugb_prover_last_witness(void* prover_object, void* out, uint64_t first, uint64_t count);
/* create-time breakdown in ms: base uploads, coefficient CSR build (host threads, overlapped), window tables, total */
int ugb_prover_create_times(void* prover_object, double* out4);
/* zero-point maps found at create: points at infinity among pointsA, among pointsB1 AND pointsB2, among the (padded) pointsC */
int ugb_prover_inf_points(void* prover_object, uint64_t* out3);
/* per-stage CUDA-event times (ms) of the last proof: upload, round, lookup, sort_w, msm_a, msm_b1, msm_b2,
 * msm_c, spmv, intt, ntt, hfinal, sort_h, msm_h, total */
int ugb_prover_set_timing(void* prover_object, int enable);
int ugb_prover_last_times(void* prover_object, float* out, int count);
uint64_t ugb_prover_kernel_launches(void* prover_object);
/* One WHOLE proof from the witness (UltraGroth: and lookup arrays) left in HBM by the previous *_prover_prove call: both rounds,
 * challenge, lookup fill, every MSM, the H pipeline, blinding, assembly, JSON -- everything except the host->device input copy
 * (bench.py's device-resident leg).  ugb_prover_last_proof reads the proof back (NUL terminated, truncated to the capacities). */
int ugb_prover_prove_resident(void* prover_object);
int ugb_prover_last_proof(void* prover_object, char* proof_buffer, uint64_t proof_cap, char* public_buffer, uint64_t public_cap);
/* per-kernel device time from CUDA events on the launching stream; stats text is "name\tlaunches\tms\n" */
int ugb_prover_set_profile(void* prover_object, int enable);
uint64_t ugb_prover_kernel_stats(void* prover_object, char* out, uint64_t cap);
void* ugb_prover_stream(void* prover_object);

/* ---- the reference's C++ class API behind a C ABI (include/groth16.hpp, include/ultra_groth.hpp wrap these) ----
 * The zkey sections as HOST pointers, exactly the arguments of Groth16::makeProver<Engine> (src/groth16.hpp:118-136,
 * src/groth16.cpp:17-45; called from src/prover.cpp:160-189) and UltraGroth::makeProver (src/ultra_groth.hpp:218-247,
 * src/ultra_groth.cpp:108-159; src/prover.cpp:237-260).  Everything is copied to HBM during the call; the pointers are not kept. */
typedef struct ugb_zkey_sections {
    uint32_t nVars, nPublic, domainSize;
    uint64_t nCoefs;
    const void *vk_alpha1, *vk_beta1, *vk_beta2, *vk_delta1, *vk_delta2; /* G1 64 B / G2 128 B; UltraGroth: delta = final_delta */
    const void* coefs;    /* nCoefs packed 44-byte Coef records (zkey section 4 + 4) */
    const void *pointsA, *pointsB1, *pointsB2, *pointsC /* UltraGroth: final_pointsC */, *pointsH;
    /* UltraGroth only (ignored for Groth16) */
    const void* round_delta1;
    const void* round_pointsC;
    const uint32_t *round_indexes, *final_round_indexes;
    uint32_t round_indexes_count, final_round_indexes_count, challenge_index;
} ugb_zkey_sections;
int ugb_prover_create_from_sections(void** prover_object, int ultra, const ugb_zkey_sections* sections, char* error_msg,
                                    unsigned long long error_msg_maxsize);
/* Groth16::Prover<Engine>::prove(FrElement* wtns) (src/groth16.cpp:48-203): nVars raw 32-byte signals -> proof points A, B, C in the
 * reference's G1PointAffine / G2PointAffine layout.  Return codes as include/prover.h. */
int ugb_groth16_prove_raw(void* prover_object, const void* wtns_signals, void* out_A64, void* out_B128, void* out_C64, char* error_msg,
                          unsigned long long error_msg_maxsize);
/* UltraGroth::Prover<Engine>::prove(FrElement* wtns, LookupInfo&) (src/ultra_groth.cpp:401-462); wtns_signals is in/out like the
 * reference's (compute_lookup writes the lookup signals into it, src/ultra_groth.cpp:62-106). */
int ugb_ultra_groth_prove_raw(void* prover_object, void* wtns_signals, const uint32_t* chunks, uint32_t chunks_len, const uint32_t* frequencies,
                              uint32_t frequencies_len, const uint32_t* wtns_indxs, const uint32_t* push_indxs, uint32_t wtns_indxs_len,
                              void* out_A64, void* out_B128, void* out_final_commitment64, void* out_round_commitment64, char* error_msg,
                              unsigned long long error_msg_maxsize);

/* ---- batched / pipelined proving: `count` witnesses against one resident prover ----
 * The same outputs as `count` calls of groth16_prover_prove / ultra_groth_prover_prove (src/prover.cpp:501-579 called repeatedly on one
 * object), but witness k+1 is uploaded on a copy stream while proof k computes and the host tail of proof k (partial-sum combine,
 * blinding, assembly, JSON) runs while the GPU already works on proof k+1.  Witness buffers should be pinned host memory (a pageable
 * buffer makes the upload synchronous).  proof_sizes / public_sizes: capacities in, as in include/prover.h.  Return codes as prover.h;
 * on error no output is valid. */
int ugb_prover_prove_batch(void* prover_object, uint32_t count, const void* const* wtns_buffers, const unsigned long long* wtns_sizes,
                           char* const* proof_buffers, unsigned long long* proof_sizes, char* const* public_buffers, unsigned long long* public_sizes,
                           char* error_msg, unsigned long long error_msg_maxsize);

/* ---- sharded proving, phase by phase (what ultragroth_b200/sharded.py drives; SURVEY.md 8(e) both bullets) ----
 * Nothing is replicated that does not have to be: every rank uploads only ITS slice of the witness over PCIe and the slices are
 * all-gathered between the GPUs (NVLink) in place; the three transforms of the H pipeline (src/groth16.cpp:110-140) run on up to
 * three different ranks and every rank receives just its point range of a', b', c' for the pointwise step and its share of the H MSM.
 * The exchanges themselves (all-gather of w, slices of a/b/c, the 896-byte partial sums) are the CALLER's: the layout call hands out the
 * device buffers; chunks are equal-sized (ceil(n / world), buffers padded) so that they map 1:1 onto collectives.  Order per proof:
 *   upload_slice -> [all-gather w] -> (UltraGroth: round_partial -> [gather] -> lookup) -> h_transform(mask) -> [send slices of a,b,c]
 *   -> witness_msms -> [recv] -> h_msm -> finish -> [gather partials] -> ugb_groth16_prove_assemble / ugb_ultra_prove_assemble.
 * Return codes as include/prover.h. */
typedef struct ugb_shard_layout {
    void *w, *a, *b, *c;             /* device buffers: witness (world * w_chunk signals), a / b / c (world * h_chunk elements), 32 B each */
    uint64_t nVars, domainSize;
    uint64_t w_chunk, h_chunk;       /* rank r UPLOADS signals [r * w_chunk, min(nVars, (r+1) * w_chunk)) and owns H points likewise; the
                                      * witness-paired bases (A, B1, B2, C) it owns are those of signals r, r + world, r + 2 world, ...
                                      * (interleaved: balanced bucket work whatever the circuit's signal order), so
                                      * ugb_shard_witness_msms needs the whole witness resident (after the all-gather) */
    uint32_t rank, world;
    int device;
    void* stream;                    /* cudaStream_t every phase is enqueued on */
} ugb_shard_layout;
int ugb_shard_layout_get(void* prover_object, ugb_shard_layout* out);
int ugb_shard_upload_slice(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, char* error_msg,
                           unsigned long long error_msg_maxsize);
int ugb_shard_round_partial(void* prover_object, void* round_partial_out128, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_shard_lookup(void* prover_object, const void* round_partials, uint32_t count, const void* rk32, char* error_msg,
                     unsigned long long error_msg_maxsize);
int ugb_shard_h_transform(void* prover_object, int vector_mask, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_shard_witness_msms(void* prover_object, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_shard_h_msm(void* prover_object, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_shard_finish(void* prover_object, void* partial_out, char* error_msg, unsigned long long error_msg_maxsize);

/* destroys a prover made by ugb_prover_create_from_sections (same as groth16_prover_destroy / ultra_groth_prover_destroy) */
void ugb_prover_destroy(void* prover_object);
/* Proof<Engine>::toJson().dump() (src/groth16.cpp:216-250, src/ultra_groth.cpp:476-513): proof points in the reference's
 * G1PointAffine / G2PointAffine layout -> the JSON text the C API emits.  ultra == 0: A, B, C -> pi_a, pi_b, pi_c, "groth16";
 * ultra != 0: A, B, C = final commitment, R = round commitment -> pi_a, pi_b, pi_f, pi_r, "ultragroth".  NUL terminated. */
int ugb_proof_points_to_json(int ultra, const void* A64, const void* B128, const void* C64, const void* R64, char* out, uint64_t cap);

/* ---- point-range sharded proving ("MSMs shard by point range across GPUs, partial sums reduced
 * on the host or via NCCL") ----
 * One prover object per GPU / process keeps share `rank` of `world` of every zkey base array resident (H and UltraGroth's
 * round C: a contiguous range; A, B1, B2, C: the points of signals rank, rank + world, ... -- see ugb_shard_layout); witness,
 * coefficients and the H pipeline are replicated on this first protocol (the phase-wise ugb_shard_* protocol below shards them too).  The MSM partial sums are opaque
 * host blobs of ugb_prover_partial_size() bytes; the caller exchanges them (one all-gather) and any rank assembles.
 * They replace, per shard, the MSM calls of src/groth16.cpp:53-64,153-154 / src/ultra_groth.cpp:168,203-234,329; the
 * assemble calls are src/groth16.cpp:158-250 / src/ultra_groth.cpp:337-399,476-513.  Return codes as include/prover.h. */
uint64_t ugb_prover_partial_size(void);
int ugb_prover_create_sharded(void** prover_object, const void* zkey_buffer, unsigned long long zkey_size, int ultra, uint32_t rank,
                              uint32_t world, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_groth16_prove_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, void* partial_out,
                              char* error_msg, unsigned long long error_msg_maxsize);
int ugb_groth16_prove_assemble(void* prover_object, const void* partials, uint32_t count, const void* wtns_buffer,
                               unsigned long long wtns_size, char* proof_buffer, unsigned long long* proof_size, char* public_buffer,
                               unsigned long long* public_size, char* error_msg, unsigned long long error_msg_maxsize);
/* UltraGroth: round partial (128 bytes) -> exchange -> final partial (needs ALL round partials and the common r_k) ->
 * exchange -> assemble.  rk32: 32-byte little-endian raw scalar, identical on every shard. */
int ugb_ultra_round_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, void* round_partial_out,
                            char* error_msg, unsigned long long error_msg_maxsize);
int ugb_ultra_final_partial(void* prover_object, const void* wtns_buffer, unsigned long long wtns_size, const void* round_partials,
                            uint32_t count, const void* rk32, void* partial_out, char* error_msg, unsigned long long error_msg_maxsize);
int ugb_ultra_prove_assemble(void* prover_object, const void* partials, uint32_t count, const void* rk32, char* proof_buffer,
                             unsigned long long* proof_size, char* public_buffer, unsigned long long* public_size, char* error_msg,
                             unsigned long long error_msg_maxsize);

/* The MSM planner's choice for n points (host arithmetic only, no device needed): window bits c, windows per scalar W = ceil(255 / c)
 * and the running-sum chunk 2^chunk_log2 buckets.  resident_table != 0: bases with their 2^(c k) multiples resident (one shared
 * bucket set, what a prover object uses); 0: ad-hoc bases (ugb_msm_*: one bucket set per window).  Honours UGB_MSM_PRE_C / UGB_MSM_C /
 * UGB_RS_MU like the MSMs themselves.  For tests and for sizing HBM (tables take W times the base array). */
int ugb_msm_plan(uint64_t n, int resident_table, int* c, int* windows, int* chunk_log2);

/* ---- microbenchmarks for the roofline denominators ----
 * kind 0: independent 32-bit IMAD chains; 1: IMAD.WIDE (mad.wide.u32) chains; 2: dependent Fq Montgomery
 * products (4 chains per thread); 3 / 4: IMAD.WIDE rows with carry-in+out / carry-out only; 5: kind 2 on the Karatsuba product
 * fp_mul_ks; 6: independent DFMA (fma.rz.f64) chains -- the FP64 pipe's peak; 7: kind 2 on the FP64-assisted product fp_mul_dfma.
 * Returns operations per second in *ops_per_s. */
int ugb_bench_int_pipe(ugb_ctx* ctx, int kind, double* ops_per_s);

/* ---- harness tooling: synthetic circuits with a VALID trusted setup ----
 * The reference has no setup tool (zkeys come from snarkjs); benchmarks at 2^20..2^24 constraints need
 * keys whose proofs verify, and the ~5n fixed-base scalar multiplications of a setup need the GPU.
 * ultra == 0: Groth16 chain circuit with domainSize 2^logn (zkey protocol 1, wtns v2);
 * ultra != 0: UltraGroth chain + range-lookup circuit (protocol 1337, .uwtns sections 3..6), table size
 * 2^lookup_bits, n_chunks looked-up signals.  Buffers are malloc'ed here; release with ugb_host_free. */
int ugb_synth_build(ugb_ctx* ctx, int ultra, int logn, uint64_t seed, int lookup_bits, uint32_t n_chunks, void** zkey, uint64_t* zkey_size,
                    void** wtns, uint64_t* wtns_size, char** vk_json);
void ugb_host_free(void* p);
/* h_out[i] = h_scalars[i] * G (group 1: G1 generator (1,2), 64-byte affine; group 2: G2 generator, 128 bytes);
 * scalars are raw 32-byte little-endian integers on the host */
int ugb_fixed_base_mul(ugb_ctx* ctx, int group, const void* h_scalars, uint64_t n, void* h_out);

#ifdef __cplusplus
}
#endif
#endif
