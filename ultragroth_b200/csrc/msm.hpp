// Host-side interface of the MSM pipeline (implemented in msm.cu).
#pragma once
#include "curve.cuh"
#include "device.hpp"
#include "host_ff.hpp"

#include <vector>

namespace ug {

using HG1Affine = Affine<HFq>;
using HG2Affine = Affine<HFq2>;
using HG1XYZZ = XYZZ<HFq>;
using HG2XYZZ = XYZZ<HFq2>;

struct MsmConfig {
    int c = 0;        // window bits
    int W = 0;        // windows = ceil(255 / c)
    uint32_t B = 0;   // buckets per window = 2^(c-1)
    uint32_t NB = 0;  // bucket count: W * B, or B when `pre`
    bool pre = false; // bases come with precomputed 2^(c*w) multiples (msm_build_table): ONE bucket set for all windows
    int nwin() const { return pre ? 1 : W; }
};

// pre = false: per-window bucket sets (any base array); pre = true: the window size for a resident table over n bases
MsmConfig msm_pick_config(uint64_t n, int force_c = 0, bool pre = false);

// Bucket-reduction schedule (msm.cuh): ONE level of chunked running sums turns the B buckets of a window into K chunk
// sums S[k] and K weighted sums R[k]; the tail  sum_k k*S[k]  (k < K <= ~64Ki) is taken bit by bit:
// T[b] = sum of S[k] over the k with bit b set -- plain, fully parallel tree sums -- and the host applies the 2^b.
// (A recursive running sum over the chunk sums was latency-bound: every level is a serial chain of 2*2^mu additions.)
struct MsmReducePlan {
    int mu = 0;        // chunk = 2^mu buckets
    uint32_t K = 0;    // chunk sums per window
    int nbits = 0;     // bits of K-1 (0 when K == 1)
    int points() const { return 1 + nbits; }  // partial sums per window handed to the host: R, T[0..nbits)
};
MsmReducePlan msm_reduce_plan(const MsmConfig& cfg);

// Resident bases with their window multiples: table[w*n + i] = 2^(c*w) * base[i], affine, zkey layout.
// Costs W times the memory of the base array (HBM is 180 GB) and removes W-1 of the W bucket reductions; with the
// reduction this cheap the window grows from c = 16 to c = 20 at n = 2^22 (22 at 2^24), i.e. 13 instead of 16 additions per point.
struct MsmTable {
    void* d_table = nullptr;  // cudaMalloc'ed, owned by the caller (ProverDevice)
    uint64_t n = 0;
    MsmConfig cfg;
};
// Builds the table on the device from n affine bases (device pointer).  group 1 = G1 (64 B), 2 = G2 (128 B).
MsmTable msm_build_table(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, int force_c = 0);

// Sorted signed digits of one scalar vector; reusable for every base array that pairs with it
// (A, B1, B2 and the shifted C all multiply the same witness, src/groth16.cpp:55-64).
struct MsmDigits {
    MsmConfig cfg;
    uint64_t n = 0;
    uint32_t* entries = nullptr;  // [<= n*W] point (or table) index | sign<<31, grouped by bucket key
    uint32_t* starts = nullptr;   // [NB+1] offsets, starts[NB] = number of entries
};

// d_scalars: n (or max(index)+1) raw 32-byte little-endian scalars on the device; d_index optional gather
// list (src/ultra_groth.cpp:415-419,439-445 folded into the scalar fetch).  `slot` names the workspace.
// `table_cfg`: when the digits are meant for MsmTable bases, the table's config (window size and layout must match).
// `d_skip`: optional bitmap of points that get NO entries (points at infinity of the arrays this sort serves, msm_inf_bitmap).
void msm_sort_digits(DeviceCtx& ctx, const uint8_t* d_scalars, const uint32_t* d_index, uint64_t n, int force_c,
                     const std::string& slot, MsmDigits& out, const MsmConfig* table_cfg = nullptr, const uint32_t* d_skip = nullptr);
// bitmap of the points at infinity among n affine bases (group 1: G1, 2: G2); combine = AND into the existing bitmap (points that
// are infinite in ALL the arrays combined).  Returns the number of set bits.  d_bitmap: (n + 31) / 32 words, device.
uint64_t msm_inf_bitmap(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, uint32_t* d_bitmap, bool combine);
// d_out = d_x & d_y (bitmaps over n_points points); returns the number of set bits
uint64_t msm_bitmap_and(DeviceCtx& ctx, const uint32_t* d_x, const uint32_t* d_y, uint64_t n_points, uint32_t* d_out);
// the same, enqueued only: the number of set bits is ADDED to *d_count (device) -- zero it first; nothing is synchronised
void msm_inf_bitmap_async(DeviceCtx& ctx, int group, const void* d_bases, uint64_t n, uint32_t* d_bitmap, bool combine, uint32_t* d_count);

// Bucket accumulation + bucket reduction on the device for one base array, final Horner on the host.
void msm_accumulate_g1(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG1XYZZ* out);
void msm_accumulate_g2(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG2XYZZ* out);

// Asynchronous halves (so that several MSMs can be queued before the first host combine):
// *_launch enqueues everything up to the D2H copy of the per-level partial sums into pinned slot `slot`;
// *_finish waits for the stream and runs the Horner combine.  d_bases is the plain base array, or MsmTable::d_table
// when the digits were sorted for that table.
void msm_g1_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot);
void msm_g1_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG1XYZZ* out);
// `count` base arrays (tables) that pair with the same digits: separate accumulations, ONE shared reduction tail; finish per slot
void msm_g1_launch_batch(DeviceCtx& ctx, const void* const* d_bases, int count, const MsmDigits& dg, const std::string* slots);
// the same with one digit set PER base array (same window config): arrays whose infinity patterns differ are sorted separately
void msm_g1_launch_batch(DeviceCtx& ctx, const void* const* d_bases, int count, const MsmDigits* const* dgs, const std::string* slots);
void msm_g2_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot);
void msm_g2_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG2XYZZ* out);

// G1 MSM that allocates nothing (msm_noalloc.cu): all scratch in one file-scope __device__ array of MSM_NOALLOC_SCRATCH_BYTES, so at
// most MSM_NOALLOC_MAX_POINTS points; result XYZZ (128 B) to host memory.  Calls that overlap on one GPU race on the shared scratch.
constexpr uint64_t MSM_NOALLOC_MAX_POINTS = 16384;
constexpr uint64_t MSM_NOALLOC_SCRATCH_BYTES = 32ull * (MSM_NOALLOC_MAX_POINTS / 2048) * 256 * 128 + 33 * 128;  // 8 MiB + 4224 B
void msm_g1_noalloc(DeviceCtx& ctx, const void* d_bases, const void* d_scalars, uint64_t n, void* h_out_xyzz);

}  // namespace ug
