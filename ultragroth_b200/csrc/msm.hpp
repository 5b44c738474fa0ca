// Host-side interface of the MSM pipeline (implemented in msm.cu).
#pragma once
#include "curve.cuh"
#include "device.hpp"
#include "host_ff.hpp"

namespace ug {

using HG1Affine = Affine<HFq>;
using HG2Affine = Affine<HFq2>;
using HG1XYZZ = XYZZ<HFq>;
using HG2XYZZ = XYZZ<HFq2>;

struct MsmConfig {
    int c = 0;        // window bits
    int W = 0;        // windows = ceil(255 / c)
    uint32_t B = 0;   // buckets per window = 2^(c-1)
    uint32_t NB = 0;  // W * B
};

MsmConfig msm_pick_config(uint64_t n, int force_c = 0);

// Sorted signed digits of one scalar vector; reusable for every base array that pairs with it
// (A, B1, B2 and the shifted C all multiply the same witness, src/groth16.cpp:55-64).
struct MsmDigits {
    MsmConfig cfg;
    uint64_t n = 0;
    uint32_t* entries = nullptr;  // [<= n*W] point index | sign<<31, grouped by (window, |digit|-1)
    uint32_t* starts = nullptr;   // [NB+1] offsets, starts[NB] = number of entries
};

// d_scalars: n (or max(index)+1) raw 32-byte little-endian scalars on the device; d_index optional gather
// list (src/ultra_groth.cpp:415-419,439-445 folded into the scalar fetch).  `slot` names the workspace.
void msm_sort_digits(DeviceCtx& ctx, const uint8_t* d_scalars, const uint32_t* d_index, uint64_t n, int force_c,
                     const std::string& slot, MsmDigits& out);

// Bucket accumulation + bucket reduction on the device for one base array, final Horner on the host.
void msm_accumulate_g1(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG1XYZZ* out);
void msm_accumulate_g2(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG2XYZZ* out);

// Asynchronous halves (so that several MSMs can be queued before the first host combine):
// *_launch enqueues everything up to the D2H copy of the per-bit sums into pinned slot `slot`;
// *_finish waits for the stream and runs the Horner combine.
void msm_g1_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot);
void msm_g1_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG1XYZZ* out);
void msm_g2_launch(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, const std::string& slot);
void msm_g2_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG2XYZZ* out);

}  // namespace ug
