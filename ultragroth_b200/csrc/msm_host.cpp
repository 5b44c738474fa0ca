// Host half of the MSM: wait for the per-bit partial sums and combine them, sum_p 2^p U[p], by Horner.
// O(W*c) = 255 doublings + additions on the CPU -- the serial tail the reference also runs on the CPU.
#include "msm.hpp"

#include <cstring>

namespace ug {

template <class HF>
static void msm_finish_impl(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, XYZZ<HF>* out) {
    ctx.wait("msm." + slot);
    int nbits = dg.cfg.W * dg.cfg.c;
    const XYZZ<HF>* U = (const XYZZ<HF>*)ctx.pinned("msm.bits." + slot, (size_t)nbits * sizeof(XYZZ<HF>) + 16);
    uint32_t entries = 0;
    memcpy(&entries, (const uint8_t*)U + (size_t)nbits * sizeof(XYZZ<HF>), 4);
    if (ctx.profile) ctx.counters[sizeof(HF) == 32 ? "#msm_entries<G1>" : "#msm_entries<G2>"] += entries;
    XYZZ<HF> acc = XYZZ<HF>::inf();
    for (int p = nbits - 1; p >= 0; p--) {
        acc = pt_dbl(acc);
        pt_add(acc, U[p]);
    }
    *out = acc;
}

void msm_g1_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG1XYZZ* out) {
    msm_finish_impl<HFq>(ctx, dg, slot, out);
}
void msm_g2_finish(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, HG2XYZZ* out) {
    msm_finish_impl<HFq2>(ctx, dg, slot, out);
}
void msm_accumulate_g1(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG1XYZZ* out) {
    msm_g1_launch(ctx, d_bases, dg, "sync");
    msm_g1_finish(ctx, dg, "sync", out);
}
void msm_accumulate_g2(DeviceCtx& ctx, const void* d_bases, const MsmDigits& dg, HG2XYZZ* out) {
    msm_g2_launch(ctx, d_bases, dg, "sync");
    msm_g2_finish(ctx, dg, "sync", out);
}


}  // namespace ug
