// Host half of the MSM: wait for the per-level / per-window partial sums of the running-sum reduction and combine
// them by Horner: a few hundred doublings on the CPU -- the serial tail the reference also runs on the CPU.
#include "msm.hpp"

#include <cstring>

namespace ug {

template <class HF>
static void msm_finish_impl(DeviceCtx& ctx, const MsmDigits& dg, const std::string& slot, XYZZ<HF>* out) {
    ctx.wait("msm." + slot);
    const MsmReducePlan plan = msm_reduce_plan(dg.cfg);
    const int nwin = dg.cfg.nwin(), L = plan.points();
    const XYZZ<HF>* U = (const XYZZ<HF>*)ctx.pinned("msm.sums." + slot, (size_t)64 * 32 * sizeof(XYZZ<HF>) + 16);
    uint32_t entries = 0;
    memcpy(&entries, (const uint8_t*)U + (size_t)L * nwin * sizeof(XYZZ<HF>), 4);
    if (ctx.profile) ctx.counters[sizeof(HF) == 32 ? "#msm_entries<G1>" : "#msm_entries<G2>"] += entries;
    // window w: U[0][w] + 2^mu * sum_b 2^b * U[1+b][w]; windows: sum_w 2^(c*w) * (...), both by Horner
    XYZZ<HF> acc = XYZZ<HF>::inf();
    for (int w = nwin - 1; w >= 0; w--) {
        for (int k = 0; k < dg.cfg.c; k++) acc = pt_dbl(acc);
        XYZZ<HF> win = XYZZ<HF>::inf();
        for (int b = plan.nbits - 1; b >= 0; b--) {
            win = pt_dbl(win);
            pt_add(win, U[(size_t)(1 + b) * nwin + w]);
        }
        for (int k = 0; k < plan.mu; k++) win = pt_dbl(win);
        pt_add(win, U[(size_t)w]);
        pt_add(acc, win);
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
