// Host-side interface of the Fr NTT (implemented in ntt.cu).
#pragma once
#include <map>
#include <vector>

#include "device.hpp"
#include "ff.cuh"
#include "host_ff.hpp"

namespace ug {

// omega_{2^k} = 5^((r-1)/2^k)  (root convention of ffiasm's FFT<Fr>), Montgomery form
HFr fr_root_of_unity(int k);
HFr hfr_pow(HFr base, uint64_t e);
HFr hfr_from_u64(uint64_t x);  // Montgomery form of a small integer

// Twiddles resident in HBM (north star item 4): built once per device / domain size, reused by every
// proof.  S tables are size independent; W and the coset table are per log2(n).
class NttTables {
public:
    explicit NttTables(DeviceCtx& ctx) : ctx_(ctx) {}
    ~NttTables();
    const Fr* stage_roots(bool inverse);           // S[2^u + j] = omega_{2^(u+1)}^(+-j), u < 12
    const Fr* full_roots(int logn);                // W[e] = omega_n^e, e < n
    const Fr* coset_inv_bitrev(int logn);          // T[p] = omega_2n^rev(p) / n
    // out[i] = base^(bitrev ? rev(i) : i) * factor, i < 2^logn  (device table generator)
    void fill_powers(Fr* d_out, int logn, const HFr& base, const HFr& factor, bool bitrev);

private:
    DeviceCtx& ctx_;
    Fr* S_[2] = {nullptr, nullptr};
    std::map<int, Fr*> W_;
    std::map<int, Fr*> coset_;
};

enum NttOrder { NTT_NATURAL_TO_BITREV = 0 /* DIF */, NTT_BITREV_TO_NATURAL = 1 /* DIT */ };

// In-place transform of `count` (<= 3) device vectors of 2^logn Fr elements (Montgomery form).
//   order   DIF (natural in, bit-reversed out) or DIT (bit-reversed in, natural out)
//   inverse use omega^-1 (no 1/n unless scale_const / scale carries it)
//   scale   optional per-position factor, fused: DIF multiplies the outputs, DIT multiplies the inputs
//   cst     optional constant factor applied to every output
void ntt_run(DeviceCtx& ctx, NttTables& tabs, Fr* const* d_vecs, int count, int logn, NttOrder order, bool inverse,
             const Fr* d_scale, const HFr* cst);

// FFT<Fr>::fft / ifft semantics: natural order in and out, ifft includes 1/n.
void ntt_natural(DeviceCtx& ctx, NttTables& tabs, Fr* d_vec, int logn, bool inverse);

}  // namespace ug
