// Fr number-theoretic transform kernels for sm_100a.
//
// Replaces ffiasm's FFT<Fr>::fft / ifft (depends/ffiasm, absent; call sites src/groth16.cpp:110-140,
// src/ultra_groth.cpp:288-309): size-n radix-2 transform over BN254 Fr with
// omega_n = 5^((r-1)/n), natural order in and out, ifft scaled by 1/n.
//
// Structure: the log2(n) radix-2 stages are grouped into at most three passes over HBM.  A pass owns
// the index bits [lo, lo+T): one CTA stages a tile of 2^T rows x C contiguous columns in shared
// memory (two uint4 planes, conflict-free 128-bit accesses), runs the T stages on it and writes it
// back.  Passes are decoupled the four-step way: element (row p, column col) of a pass is multiplied
// once by omega_{2^(lo+T)}^(col * bitrev_T(p)) (a gather from the HBM-resident table W[e] = omega_n^e),
// so the stages themselves only need the 2^T-th roots, read from a 128 KiB stage-compact table
// S[2^u + j] = omega_{2^(u+1)}^j that is the same for every n.
//   DIF (Gentleman-Sande): natural in  -> bit-reversed out, passes from the high bits down
//   DIT (Cooley-Tukey):    bit-reversed in -> natural out, passes from the low bits up
// The H pipeline (src/groth16.cpp:110-140) runs iNTT as DIF, multiplies by the bit-reversed coset
// table (omega_2n^i / n) in the store of the last DIF pass, and runs the forward NTT as DIT: no
// bit-reversal permutation ever touches HBM.  Up to three vectors (a, b, c) share one launch.
//
// Cost model: per element and pass 64 B of HBM traffic but T/2 Montgomery products of
// 136 IMAD.WIDE each -- on B200 the transform is bound by the integer pipe, not by HBM.
#pragma once
#include "ff.cuh"
#include "memops.cuh"

namespace ug {

constexpr int NTT_MAX_STAGE_BITS = 12;  // S tables hold stages u < 12 (4096 entries)

struct NttPass {
    Fr* data[3];        // vectors transformed by this launch (blockIdx.y)
    const Fr* S;        // stage-compact roots (forward or inverse)
    const Fr* W;        // omega_n^e, e < n (only read when lo > 0)
    const Fr* scale;    // optional per-position factor (DIF: applied at the store, DIT: at the load)
    uint32_t n;         // transform size
    int logn, lo, T, logC;
    int inverse;        // W is indexed with (n - e) mod n
    int use_const;      // multiply every output by cst (standalone ifft: 1/n)
    Fr cst;
};

#if defined(__CUDACC__)

__device__ __forceinline__ Fr sm_load(const uint4* sm, uint32_t E, uint32_t i) {
    Fr r;
    uint4 a = sm[i], b = sm[E + i];
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void sm_store(uint4* sm, uint32_t E, uint32_t i, const Fr& r) {
    sm[i] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
    sm[E + i] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}

template <bool DIF>
__global__ void __launch_bounds__(1024, 1) k_ntt_pass(const NttPass a) {
    extern __shared__ uint4 sm[];
    const int T = a.T, logC = a.logC, lo = a.lo;
    const uint32_t C = 1u << logC;
    const uint32_t E = 1u << (T + logC);
    Fr* __restrict__ data = a.data[blockIdx.y];
    // tile -> fixed index bits
    const uint32_t colblocks = (1u << lo) >> logC;
    const uint32_t colblock = blockIdx.x % colblocks;
    const uint32_t hi = blockIdx.x / colblocks;
    const uint32_t base = (hi << (lo + T)) | (colblock << logC);
    const uint32_t tw_shift = a.logn - (lo + T);  // exponent scale n / 2^(lo+T)

    // ---- load (DIT: inter-pass twiddle and scale applied here) ----
    for (uint32_t idx = threadIdx.x; idx < E; idx += blockDim.x) {
        const uint32_t c = idx & (C - 1), p = idx >> logC;
        const uint32_t g = base + ((p << lo) | c);
        Fr v = ld_g(&data[g]);
        if (!DIF) {
            if (a.scale) v = fp_mul(v, ld_nc(&a.scale[g]));
            if (lo > 0) {
                const uint32_t col = (colblock << logC) | c;
                uint32_t e = (col * (__brev(p) >> (32 - T))) << tw_shift;
                if (a.inverse) e = (a.n - e) & (a.n - 1);
                v = fp_mul(v, ld_nc(&a.W[e]));
            }
        }
        sm_store(sm, E, idx, v);
    }
    __syncthreads();

    // ---- T radix-2 stages in shared memory ----
    const uint32_t nb = E >> 1;
    for (int s = 0; s < T; s++) {
        const int u = DIF ? (T - 1 - s) : s;
        const Fr* __restrict__ Su = a.S + (1u << u);
        for (uint32_t b = threadIdx.x; b < nb; b += blockDim.x) {
            const uint32_t c = b & (C - 1), pb = b >> logC;
            const uint32_t j = pb & ((1u << u) - 1), blk = pb >> u;
            const uint32_t p0 = (blk << (u + 1)) | j;
            const uint32_t i0 = (p0 << logC) | c, i1 = i0 + (1u << (u + logC));
            Fr U = sm_load(sm, E, i0), V = sm_load(sm, E, i1);
            if (DIF) {
                Fr d = fp_sub(U, V);
                if (u > 0) d = fp_mul(d, ld_nc(&Su[j]));
                sm_store(sm, E, i0, fp_add(U, V));
                sm_store(sm, E, i1, d);
            } else {
                if (u > 0) V = fp_mul(V, ld_nc(&Su[j]));
                sm_store(sm, E, i0, fp_add(U, V));
                sm_store(sm, E, i1, fp_sub(U, V));
            }
        }
        __syncthreads();
    }

    // ---- store (DIF: inter-pass twiddle and scale applied here) ----
    for (uint32_t idx = threadIdx.x; idx < E; idx += blockDim.x) {
        const uint32_t c = idx & (C - 1), p = idx >> logC;
        const uint32_t g = base + ((p << lo) | c);
        Fr v = sm_load(sm, E, idx);
        if (DIF) {
            if (lo > 0) {
                const uint32_t col = (colblock << logC) | c;
                uint32_t e = (col * (__brev(p) >> (32 - T))) << tw_shift;
                if (a.inverse) e = (a.n - e) & (a.n - 1);
                v = fp_mul(v, ld_nc(&a.W[e]));
            }
            if (a.scale) v = fp_mul(v, ld_nc(&a.scale[g]));
        }
        if (a.use_const) v = fp_mul(v, a.cst);
        st_g(&data[g], v);
    }
}

// ---- contiguous pass with TMA bulk copies (sm_90+ / sm_100a only) ------------------------------------------------------------
// The same pass as k_ntt_pass for lo == 0 (a tile is 2^T CONTIGUOUS elements), staged the Blackwell way: a persistent CTA per SM
// walks its tiles; one thread issues cp.async.bulk global -> shared for tile i+1 (completion on an mbarrier, expect_tx = tile bytes)
// while all threads run the T butterfly stages of tile i in the other buffer, and the finished tile leaves with cp.async.bulk
// shared -> global (bulk_group).  No thread ever issues a global load or store for the data; latency is hidden by the copy engine
// instead of by co-resident CTAs.  Elements stay in their memory layout in shared memory (32 B apart), so 128-bit accesses are
// 2-way bank conflicted -- the price of letting the copy engine place them.  Selected by UGB_NTT_TMA=1 (ntt.cu); measured against
// the default kernel in profiles/ (north star item 3 asks for TMA staging; DESIGN.md section 4 has the verdict).
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ Fr raw_load(const uint4* sm, uint32_t i) {
    Fr r;
    uint4 a = sm[2 * i], b = sm[2 * i + 1];
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
    r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
    return r;
}
__device__ __forceinline__ void raw_store(uint4* sm, uint32_t i, const Fr& r) {
    sm[2 * i] = make_uint4(r.v[0], r.v[1], r.v[2], r.v[3]);
    sm[2 * i + 1] = make_uint4(r.v[4], r.v[5], r.v[6], r.v[7]);
}

template <bool DIF>
__global__ void __launch_bounds__(1024, 1) k_ntt_pass_bulk(const NttPass a, uint32_t tiles_per_vec, uint32_t total_tiles) {
    extern __shared__ __align__(128) uint8_t smraw[];
    const int T = a.T;
    const uint32_t E = 1u << T, bytes = E * 32u;
    uint4* buf[2] = {reinterpret_cast<uint4*>(smraw), reinterpret_cast<uint4*>(smraw + bytes)};
    uint64_t* full = reinterpret_cast<uint64_t*>(smraw + 2 * (size_t)bytes);
    const uint32_t mb[2] = {smem_u32(&full[0]), smem_u32(&full[1])};
    const uint32_t sb[2] = {smem_u32(buf[0]), smem_u32(buf[1])};
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb[0]));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb[1]));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto tile_ptr = [&](uint32_t id) -> Fr* { return a.data[id / tiles_per_vec] + (size_t)(id % tiles_per_vec) * E; };
    auto issue_load = [&](uint32_t id, int b) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb[b]), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sb[b]), "l"(tile_ptr(id)),
                     "r"(bytes), "r"(mb[b])
                     : "memory");
    };
    if (threadIdx.x == 0 && blockIdx.x < total_tiles) issue_load(blockIdx.x, 0);
    uint32_t it = 0;
    for (uint32_t id = blockIdx.x; id < total_tiles; id += gridDim.x, it++) {
        const int b = it & 1;
        const uint32_t next = id + gridDim.x;
        if (threadIdx.x == 0 && next < total_tiles) {
            // the other buffer still feeds the store of the previous tile: wait until the copy engine has READ it
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            issue_load(next, b ^ 1);
        }
        {  // wait for this tile to land
            const uint32_t parity = (it >> 1) & 1u;
            uint32_t done = 0;
            while (!done) {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(done)
                             : "r"(mb[b]), "r"(parity)
                             : "memory");
            }
        }
        uint4* sm = buf[b];
        const uint32_t base = (id % tiles_per_vec) * E;  // position of the tile inside its vector (for the scale table)
        if (!DIF && a.scale) {
            for (uint32_t idx = threadIdx.x; idx < E; idx += blockDim.x) raw_store(sm, idx, fp_mul(raw_load(sm, idx), ld_nc(&a.scale[base + idx])));
            __syncthreads();
        }
        const uint32_t nb = E >> 1;
        for (int s = 0; s < T; s++) {
            const int u = DIF ? (T - 1 - s) : s;
            const Fr* __restrict__ Su = a.S + (1u << u);
            for (uint32_t q = threadIdx.x; q < nb; q += blockDim.x) {
                const uint32_t j = q & ((1u << u) - 1), blk = q >> u;
                const uint32_t i0 = (blk << (u + 1)) | j, i1 = i0 + (1u << u);
                Fr U = raw_load(sm, i0), V = raw_load(sm, i1);
                if (DIF) {
                    Fr d = fp_sub(U, V);
                    if (u > 0) d = fp_mul(d, ld_nc(&Su[j]));
                    raw_store(sm, i0, fp_add(U, V));
                    raw_store(sm, i1, d);
                } else {
                    if (u > 0) V = fp_mul(V, ld_nc(&Su[j]));
                    raw_store(sm, i0, fp_add(U, V));
                    raw_store(sm, i1, fp_sub(U, V));
                }
            }
            __syncthreads();
        }
        if ((DIF && a.scale) || a.use_const) {
            for (uint32_t idx = threadIdx.x; idx < E; idx += blockDim.x) {
                Fr v = raw_load(sm, idx);
                if (DIF && a.scale) v = fp_mul(v, ld_nc(&a.scale[base + idx]));
                if (a.use_const) v = fp_mul(v, a.cst);
                raw_store(sm, idx, v);
            }
        }
        // generic-proxy writes to shared memory must be visible to the async proxy (the copy engine) before the bulk store reads them
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(tile_ptr(id)), "r"(sb[b]), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// out[i] = hi[e >> lo_bits] * lo[e & mask], e = bitrev ? rev_logn(i) : i   (root tables, coset table)
__global__ void __launch_bounds__(256) k_fill_powers(Fr* __restrict__ out, uint32_t n, int logn, const Fr* __restrict__ hi_tab,
                                                     const Fr* __restrict__ lo_tab, int lo_bits, int bitrev) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t e = bitrev ? (logn ? (__brev(i) >> (32 - logn)) : 0) : i;
        Fr v = fp_mul(ld_nc(&hi_tab[e >> lo_bits]), ld_nc(&lo_tab[e & ((1u << lo_bits) - 1)]));
        st_g(&out[i], v);
    }
}

// in-place bit-reversal permutation (standalone natural-order API only)
__global__ void __launch_bounds__(256) k_bitrev_swap(Fr* __restrict__ data, uint32_t n, int logn) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t r = __brev(i) >> (32 - logn);
        if (i < r) {
            Fr x = ld_g(&data[i]), y = ld_g(&data[r]);
            st_g(&data[i], y);
            st_g(&data[r], x);
        }
    }
}

#endif  // __CUDACC__
}  // namespace ug
