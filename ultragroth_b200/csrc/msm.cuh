// Pippenger multi-scalar multiplication kernels for sm_100a (G1 and G2 via the field template F).
//
// Replaces ffiasm's Curve<F>::multiMulByScalarMSM (depends/ffiasm, absent), call sites
// src/groth16.cpp:55,58,61,64,154 and src/ultra_groth.cpp:168,203,214,224,234,329:
//   result = sum_i scalar_i * base_i,  bases affine Montgomery (all-zero = infinity),
//   scalars raw (non-Montgomery) 32-byte little-endian integers.
//
// Pipeline (all on one stream, no host round trip until the final 256 partial sums):
//   1. k_digits_count     signed c-bit digits of every scalar, histogram per (window, |digit|)
//   2. scan               exclusive prefix sum of the histogram -> bucket start offsets
//   3. k_digits_scatter   counting-sort scatter of (point index | sign) into bucket order
//   4. k_accumulate       atomics-free segmented bucket accumulation: every thread owns an equal
//                         slice of the SORTED entry list (so skewed scalar distributions cannot
//                         unbalance it), adds affine points into an XYZZ accumulator and flushes at
//                         bucket boundaries; runs cut by a slice boundary become (key, partial) pairs
//   5. k_seg_reduce       segmented scan over those pairs (recursively, 128:2 per level)
//   6. k_running_sum      chunked running sums over every window's buckets (2 additions per bucket) -> chunk sums S, R
//      k_sum_points       sum of the R;   k_bit_sums: sum_k k*S[k] bit by bit (plain tree sums, no serial chain)
//   7. host               Horner over the bits and the windows (a few hundred doublings) -- O(1) serial
//                         work, as the reference's own final double-and-add (north star: "reduced on the host")
#pragma once
#include "curve.cuh"
#include "memops.cuh"

namespace ug {

// ---- scalar digits ----------------------------------------------------------------------------
struct Scalar256 { uint32_t v[8]; };

// reduce an arbitrary 256-bit scalar below r (s*P == (s mod r)*P on the r-torsion); at most 5 steps
UG_HD void scalar_reduce(Scalar256& s) {
    for (int it = 0; it < 6; it++) {
        uint32_t t[8];
        t[0] = sub_cc(s.v[0], FrParams::mod(0));
#pragma unroll
        for (int i = 1; i < 8; i++) t[i] = subc_cc(s.v[i], FrParams::mod(i));
        uint32_t borrow = subc(0, 0);
        if (borrow) break;
#pragma unroll
        for (int i = 0; i < 8; i++) s.v[i] = t[i];
    }
}

UG_HD uint32_t scalar_bits(const Scalar256& s, int pos, int c) {
    // bits [pos, pos+c) ; c <= 24 ; bits beyond 256 read as zero
    int w = pos >> 5, off = pos & 31;
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        if (i == w) lo = s.v[i];
        if (i == w + 1) hi = s.v[i];
    }
    uint64_t x = ((uint64_t)hi << 32) | lo;
    return (uint32_t)(x >> off) & ((1u << c) - 1u);
}

// signed-digit recoding: digit in [-(B-1), B], B = 2^(c-1); emits (window, |d|-1, sign) via f
template <class Fn>
UG_HD void scalar_digits(const Scalar256& s, int c, int W, Fn f) {
    uint32_t carry = 0;
    const uint32_t B = 1u << (c - 1);
    for (int w = 0; w < W; w++) {
        uint32_t d = scalar_bits(s, w * c, c) + carry;
        carry = 0;
        bool neg = false;
        if (d > B) { d = (1u << c) - d; neg = true; carry = 1; }
        if (d != 0) f(w, d - 1, neg);
    }
}

// The next non-empty bucket after `key`: the smallest k > key with starts[k + 1] > pos.  Dense digit distributions find it at the
// first or second probe; a shard that holds only small scalars leaves ~2^21 empty buckets between two windows, and a linear walk
// over them stalls one thread (and with it its whole launch: 5 ms kernels at 2 % pipe use on shard 0 of 8) -- so after two probes
// the search is binary.  Caller guarantees pos < starts[NB].  (Host-callable so that tests/emul can check it without a GPU.)
UG_HD void next_bucket(const uint32_t* starts, uint32_t NB, uint32_t& key, uint32_t pos, uint32_t& end) {
    key++; end = starts[key + 1];
    if (end > pos) return;
    key++; end = starts[key + 1];
    if (end > pos) return;
    uint32_t a = key + 1, b = NB - 1;
    while (a < b) {
        uint32_t mid = (a + b) >> 1;
        if (starts[mid + 1] > pos) b = mid; else a = mid + 1;
    }
    key = a; end = starts[key + 1];
}

#if defined(__CUDACC__)

UG_D Scalar256 load_scalar(const uint8_t* scalars, const uint32_t* index, uint64_t i) {
    uint64_t src = index ? (uint64_t)index[i] : i;
    const uint4* p = reinterpret_cast<const uint4*>(scalars + src * 32);
    uint4 a = __ldg(p), b = __ldg(p + 1);
    Scalar256 s;
    s.v[0] = a.x; s.v[1] = a.y; s.v[2] = a.z; s.v[3] = a.w;
    s.v[4] = b.x; s.v[5] = b.y; s.v[6] = b.z; s.v[7] = b.w;
    return s;
}

// key_stride / entry_stride select the bucket layout:
//   per-window bucket sets (ad-hoc bases):   key = w*B + |d|-1,  entry = i                 (key_stride = B, entry_stride = 0)
//   one shared bucket set (precomputed 2^(c*w) multiples of every base, msm_build_table):
//                                            key = |d|-1,        entry = w*n_bases + i     (key_stride = 0, entry_stride = n_bases)
// skip: optional bitmap over the points (bit i set = base i is the point at infinity in EVERY array this sort serves): such points
// get no digit entries at all, so the accumulation never sees them (a lane that meets an infinity base in k_accumulate idles
// for a whole mixed addition while the rest of its warp works)
__global__ void k_digits_count(const uint8_t* __restrict__ scalars, const uint32_t* __restrict__ index,
                               uint64_t n, int c, int W, uint32_t key_stride, uint32_t* __restrict__ counts,
                               const uint32_t* __restrict__ skip) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (skip && ((skip[i >> 5] >> (i & 31)) & 1u)) continue;
        Scalar256 s = load_scalar(scalars, index, i);
        scalar_reduce(s);
        scalar_digits(s, c, W, [&](int w, uint32_t b, bool) { atomicAdd(&counts[(uint32_t)w * key_stride + b], 1u); });
    }
}

__global__ void k_digits_scatter(const uint8_t* __restrict__ scalars, const uint32_t* __restrict__ index,
                                 uint64_t n, int c, int W, uint32_t key_stride, uint32_t entry_stride,
                                 uint32_t* __restrict__ cursor, uint32_t* __restrict__ entries, const uint32_t* __restrict__ skip) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        if (skip && ((skip[i >> 5] >> (i & 31)) & 1u)) continue;
        Scalar256 s = load_scalar(scalars, index, i);
        scalar_reduce(s);
        scalar_digits(s, c, W, [&](int w, uint32_t b, bool neg) {
            uint32_t pos = atomicAdd(&cursor[(uint32_t)w * key_stride + b], 1u);
            entries[pos] = ((uint32_t)i + (uint32_t)w * entry_stride) | (neg ? 0x80000000u : 0u);
        });
    }
}

// bitmap[i / 32] bit (i % 32) = 1 iff bases[i] is the point at infinity (AND-ed into an existing bitmap when `combine`); *count += set bits
template <class F>
__global__ void __launch_bounds__(256) k_inf_bitmap(const Affine<F>* __restrict__ bases, uint32_t n, uint32_t* __restrict__ bitmap, int combine,
                                                    uint32_t* __restrict__ count) {
    // one warp owns 32 consecutive points = one bitmap word
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5, wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    uint32_t local = 0;
    for (uint32_t w = wid; w * 32 < n; w += warps) {
        const uint32_t i = w * 32 + lane;
        bool inf = false;
        if (i < n) {
            // the x and y coordinates are all-zero exactly for the point at infinity (zkey convention)
            const uint4* p = reinterpret_cast<const uint4*>(&bases[i]);
            uint32_t o = 0;
#pragma unroll
            for (int k = 0; k < (int)(sizeof(Affine<F>) / 16); k++) { uint4 v = __ldg(p + k); o |= v.x | v.y | v.z | v.w; }
            inf = o == 0;
        }
        uint32_t word = __ballot_sync(0xffffffffu, inf);
        if (lane == 0) {
            if (combine) word &= bitmap[w];
            bitmap[w] = word;
            local += __popc(word);
        }
    }
    if (lane == 0 && local) atomicAdd(count, local);
}

// out = x & y over `words` bitmap words; *count += set bits of the result
__global__ void __launch_bounds__(256) k_bitmap_and(const uint32_t* __restrict__ x, const uint32_t* __restrict__ y, uint32_t words,
                                                    uint32_t* __restrict__ out, uint32_t* __restrict__ count) {
    uint32_t local = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) {
        const uint32_t v = x[i] & y[i];
        out[i] = v;
        local += __popc(v);
    }
    if (local) atomicAdd(count, local);
}

// ---- exclusive scan of the histogram (3 small launches; tile = 2048 counters) -----------------
constexpr int SCAN_BS = 256;
constexpr int SCAN_TILE = 2048;

__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[8];
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < 8 ? warp_sums[lane] : 0;
#pragma unroll
        for (int d = 1; d < 8; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += y;
        }
        if (lane < 8) warp_sums[lane] = s;
    }
    __syncthreads();
    uint32_t base = wid ? warp_sums[wid - 1] : 0;
    if (total) *total = warp_sums[7];
    uint32_t r = base + x - v;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_BS) k_scan_tile_sums(const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ tile_sums) {
    uint32_t base = blockIdx.x * SCAN_TILE;
    uint32_t s = 0;
    for (int k = 0; k < SCAN_TILE / SCAN_BS; k++) {
        uint32_t i = base + k * SCAN_BS + threadIdx.x;
        if (i < n) s += in[i];
    }
    uint32_t tot;
    block_exclusive_scan_256(s, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// single block: exclusive scan of up to 256*per tile sums, also writes the grand total to out_total
__global__ void __launch_bounds__(SCAN_BS) k_scan_tiles(uint32_t* tile_sums, uint32_t ntiles, uint32_t* out_total) {
    uint32_t per = (ntiles + SCAN_BS - 1) / SCAN_BS;
    uint32_t lo = threadIdx.x * per, hi = min(lo + per, ntiles);
    uint32_t s = 0;
    for (uint32_t i = lo; i < hi; i++) s += tile_sums[i];
    uint32_t tot;
    uint32_t ex = block_exclusive_scan_256(s, &tot);
    for (uint32_t i = lo; i < hi; i++) { uint32_t v = tile_sums[i]; tile_sums[i] = ex; ex += v; }
    if (threadIdx.x == 0) *out_total = tot;
}

__global__ void __launch_bounds__(SCAN_BS) k_scan_apply(const uint32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ tile_offs,
                                                       uint32_t* __restrict__ out) {
    // thread t owns 8 consecutive counters of the tile
    uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * 8;
    uint32_t v[8];
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
    uint32_t ex = block_exclusive_scan_256(s, nullptr) + tile_offs[blockIdx.x];
#pragma unroll
    for (int k = 0; k < 8; k++) { if (base + k < n) out[base + k] = ex; ex += v[k]; }
}

// ---- bucket accumulation ----------------------------------------------------------------------
// starts[0..NB] = bucket offsets into entries (starts[NB] = total M).  Thread t owns entries
// [t*L, (t+1)*L), L = ceil(M/T).  Outputs: buckets[key] for runs strictly inside the slice,
// (pkeys, ppts)[2t], [2t+1] for the first and last run of the slice.  key_off / key_none: several MSMs that share one digit sort
// are reduced together (msm.cu: batch): pkeys carry key + key_off (position in the concatenated bucket array), key_none = "no run".
template <class F>
__global__ void __launch_bounds__(128, sizeof(F) == 32 ? 4 : 1) k_accumulate(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ entries,
                                                    const uint32_t* __restrict__ starts, uint32_t NB,
                                                    XYZZ<F>* __restrict__ buckets, uint32_t* __restrict__ pkeys,
                                                    XYZZ<F>* __restrict__ ppts, uint32_t key_off, uint32_t key_none) {
    const uint32_t T = gridDim.x * blockDim.x;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t M = starts[NB];
    const uint32_t L = (uint32_t)(((uint64_t)M + T - 1) / T);
    const uint64_t lo64 = (uint64_t)t * L;
    if (L == 0 || lo64 >= M) {
        pkeys[2 * t] = key_none; pkeys[2 * t + 1] = key_none;
        st_g(&ppts[2 * t], XYZZ<F>::inf());
        st_g(&ppts[2 * t + 1], XYZZ<F>::inf());
        return;
    }
    const uint32_t lo = (uint32_t)lo64;
    const uint32_t hi = (uint32_t)min((uint64_t)M, lo64 + L);
    // smallest key with starts[key+1] > lo
    uint32_t a = 0, b = NB - 1;
    while (a < b) {
        uint32_t mid = (a + b) >> 1;
        if (starts[mid + 1] > lo) b = mid; else a = mid + 1;
    }
    uint32_t key = a;
    uint32_t end = starts[key + 1];
    XYZZ<F> acc = XYZZ<F>::inf();
    bool first = true;
    uint32_t e = entries[lo];
    Affine<F> pt = ld_nc(&bases[e & 0x7fffffffu]);
    for (uint32_t pos = lo; pos < hi; pos++) {
        if (pos == end) {
            if (first) { pkeys[2 * t] = key + key_off; st_g(&ppts[2 * t], acc); first = false; }
            else st_g(&buckets[key], acc);
            acc = XYZZ<F>::inf();
            next_bucket(starts, NB, key, pos, end);
        }
        // prefetch the next point while this one is being added
        uint32_t e_next = 0;
        Affine<F> pt_next;
        bool has_next = pos + 1 < hi;
        if (has_next) { e_next = entries[pos + 1]; pt_next = ld_nc(&bases[e_next & 0x7fffffffu]); }
        pt_madd(acc, pt, (e >> 31) != 0);
        if (has_next) { e = e_next; pt = pt_next; }
    }
    if (first) {
        pkeys[2 * t] = key + key_off; st_g(&ppts[2 * t], acc);
        pkeys[2 * t + 1] = key + key_off; st_g(&ppts[2 * t + 1], XYZZ<F>::inf());
    } else {
        pkeys[2 * t + 1] = key + key_off; st_g(&ppts[2 * t + 1], acc);
    }
}

// Same contract as k_accumulate, with the XYZZ accumulator in shared memory (SmemPoint, curve.cuh) and the next affine
// point prefetched into registers: MINB CTAs of 128 threads per SM instead of the 2 that 255 registers allow for G2.
template <class F, int MINB>
__global__ void __launch_bounds__(128, MINB) k_accumulate_s(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ entries,
                                                            const uint32_t* __restrict__ starts, uint32_t NB,
                                                            XYZZ<F>* __restrict__ buckets, uint32_t* __restrict__ pkeys,
                                                            XYZZ<F>* __restrict__ ppts, uint32_t key_off, uint32_t key_none) {
    extern __shared__ uint4 slab[];
    const SmemPoint<F, 128> acc{slab + threadIdx.x};
    const uint32_t T = gridDim.x * blockDim.x;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t M = starts[NB];
    const uint32_t L = (uint32_t)(((uint64_t)M + T - 1) / T);
    const uint64_t lo64 = (uint64_t)t * L;
    if (L == 0 || lo64 >= M) {
        pkeys[2 * t] = key_none; pkeys[2 * t + 1] = key_none;
        st_g(&ppts[2 * t], XYZZ<F>::inf());
        st_g(&ppts[2 * t + 1], XYZZ<F>::inf());
        return;
    }
    const uint32_t lo = (uint32_t)lo64;
    const uint32_t hi = (uint32_t)min((uint64_t)M, lo64 + L);
    uint32_t a = 0, b = NB - 1;
    while (a < b) {
        uint32_t mid = (a + b) >> 1;
        if (starts[mid + 1] > lo) b = mid; else a = mid + 1;
    }
    uint32_t key = a;
    uint32_t end = starts[key + 1];
    acc.set_inf();
    bool first = true;
    uint32_t e = entries[lo];
    Affine<F> pt = ld_nc(&bases[e & 0x7fffffffu]);
    for (uint32_t pos = lo; pos < hi; pos++) {
        if (pos == end) {
            if (first) { pkeys[2 * t] = key + key_off; st_g(&ppts[2 * t], acc.get()); first = false; }
            else st_g(&buckets[key], acc.get());
            acc.set_inf();
            next_bucket(starts, NB, key, pos, end);
        }
        const bool neg = (e >> 31) != 0;
        const Affine<F> cur = pt;
        if (pos + 1 < hi) { e = entries[pos + 1]; pt = ld_nc(&bases[e & 0x7fffffffu]); }
        pt_madd_at<F>(acc, cur, neg);
    }
    if (first) {
        pkeys[2 * t] = key + key_off; st_g(&ppts[2 * t], acc.get());
        pkeys[2 * t + 1] = key + key_off; st_g(&ppts[2 * t + 1], XYZZ<F>::inf());
    } else {
        pkeys[2 * t + 1] = key + key_off; st_g(&ppts[2 * t + 1], acc.get());
    }
}

// (the measured-and-rejected batched-affine accumulation, k_accumulate_affine, lives in msm_variants.cuh)

// ---- segmented reduction of (key, point) pairs sorted by key -----------------------------------
constexpr int SEG_BS = 128;

template <class F>
__global__ void __launch_bounds__(SEG_BS) k_seg_reduce(const uint32_t* __restrict__ keys, const XYZZ<F>* __restrict__ pts, uint32_t K,
                                                       uint32_t NB, XYZZ<F>* __restrict__ buckets,
                                                       uint32_t* __restrict__ out_keys, XYZZ<F>* __restrict__ out_pts, int final_pass) {
    __shared__ XYZZ<F> sh[SEG_BS];
    __shared__ uint32_t shk[SEG_BS];
    const int tid = threadIdx.x;
    const uint32_t i = blockIdx.x * SEG_BS + tid;
    uint32_t key = NB;
    XYZZ<F> val = XYZZ<F>::inf();
    if (i < K) { key = keys[i]; val = ld_g(&pts[i]); }
    shk[tid] = key;
    for (int d = 1; d < SEG_BS; d <<= 1) {
        sh[tid] = val;
        __syncthreads();
        if (tid >= d && shk[tid - d] == key) {
            XYZZ<F> other = sh[tid - d];
            pt_add(val, other);
        }
        __syncthreads();
    }
    const bool is_last = (tid == SEG_BS - 1) || (shk[tid + 1] != key);
    if (!is_last) return;
    const bool touches_start = (shk[0] == key);
    const bool touches_end = (tid == SEG_BS - 1);
    if (final_pass || !(touches_start || touches_end)) {
        if (key < NB) st_g(&buckets[key], val);
        return;
    }
    const uint32_t s0 = 2 * blockIdx.x, s1 = s0 + 1;
    if (touches_start) {
        out_keys[s0] = key; st_g(&out_pts[s0], val);
        if (touches_end) { out_keys[s1] = key; st_g(&out_pts[s1], XYZZ<F>::inf()); }
    } else {  // touches_end only
        out_keys[s1] = key; st_g(&out_pts[s1], val);
    }
}

// ---- bucket reduction by chunked running sums ---------------------------------------------------
// One window's result is sum_j (j+1) * X[j] over its N buckets.  Thread k owns the chunk [k*m, (k+1)*m) and walks it
// from the top with the classic running sum (2 additions per bucket):
//     S[k] = sum X[j],      R[k] = sum (j - k*m + 1) * X[j]
// so that   sum_j (j+1) X[j] = sum_k R[k] + m * sum_k k * S[k];   the second term goes to k_bit_sums and the host
// applies m = 2^mu and the bit weights by doublings at the very end.
// blockIdx.y selects the window (independent bucket arrays `strideX` apart).
template <class F>
__global__ void __launch_bounds__(128) k_running_sum(const XYZZ<F>* __restrict__ X, uint32_t N, uint32_t strideX, uint32_t m,
                                                     uint32_t K, XYZZ<F>* __restrict__ S, XYZZ<F>* __restrict__ R) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const XYZZ<F>* x = X + (size_t)blockIdx.y * strideX;
    const uint32_t lo = k * m, hi = min(N, lo + m);
    XYZZ<F> acc = XYZZ<F>::inf(), sum = XYZZ<F>::inf();
    for (uint32_t j = hi; j-- > lo;) {
        XYZZ<F> p = ld_g(&x[j]);
        pt_add(acc, p);
        pt_add(sum, acc);
    }
    st_g(&S[(size_t)blockIdx.y * K + k], acc);
    st_g(&R[(size_t)blockIdx.y * K + k], sum);
}

// k_running_sum with both accumulators in shared memory and the bucket read coordinate by coordinate from global.
template <class F, int MINB>
__global__ void __launch_bounds__(128, MINB) k_running_sum_s(const XYZZ<F>* __restrict__ X, uint32_t N, uint32_t strideX, uint32_t m,
                                                             uint32_t K, XYZZ<F>* __restrict__ S, XYZZ<F>* __restrict__ R) {
    extern __shared__ uint4 slab[];
    constexpr int PT = sizeof(XYZZ<F>) / 16;  // uint4 words per point
    const SmemPoint<F, 128> acc{slab + threadIdx.x};
    const SmemPoint<F, 128> sum{slab + PT * 128 + threadIdx.x};
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const XYZZ<F>* x = X + (size_t)blockIdx.y * strideX;
    const uint32_t lo = k * m, hi = min(N, lo + m);
    acc.set_inf();
    sum.set_inf();
    for (uint32_t j = hi; j-- > lo;) {
        pt_add_at<F>(acc, GlobalPoint<F>{x + j});
        pt_add_at<F>(sum, acc);
    }
    st_g(&S[(size_t)blockIdx.y * K + k], acc.get());
    st_g(&R[(size_t)blockIdx.y * K + k], sum.get());
}

// out[w*G + g] = sum of in[w*strideIn + i] over i = g*BS + tid, stepping by G*BS  (plain sums of the R arrays)
constexpr int SUM_BS = 128;
template <class F>
__global__ void __launch_bounds__(SUM_BS) k_sum_points(const XYZZ<F>* __restrict__ in, uint32_t N, uint32_t strideIn,
                                                       XYZZ<F>* __restrict__ out) {
    __shared__ XYZZ<F> sh[SUM_BS];
    const XYZZ<F>* x = in + (size_t)blockIdx.y * strideIn;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t i = blockIdx.x * SUM_BS + threadIdx.x; i < N; i += gridDim.x * SUM_BS) {
        XYZZ<F> p = ld_g(&x[i]);
        pt_add(acc, p);
    }
    for (int d = SUM_BS / 2; d > 0; d >>= 1) {
        sh[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x < d) { XYZZ<F> o = sh[threadIdx.x + d]; pt_add(acc, o); }
        __syncthreads();
    }
    if (threadIdx.x == 0) st_g(&out[(size_t)blockIdx.y * gridDim.x + blockIdx.x], acc);
}

// T[bit][w][g] = sum of S[w*strideS + k] over the k in [1, K) with bit `bit` set (blockIdx.z = bit, .y = window, .x = g).
// The j-th such k is enumerated directly, so every lane adds: k = ((j >> bit) << (bit+1)) | 1<<bit | (j & (2^bit - 1)).
// Output index ((bit*nwin + w)*G + g): with G == 1 this is the final [bit][w] layout.
template <class F>
__global__ void __launch_bounds__(SUM_BS) k_bit_sums(const XYZZ<F>* __restrict__ S, uint32_t K, uint32_t strideS, XYZZ<F>* __restrict__ out) {
    __shared__ XYZZ<F> sh[SUM_BS];
    const uint32_t bit = blockIdx.z;
    const XYZZ<F>* x = S + (size_t)blockIdx.y * strideS;
    const uint32_t lowmask = (1u << bit) - 1u;
    const uint32_t rem = K & ((2u << bit) - 1u);
    const uint32_t count = ((K >> (bit + 1)) << bit) + (rem > (1u << bit) ? rem - (1u << bit) : 0u);
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t j = blockIdx.x * SUM_BS + threadIdx.x; j < count; j += gridDim.x * SUM_BS) {
        const uint32_t k = ((j >> bit) << (bit + 1)) | (1u << bit) | (j & lowmask);
        XYZZ<F> p = ld_g(&x[k]);
        pt_add(acc, p);
    }
    for (int d = SUM_BS / 2; d > 0; d >>= 1) {
        sh[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x < d) { XYZZ<F> o = sh[threadIdx.x + d]; pt_add(acc, o); }
        __syncthreads();
    }
    if (threadIdx.x == 0) st_g(&out[((size_t)bit * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x], acc);
}

// ---- precomputed window tables (resident bases only) ---------------------------------------------
// level k of the table holds 2^(c*k) * P_i in affine form.  k_table_shift: out = 2^c * in (XYZZ, no inversion);
// k_batch_to_affine: RUN points per thread share one field inversion (Montgomery's trick).
template <class F>
__global__ void __launch_bounds__(128) k_table_shift(const Affine<F>* __restrict__ in, uint32_t n, int c, Jac<F>* __restrict__ out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        Affine<F> a = ld_nc(&in[i]);
        Jac<F> p = jac_dbl_affine(a);
        for (int k = 1; k < c; k++) p = jac_dbl(p);
        st_g(&out[i], p);
    }
}

constexpr int AFF_RUN = 8;
template <class F>
__global__ void __launch_bounds__(128) k_batch_to_affine(const Jac<F>* __restrict__ in, uint32_t n, Affine<F>* __restrict__ out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t T = gridDim.x * blockDim.x;
    // thread t owns points t, t+T, t+2T, ... (coalesced), AFF_RUN at a time: one field inversion per run (Montgomery's trick)
    for (uint32_t base = t; base < n; base += T * AFF_RUN) {
        F pre[AFF_RUN];
        F acc = F::one();
#pragma unroll
        for (int k = 0; k < AFF_RUN; k++) {
            const uint32_t i = base + k * T;
            pre[k] = acc;
            if (i < n) {
                F z = ld_g(&in[i].Z);
                if (!z.is_zero()) acc = fmul(acc, z);
            }
        }
        F inv = finv(acc);
#pragma unroll
        for (int k = AFF_RUN - 1; k >= 0; k--) {
            const uint32_t i = base + k * T;
            if (i >= n) continue;
            Jac<F> p = ld_g(&in[i]);
            Affine<F> a = Affine<F>::inf();
            if (!p.Z.is_zero()) {
                F zi = fmul(inv, pre[k]);
                inv = fmul(inv, p.Z);
                F zi2 = fsqr(zi);
                a.x = fmul(p.X, zi2);
                a.y = fmul(p.Y, fmul(zi2, zi));
            }
            st_g(&out[i], a);
        }
    }
}

#endif  // __CUDACC__

}  // namespace ug
