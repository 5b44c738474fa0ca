// Measured-and-rejected MSM kernel variants (opt-in through UGB_ACC_G1=24 / UGB_ACC_G2=33, used by NO default path):
//   k_accumulate_affine ... batched-affine bucket accumulation with a CTA-wide shared inversion: bit-exact, 12.0 vs 9.0 ms (G1) and
//                           29.5 vs 28.0 ms (G2) in round 1 (profiles/r01_ncu_affine.txt; DESIGN.md section 4 for why the
//                           level-synchronous formulation was costed and not built either)
#pragma once
#include "msm.cuh"

namespace ug {
#if defined(__CUDACC__)

// ---- bucket accumulation in AFFINE coordinates with batched inversion -------------------------------------------
// An affine addition costs 3 products once 1/(x2-x1) is known, and Montgomery's trick shares ONE field inversion among
// all additions of a batch for 3 more products each: 6 products per addition instead of the 10 of madd-2008-s (Fq2: 17
// instead of 28).  Additions inside a bucket are dependent, so the batch is built level by level: every thread owns a
// slice of the sorted entry list (as k_accumulate), pairs up neighbours of the same bucket, and one level halves every
// run.  Per level and CTA:
//   pre-pass   pairs[p] = index of the first operand of pair p                         (keys only)
//   phase A    p descending: d_p = x2-x1 (2y for a doubling, 1 for the degenerate cases), suf[p] = prod_{q>p} d_q
//   inversion  CTA-wide product tree in shared memory, ONE Fermat inversion, inverses pushed back down the tree
//   phase B    p ascending: 1/d_p = u*suf[p], u *= d_p; the sum replaces its operands IN PLACE (output index <= input)
// until no CTA thread has a pair left.  What remains is one affine point per run, written out under k_accumulate's
// contract (XYZZ buckets, first/last run of the slice as (key, partial) pairs for k_seg_reduce).
// STATUS (round 1, B200, 2^22 UltraGroth): correct (all MSM / prover parity tests pass with UGB_ACC_G1=24 UGB_ACC_G2=33)
// but SLOWER than k_accumulate -- 12.0 vs 9.0 ms per G1 MSM, 29.5 vs 28.0 ms for G2 -- so it is NOT the default: ncu
// shows the integer pipe only 48 % busy, long-scoreboard stalls on the dependent loads (pairs -> entries -> bases) and
// 50 % more instructions than the modmul count explains (profiles/r01_ncu_affine.txt).  A software-pipelined variant
// spilled and was slower still.  Kept as an opt-in variant; see DESIGN.md section 4 for what it would take.
// Scratch is laid out word-interleaved per CTA (word w of element j of thread t at ((j*NW + w)*BS + t)), so a warp's
// access to "its" element j is one contiguous 512-byte transaction.
constexpr int BA_BS = 128;

template <class T, int BS>
struct Interleaved {  // per-CTA array of per-thread lists of T (sizeof(T) % 16 == 0)
    uint4* base;      // CTA region + threadIdx.x
    static constexpr int NW = sizeof(T) / 16;
    __device__ __forceinline__ T ld(uint32_t j) const {
        T r;
        uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
        for (int w = 0; w < NW; w++) d[w] = base[((size_t)j * NW + w) * BS];
        return r;
    }
    __device__ __forceinline__ void st(uint32_t j, const T& v) const {
        const uint4* s = reinterpret_cast<const uint4*>(&v);
#pragma unroll
        for (int w = 0; w < NW; w++) base[((size_t)j * NW + w) * BS] = s[w];
    }
};

// denominator of the affine addition P + Q and its kind: 0 chord (x2 - x1), 1 tangent (2*y1), 2 degenerate (1)
template <class F>
__device__ __forceinline__ int ba_denominator(const F& x1, const F& x2, const F& y1, const F& y2, F& d) {
    if (!(x1 == x2)) { d = fsub(x2, x1); return 0; }
    if (y1 == y2 && !y1.is_zero()) { d = fdbl(y1); return 1; }
    d = F::one();
    return 2;
}

template <class F>
__device__ __forceinline__ Affine<F> ba_add(const Affine<F>& P, const Affine<F>& Q, int kind, const F& inv) {
    if (P.is_inf()) return Q;
    if (Q.is_inf()) return P;
    if (kind == 2) return Affine<F>::inf();
    F lam;
    if (kind == 0) {
        lam = fmul(fsub(Q.y, P.y), inv);
    } else {
        F xx = fsqr(P.x);
        lam = fmul(fadd(fdbl(xx), xx), inv);
    }
    Affine<F> R;
    R.x = fsub(fsub(fsqr(lam), P.x), Q.x);
    R.y = fsub(fmul(lam, fsub(P.x, R.x)), P.y);
    return R;
}

// CTA-wide batched inversion: every thread contributes v != 0 and gets 1/v back.  tree: 2*BA_BS elements of shared memory.
template <class F>
__device__ __forceinline__ F ba_block_invert(const F& v, F* tree) {
    const int t = threadIdx.x;
    tree[BA_BS + t] = v;
    __syncthreads();
    for (int s = BA_BS / 2; s >= 1; s >>= 1) {
        if (t < s) tree[s + t] = fmul(tree[2 * (s + t)], tree[2 * (s + t) + 1]);
        __syncthreads();
    }
    if (t == 0) tree[1] = finv(tree[1]);
    __syncthreads();
    for (int s = 1; s < BA_BS; s <<= 1) {
        if (t < s) {
            const int k = s + t;
            const F ik = tree[k], a = tree[2 * k], b = tree[2 * k + 1];
            tree[2 * k] = fmul(ik, b);
            tree[2 * k + 1] = fmul(ik, a);
        }
        __syncthreads();
    }
    F r = tree[BA_BS + t];
    __syncthreads();
    return r;
}

template <class F, int MINB>
__global__ void __launch_bounds__(BA_BS, MINB) k_accumulate_affine(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ entries,
                                                                   const uint32_t* __restrict__ starts, uint32_t NB,
                                                                   XYZZ<F>* __restrict__ buckets, uint32_t* __restrict__ pkeys,
                                                                   XYZZ<F>* __restrict__ ppts, uint32_t key_off, uint32_t key_none,
                                                                   uint32_t Lmax, uint4* __restrict__ list_mem,
                                                                   uint4* __restrict__ suf_mem, uint32_t* __restrict__ keys_mem,
                                                                   uint32_t* __restrict__ pairs_mem) {
    __shared__ F tree[2 * BA_BS];
    const uint32_t T = gridDim.x * BA_BS;
    const uint32_t t = blockIdx.x * BA_BS + threadIdx.x;
    const uint32_t M = starts[NB];
    const uint32_t L = (uint32_t)(((uint64_t)M + T - 1) / T);  // <= Lmax by construction (host sizes scratch for max M)
    const uint64_t lo64 = (uint64_t)t * L;
    const uint32_t lo = (uint32_t)min((uint64_t)M, lo64);
    const uint32_t hi = (uint32_t)min((uint64_t)M, lo64 + L);
    uint32_t n = hi - lo;  // may be 0: the thread still takes part in the CTA-wide steps

    constexpr int NWP = sizeof(Affine<F>) / 16, NWF = sizeof(F) / 16;
    const Interleaved<Affine<F>, BA_BS> list{list_mem + (size_t)blockIdx.x * Lmax * NWP * BA_BS + threadIdx.x};
    const Interleaved<F, BA_BS> suf{suf_mem + (size_t)blockIdx.x * (Lmax / 2 + 1) * NWF * BA_BS + threadIdx.x};
    uint32_t* const K = keys_mem + (size_t)blockIdx.x * Lmax * BA_BS + threadIdx.x;        // K[j*BA_BS]
    uint32_t* const PR = pairs_mem + (size_t)blockIdx.x * (Lmax / 2 + 1) * BA_BS + threadIdx.x;

    // keys of the slice (bucket of every entry), from the bucket offsets
    if (n) {
        uint32_t a = 0, b = NB - 1;
        while (a < b) {
            uint32_t mid = (a + b) >> 1;
            if (starts[mid + 1] > lo) b = mid; else a = mid + 1;
        }
        uint32_t key = a, end = starts[key + 1];
        for (uint32_t i = 0; i < n; i++) {
            while (end <= lo + i) { key++; end = starts[key + 1]; }
            K[(size_t)i * BA_BS] = key;
        }
    }

    bool first_level = true;
    // element i of the current level: level 1 reads the signed base point, later levels the in-place list
    auto fetch = [&](uint32_t i) -> Affine<F> {
        if (first_level) {
            const uint32_t e = entries[lo + i];
            Affine<F> p = ld_nc(&bases[e & 0x7fffffffu]);
            if (e >> 31) p.y = fneg(p.y);
            return p;
        }
        return list.ld(i);
    };

    while (true) {
        // pre-pass: neighbours with equal keys pair up
        uint32_t np = 0;
        for (uint32_t i = 0; i + 1 < n;) {
            if (K[(size_t)i * BA_BS] == K[(size_t)(i + 1) * BA_BS]) { PR[(size_t)np * BA_BS] = i; np++; i += 2; }
            else i++;
        }
        if (!__syncthreads_or(np > 0)) break;
        // phase A: suffix products of the denominators
        F run = F::one();
        for (uint32_t p = np; p-- > 0;) {
            const uint32_t i = PR[(size_t)p * BA_BS];
            const Affine<F> P = fetch(i), Q = fetch(i + 1);
            F d;
            ba_denominator(P.x, Q.x, P.y, Q.y, d);
            suf.st(p, run);
            run = fmul(run, d);
        }
        F u = ba_block_invert(run, tree);
        // phase B: additions, in place (output index <= input index)
        uint32_t o = 0, p = 0;
        for (uint32_t i = 0; i < n;) {
            const uint32_t key = K[(size_t)i * BA_BS];
            const bool paired = p < np && PR[(size_t)p * BA_BS] == i;
            Affine<F> R;
            if (paired) {
                const Affine<F> P = fetch(i), Q = fetch(i + 1);
                F d;
                const int kind = ba_denominator(P.x, Q.x, P.y, Q.y, d);
                const F inv = fmul(u, suf.ld(p));
                u = fmul(u, d);
                R = ba_add(P, Q, kind, inv);
                p++;
                i += 2;
            } else {
                R = fetch(i);
                i++;
            }
            list.st(o, R);
            K[(size_t)o * BA_BS] = key;
            o++;
        }
        n = o;
        first_level = false;
    }
    // no level ran at all (every run of every thread had one entry): materialise the list once
    if (first_level) {
        for (uint32_t i = 0; i < n; i++) list.st(i, fetch(i));
        first_level = false;
    }

    // one point per run: interior runs are complete buckets, the first and last run may continue in a neighbour's slice
    if (n == 0) {
        pkeys[2 * t] = key_none; pkeys[2 * t + 1] = key_none;
        st_g(&ppts[2 * t], XYZZ<F>::inf());
        st_g(&ppts[2 * t + 1], XYZZ<F>::inf());
        return;
    }
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t key = K[(size_t)i * BA_BS];
        const XYZZ<F> v = XYZZ<F>::from_affine(list.ld(i));
        if (i == 0) { pkeys[2 * t] = key + key_off; st_g(&ppts[2 * t], v); }
        else if (i == n - 1) { pkeys[2 * t + 1] = key + key_off; st_g(&ppts[2 * t + 1], v); }
        else st_g(&buckets[key], v);
    }
    if (n == 1) { pkeys[2 * t + 1] = K[0] + key_off; st_g(&ppts[2 * t + 1], XYZZ<F>::inf()); }
}

#endif  // __CUDACC__
}  // namespace ug
