// K3a: importance-ranked bit-width selection (block radix select).
//
// Reference: torch.argsort(importance_values)[:int(ratio*S)] -- a full device sort used only to
// pick the k least important tokens (Experiments/Pythia-70M/pythia_model.py:169-170,
// Experiments/Qwen2-0.5B/qwen_layer_wise.py:57).  Here one CTA per window finds, for every rung
// of the ladder at once, the pivot key with an MSB-first 8-bit radix select over the order-
// preserving integer image of the fp32 scores (4 histogram passes, no sort), resolves ties in
// favour of the lower token index with a block scan, and emits the per-token width map plus the
// exclusive prefix sum of row sizes that the pack kernel uses as row offsets.
//
// What keeps the pass chain short (it is pure latency: S floats per window):
//   * the four histograms of every rung live in shared memory side by side and are cleared once,
//     so a pass is "atomics, barrier, one warp per rung scans 256 bins, barrier";
//   * pass 0 is the same for every rung (no prefix yet): one histogram, one warp-aggregated
//     atomic per key (importance scores share a handful of exponents, plain atomics would
//     serialise); later passes touch only the keys inside a rung's pivot bin -- a whole warp
//     skips them with one vote;
//   * the tie scan is skipped when the last histogram shows that every key equal to a pivot is
//     selected anyway (always, unless scores repeat exactly at a rung boundary).
//
// Second mode ("top-rho", Experiments/Pythia-70M/pythia_model.py:93-109 `batched_top_rho_
// quantization`): keep the fewest most-important tokens whose importance mass reaches a
// threshold and quantise the rest.  The kept count comes from the same radix machinery run from
// the top with a mass histogram beside the count histogram; masses are 2^-47 fixed-point
// integers (values clamped to [0, 1]), so the sums -- and the count -- do not depend on the order of
// the atomics and equal the oracle's integer arithmetic exactly.
#include <mutex>

#include "common.cuh"

namespace splitq {

#ifndef SPLITQ_SEL_THREADS
#define SPLITQ_SEL_THREADS 1024
#endif
// 1024 threads per window: with 512 / 256 the kernel takes 13.8 / 15.0 us instead of 10.4 at S = 4096 (ncu) -- the
// per-key work dominates the per-thread overhead, so more threads per window win as long as they fit one CTA
constexpr int kSelThreads = SPLITQ_SEL_THREADS;
constexpr int kSelWarps = kSelThreads / 32;
static_assert(32 * kSelThreads >= SPLITQ_MAX_SELECT_S, "32 keys per thread must cover the longest window");
constexpr int kSelMaxWindows = 32;  // windows per launch of the per-window modes (kernel-parameter space)

struct Ladder {
    int n;
    int thresh[SPLITQ_MAX_LADDER];       // cumulative token counts
    int bits[SPLITQ_MAX_LADDER];
    uint32_t rb16[SPLITQ_MAX_LADDER + 1];  // row bytes / 16 per rung; last = raw fp32
};
// per-window overrides of the single rung's count (mode 1) or the mass threshold it is derived from (mode 2)
struct PerWindow {
    int mode;                             // 0: ladder as given; 1: k[w] tokens; 2: top-rho with threshold mass[w]
    int k[kSelMaxWindows];
    unsigned long long mass[kSelMaxWindows];  // ceil(threshold * 2^47)
};

// fp32 -> uint32 such that a < b  <=>  key(a) < key(b); -0.0 and +0.0 map to the same key
__device__ __forceinline__ uint32_t order_key(float f) {
    uint32_t u = __float_as_uint(f);
    if (u == 0x80000000u) u = 0u;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_value(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// importance mass of one token as a 2^-47 fixed-point integer: values are clamped to [0, 1] (NaN -> 0)
__device__ __forceinline__ unsigned long long mass_of(float v) {
    v = fminf(fmaxf(v, 0.f), 1.f);                      // fmaxf(NaN, 0) = 0
    return (unsigned long long)(v * 140737488355328.0f);  // exact: a power-of-two scaling, then truncation
}

// exclusive prefix sum of one counter over the whole CTA (thread order); `total` gets the block sum
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t x, uint32_t* warp_tot /* smem[32] */, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t w = lane < kSelWarps ? warp_tot[lane] : 0u;  // every warp scans the warp totals itself: no second barrier
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += t;
    }
    const uint32_t base = __shfl_sync(0xffffffffu, wi - w, warp);
    if (total) *total = __shfl_sync(0xffffffffu, wi, 31);  // lanes past the last warp carry the running total
    __syncthreads();  // warp_tot may be reused by the caller
    return base + inc - x;
}

struct SelShared {
    // Two ways to count, each where it is cheap (cycle counts from a timing build, S = 4096):
    //   pass 0 -- scores share a handful of exponents, so atomics would pile onto a few bins; a warp
    //   finds its few distinct digits with match.any (cost grows with the number of distinct values:
    //   ~2 k cycles here) and one leader lane per digit adds the count to the warp's OWN histogram with
    //   a plain store; the 32 warp histograms are summed column-wise afterwards (counts fit 16 bits);
    //   passes 1-3 -- the digits inside a pivot bin are spread over all 256 values, where match.any costs
    //   ~300 cycles per call (13 k cycles for pass 1 alone) and plain shared atomics hardly collide.
    uint16_t whist[32][256];                    // pass 0: [warp][digit]
    uint32_t hist[4][SPLITQ_MAX_LADDER][256];   // [pass][rung][digit], cleared once; pass 0 uses rung slot 0 for all
    uint32_t tcnt[4][256];                      // top-rho: key counts per digit, per pass
    unsigned long long tmass[4][256];           // top-rho: fixed-point mass per digit, per pass
    uint32_t prefix[SPLITQ_MAX_LADDER], rem[SPLITQ_MAX_LADDER], eqcnt[SPLITQ_MAX_LADDER];
    uint32_t cand_n[SPLITQ_MAX_LADDER], cand_fill[SPLITQ_MAX_LADDER], cand[SPLITQ_MAX_LADDER][32];
    uint32_t warp_tot[32];
    uint32_t t_prefix, t_nhi, t_done, t_keep;
    unsigned long long t_mhi;
};

// one warp: find the digit of `hist` (256 bins, ascending) that holds rank `rem`; returns through lane-uniform refs
__device__ __forceinline__ void scan_bins_ascending(const uint32_t* hist, uint32_t rem, int lane, uint32_t& digit,
                                                    uint32_t& rem_out) {
    uint32_t loc[8], sum = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        loc[q] = hist[lane * 8 + q];
        sum += loc[q];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    uint32_t before = inc - sum;
    const bool mine = before <= rem && rem < inc;
    uint32_t d = 0, r = 0;
    if (mine) {
        int q = 0;
        while (before + loc[q] <= rem) before += loc[q++];
        d = (uint32_t)(lane * 8 + q);
        r = rem - before;
    }
    const uint32_t who = __ballot_sync(0xffffffffu, mine);
    const int src = who ? __ffs(who) - 1 : 0;
    digit = __shfl_sync(0xffffffffu, d, src);
    rem_out = __shfl_sync(0xffffffffu, r, src);
}

template <int ITEMS>
__global__ void __launch_bounds__(kSelThreads)
select_bits_kernel(const float* __restrict__ imp, int S, Ladder lad, const PerWindow pw, uint8_t* __restrict__ bits_out,
                   uint32_t* __restrict__ row_off_out, int* __restrict__ k_out) {
#ifdef SPLITQ_SEL_TIMING
    long long tk[12]; int nt = 0;
#define SEL_TICK() do { __syncthreads(); tk[nt++] = clock64(); } while (0)
#else
#define SEL_TICK() do {} while (0)
#endif
    SEL_TICK();
    extern __shared__ __align__(16) uint8_t sel_smem[];
    SelShared& sh = *reinterpret_cast<SelShared*>(sel_smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t wbase = (size_t)blockIdx.x * S;
    imp += wbase;

    uint32_t key[ITEMS];
    if (ITEMS % 4 == 0 && (S & 3) == 0 && (reinterpret_cast<uintptr_t>(imp) & 15u) == 0) {
#pragma unroll
        for (int i = 0; i < ITEMS; i += 4) {
            const int idx = tid * ITEMS + i;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < S) v = __ldg(reinterpret_cast<const float4*>(imp + idx));
            key[i + 0] = idx + 0 < S ? order_key(v.x) : 0xffffffffu;
            key[i + 1] = idx + 1 < S ? order_key(v.y) : 0xffffffffu;
            key[i + 2] = idx + 2 < S ? order_key(v.z) : 0xffffffffu;
            key[i + 3] = idx + 3 < S ? order_key(v.w) : 0xffffffffu;
        }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const int idx = tid * ITEMS + i;
            key[i] = idx < S ? order_key(__ldg(imp + idx)) : 0xffffffffu;
        }
    }
    {   // clear the atomically filled histograms once (the top-rho ones only when they are used)
        uint32_t* z = &sh.hist[0][0][0];
        const int nz = (int)((sizeof(sh.hist) + (pw.mode == 2 ? sizeof(sh.tcnt) + sizeof(sh.tmass) : 0)) / 4);
        static_assert(offsetof(SelShared, tcnt) == offsetof(SelShared, hist) + sizeof(sh.hist) &&
                      offsetof(SelShared, tmass) == offsetof(SelShared, tcnt) + sizeof(sh.tcnt), "histograms must be contiguous");
        for (int i = tid; i < nz; i += kSelThreads) z[i] = 0u;
        if (tid == 0) { sh.t_prefix = 0u; sh.t_nhi = 0u; sh.t_done = 0u; sh.t_keep = 0u; sh.t_mhi = 0ull; }
        if (pw.mode == 2) __syncthreads();
    }

    SEL_TICK();  // keys loaded
    int thresh[SPLITQ_MAX_LADDER];
#pragma unroll
    for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) thresh[c] = c < lad.n ? lad.thresh[c] : 0;
    if (pw.mode == 1) thresh[0] = pw.k[blockIdx.x];

    // ---- top-rho: how many of the most important tokens reach the mass threshold T -----------------
    if (pw.mode == 2) {
        const unsigned long long T = pw.mass[blockIdx.x];
        for (int pass = 0; pass < 4; ++pass) {
            const int shift = 24 - 8 * pass;
            const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
            const uint32_t pre = sh.t_prefix;
            if (!sh.t_done) {
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    if (tid * ITEMS + i < S && (key[i] & himask) == pre) {
                        const uint32_t digit = (key[i] >> shift) & 255u;
                        atomicAdd(&sh.tcnt[pass][digit], 1u);
                        const unsigned long long m = mass_of(key_value(key[i]));
                        if (m) atomicAdd(&sh.tmass[pass][digit], m);
                    }
                }
            }
            __syncthreads();
            if (warp == 0 && !sh.t_done) {
                // descending order: lane l owns bins 255-8l .. 248-8l
                uint32_t cl[8], csum = 0;
                unsigned long long ml[8], msum = 0;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int bin = 255 - (lane * 8 + q);
                    cl[q] = sh.tcnt[pass][bin];
                    ml[q] = sh.tmass[pass][bin];
                    csum += cl[q];
                    msum += ml[q];
                }
                uint32_t cinc = csum;
                unsigned long long minc = msum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t ct = __shfl_up_sync(0xffffffffu, cinc, o);
                    const unsigned long long mt = __shfl_up_sync(0xffffffffu, minc, o);
                    if (lane >= o) { cinc += ct; minc += mt; }
                }
                const unsigned long long mhi = sh.t_mhi;
                unsigned long long mbefore = mhi + (minc - msum);
                uint32_t cbefore = cinc - csum;
                const bool mine = mbefore < T && T <= mhi + minc;   // the running mass reaches T inside my bins
                const uint32_t who = __ballot_sync(0xffffffffu, mine);
                if (who == 0u) {
                    if (lane == 0) {  // T == 0: nothing is kept; otherwise the whole distribution is lighter than T
                        sh.t_keep = T == 0ull ? 0u : (uint32_t)S;
                        sh.t_done = 1u;
                    }
                } else if (mine) {
                    int q = 0;
                    while (mbefore + ml[q] < T) { mbefore += ml[q]; cbefore += cl[q]; ++q; }
                    const uint32_t digit = (uint32_t)(255 - (lane * 8 + q));
                    sh.t_prefix = pre | (digit << shift);
                    sh.t_mhi = mbefore;
                    sh.t_nhi += cbefore;
                    if (pass == 3) {  // cl[q] keys equal to the pivot, each of mass w > 0 (their bin moved the sum)
                        const unsigned long long w = ml[q] / cl[q];
                        const unsigned long long need = (T - mbefore + w - 1ull) / w;
                        sh.t_keep = sh.t_nhi + (uint32_t)(need < cl[q] ? need : cl[q]);
                        sh.t_done = 1u;
                    }
                }
            }
            __syncthreads();
        }
        thresh[0] = S - (int)sh.t_keep;
    }
    if (k_out && tid == 0) k_out[blockIdx.x] = thresh[0];
    if (tid < SPLITQ_MAX_LADDER) {
        sh.cand_fill[tid] = 0u;
        sh.cand_n[tid] = 0xffffffffu;
        sh.prefix[tid] = 0u;
        sh.rem[tid] = thresh[tid] > 0 ? (uint32_t)(thresh[tid] - 1) : 0u;
    }

    // ---- 4 passes of 8-bit MSB radix select, all rungs at once --------------------------------
    {   // pass 0: no rung has a prefix yet -> one histogram for all rungs, counted per warp
        uint16_t* mine = sh.whist[warp];
        uint32_t* z = reinterpret_cast<uint32_t*>(mine);
#pragma unroll
        for (int i = 0; i < 4; ++i) z[lane + 32 * i] = 0u;
        __syncwarp();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint32_t digit = key[i] >> 24;
            const uint32_t peers = __match_any_sync(0xffffffffu, digit);
            if (lane == (__ffs(peers) - 1)) mine[digit] = (uint16_t)(mine[digit] + __popc(peers));
            __syncwarp();
        }
        __syncthreads();
        SEL_TICK();
        if (tid < 256) {
            uint32_t sum = 0;
#pragma unroll 8
            for (int w = 0; w < kSelWarps; ++w) sum += sh.whist[w][tid];
            sh.hist[0][0][tid] = sum;
        }
        __syncthreads();
        if (warp < lad.n) {
            uint32_t digit, rem;
            scan_bins_ascending(sh.hist[0][0], sh.rem[warp], lane, digit, rem);
            if (lane == 0) { sh.prefix[warp] = digit << 24; sh.rem[warp] = rem; }
        }
        __syncthreads();
    }
    bool finished = false;  // block-uniform: the shortcut below has already produced the pivots
    for (int pass = 1; pass < 4 && !finished; ++pass) {
        const int shift = 24 - 8 * pass;
        const uint32_t himask = 0xffffffffu << (shift + 8);
        if (pass == 2) {
            // After two passes a pivot bin usually holds a handful of keys (4096 scores over 65536 bins).  If
            // every rung's bin holds at most 32, the two remaining histogram passes are replaced by a direct
            // ranking: the candidates of rung c are gathered into a short list and warp c ranks them with
            // shuffles (count of smaller keys = rank, count of equal keys = what the tie rule needs).
            bool few = true;
            for (int c = 0; c < lad.n; ++c) few = few && sh.cand_n[c] <= 32u;
            if (few) {
                for (int c = 0; c < lad.n; ++c) {
                    const uint32_t pre = sh.prefix[c];
#pragma unroll
                    for (int i = 0; i < ITEMS; ++i)
                        if ((key[i] & himask) == pre) sh.cand[c][atomicAdd(&sh.cand_fill[c], 1u)] = key[i];
                }
                __syncthreads();
                if (warp < lad.n) {
                    const uint32_t n = sh.cand_n[warp], mine = lane < n ? sh.cand[warp][lane] : 0xffffffffu;
                    uint32_t smaller = 0, equal = 0;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const uint32_t other = __shfl_sync(0xffffffffu, mine, j);
                        if ((uint32_t)j < n) { smaller += other < mine; equal += other == mine; }
                    }
                    const uint32_t rem = sh.rem[warp];
                    if (lane < n && smaller <= rem && rem < smaller + equal) {  // all lanes holding the pivot key agree
                        sh.prefix[warp] = mine;
                        sh.rem[warp] = rem - smaller;
                        sh.eqcnt[warp] = equal;
                    }
                }
                __syncthreads();
                finished = true;
                break;
            }
        }
        for (int c = 0; c < lad.n; ++c) {
            const uint32_t pre = sh.prefix[c];
#pragma unroll
            for (int i = 0; i < ITEMS; ++i)
                if ((key[i] & himask) == pre) atomicAdd(&sh.hist[pass][c][(key[i] >> shift) & 255u], 1u);
        }
        __syncthreads();
        SEL_TICK();
        if (warp < lad.n) {
            uint32_t digit, rem;
            scan_bins_ascending(sh.hist[pass][warp], sh.rem[warp], lane, digit, rem);
            if (lane == 0) {
                sh.prefix[warp] |= digit << shift;
                sh.rem[warp] = rem;
                if (pass == 1) sh.cand_n[warp] = sh.hist[1][warp][digit];  // keys left in the pivot bin
                if (pass == 3) sh.eqcnt[warp] = sh.hist[3][warp][digit];  // keys equal to the pivot
            }
        }
        __syncthreads();
    }

    SEL_TICK();  // passes done
    // ---- ties: equal-to-pivot keys are taken in index order.  The last histogram knows how many keys
    // equal each pivot; if all of them are wanted there is nothing to break.
    uint32_t pivot[SPLITQ_MAX_LADDER], need[SPLITQ_MAX_LADDER], eq_base[SPLITQ_MAX_LADDER];
    bool ties = false;
#pragma unroll
    for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) {
        pivot[c] = sh.prefix[c];
        need[c] = (c < lad.n && thresh[c] > 0) ? sh.rem[c] + 1u : 0u;
        eq_base[c] = 0u;
        if (need[c] > 0u && sh.eqcnt[c] != need[c]) ties = true;
    }
    if (ties) {  // block-uniform
#pragma unroll
        for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) {
            uint32_t cnt = 0;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) cnt += (key[i] == pivot[c]) ? 1u : 0u;
            eq_base[c] = block_exclusive_scan(cnt, sh.warp_tot, nullptr);
        }
    }

    uint8_t cls[ITEMS];
    uint32_t mine = 0;  // row bytes / 16 of this thread's tokens
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        int k = lad.n;
#pragma unroll
        for (int c = SPLITQ_MAX_LADDER - 1; c >= 0; --c) {
            if (c < lad.n) {
                const bool eq = key[i] == pivot[c];
                const bool sel = need[c] > 0u && (key[i] < pivot[c] || (eq && eq_base[c] < need[c]));
                if (eq) eq_base[c]++;
                if (sel) k = c;
            }
        }
        cls[i] = (uint8_t)k;
        if (tid * ITEMS + i < S) mine += lad.rb16[k];
    }
    SEL_TICK();  // classified
    uint32_t o16 = block_exclusive_scan(mine, sh.warp_tot, nullptr);
    SEL_TICK();  // offsets scanned
    if (ITEMS % 4 == 0 && (S & 3) == 0 && ((wbase & 3) == 0) &&
        ((reinterpret_cast<uintptr_t>(bits_out) & 3u) | (reinterpret_cast<uintptr_t>(row_off_out) & 15u)) == 0) {
#pragma unroll
        for (int i = 0; i < ITEMS; i += 4) {
            const int idx = tid * ITEMS + i;
            uint32_t b4 = 0;
            uint4 o4;
            uint32_t* o = &o4.x;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = cls[i + j];
                b4 |= (uint32_t)(k < lad.n ? lad.bits[k] : SPLITQ_BITS_RAW) << (8 * j);
                o[j] = o16;
                o16 += lad.rb16[k];
            }
            if (idx < S) {
                *reinterpret_cast<uint32_t*>(bits_out + wbase + idx) = b4;
                *reinterpret_cast<uint4*>(row_off_out + wbase + idx) = o4;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const int idx = tid * ITEMS + i;
            if (idx < S) {
                const int k = cls[i];
                bits_out[wbase + idx] = (uint8_t)(k < lad.n ? lad.bits[k] : SPLITQ_BITS_RAW);
                row_off_out[wbase + idx] = o16;
                o16 += lad.rb16[k];
            }
        }
    }
#ifdef SPLITQ_SEL_TIMING
    SEL_TICK();
    if (blockIdx.x == 0 && tid == 0) {
        printf("select_bits timing (cycles since start):");
        for (int i = 1; i < nt; ++i) printf(" %lld", tk[i] - tk[0]);
        printf("\n");
    }
#endif
}

}  // namespace splitq

using namespace splitq;

static int launch_select(const float* imp, int nwin, int S, const Ladder& lad, const PerWindow& pw, uint8_t* bits_out,
                         uint32_t* row_off_out, int* k_out, cudaStream_t st) {
#define SPLITQ_SEL(ITEMS) \
    do {                                                                                                  \
        static std::once_flag once[64];                                                                   \
        int dev = 0;                                                                                      \
        SPLITQ_CUDA(cudaGetDevice(&dev));                                                                 \
        cudaError_t attr_err = cudaSuccess;                                                               \
        std::call_once(once[dev & 63], [&] {                                                              \
            attr_err = cudaFuncSetAttribute(select_bits_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                            (int)sizeof(SelShared));                                      \
        });                                                                                               \
        SPLITQ_CUDA(attr_err);                                                                            \
        select_bits_kernel<ITEMS><<<nwin, kSelThreads, sizeof(SelShared), st>>>(imp, S, lad, pw, bits_out, row_off_out, k_out); \
    } while (0)
    if (S <= 1 * kSelThreads) SPLITQ_SEL(1);
    else if (S <= 2 * kSelThreads) SPLITQ_SEL(2);
    else if (S <= 4 * kSelThreads) SPLITQ_SEL(4);
    else if (S <= 8 * kSelThreads) SPLITQ_SEL(8);
    else if (S <= 16 * kSelThreads) SPLITQ_SEL(16);
    else SPLITQ_SEL(32);
#undef SPLITQ_SEL
    SPLITQ_LAUNCH_CHECK("select_bits_kernel");
    return SPLITQ_OK;
}

static void fill_ladder(Ladder& lad, int D, const int* k, const int* bits, int n) {
    lad.n = n;
    int cum = 0;
    for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) {
        if (c < n) {
            cum += k[c];
            lad.thresh[c] = cum;
            lad.bits[c] = bits[c];
            lad.rb16[c] = row_bytes(D, bits[c]) / 16u;
        } else {
            lad.thresh[c] = 0;
            lad.bits[c] = SPLITQ_BITS_RAW;
            lad.rb16[c] = 0;
        }
    }
    lad.rb16[n] = row_bytes(D, SPLITQ_BITS_RAW) / 16u;
}

extern "C" int splitq_select_bits(const float* imp, int B, int S, int D, const int* ladder_k_host,
                                  const int* ladder_bits_host, int n_ladder, uint8_t* bits_out,
                                  uint32_t* row_off_out, splitq_stream_t stream) {
    SPLITQ_REQUIRE(imp && bits_out && row_off_out, SPLITQ_ERR_ARG, "select_bits: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0, SPLITQ_ERR_ARG, "select_bits: bad shape B=%d S=%d D=%d", B, S, D);
    SPLITQ_REQUIRE(S <= SPLITQ_MAX_SELECT_S, SPLITQ_ERR_ARG, "select_bits: S=%d > %d", S,
                   SPLITQ_MAX_SELECT_S);
    if (int rc = validate_ladder(S, ladder_k_host, ladder_bits_host, n_ladder)) return rc;
    if (int rc = check_arch()) return rc;
    Ladder lad;
    fill_ladder(lad, D, ladder_k_host, ladder_bits_host, n_ladder);
    PerWindow pw;
    pw.mode = 0;
    return launch_select(imp, B, S, lad, pw, bits_out, row_off_out, nullptr, (cudaStream_t)stream);
}

extern "C" int splitq_select_bits_per_window(const float* imp, int B, int S, int D, int bits, const int* k_host,
                                             const double* mass_threshold_host, uint8_t* bits_out,
                                             uint32_t* row_off_out, int* k_out, splitq_stream_t stream) {
    SPLITQ_REQUIRE(imp && bits_out && row_off_out, SPLITQ_ERR_ARG, "select_bits_per_window: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0, SPLITQ_ERR_ARG, "select_bits_per_window: bad shape B=%d S=%d D=%d", B, S, D);
    SPLITQ_REQUIRE(S <= SPLITQ_MAX_SELECT_S, SPLITQ_ERR_ARG, "select_bits_per_window: S=%d > %d", S, SPLITQ_MAX_SELECT_S);
    SPLITQ_REQUIRE((k_host != nullptr) != (mass_threshold_host != nullptr), SPLITQ_ERR_ARG,
                   "select_bits_per_window: give either per-window counts or per-window mass thresholds");
    SPLITQ_REQUIRE(bits == 2 || bits == 4 || bits == 8, SPLITQ_ERR_ARG, "select_bits_per_window: width %d not in {2,4,8}", bits);
    if (int rc = check_arch()) return rc;
    const int one_k = 0;
    Ladder lad;
    fill_ladder(lad, D, &one_k, &bits, 1);
    for (int w0 = 0; w0 < B; w0 += kSelMaxWindows) {
        const int n = B - w0 < kSelMaxWindows ? B - w0 : kSelMaxWindows;
        PerWindow pw;
        pw.mode = k_host ? 1 : 2;
        for (int i = 0; i < n; ++i) {
            pw.k[i] = 0;
            pw.mass[i] = 0ull;
            if (k_host) {
                SPLITQ_REQUIRE(k_host[w0 + i] >= 0 && k_host[w0 + i] <= S, SPLITQ_ERR_ARG,
                               "select_bits_per_window: count %d of window %d not in [0,%d]", k_host[w0 + i], w0 + i, S);
                pw.k[i] = k_host[w0 + i];
            } else {
                const double t = mass_threshold_host[w0 + i];
                SPLITQ_REQUIRE(t == t && t <= 32768.0, SPLITQ_ERR_ARG,
                               "select_bits_per_window: mass threshold of window %d is not a number <= 32768", w0 + i);
                // ceil(t * 2^47) in exact integer arithmetic on the double's value
                const double scaled = t <= 0.0 ? 0.0 : t * 140737488355328.0;
                unsigned long long q = (unsigned long long)scaled;
                if ((double)q < scaled) ++q;
                pw.mass[i] = q;
            }
        }
        if (int rc = launch_select(imp + (size_t)w0 * S, n, S, lad, pw, bits_out + (size_t)w0 * S,
                                   row_off_out + (size_t)w0 * S, k_out ? k_out + w0 : nullptr, (cudaStream_t)stream))
            return rc;
    }
    return SPLITQ_OK;
}
