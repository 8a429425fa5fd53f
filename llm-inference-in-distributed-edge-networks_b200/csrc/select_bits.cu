// K3a: importance-ranked bit-width selection (block radix select).
//
// Reference: torch.argsort(importance_values)[:int(ratio*S)] -- a full device sort used only to
// pick the k least important tokens (Experiments/Pythia-70M/pythia_model.py:169-170,
// Experiments/Qwen2-0.5B/qwen_layer_wise.py:57).  Here one CTA per window finds, for every rung
// of the ladder at once, the pivot key with an MSB-first 8-bit radix select over the order-
// preserving integer image of the fp32 scores (4 histogram passes, no sort), resolves ties in
// favour of the lower token index with a block scan, and emits the per-token width map plus the
// exclusive prefix sum of row sizes that the pack kernel uses as row offsets.
#include "common.cuh"

namespace splitq {

constexpr int kSelThreads = 1024;

struct Ladder {
    int n;
    int thresh[SPLITQ_MAX_LADDER];       // cumulative token counts
    int bits[SPLITQ_MAX_LADDER];
    uint32_t rb16[SPLITQ_MAX_LADDER + 1];  // row bytes / 16 per rung; last = raw fp32
};

// fp32 -> uint32 such that a < b  <=>  key(a) < key(b); -0.0 and +0.0 map to the same key
__device__ __forceinline__ uint32_t order_key(float f) {
    uint32_t u = __float_as_uint(f);
    if (u == 0x80000000u) u = 0u;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

struct U4 {
    uint32_t v[4];
};

// exclusive prefix sum of four counters over the whole CTA (thread order)
__device__ __forceinline__ U4 block_exclusive_scan4(U4 x, U4* warp_tot /* smem[32] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    U4 inc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc.v[c], o);
            if (lane >= o) inc.v[c] += t;
        }
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        U4 t = warp_tot[lane];
        U4 ti = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t u = __shfl_up_sync(0xffffffffu, ti.v[c], o);
                if (lane >= o) ti.v[c] += u;
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) ti.v[c] -= t.v[c];
        warp_tot[lane] = ti;
    }
    __syncthreads();
    U4 base = warp_tot[warp];
    U4 out;
#pragma unroll
    for (int c = 0; c < 4; ++c) out.v[c] = base.v[c] + inc.v[c] - x.v[c];
    __syncthreads();  // warp_tot may be reused by the caller
    return out;
}

template <int ITEMS>
__global__ void __launch_bounds__(kSelThreads)
select_bits_kernel(const float* __restrict__ imp, int S, Ladder lad, uint8_t* __restrict__ bits_out,
                   uint32_t* __restrict__ row_off_out) {
    __shared__ uint32_t hist[SPLITQ_MAX_LADDER][256];
    __shared__ uint32_t s_prefix[SPLITQ_MAX_LADDER];
    __shared__ uint32_t s_rem[SPLITQ_MAX_LADDER];
    __shared__ U4 warp_tot[32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t wbase = (size_t)blockIdx.x * S;
    imp += wbase;

    uint32_t key[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int idx = tid * ITEMS + i;
        key[i] = idx < S ? order_key(imp[idx]) : 0xffffffffu;
    }
    if (tid < SPLITQ_MAX_LADDER) {
        s_prefix[tid] = 0u;
        s_rem[tid] = (tid < lad.n && lad.thresh[tid] > 0) ? (uint32_t)(lad.thresh[tid] - 1) : 0u;
    }

    // ---- 4 passes of 8-bit MSB radix select, all rungs at once --------------------------------
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        const uint32_t himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        for (int i = tid; i < SPLITQ_MAX_LADDER * 256; i += kSelThreads) (&hist[0][0])[i] = 0u;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint32_t digit = (key[i] >> shift) & 255u;
            for (int c = 0; c < lad.n; ++c) {
                const bool in = (key[i] & himask) == s_prefix[c];
                // warp-aggregated shared atomics: scores cluster in a few exponent buckets
                const uint32_t peers = __match_any_sync(0xffffffffu, in ? digit : 0xffffffffu);
                if (in && lane == (__ffs(peers) - 1)) atomicAdd(&hist[c][digit], __popc(peers));
            }
        }
        __syncthreads();
        if (warp < lad.n) {
            const int c = warp;
            uint32_t loc[8], sum = 0;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                loc[q] = hist[c][lane * 8 + q];
                sum += loc[q];
            }
            uint32_t inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            const uint32_t rem = s_rem[c];
            uint32_t before = inc - sum;
            const bool mine = before <= rem && rem < inc;
            __syncwarp();
            if (mine) {
                int q = 0;
                while (before + loc[q] <= rem) before += loc[q++];
                s_prefix[c] |= (uint32_t)(lane * 8 + q) << shift;
                s_rem[c] = rem - before;
            }
        }
        __syncthreads();
    }

    // ---- ties: equal-to-pivot elements are taken in index order ---------------------------------
    uint32_t pivot[SPLITQ_MAX_LADDER], need[SPLITQ_MAX_LADDER];
#pragma unroll
    for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) {
        pivot[c] = s_prefix[c];
        need[c] = (c < lad.n && lad.thresh[c] > 0) ? s_rem[c] + 1u : 0u;
    }
    U4 cnt = {{0u, 0u, 0u, 0u}};
#pragma unroll
    for (int i = 0; i < ITEMS; ++i)
#pragma unroll
        for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) cnt.v[c] += (key[i] == pivot[c]) ? 1u : 0u;
    U4 eq_base = block_exclusive_scan4(cnt, warp_tot);

    uint8_t cls[ITEMS];
    U4 mine = {{0u, 0u, 0u, 0u}};  // v[3] = row bytes / 16 of this thread's tokens
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        int k = lad.n;
#pragma unroll
        for (int c = SPLITQ_MAX_LADDER - 1; c >= 0; --c) {
            if (c < lad.n) {
                const bool eq = key[i] == pivot[c];
                const bool sel = need[c] > 0u && (key[i] < pivot[c] || (eq && eq_base.v[c] < need[c]));
                if (eq) eq_base.v[c]++;
                if (sel) k = c;
            }
        }
        cls[i] = (uint8_t)k;
        if (tid * ITEMS + i < S) mine.v[3] += lad.rb16[k];
    }
    U4 off = block_exclusive_scan4(mine, warp_tot);
    uint32_t o16 = off.v[3];
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

}  // namespace splitq

using namespace splitq;

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
    lad.n = n_ladder;
    int cum = 0;
    for (int c = 0; c < SPLITQ_MAX_LADDER; ++c) {
        if (c < n_ladder) {
            cum += ladder_k_host[c];
            lad.thresh[c] = cum;
            lad.bits[c] = ladder_bits_host[c];
            lad.rb16[c] = row_bytes(D, ladder_bits_host[c]) / 16u;
        } else {
            lad.thresh[c] = 0;
            lad.bits[c] = SPLITQ_BITS_RAW;
            lad.rb16[c] = 0;
        }
    }
    lad.rb16[n_ladder] = row_bytes(D, SPLITQ_BITS_RAW) / 16u;
    cudaStream_t st = (cudaStream_t)stream;
#define SPLITQ_SEL(ITEMS) \
    select_bits_kernel<ITEMS><<<B, kSelThreads, 0, st>>>(imp, S, lad, bits_out, row_off_out)
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
