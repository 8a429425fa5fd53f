// K1: per-token attention importance from materialised attention probabilities.
//
//   col[b][j]  = (1/Sq) * sum_h w_h * sum_i A[b,h,i,j]     ("regular_importance" with w_h = 1/H,
//                                                            "weighted_importance" with given w_h)
//   last[b][j] = sum_h w_h * A[b,h,Sq-1,j]                  ("last_row")
//
// Reference: Experiments/Pythia-70M/last_row_exp.py:13-30, Experiments/Qwen2-0.5B/main.py:46-86,
// which run two torch.mean reductions (and a Python loop over heads for the weighted form).
//
// HBM-bound: A is read exactly once with 128-bit streaming loads, a warp covering 512 contiguous
// bytes of one attention row; partial column sums live in registers, are combined per CTA through
// shared memory and written as one row-stripe partial; a second tiny kernel adds the stripes in a
// fixed order (deterministic, no float atomics) and applies 1/Sq.
#include "common.cuh"

namespace splitq {

struct HeadWeights {
    float w[SPLITQ_MAX_HEADS];
};

constexpr int kColTile = 128;  // columns per CTA: 32 lanes x float4
constexpr int kWarps = 8;
constexpr int kUnroll = 8;

template <bool VEC>
__global__ void __launch_bounds__(kWarps * 32)
colsum_partial_kernel(const float* __restrict__ A, int H, int Sq, int Sk, int rows_per_chunk, int nchunks,
                      HeadWeights hw, float* __restrict__ partial, unsigned flags) {
    const int tile = blockIdx.x;
    const int h = blockIdx.y / nchunks;
    const int chunk = blockIdx.y - h * nchunks;
    const int b = blockIdx.z;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int j0 = tile * kColTile + lane * 4;

    int r0 = chunk * rows_per_chunk;
    const int r1 = min(Sq, r0 + rows_per_chunk);
    // causal hint: rows above the diagonal block hold only zeros for this column tile
    if ((flags & SPLITQ_FLAG_CAUSAL_HINT) && r0 < tile * kColTile) r0 = min(r1, tile * kColTile);

    const float* base = A + ((size_t)(b * H + h) * Sq) * Sk;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);

    if (VEC) {
        if (j0 < Sk) {
            const float4* p = reinterpret_cast<const float4*>(base + (size_t)(r0 + warp) * Sk + j0);
            const size_t step = (size_t)kWarps * Sk / 4;  // in float4
            int r = r0 + warp;
            for (; r + (kUnroll - 1) * kWarps < r1; r += kUnroll * kWarps) {
                float4 v[kUnroll];
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) v[u] = ld_stream_f4(p + u * step);
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
                }
                p += kUnroll * step;
            }
            for (; r < r1; r += kWarps) {
                float4 v = ld_stream_f4(p);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                p += step;
            }
        }
    } else {
        for (int r = r0 + warp; r < r1; r += kWarps) {
            const float* row = base + (size_t)r * Sk;
            if (j0 + 0 < Sk) acc.x += __ldg(row + j0 + 0);
            if (j0 + 1 < Sk) acc.y += __ldg(row + j0 + 1);
            if (j0 + 2 < Sk) acc.z += __ldg(row + j0 + 2);
            if (j0 + 3 < Sk) acc.w += __ldg(row + j0 + 3);
        }
    }

    __shared__ float4 red[kWarps][32];
    red[warp][lane] = acc;
    __syncthreads();
    if (warp == 0) {
        float4 s = red[0][lane];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) {
            float4 v = red[w][lane];
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        const float wh = hw.w[h];
        float* out = partial + ((size_t)b * gridDim.y + blockIdx.y) * Sk;
        if (j0 + 0 < Sk) out[j0 + 0] = s.x * wh;
        if (j0 + 1 < Sk) out[j0 + 1] = s.y * wh;
        if (j0 + 2 < Sk) out[j0 + 2] = s.z * wh;
        if (j0 + 3 < Sk) out[j0 + 3] = s.w * wh;
    }
}

__global__ void __launch_bounds__(256)
importance_finalize_kernel(const float* __restrict__ A, const float* __restrict__ partial, int H, int Sq,
                           int Sk, int nparts, HeadWeights hw, float* __restrict__ out_col,
                           float* __restrict__ out_last) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (j >= Sk) return;
    if (out_col) {
        const float* p = partial + (size_t)b * nparts * Sk + j;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int q = 0;
        for (; q + 3 < nparts; q += 4) {
            s0 += p[(size_t)(q + 0) * Sk];
            s1 += p[(size_t)(q + 1) * Sk];
            s2 += p[(size_t)(q + 2) * Sk];
            s3 += p[(size_t)(q + 3) * Sk];
        }
        for (; q < nparts; ++q) s0 += p[(size_t)q * Sk];
        out_col[(size_t)b * Sk + j] = ((s0 + s1) + (s2 + s3)) / (float)Sq;
    }
    if (out_last) {
        float s = 0.f;
        for (int h = 0; h < H; ++h)
            s += hw.w[h] * __ldg(A + (((size_t)(b * H + h) * Sq) + (Sq - 1)) * Sk + j);
        out_last[(size_t)b * Sk + j] = s;
    }
}

// rows per stripe: as large as possible while the grid still fills the 148 SMs several times over
static int pick_rows_per_chunk(int B, int H, int Sq, int Sk) {
    const long long tiles = (long long)B * H * cdiv(Sk, kColTile);
    int R = 512;
    while (R > 32 && tiles * cdiv(Sq, R) < 148LL * 8) R >>= 1;
    return R;
}

}  // namespace splitq

using namespace splitq;

extern "C" size_t splitq_importance_colmean_workspace_bytes(int B, int H, int Sq, int Sk) {
    if (B <= 0 || H <= 0 || Sq <= 0 || Sk <= 0) return 0;
    const int R = pick_rows_per_chunk(B, H, Sq, Sk);
    return (size_t)B * H * cdiv(Sq, R) * Sk * sizeof(float);
}

extern "C" int splitq_importance_colmean(const float* A, int B, int H, int Sq, int Sk,
                                         const float* head_w_host, float* out_col, float* out_last,
                                         void* ws, size_t ws_bytes, unsigned flags,
                                         splitq_stream_t stream) {
    SPLITQ_REQUIRE(A && (out_col || out_last), SPLITQ_ERR_ARG, "importance_colmean: NULL input/output");
    SPLITQ_REQUIRE(B > 0 && H > 0 && Sq > 0 && Sk > 0, SPLITQ_ERR_ARG,
                   "importance_colmean: bad shape B=%d H=%d Sq=%d Sk=%d", B, H, Sq, Sk);
    SPLITQ_REQUIRE(H <= SPLITQ_MAX_HEADS, SPLITQ_ERR_ARG, "importance_colmean: H=%d > %d", H,
                   SPLITQ_MAX_HEADS);
    SPLITQ_REQUIRE(B <= 65535, SPLITQ_ERR_ARG, "importance_colmean: B=%d > 65535", B);
    SPLITQ_REQUIRE(!(flags & SPLITQ_FLAG_CAUSAL_HINT) || Sq == Sk, SPLITQ_ERR_ARG,
                   "importance_colmean: causal hint needs a square map");
    if (int rc = check_arch()) return rc;
    cudaStream_t st = (cudaStream_t)stream;

    HeadWeights hw;
    for (int h = 0; h < H; ++h) hw.w[h] = head_w_host ? head_w_host[h] : 1.0f / (float)H;

    const int R = pick_rows_per_chunk(B, H, Sq, Sk);
    const int nchunks = cdiv(Sq, R);
    SPLITQ_REQUIRE((long long)H * nchunks <= 65535, SPLITQ_ERR_ARG, "importance_colmean: grid too large");
    if (out_col) {
        const size_t need = (size_t)B * H * nchunks * Sk * sizeof(float);
        SPLITQ_REQUIRE(ws && ws_bytes >= need, SPLITQ_ERR_WORKSPACE,
                       "importance_colmean: workspace %zu < %zu bytes", ws_bytes, need);
        dim3 grid(cdiv(Sk, kColTile), H * nchunks, B);
        const bool vec = (Sk % 4 == 0) && aligned16(A);
        if (vec)
            colsum_partial_kernel<true><<<grid, kWarps * 32, 0, st>>>(A, H, Sq, Sk, R, nchunks, hw,
                                                                     (float*)ws, flags);
        else
            colsum_partial_kernel<false><<<grid, kWarps * 32, 0, st>>>(A, H, Sq, Sk, R, nchunks, hw,
                                                                      (float*)ws, flags);
        SPLITQ_LAUNCH_CHECK("colsum_partial_kernel");
    }
    dim3 fgrid(cdiv(Sk, 256), B);
    importance_finalize_kernel<<<fgrid, 256, 0, st>>>(A, (const float*)ws, H, Sq, Sk, H * nchunks, hw,
                                                     out_col, out_last);
    SPLITQ_LAUNCH_CHECK("importance_finalize_kernel");
    return SPLITQ_OK;
}
