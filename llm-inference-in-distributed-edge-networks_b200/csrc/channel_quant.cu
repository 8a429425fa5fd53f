// Channel-wise quantisers of the split-layer hidden states.
//
// Reference: QwenPointFiveBModel.channel_wise (Experiments/Qwen2-0.5B/qwen_layer_wise.py:106-167),
// a Python loop over the D channels with ~6 ATen launches each:
//   channel_8 / channel_4 : m = max|x_c|,  q = round(x / m * L),             x' = q * m / L   (L = 127 / 7)
//   channel_1_mean        : u = mean(x_c) + 1e-8, q = clamp(round(x / u), -1, 1), x' = q * u
//   channel_1_max         : m = max|x_c|,  q = clamp(round(x / m), -1, 1),   x' = q * m
// Here: one column reduction over the B*S rows (same stripe/partial scheme as K1) and one
// elementwise pass with IEEE division/multiplication in the reference's order.
#include <algorithm>

#include "common.cuh"

namespace splitq {

constexpr int kCWarps = 8;

// per-(stripe, channel) partials: absmax bit pattern or fp32 sum
template <bool MEAN>
__global__ void __launch_bounds__(kCWarps * 32)
channel_partial_kernel(const float* __restrict__ h, long long R, int D, int rows_per_chunk,
                       float* __restrict__ partial) {
    // blockIdx.z = window when the statistic is taken per window (R = rows of one window)
    h += (size_t)blockIdx.z * R * D;
    partial += (size_t)blockIdx.z * gridDim.y * D;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = blockIdx.x * 128 + lane * 4;
    const long long r0 = (long long)blockIdx.y * rows_per_chunk;
    const long long r1 = min(R, r0 + rows_per_chunk);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool vec = (D & 3) == 0 && (reinterpret_cast<uintptr_t>(h) & 15u) == 0;
    for (long long r = r0 + warp; r < r1; r += kCWarps) {
        float v[4];
        if (vec) {
            const float4 t = c0 < D ? ld_stream_f4(reinterpret_cast<const float4*>(h + r * D + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = (c0 + i < D) ? __ldg(h + r * D + c0 + i) : 0.f;
        }
        if (MEAN) {
            acc.x += v[0]; acc.y += v[1]; acc.z += v[2]; acc.w += v[3];
        } else {
            acc.x = fmaxf(acc.x, fabsf(v[0])); acc.y = fmaxf(acc.y, fabsf(v[1]));
            acc.z = fmaxf(acc.z, fabsf(v[2])); acc.w = fmaxf(acc.w, fabsf(v[3]));
        }
    }
    __shared__ float4 red[kCWarps][32];
    red[warp][lane] = acc;
    __syncthreads();
    if (warp == 0) {
        float4 s = red[0][lane];
        for (int w = 1; w < kCWarps; ++w) {
            const float4 v = red[w][lane];
            if (MEAN) { s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
            else { s.x = fmaxf(s.x, v.x); s.y = fmaxf(s.y, v.y); s.z = fmaxf(s.z, v.z); s.w = fmaxf(s.w, v.w); }
        }
        float* out = partial + (size_t)blockIdx.y * D;
        if (c0 + 0 < D) out[c0 + 0] = s.x;
        if (c0 + 1 < D) out[c0 + 1] = s.y;
        if (c0 + 2 < D) out[c0 + 2] = s.z;
        if (c0 + 3 < D) out[c0 + 3] = s.w;
    }
}

template <bool MEAN>
__global__ void channel_finalize_kernel(const float* __restrict__ partial, int nparts, int D, long long R,
                                        float* __restrict__ stats, size_t stats_stride_bytes) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= D) return;
    partial += (size_t)blockIdx.y * nparts * D;                                   // blockIdx.y = window
    stats = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(stats) + (size_t)blockIdx.y * stats_stride_bytes);
    if (MEAN) {
        double s = 0.0;
        for (int p = 0; p < nparts; ++p) s += (double)partial[(size_t)p * D + c];
        stats[c] = __fadd_rn((float)(s / (double)R), 1e-8f);  // torch.mean(...) + 1e-8
    } else {
        float m = 0.f;
        for (int p = 0; p < nparts; ++p) m = fmaxf(m, partial[(size_t)p * D + c]);
        stats[c] = m;
    }
}

__device__ __forceinline__ float channel_fake(float x, float st, int method, int8_t* code) {
    float q, out;
    if (method == SPLITQ_CHANNEL_8 || method == SPLITQ_CHANNEL_4) {
        const float L = method == SPLITQ_CHANNEL_8 ? 127.f : 7.f;
        q = rintf(__fmul_rn(__fdiv_rn(x, st), L));          // round(x / m * L)
        out = __fdiv_rn(__fmul_rn(q, st), L);               // q * m / L
    } else {
        q = rintf(__fdiv_rn(x, st));                        // round(x / u)
        q = fminf(fmaxf(q, -1.f), 1.f);                     // clamp (NaN stays NaN in ATen; see below)
        if (!(__fdiv_rn(x, st) == __fdiv_rn(x, st))) q = __uint_as_float(0x7fc00000u);
        out = __fmul_rn(q, st);                             // q * u
    }
    *code = (q == q) ? (int8_t)q : (int8_t)0;
    return out;
}

__global__ void __launch_bounds__(256)
channel_quant_kernel(const float* __restrict__ h, long long n, int D, int method, const float* __restrict__ stats,
                     float* __restrict__ out, int8_t* __restrict__ codes) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % D);
        int8_t q;
        const float y = channel_fake(h[i], __ldg(stats + c), method, &q);
        out[i] = y;
        if (codes) codes[i] = q;
    }
}

// vector form: a thread owns four adjacent channels (their scales stay in registers) and walks down
// the rows; a 128-thread row segment is one contiguous 2 KB access.  D % 4 == 0, 16-byte aligned.
__global__ void __launch_bounds__(128)
channel_quant_vec_kernel(const float* __restrict__ h, long long R, int D, int method, const float* __restrict__ stats,
                         float* __restrict__ out, int8_t* __restrict__ codes) {
    const int c = (blockIdx.x * 128 + threadIdx.x) * 4;
    if (c >= D) return;
    const float4 st = __ldg(reinterpret_cast<const float4*>(stats + c));
    for (long long r = blockIdx.y; r < R; r += gridDim.y) {
        const float4 x = *reinterpret_cast<const float4*>(h + r * D + c);  // h may alias out: plain load
        int8_t q0, q1, q2, q3;
        float4 y;
        y.x = channel_fake(x.x, st.x, method, &q0);
        y.y = channel_fake(x.y, st.y, method, &q1);
        y.z = channel_fake(x.z, st.z, method, &q2);
        y.w = channel_fake(x.w, st.w, method, &q3);
        *reinterpret_cast<float4*>(out + r * D + c) = y;
        if (codes)
            *reinterpret_cast<uint32_t*>(codes + r * D + c) = (uint32_t)(uint8_t)q0 | ((uint32_t)(uint8_t)q1 << 8) |
                                                              ((uint32_t)(uint8_t)q2 << 16) | ((uint32_t)(uint8_t)q3 << 24);
    }
}

// ---- payload form: one warp per token packs the channel codes (statistics already in the payload) ----
__global__ void __launch_bounds__(128)
channel_pack_kernel(const float* __restrict__ h, int S, int D, int method, uint8_t* __restrict__ payload, size_t stride) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * 4 + warp, b = blockIdx.y;
    if (t >= S) return;
    const int mode = SPLITQ_SCALE_CHANNEL_BASE + method, nb = channel_bits(method);
    const PayloadLayout lay = payload_layout((uint32_t)S, mode, (uint32_t)D);
    uint8_t* P = payload + (size_t)b * stride;
    const uint32_t rb = row_bytes((uint32_t)D, (uint32_t)nb);
    const uint32_t off16 = (uint32_t)t * (rb / 16u);
    if (t == 0) {
        if (lane < 8) {
            const uint32_t w[8] = {SPLITQ_MAGIC, (uint32_t)mode, (uint32_t)S, (uint32_t)D, lay.n_scales, lay.codes_off,
                                   lay.codes_off + (uint32_t)S * rb, 0u};
            reinterpret_cast<uint32_t*>(P)[lane] = w[lane];
        }
        for (uint32_t i = (uint32_t)S + lane; i < pad16((uint32_t)S); i += 32) P[lay.bits_off + i] = 0;
        for (uint32_t i = 4u * S + lane; i < pad16(4u * S); i += 32) P[lay.rowoff_off + i] = 0;
        for (uint32_t i = 4u * D + lane; i < pad16(4u * D); i += 32) P[lay.scales_off + i] = 0;
    }
    if (lane == 0) {
        P[lay.bits_off + t] = (uint8_t)nb;
        reinterpret_cast<uint32_t*>(P + lay.rowoff_off)[t] = off16;
    }
    const float4* src = reinterpret_cast<const float4*>(h + ((size_t)b * S + t) * D);
    const float4* sc = reinterpret_cast<const float4*>(P + lay.scales_off);
    uint8_t* dst = P + lay.codes_off + (size_t)off16 * 16u;
    const uint32_t mask = (1u << nb) - 1u;
    const int nch = D >> 2;
    for (int c = lane; (uint32_t)c * (uint32_t)nb / 2u < rb; c += 32) {  // chunks past D pad the row with zeros
        uint32_t w = 0u;
        if (c < nch) {
            const float4 x = ld_stream_f4(src + c);
            const float4 m = sc[c];  // written by the statistics kernel of this call: a plain load
            int8_t q0, q1, q2, q3;
            channel_fake(x.x, m.x, method, &q0);
            channel_fake(x.y, m.y, method, &q1);
            channel_fake(x.z, m.z, method, &q2);
            channel_fake(x.w, m.w, method, &q3);
            w = ((uint32_t)q0 & mask) | (((uint32_t)q1 & mask) << nb) | (((uint32_t)q2 & mask) << (2 * nb)) |
                (((uint32_t)q3 & mask) << (3 * nb));
        }
        if (nb == 8) reinterpret_cast<uint32_t*>(dst)[c] = w;
        else if (nb == 4) reinterpret_cast<uint16_t*>(dst)[c] = (uint16_t)w;
        else dst[c] = (uint8_t)w;
    }
}

static int pick_rows(long long R, int D) {
    const long long tiles = cdiv(D, 128);
    int rows = 512;
    while (rows > 32 && tiles * cdiv(R, rows) < 148LL * 4) rows >>= 1;
    return rows;
}

}  // namespace splitq

using namespace splitq;

extern "C" size_t splitq_channel_workspace_bytes(int B, int S, int D) {
    if (B <= 0 || S <= 0 || D <= 0) return 0;
    const long long R = (long long)B * S;
    const size_t whole = (size_t)cdiv(R, pick_rows(R, D)) * D * sizeof(float);            // statistic over all B*S rows
    const size_t per_window = (size_t)B * cdiv(S, pick_rows(S, D)) * D * sizeof(float);   // splitq_channel_quant_pack
    return std::max(whole, per_window);
}

extern "C" int splitq_channel_fake_quant(const float* h, int B, int S, int D, int method, float* stats_out,
                                         float* h_out, int8_t* codes_out, void* ws, size_t ws_bytes,
                                         splitq_stream_t stream) {
    SPLITQ_REQUIRE(h && stats_out && h_out, SPLITQ_ERR_ARG, "channel_fake_quant: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0, SPLITQ_ERR_ARG, "channel_fake_quant: bad shape B=%d S=%d D=%d", B, S, D);
    SPLITQ_REQUIRE(method >= SPLITQ_CHANNEL_8 && method <= SPLITQ_CHANNEL_1_MAX, SPLITQ_ERR_ARG,
                   "channel_fake_quant: unknown method %d", method);
    const long long R = (long long)B * S;
    const int rows = pick_rows(R, D), nparts = cdiv(R, rows);
    const size_t need = (size_t)nparts * D * sizeof(float);
    SPLITQ_REQUIRE(ws && ws_bytes >= need, SPLITQ_ERR_WORKSPACE, "channel_fake_quant: workspace %zu < %zu bytes",
                   ws_bytes, need);
    SPLITQ_REQUIRE(nparts <= 65535, SPLITQ_ERR_ARG, "channel_fake_quant: too many rows");
    if (int rc = check_arch()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(cdiv(D, 128), nparts);
    const bool mean = method == SPLITQ_CHANNEL_1_MEAN;
    if (mean) {
        channel_partial_kernel<true><<<grid, kCWarps * 32, 0, st>>>(h, R, D, rows, (float*)ws);
        channel_finalize_kernel<true><<<cdiv(D, 256), 256, 0, st>>>((const float*)ws, nparts, D, R, stats_out, 0);
    } else {
        channel_partial_kernel<false><<<grid, kCWarps * 32, 0, st>>>(h, R, D, rows, (float*)ws);
        channel_finalize_kernel<false><<<cdiv(D, 256), 256, 0, st>>>((const float*)ws, nparts, D, R, stats_out, 0);
    }
    SPLITQ_LAUNCH_CHECK("channel stats kernels");
    const long long n = R * D;
    const bool vec = (D & 3) == 0 && aligned16(h) && aligned16(h_out) && aligned16(stats_out) &&
                     (!codes_out || (reinterpret_cast<uintptr_t>(codes_out) & 3u) == 0);
    if (vec) {
        const int gx = cdiv(D / 4, 128);
        const int gy = (int)std::min<long long>(R, std::max<long long>(1, 148LL * 16 / gx));
        channel_quant_vec_kernel<<<dim3(gx, gy), 128, 0, st>>>(h, R, D, method, stats_out, h_out, codes_out);
        SPLITQ_LAUNCH_CHECK("channel_quant_vec_kernel");
        return SPLITQ_OK;
    }
    const int blocks = (int)std::min<long long>((n + 255) / 256, 148LL * 16);
    channel_quant_kernel<<<blocks, 256, 0, st>>>(h, n, D, method, stats_out, h_out, codes_out);
    SPLITQ_LAUNCH_CHECK("channel_quant_kernel");
    return SPLITQ_OK;
}

extern "C" size_t splitq_channel_payload_bytes(int S, int D, int method) {
    if (S <= 0 || D <= 0 || method < SPLITQ_CHANNEL_8 || method > SPLITQ_CHANNEL_1_MAX) return 0;
    const PayloadLayout lay = payload_layout((uint32_t)S, SPLITQ_SCALE_CHANNEL_BASE + method, (uint32_t)D);
    return (size_t)lay.codes_off + (size_t)S * row_bytes((uint32_t)D, (uint32_t)channel_bits(method));
}

extern "C" int splitq_channel_quant_pack(const float* h, int B, int S, int D, int method, uint8_t* payload,
                                         size_t payload_stride, void* ws, size_t ws_bytes, splitq_stream_t stream) {
    SPLITQ_REQUIRE(h && payload, SPLITQ_ERR_ARG, "channel_quant_pack: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0 && B <= 65535, SPLITQ_ERR_ARG, "channel_quant_pack: bad shape B=%d S=%d D=%d", B, S, D);
    SPLITQ_REQUIRE(method >= SPLITQ_CHANNEL_8 && method <= SPLITQ_CHANNEL_1_MAX, SPLITQ_ERR_ARG,
                   "channel_quant_pack: unknown method %d", method);
    SPLITQ_REQUIRE(D % 4 == 0 && aligned16(h), SPLITQ_ERR_ALIGN, "channel_quant_pack: D must be a multiple of 4 and h 16-byte aligned");
    SPLITQ_REQUIRE(aligned16(payload) && payload_stride % 16 == 0, SPLITQ_ERR_ALIGN,
                   "channel_quant_pack: payload and payload_stride must be 16-byte aligned");
    const size_t nbytes = splitq_channel_payload_bytes(S, D, method);
    SPLITQ_REQUIRE(B == 1 || payload_stride >= nbytes, SPLITQ_ERR_WORKSPACE, "channel_quant_pack: payload_stride %zu < %zu",
                   payload_stride, nbytes);
    const int rows = pick_rows(S, D), nparts = cdiv(S, rows);
    const size_t need = (size_t)B * nparts * D * sizeof(float);
    SPLITQ_REQUIRE(ws && ws_bytes >= need, SPLITQ_ERR_WORKSPACE, "channel_quant_pack: workspace %zu < %zu bytes", ws_bytes, need);
    SPLITQ_REQUIRE(nparts <= 65535, SPLITQ_ERR_ARG, "channel_quant_pack: too many rows");
    if (int rc = check_arch()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const PayloadLayout lay = payload_layout((uint32_t)S, SPLITQ_SCALE_CHANNEL_BASE + method, (uint32_t)D);
    float* scales0 = reinterpret_cast<float*>(payload + lay.scales_off);  // window 0; the others payload_stride apart
    dim3 grid(cdiv(D, 128), nparts, B);
    if (method == SPLITQ_CHANNEL_1_MEAN) {
        channel_partial_kernel<true><<<grid, kCWarps * 32, 0, st>>>(h, S, D, rows, (float*)ws);
        channel_finalize_kernel<true><<<dim3(cdiv(D, 256), B), 256, 0, st>>>((const float*)ws, nparts, D, S, scales0, payload_stride);
    } else {
        channel_partial_kernel<false><<<grid, kCWarps * 32, 0, st>>>(h, S, D, rows, (float*)ws);
        channel_finalize_kernel<false><<<dim3(cdiv(D, 256), B), 256, 0, st>>>((const float*)ws, nparts, D, S, scales0, payload_stride);
    }
    SPLITQ_LAUNCH_CHECK("channel stats kernels");
    channel_pack_kernel<<<dim3(cdiv(S, 4), B), 128, 0, st>>>(h, S, D, method, payload, payload_stride);
    SPLITQ_LAUNCH_CHECK("channel_pack_kernel");
    return SPLITQ_OK;
}
