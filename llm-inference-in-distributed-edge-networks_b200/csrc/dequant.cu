// K4: dequantise the boundary payload, optionally fused with the downstream half's first norm.
//
// Reference: `quantized_values / 7 * max_val` scattered back with index_put
// (Experiments/Pythia-70M/pythia_model.py:186-188, Experiments/Qwen2-0.5B/qwen_layer_wise.py:68-70),
// after which the next decoder layer immediately re-reads the tensor to normalise it
// (input_layernorm / final norm, reached through pythia_model.py:165,193 and
// qwen_layer_wise.py:52,75).  Here one warp owns one token: it reads the packed row once,
// rebuilds the fp32 row in registers as (q / L) * m (IEEE division and multiplication, same
// order as ATen, so the values are bit-identical), writes the residual stream, and -- in the
// fused variants -- computes LayerNorm / RMSNorm statistics with warp shuffles and writes the
// normalised activations without a second trip to HBM.
#include <math.h>

#include "common.cuh"

namespace splitq {

// like the pack kernel: small CTAs and a register cap for many resident row-warps (plain mode, bench
// shape: 8 warps x default occupancy 27.8 us, 4 warps x 10 CTAs/SM 25.2 us = 0.97 of the measured copy rate)
constexpr int kDWarps = 4;
constexpr int dq_min_blocks(int maxch, int mode) { return maxch <= 8 ? (mode == 0 ? 10 : 8) : 1; }
enum { MODE_PLAIN = 0, MODE_LAYERNORM = 1, MODE_RMSNORM = 2 };

struct NormArgs {
    const float* gamma;
    const float* beta;
    const float* gamma2;
    const float* beta2;
    float eps;
    float* normed;
    float* normed2;
};

// q / L for the three level counts, without the general division: one multiplication by the
// correctly rounded reciprocal plus one Newton correction through FMA gives the correctly rounded
// quotient for every code q in [-128, 127] and L in {1, 7, 127} (checked exhaustively against
// __fdiv_rn by tests/test_kernels_gpu.py::test_every_code_dequantises_like_ieee_division).
__device__ __forceinline__ float div_levels(float q, float L, float rL) {
    const float y = __fmul_rn(q, rL);
    const float e = __fmaf_rn(-L, y, q);
    return __fmaf_rn(e, rL, y);
}
// exact int -> float for |q| < 2^22 on the integer and FMA pipes (I2F runs on the quarter-rate conversion pipe,
// one per element here): 1.5 * 2^23 + q is representable, and subtracting the constant leaves q
__device__ __forceinline__ float small_int_to_float(int q) {
    return __fsub_rn(__int_as_float(0x4B400000 + q), 12582912.f);
}
// recip: the flavour CUDA ATen computes for `q / 7 * m` -- a tensor divided by a Python scalar is
// multiplied by fl(1/7) there (BinaryDivTrueKernel.cu), which differs from the IEEE quotient by up
// to one ulp for some codes; the default is the correctly rounded quotient (CPU ATen).
__device__ __forceinline__ float decode_field(uint32_t w, int i, int nb, float L, float rL, float m, bool recip) {
    const int q = (int)(w << (32 - nb * (i + 1))) >> (32 - nb);  // sign-extend field i
    const float qf = small_int_to_float(q);
    return __fmul_rn(recip ? __fmul_rn(qf, rL) : div_levels(qf, L, rL), m);
}

struct RowInfo {
    int nb;
    float m, L, rL;
    bool nan_row, bad, recip;
    int channel;              // -1, or the channel method of a per-channel payload (scales: one per channel)
    const float* ch_scales;
    const uint8_t* src;
};
// per-channel payloads (qwen_layer_wise.py:132-134,139-150): q * m / L for the absmax widths (multiply first,
// then divide -- not the token path's q / L * m), q * u for the ternary methods; a zero statistic is the
// reference's 0/0
__device__ __forceinline__ float channel_decode(uint32_t w, int i, int nb, float st, int method) {
    const int q = (int)(w << (32 - nb * (i + 1))) >> (32 - nb);
    if (st == 0.f) return __uint_as_float(0x7fc00000u);
    if (method == SPLITQ_CHANNEL_8) return __fdiv_rn(__fmul_rn((float)q, st), 127.f);
    if (method == SPLITQ_CHANNEL_4) return __fdiv_rn(__fmul_rn((float)q, st), 7.f);
    return __fmul_rn((float)q, st);
}

// `limit` = bytes the caller says one window's payload may span (payload_stride; 0 = unknown).  The
// payload crosses process and device boundaries, so nothing read from it is trusted: a header that
// does not describe this (S, D, mode) layout, or a row that would end past the declared size or
// past `limit`, marks the row bad (NaN output) instead of being dereferenced.
__device__ __forceinline__ RowInfo row_info(const uint8_t* P, int S, int D, int t, size_t limit) {
    RowInfo r;
    // every index load below is independent of the others (section offsets depend only on S), so
    // they are all in flight at once; only the row data depends on them
    const uint4 h0 = __ldg(reinterpret_cast<const uint4*>(P));
    const uint4 h1 = __ldg(reinterpret_cast<const uint4*>(P) + 1);
    const PayloadLayout lt = payload_layout((uint32_t)S, SPLITQ_SCALE_TOKEN);
    const int nb = P[lt.bits_off + t];
    const uint32_t off16 = __ldg(reinterpret_cast<const uint32_t*>(P + lt.rowoff_off) + t);
    const float4 gs = __ldg(reinterpret_cast<const float4*>(P + lt.scales_off));  // group slots (or tokens 0..3)
    const float ts = __ldg(reinterpret_cast<const float*>(P + lt.scales_off) + t);  // token slot (inside the payload in either mode)
    const int mode = (int)h0.y;
    r.channel = is_channel_mode(mode) ? mode - SPLITQ_SCALE_CHANNEL_BASE : -1;
    r.ch_scales = reinterpret_cast<const float*>(P + lt.scales_off);
    r.bad = h0.x != SPLITQ_MAGIC || h0.z != (uint32_t)S || h0.w != (uint32_t)D ||
            (mode != SPLITQ_SCALE_GROUP && mode != SPLITQ_SCALE_TOKEN && r.channel < 0);
    if (!r.bad) {
        const uint32_t nbw = (nb == 2 || nb == 4 || nb == 8 || nb == SPLITQ_BITS_RAW) ? (uint32_t)nb : 8u;
        const unsigned long long end = (unsigned long long)h1.y + 16ull * off16 + row_bytes((uint32_t)D, nbw);
        r.bad = h1.y != payload_layout((uint32_t)S, mode, (uint32_t)D).codes_off || end > h1.z || (limit != 0 && h1.z > limit) ||
                (r.channel >= 0 && nb != channel_bits(r.channel));
    }
    r.nb = nb;
    r.recip = (h1.w & SPLITQ_HDR_DEQUANT_RECIP) != 0u;
    r.src = P + h1.y /* codes_offset */ + (size_t)off16 * 16u;
    r.m = 0.f;
    r.L = 1.f;
    r.rL = 1.f;
    r.nan_row = false;
    if (nb != SPLITQ_BITS_RAW) {
        if (nb != 2 && nb != 4 && nb != 8) {
            r.bad = true;
            r.nb = 8;
        }
        r.m = mode == SPLITQ_SCALE_GROUP ? (r.nb == 8 ? gs.x : (r.nb == 4 ? gs.y : gs.z)) : ts;
        r.L = (float)((1 << (r.nb - 1)) - 1);
        r.rL = r.nb == 8 ? 0x1.020408p-7f : (r.nb == 4 ? 0x1.24924ap-3f : 1.0f);  // fl(1/127), fl(1/7), 1
        // reference semantics: one scale for the set, all zeros -> 0/0 -> NaN rows
        r.nan_row = mode == SPLITQ_SCALE_GROUP && r.m == 0.f;
    }
    if (r.bad) r.src = P;  // never dereference an offset taken from a corrupt header
    return r;
}

template <int MAXCH, int MODE>
__global__ void __launch_bounds__(kDWarps * 32, dq_min_blocks(MAXCH, MODE))
dequant_kernel(const uint8_t* __restrict__ payload, size_t stride, int S, int D, float* __restrict__ h_out,
               NormArgs na) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * kDWarps + warp;
    const int b = blockIdx.y;
    if (t >= S) return;
    const uint8_t* P = payload + (size_t)b * stride;
    const RowInfo r = row_info(P, S, D, t, stride);
    const int nch = D >> 2;
    const size_t tok = (size_t)b * S + t;
    const float qnan = __uint_as_float(0x7fc00000u);

    float4 v[MAXCH];
    if (r.bad) {
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) v[u] = make_float4(qnan, qnan, qnan, qnan);
    } else if (r.nb == SPLITQ_BITS_RAW) {
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            v[u] = c < nch ? ld_stream_f4(reinterpret_cast<const float4*>(r.src) + c)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    } else {
        uint32_t w[MAXCH];
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            w[u] = 0u;
            if (c < nch) {
                if (r.nb == 8) w[u] = __ldg(reinterpret_cast<const uint32_t*>(r.src) + c);
                else if (r.nb == 4) w[u] = __ldg(reinterpret_cast<const uint16_t*>(r.src) + c);
                else w[u] = __ldg(r.src + c);
            }
        }
        if (r.channel >= 0) {
#pragma unroll
            for (int u = 0; u < MAXCH; ++u) {
                const int c = lane + 32 * u;
                if (c < nch) {
                    const float4 st = __ldg(reinterpret_cast<const float4*>(r.ch_scales) + c);
                    v[u].x = channel_decode(w[u], 0, r.nb, st.x, r.channel);
                    v[u].y = channel_decode(w[u], 1, r.nb, st.y, r.channel);
                    v[u].z = channel_decode(w[u], 2, r.nb, st.z, r.channel);
                    v[u].w = channel_decode(w[u], 3, r.nb, st.w, r.channel);
                } else {
                    v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
        } else {
#pragma unroll
            for (int u = 0; u < MAXCH; ++u) {
                v[u].x = decode_field(w[u], 0, r.nb, r.L, r.rL, r.m, r.recip);
                v[u].y = decode_field(w[u], 1, r.nb, r.L, r.rL, r.m, r.recip);
                v[u].z = decode_field(w[u], 2, r.nb, r.L, r.rL, r.m, r.recip);
                v[u].w = decode_field(w[u], 3, r.nb, r.L, r.rL, r.m, r.recip);
                if (r.nan_row) v[u] = make_float4(qnan, qnan, qnan, qnan);
            }
        }
    }

    if (h_out) {
        float4* dst = reinterpret_cast<float4*>(h_out + tok * D);
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if (c < nch) dst[c] = v[u];
        }
    }
    if (MODE == MODE_PLAIN) return;

    const float invD = 1.0f / (float)D;
    float mean = 0.f;
    if (MODE == MODE_LAYERNORM) {
        float s = 0.f;
#pragma unroll
        for (int u = 0; u < MAXCH; ++u)
            if (lane + 32 * u < nch) s += (v[u].x + v[u].y) + (v[u].z + v[u].w);
        mean = warp_sum(s) * invD;
    }
    float ss = 0.f;
#pragma unroll
    for (int u = 0; u < MAXCH; ++u) {
        if (lane + 32 * u < nch) {
            const float a = v[u].x - mean, bb = v[u].y - mean, cc = v[u].z - mean, d = v[u].w - mean;
            ss += (a * a + bb * bb) + (cc * cc + d * d);
        }
    }
    const float var = warp_sum(ss) * invD;
    const float rstd = 1.0f / sqrtf(var + na.eps);

#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        const float* gamma = pass == 0 ? na.gamma : na.gamma2;
        const float* beta = pass == 0 ? na.beta : na.beta2;
        float* out = pass == 0 ? na.normed : na.normed2;
        if (!out) continue;
        float4* dst = reinterpret_cast<float4*>(out + tok * D);
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if (c < nch) {
                const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + c);
                float4 o;
                if (MODE == MODE_LAYERNORM) {
                    const float4 be = beta ? __ldg(reinterpret_cast<const float4*>(beta) + c)
                                           : make_float4(0.f, 0.f, 0.f, 0.f);
                    o.x = (v[u].x - mean) * rstd * g.x + be.x;
                    o.y = (v[u].y - mean) * rstd * g.y + be.y;
                    o.z = (v[u].z - mean) * rstd * g.z + be.z;
                    o.w = (v[u].w - mean) * rstd * g.w + be.w;
                } else {
                    o.x = g.x * (v[u].x * rstd);
                    o.y = g.y * (v[u].y * rstd);
                    o.z = g.z * (v[u].z * rstd);
                    o.w = g.w * (v[u].w * rstd);
                }
                dst[c] = o;
            }
        }
    }
}

// any D / unaligned outputs: scalar accesses, row re-decoded per use (stays in L1)
template <int MODE>
__global__ void __launch_bounds__(kDWarps * 32)
dequant_generic_kernel(const uint8_t* __restrict__ payload, size_t stride, int S, int D,
                       float* __restrict__ h_out, NormArgs na) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * kDWarps + warp;
    const int b = blockIdx.y;
    if (t >= S) return;
    const uint8_t* P = payload + (size_t)b * stride;
    const RowInfo r = row_info(P, S, D, t, stride);
    const size_t tok = (size_t)b * S + t;
    const float qnan = __uint_as_float(0x7fc00000u);
    auto value = [&](int d) -> float {
        if (r.bad || r.nan_row) return qnan;
        if (r.nb == SPLITQ_BITS_RAW) return reinterpret_cast<const float*>(r.src)[d];
        const int per = 8 / r.nb;
        const uint32_t byte = r.src[d / per];
        if (r.channel >= 0) return channel_decode(byte, d % per, r.nb, r.ch_scales[d], r.channel);
        return decode_field(byte, d % per, r.nb, r.L, r.rL, r.m, r.recip);
    };
    if (h_out)
        for (int d = lane; d < D; d += 32) h_out[tok * D + d] = value(d);
    if (MODE == MODE_PLAIN) return;
    const float invD = 1.0f / (float)D;
    float mean = 0.f;
    if (MODE == MODE_LAYERNORM) {
        float s = 0.f;
        for (int d = lane; d < D; d += 32) s += value(d);
        mean = warp_sum(s) * invD;
    }
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) {
        const float a = value(d) - mean;
        ss += a * a;
    }
    const float rstd = 1.0f / sqrtf(warp_sum(ss) * invD + na.eps);
    for (int pass = 0; pass < 2; ++pass) {
        const float* gamma = pass == 0 ? na.gamma : na.gamma2;
        const float* beta = pass == 0 ? na.beta : na.beta2;
        float* out = pass == 0 ? na.normed : na.normed2;
        if (!out) continue;
        for (int d = lane; d < D; d += 32) {
            const float x = value(d);
            out[tok * D + d] = MODE == MODE_LAYERNORM ? (x - mean) * rstd * gamma[d] + (beta ? beta[d] : 0.f)
                                                      : gamma[d] * (x * rstd);
        }
    }
}

template <int MODE>
static int launch_dequant(const uint8_t* payload, size_t stride, int B, int S, int D, float* h_out,
                          const NormArgs& na, bool vec, cudaStream_t st) {
    dim3 grid(cdiv(S, kDWarps), B);
    if (!vec) {
        dequant_generic_kernel<MODE><<<grid, kDWarps * 32, 0, st>>>(payload, stride, S, D, h_out, na);
    } else {
        const int nch = D / 4;
        if (nch <= 32 * 4) dequant_kernel<4, MODE><<<grid, kDWarps * 32, 0, st>>>(payload, stride, S, D, h_out, na);
        else if (nch <= 32 * 8) dequant_kernel<8, MODE><<<grid, kDWarps * 32, 0, st>>>(payload, stride, S, D, h_out, na);
        else if (nch <= 32 * 16) dequant_kernel<16, MODE><<<grid, kDWarps * 32, 0, st>>>(payload, stride, S, D, h_out, na);
        else dequant_kernel<32, MODE><<<grid, kDWarps * 32, 0, st>>>(payload, stride, S, D, h_out, na);
    }
    SPLITQ_LAUNCH_CHECK("dequant_kernel");
    return SPLITQ_OK;
}

static int check_payload_args(const char* who, const uint8_t* payload, size_t stride, int B, int S, int D) {
    SPLITQ_REQUIRE(payload, SPLITQ_ERR_ARG, "%s: NULL payload", who);
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0 && B <= 65535, SPLITQ_ERR_ARG, "%s: bad shape B=%d S=%d D=%d", who,
                   B, S, D);
    SPLITQ_REQUIRE(aligned16(payload) && stride % 16 == 0, SPLITQ_ERR_ALIGN,
                   "%s: payload and payload_stride must be 16-byte aligned", who);
    return check_arch();
}

}  // namespace splitq

using namespace splitq;

extern "C" int splitq_dequant(const uint8_t* payload, size_t payload_stride, int B, int S, int D,
                              float* h_out, splitq_stream_t stream) {
    if (int rc = check_payload_args("dequant", payload, payload_stride, B, S, D)) return rc;
    SPLITQ_REQUIRE(h_out, SPLITQ_ERR_ARG, "dequant: NULL output");
    NormArgs na = {};
    const bool vec = (D % 4 == 0) && D <= 4096 && aligned16(h_out);
    return launch_dequant<MODE_PLAIN>(payload, payload_stride, B, S, D, h_out, na, vec, (cudaStream_t)stream);
}

extern "C" int splitq_dequant_norm(const uint8_t* payload, size_t payload_stride, int B, int S, int D,
                                   int norm_kind, const float* gamma, const float* beta, float eps,
                                   const float* gamma2, const float* beta2, float* resid_out,
                                   float* normed_out, float* normed2_out, splitq_stream_t stream) {
    if (int rc = check_payload_args("dequant_norm", payload, payload_stride, B, S, D)) return rc;
    SPLITQ_REQUIRE(norm_kind == SPLITQ_NORM_LAYERNORM || norm_kind == SPLITQ_NORM_RMSNORM, SPLITQ_ERR_ARG,
                   "dequant_norm: unknown norm_kind %d", norm_kind);
    SPLITQ_REQUIRE(gamma && normed_out, SPLITQ_ERR_ARG, "dequant_norm: gamma / normed_out are NULL");
    SPLITQ_REQUIRE(!normed2_out || gamma2, SPLITQ_ERR_ARG, "dequant_norm: normed2_out given without gamma2");
    NormArgs na = {gamma, beta, gamma2, beta2, eps, normed_out, normed2_out};
    const bool vec = (D % 4 == 0) && D <= 4096 && aligned16(normed_out) && aligned16(gamma) &&
                     (!resid_out || aligned16(resid_out)) && (!beta || aligned16(beta)) &&
                     (!normed2_out || (aligned16(normed2_out) && aligned16(gamma2) && (!beta2 || aligned16(beta2))));
    cudaStream_t st = (cudaStream_t)stream;
    if (norm_kind == SPLITQ_NORM_LAYERNORM)
        return launch_dequant<MODE_LAYERNORM>(payload, payload_stride, B, S, D, resid_out, na, vec, st);
    return launch_dequant<MODE_RMSNORM>(payload, payload_stride, B, S, D, resid_out, na, vec, st);
}
