// K3: absmax scale, quantise, bit-pack the split-layer hidden states into the boundary payload.
//
// Reference: the fake-quant block of activation_quantization
// (Experiments/Pythia-70M/pythia_model.py:176-184, Experiments/Qwen2-0.5B/qwen_layer_wise.py:60-66):
//     m = max|h[:, idx, :]|;  q = round(clamp(h / m * 7, -8, 7))
// run there as gather + 5 elementwise ATen kernels and never stored as integers.  Here one warp
// owns one token row: the row is read once with 128-bit streaming loads and kept in registers,
// absmax is a single integer warp reduction on the |x| bit patterns, codes are produced with
// IEEE division / multiplication (no reciprocal, no FMA contraction) so they equal the ATen
// result bit for bit, and fields are packed in registers and stored straight into the payload.
// The reference's single scale for the whole selected set ("group" mode) needs the class maximum
// before any code can be emitted, hence a small first phase over the selected rows only.
#include <mutex>

#include "common.cuh"

namespace splitq {

// 4 warps per CTA and (for rows of up to 1024 floats) a register cap that lets 10 CTAs share an SM: 40
// independent row-warps per SM instead of 32 was worth 15 % at the bench shape (37.2 -> 31.5 us; 8 warps x
// 5 CTAs 32.2, 16 x 2 33.3), the kernel lives on bytes in flight and has no barrier to amortise.  What it
// lived on even more was its instruction footprint: see pack4_exact (31.5 -> 23.1 us).
constexpr int kQWarps = 4;
constexpr int qp_min_blocks(int maxch) { return maxch <= 8 ? 10 : (maxch == 16 ? 5 : 2); }

__device__ __forceinline__ int code_of(float x, float m, float L, float lo) {
    // (x / m) * L with round-to-nearest at each step == ATen `x / m * L`
    float s = __fmul_rn(__fdiv_rn(x, m), L);
    // 0/0 (all-zero set): ATen propagates NaN through clamp/round; an integer field cannot, store 0
    if (!(s == s)) return 0;
    s = fminf(fmaxf(s, lo), L);
    return (int)rintf(s);  // round half to even, like torch.round
}

// Four codes of one float4, packed, for a width known at compile time.  Same result as code_of() for every
// input at a fraction of the instruction count (with one code at a time through rintf / F2I the kernel was
// issue-bound: 64 % of the issue slots, 37 % of the conversion pipe, 48 % of DRAM).  The IEEE division is
// replaced by a multiplication with rcp = 1/m, which can differ from the exact sequence by at most
// ~3e-7 * |s| <= 4e-5 in s.  Rounding is the 1.5 * 2^23 trick:
// t = s + 12582912 rounds s to the nearest integer, ties to even, like rintf, and leaves the two's-complement
// code in t's low mantissa bits -- no FRND, no F2I, no clamp (|x| <= m bounds |s| by L, so the approximate s
// cannot round past L either).  The four distances d = s - round(s) decide together whether the reciprocal
// could have changed a rounding (some |d| within 2e-4 of 1/2, or NaN anywhere): then, about once in 600
// chunks, all four codes come from the exact sequence.
__device__ __forceinline__ float max_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ uint32_t bfi(uint32_t field, uint32_t into, int pos, int len) {
    uint32_t r;
    asm("bfi.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(field), "r"(into), "r"(pos), "r"(len));
    return r;
}
// the exact sequence for one chunk, out of line: it runs about once in 600 chunks, and inlined into every chunk of
// every width it made the row loop ~80 KB of code -- the warps of an SM, each in another width's copy, then stalled
// on instruction fetch more than on anything else (ncu: "no instructions" the top stall reason)
__device__ __noinline__ uint32_t pack4_exact(float4 v, float m, float L, float lo, int nb) {
    const uint32_t mask = (1u << nb) - 1u;
    return ((uint32_t)code_of(v.x, m, L, lo) & mask) | (((uint32_t)code_of(v.y, m, L, lo) & mask) << nb) |
           (((uint32_t)code_of(v.z, m, L, lo) & mask) << (2 * nb)) | (((uint32_t)code_of(v.w, m, L, lo) & mask) << (3 * nb));
}
template <int NB>
__device__ __forceinline__ uint32_t pack4_w(float4 v, float m, float rcp, float L, float lo) {
    constexpr float kMagic = 12582912.f;
    const float s0 = __fmul_rn(__fmul_rn(v.x, rcp), L), s1 = __fmul_rn(__fmul_rn(v.y, rcp), L);
    const float s2 = __fmul_rn(__fmul_rn(v.z, rcp), L), s3 = __fmul_rn(__fmul_rn(v.w, rcp), L);
    const float t0 = __fadd_rn(s0, kMagic), t1 = __fadd_rn(s1, kMagic), t2 = __fadd_rn(s2, kMagic), t3 = __fadd_rn(s3, kMagic);
    const float d0 = __fsub_rn(s0, __fsub_rn(t0, kMagic)), d1 = __fsub_rn(s1, __fsub_rn(t1, kMagic));
    const float d2 = __fsub_rn(s2, __fsub_rn(t2, kMagic)), d3 = __fsub_rn(s3, __fsub_rn(t3, kMagic));
    const float worst = max_nan(max_nan(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), max_nan(__fmul_rn(d2, d2), __fmul_rn(d3, d3)));
    // |d| > 0.4998 somewhere (the bound that matters is 0.5 - 1.2e-4), or NaN / Inf
    if (!(worst <= 0.2498f)) return pack4_exact(v, m, L, lo, NB);
    const uint32_t q0 = __float_as_uint(t0), q1 = __float_as_uint(t1), q2 = __float_as_uint(t2), q3 = __float_as_uint(t3);
    if (NB == 8) return __byte_perm(__byte_perm(q0, q1, 0x0040), __byte_perm(q2, q3, 0x0040), 0x5410);
    return bfi(q3, bfi(q2, bfi(q1, q0 & ((1u << NB) - 1u), NB, NB), 2 * NB, NB), 3 * NB, NB);
}

// Window index written by the warp that owns token 0: header words (all but the total size, which
// only the owner of the last token knows), zeroed section padding, and the group-scale slots.
template <bool GROUP>
__device__ __forceinline__ void write_window_index(uint8_t* P, int S, int D, const PayloadLayout& lay,
                                                   const uint32_t* class_max_w, int lane, uint32_t hdr_flags) {
    if (lane < 8 && lane != 6) {
        uint32_t w = 0;
        switch (lane) {
            case 0: w = SPLITQ_MAGIC; break;
            case 1: w = GROUP ? SPLITQ_SCALE_GROUP : SPLITQ_SCALE_TOKEN; break;
            case 2: w = (uint32_t)S; break;
            case 3: w = (uint32_t)D; break;
            case 4: w = lay.n_scales; break;
            case 5: w = lay.codes_off; break;
            default: w = hdr_flags; break;  // word 7: SPLITQ_HDR_DEQUANT_RECIP
        }
        reinterpret_cast<uint32_t*>(P)[lane] = w;
    }
    for (uint32_t i = (uint32_t)S + lane; i < pad16((uint32_t)S); i += 32) P[lay.bits_off + i] = 0;
    for (uint32_t i = 4u * S + lane; i < pad16(4u * S); i += 32) P[lay.rowoff_off + i] = 0;
    if (GROUP) {
        if (lane < 4) reinterpret_cast<float*>(P + lay.scales_off)[lane] = __uint_as_float(class_max_w[lane]);
    } else {
        for (uint32_t i = 4u * S + lane; i < pad16(4u * S); i += 32) P[lay.scales_off + i] = 0;
    }
}

// ---- phase 1 of group mode: absmax per width class over the selected rows ---------------------
template <bool VEC>
__global__ void __launch_bounds__(kQWarps * 32)
class_absmax_kernel(const float* __restrict__ h, int S, int D, const uint8_t* __restrict__ bits,
                    uint32_t* __restrict__ class_max /* (B,4) |x| bit patterns, pre-zeroed */) {
    __shared__ uint32_t smax[4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    if (threadIdx.x < 4) smax[threadIdx.x] = 0u;
    __syncthreads();
    const int t0 = blockIdx.x * (kQWarps * 4);
    for (int q = 0; q < 4; ++q) {
        const int t = t0 + q * kQWarps + warp;
        if (t >= S) break;
        const int nb = bits[(size_t)b * S + t];
        if (nb == SPLITQ_BITS_RAW) continue;
        const float* row = h + ((size_t)b * S + t) * D;
        uint32_t mx = 0u;
        if (VEC) {
            const float4* src = reinterpret_cast<const float4*>(row);
            for (int c = lane; c < D / 4; c += 32) {
                float4 v = __ldg(src + c);  // stays in L2 for phase 2
                mx = max(mx, __float_as_uint(v.x) & 0x7fffffffu);
                mx = max(mx, __float_as_uint(v.y) & 0x7fffffffu);
                mx = max(mx, __float_as_uint(v.z) & 0x7fffffffu);
                mx = max(mx, __float_as_uint(v.w) & 0x7fffffffu);
            }
        } else {
            for (int d = lane; d < D; d += 32) mx = max(mx, __float_as_uint(__ldg(row + d)) & 0x7fffffffu);
        }
        mx = __reduce_max_sync(0xffffffffu, mx);
        if (lane == 0) atomicMax(&smax[slot_of_bits(nb)], mx);
    }
    __syncthreads();
    if (threadIdx.x < 3 && smax[threadIdx.x] != 0u) atomicMax(&class_max[b * 4 + threadIdx.x], smax[threadIdx.x]);
}

// ---- one token row, resident in registers as MAXCH float4 per lane (chunk c = lane + 32 u) -> payload ----
template <int MAXCH, bool GROUP>
__device__ __forceinline__ void emit_row(const float4 (&v)[MAXCH], uint8_t* __restrict__ P, const PayloadLayout& lay, int t,
                                         int S, int D, int nb, uint32_t off16, int lane,
                                         const uint32_t* __restrict__ class_max_w, uint32_t hdr_flags) {
    const int nch = D >> 2;
    float* scales = reinterpret_cast<float*>(P + lay.scales_off);
    if (t == 0) write_window_index<GROUP>(P, S, D, lay, class_max_w, lane, hdr_flags);
    if (lane == 0) {
        P[lay.bits_off + t] = (uint8_t)nb;
        reinterpret_cast<uint32_t*>(P + lay.rowoff_off)[t] = off16;
        if (t == S - 1)  // total payload size = end of the last row
            reinterpret_cast<uint32_t*>(P)[6] = lay.codes_off + 16u * off16 + row_bytes((uint32_t)D, (uint32_t)nb);
    }
    uint8_t* dst = P + lay.codes_off + (size_t)off16 * 16u;

    if (nb == SPLITQ_BITS_RAW) {
        if (!GROUP && lane == 0) scales[t] = 0.f;
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if (c < nch) reinterpret_cast<float4*>(dst)[c] = v[u];
        }
        return;
    }

    float m;
    if (GROUP) {
        m = __uint_as_float(class_max_w[slot_of_bits(nb)]);
    } else {
        uint32_t mx = 0u;
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            mx = max(mx, __float_as_uint(v[u].x) & 0x7fffffffu);
            mx = max(mx, __float_as_uint(v[u].y) & 0x7fffffffu);
            mx = max(mx, __float_as_uint(v[u].z) & 0x7fffffffu);
            mx = max(mx, __float_as_uint(v[u].w) & 0x7fffffffu);
        }
        m = __uint_as_float(__reduce_max_sync(0xffffffffu, mx));
        if (lane == 0) scales[t] = m;
    }
    const float L = (float)((1 << (nb - 1)) - 1);
    const float lo = -(L + 1.f);
    const float rcp = __fdiv_rn(1.0f, m);
    const uint32_t rb = row_bytes((uint32_t)D, (uint32_t)nb);
    // chunk c (4 elements) occupies nb/2 bytes at byte c*nb/2; chunks past D pad the row with zeros
    if (nb == 8) {           // one copy of the loop per width: shifts, masks and the store width are constants
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if ((uint32_t)c * 4u < rb) reinterpret_cast<uint32_t*>(dst)[c] = pack4_w<8>(v[u], m, rcp, L, lo);
        }
    } else if (nb == 4) {
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if ((uint32_t)c * 2u < rb) reinterpret_cast<uint16_t*>(dst)[c] = (uint16_t)pack4_w<4>(v[u], m, rcp, L, lo);
        }
    } else {
#pragma unroll
        for (int u = 0; u < MAXCH; ++u) {
            const int c = lane + 32 * u;
            if ((uint32_t)c < rb) dst[c] = (uint8_t)pack4_w<2>(v[u], m, rcp, L, lo);
        }
    }
    // rows wider than the register tile never reach here (dispatch guarantees nch <= 32*MAXCH);
    // padding beyond 32*MAXCH chunks cannot exist because rb <= pad16(nch*nb/2) < (nch+32)*nb/2
}

// ---- main kernel: one warp per token, row resident in registers -------------------------------
template <int MAXCH, bool GROUP>
__global__ void __launch_bounds__(kQWarps * 32, qp_min_blocks(MAXCH))
quant_pack_kernel(const float* __restrict__ h, int S, int D, const uint8_t* __restrict__ bits,
                  const uint32_t* __restrict__ row_off, uint8_t* __restrict__ payload, size_t stride,
                  const uint32_t* __restrict__ class_max, uint32_t hdr_flags) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * kQWarps + warp;
    const int b = blockIdx.y;
    if (t >= S) return;
    const size_t tok = (size_t)b * S + t;
    const PayloadLayout lay = payload_layout((uint32_t)S, GROUP ? SPLITQ_SCALE_GROUP : SPLITQ_SCALE_TOKEN);
    const int nch = D >> 2;
    const float4* src = reinterpret_cast<const float4*>(h + tok * D);
    float4 v[MAXCH];
#pragma unroll
    for (int u = 0; u < MAXCH; ++u) {
        const int c = lane + 32 * u;
        v[u] = c < nch ? ld_stream_f4(src + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    emit_row<MAXCH, GROUP>(v, payload + (size_t)b * stride, lay, t, S, D, bits[tok], row_off[tok], lane, class_max + b * 4, hdr_flags);
}

// ---- generic path: any D / unaligned rows; one warp per token, scalar accesses ----------------
template <bool GROUP>
__global__ void __launch_bounds__(kQWarps * 32)
quant_pack_generic_kernel(const float* __restrict__ h, int S, int D, const uint8_t* __restrict__ bits,
                          const uint32_t* __restrict__ row_off, uint8_t* __restrict__ payload,
                          size_t stride, const uint32_t* __restrict__ class_max, uint32_t hdr_flags) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * kQWarps + warp;
    const int b = blockIdx.y;
    if (t >= S) return;
    const size_t tok = (size_t)b * S + t;
    const PayloadLayout lay = payload_layout((uint32_t)S, GROUP ? SPLITQ_SCALE_GROUP : SPLITQ_SCALE_TOKEN);
    uint8_t* P = payload + (size_t)b * stride;
    const int nb = bits[tok];
    const uint32_t off16 = row_off[tok];
    const float* row = h + tok * D;
    float* scales = reinterpret_cast<float*>(P + lay.scales_off);
    if (t == 0) write_window_index<GROUP>(P, S, D, lay, class_max + b * 4, lane, hdr_flags);
    if (lane == 0) {
        P[lay.bits_off + t] = (uint8_t)nb;
        reinterpret_cast<uint32_t*>(P + lay.rowoff_off)[t] = off16;
        if (t == S - 1)  // total payload size = end of the last row
            reinterpret_cast<uint32_t*>(P)[6] = lay.codes_off + 16u * off16 + row_bytes((uint32_t)D, (uint32_t)nb);
    }
    uint8_t* dst = P + lay.codes_off + (size_t)off16 * 16u;
    const uint32_t rb = row_bytes((uint32_t)D, (uint32_t)nb);
    if (nb == SPLITQ_BITS_RAW) {
        if (!GROUP && lane == 0) scales[t] = 0.f;
        for (uint32_t d = lane; d < rb / 4; d += 32)
            reinterpret_cast<float*>(dst)[d] = d < (uint32_t)D ? __ldg(row + d) : 0.f;
        return;
    }
    float m;
    if (GROUP) {
        m = __uint_as_float(class_max[b * 4 + slot_of_bits(nb)]);
    } else {
        uint32_t mx = 0u;
        for (int d = lane; d < D; d += 32) mx = max(mx, __float_as_uint(__ldg(row + d)) & 0x7fffffffu);
        m = __uint_as_float(__reduce_max_sync(0xffffffffu, mx));
        if (lane == 0) scales[t] = m;
    }
    const float L = (float)((1 << (nb - 1)) - 1);
    const float lo = -(L + 1.f);
    const int per = 8 / nb;
    const uint32_t mask = (1u << nb) - 1u;
    for (uint32_t y = lane; y < rb; y += 32) {
        uint32_t byte = 0u;
        for (int i = 0; i < per; ++i) {
            const uint32_t e = y * per + i;
            if (e < (uint32_t)D) byte |= ((uint32_t)code_of(__ldg(row + e), m, L, lo) & mask) << (nb * i);
        }
        dst[y] = (uint8_t)byte;
    }
}

template <bool GROUP> constexpr auto quant_pack_kernel_4 = quant_pack_kernel<4, GROUP>;
template <bool GROUP> constexpr auto quant_pack_kernel_8 = quant_pack_kernel<8, GROUP>;
template <bool GROUP> constexpr auto quant_pack_kernel_16 = quant_pack_kernel<16, GROUP>;
template <bool GROUP> constexpr auto quant_pack_kernel_32 = quant_pack_kernel<32, GROUP>;

// Both entry points launch through here: the optional class-maximum phase (group mode; `class_max`
// = B x 4 words that this function clears first) and the pack kernel sized for the row width.
static int launch_quant_pack(const char* who, const float* h, int B, int S, int D, const uint8_t* bits,
                             const uint32_t* row_off, bool group, uint8_t* payload, size_t payload_stride,
                             uint32_t* class_max, uint32_t hdr_flags, cudaStream_t st) {
    SPLITQ_REQUIRE(aligned16(payload) && payload_stride % 16 == 0, SPLITQ_ERR_ALIGN,
                   "%s: payload and payload_stride must be 16-byte aligned", who);
    // the payload cannot be larger than an all-raw window; the caller sizes it with
    // splitq_payload_bytes(), we can only bound-check the stride against the smallest legal size
    const PayloadLayout lay = payload_layout((uint32_t)S, group ? SPLITQ_SCALE_GROUP : SPLITQ_SCALE_TOKEN);
    SPLITQ_REQUIRE(B == 1 || payload_stride >= lay.codes_off, SPLITQ_ERR_WORKSPACE,
                   "%s: payload_stride %zu smaller than the payload index (%u)", who, payload_stride, lay.codes_off);
    if (int rc = check_arch()) return rc;
    const bool vec = (D % 4 == 0) && aligned16(h) && D <= 4096;
    if (group) {
        // the scratch keeps whatever the previous call left: the maxima must start from zero
        SPLITQ_CUDA(cudaMemsetAsync(class_max, 0, (size_t)B * 4 * sizeof(uint32_t), st));
        dim3 g1(cdiv(S, kQWarps * 4), B);
        if (vec) class_absmax_kernel<true><<<g1, kQWarps * 32, 0, st>>>(h, S, D, bits, class_max);
        else class_absmax_kernel<false><<<g1, kQWarps * 32, 0, st>>>(h, S, D, bits, class_max);
        SPLITQ_LAUNCH_CHECK("class_absmax_kernel");
    }
    dim3 grid(cdiv(S, kQWarps), B);
#define SPLITQ_QP(KERNEL)                                                                                      \
    do {                                                                                                       \
        if (group)                                                                                             \
            KERNEL<true><<<grid, kQWarps * 32, 0, st>>>(h, S, D, bits, row_off, payload, payload_stride, class_max, hdr_flags); \
        else                                                                                                   \
            KERNEL<false><<<grid, kQWarps * 32, 0, st>>>(h, S, D, bits, row_off, payload, payload_stride, class_max, hdr_flags); \
    } while (0)
    const int nch = D / 4;
    if (!vec) SPLITQ_QP(quant_pack_generic_kernel);
    else if (nch <= 32 * 4) SPLITQ_QP(quant_pack_kernel_4);
    else if (nch <= 32 * 8) SPLITQ_QP(quant_pack_kernel_8);
    else if (nch <= 32 * 16) SPLITQ_QP(quant_pack_kernel_16);
    else SPLITQ_QP(quant_pack_kernel_32);
#undef SPLITQ_QP
    SPLITQ_LAUNCH_CHECK("quant_pack_kernel");
    return SPLITQ_OK;
}

}  // namespace splitq

using namespace splitq;

extern "C" size_t splitq_quant_pack_workspace_bytes(int B, int S, int D, int scale_mode) {
    (void)S; (void)D;
    if (B <= 0) return 0;
    return (scale_mode & 0xff) == SPLITQ_SCALE_GROUP ? (size_t)B * 4 * sizeof(uint32_t) : 0;
}

extern "C" int splitq_quant_pack(const float* h, int B, int S, int D, const uint8_t* bits,
                                 const uint32_t* row_off, int scale_mode, uint8_t* payload,
                                 size_t payload_stride, void* ws, size_t ws_bytes,
                                 splitq_stream_t stream) {
    SPLITQ_REQUIRE(h && bits && row_off && payload, SPLITQ_ERR_ARG, "quant_pack: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0, SPLITQ_ERR_ARG, "quant_pack: bad shape B=%d S=%d D=%d", B, S, D);
    SPLITQ_REQUIRE(B <= 65535, SPLITQ_ERR_ARG, "quant_pack: B=%d > 65535", B);
    const uint32_t hdr_flags = (scale_mode & SPLITQ_FLAG_DEQUANT_RECIP) ? SPLITQ_HDR_DEQUANT_RECIP : 0u;
    scale_mode &= ~SPLITQ_FLAG_DEQUANT_RECIP;
    SPLITQ_REQUIRE(scale_mode == SPLITQ_SCALE_GROUP || scale_mode == SPLITQ_SCALE_TOKEN, SPLITQ_ERR_ARG,
                   "quant_pack: unknown scale_mode %d", scale_mode);
    const bool group = scale_mode == SPLITQ_SCALE_GROUP;
    if (group) {
        const size_t need = (size_t)B * 4 * sizeof(uint32_t);
        SPLITQ_REQUIRE(ws && ws_bytes >= need, SPLITQ_ERR_WORKSPACE, "quant_pack: workspace %zu < %zu bytes",
                       ws_bytes, need);
    }
    return launch_quant_pack("quant_pack", h, B, S, D, bits, row_off, group, payload, payload_stride, (uint32_t*)ws,
                             hdr_flags, (cudaStream_t)stream);
}

// ---- allocation-free group-mode variant -------------------------------------------------------
// Group mode needs a few words of scratch for the per-class maxima.  Some hosts cannot supply
// it: a serving runtime that handed every byte of HBM to its own pool at start-up (a late
// cudaMalloc fails against that pool, and it also stalls every stream on the device), or a
// plugin hook whose contract forbids allocating while it enqueues work.  For them the scratch
// lives in the library image instead, once per GPU that loads it, for the life of the process.
namespace splitq {

static __device__ __align__(256) uint32_t g_static_class_max[SPLITQ_NOALLOC_MAX_WINDOWS * 4];

// In host code `g_static_class_max` is only a handle; the usable device address has to be asked
// for, with the right device current, once per device.
static int static_class_max_ptr(uint32_t** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static uint32_t* ptrs[kMaxDevices] = {};
    int dev = -1;
    SPLITQ_CUDA(cudaGetDevice(&dev));
    SPLITQ_REQUIRE(dev >= 0 && dev < kMaxDevices, SPLITQ_ERR_ARG, "quant_pack_group_noalloc: device %d out of range",
                   dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        SPLITQ_CUDA(cudaGetSymbolAddress(&p, g_static_class_max));
        SPLITQ_REQUIRE(p != nullptr, SPLITQ_ERR_CUDA, "quant_pack_group_noalloc: no address for the static scratch");
        ptrs[dev] = static_cast<uint32_t*>(p);
    }
    *out = ptrs[dev];
    return SPLITQ_OK;
}

}  // namespace splitq

extern "C" int splitq_quant_pack_group_noalloc(const float* h, int B, int S, int D, const uint8_t* bits,
                                               const uint32_t* row_off, uint8_t* payload,
                                               size_t payload_stride, splitq_stream_t stream) {
    SPLITQ_REQUIRE(h && bits && row_off && payload, SPLITQ_ERR_ARG, "quant_pack_group_noalloc: NULL buffer");
    SPLITQ_REQUIRE(B > 0 && S > 0 && D > 0, SPLITQ_ERR_ARG, "quant_pack_group_noalloc: bad shape B=%d S=%d D=%d",
                   B, S, D);
    SPLITQ_REQUIRE(B <= SPLITQ_NOALLOC_MAX_WINDOWS, SPLITQ_ERR_ARG,
                   "quant_pack_group_noalloc: B=%d windows exceeds the built-in scratch (max %d per call); "
                   "split the batch or use splitq_quant_pack with a caller workspace",
                   B, SPLITQ_NOALLOC_MAX_WINDOWS);
    if (int rc = check_arch()) return rc;
    uint32_t* class_max = nullptr;
    if (int rc = static_class_max_ptr(&class_max)) return rc;
    return launch_quant_pack("quant_pack_group_noalloc", h, B, S, D, bits, row_off, true, payload, payload_stride,
                             class_max, 0u, (cudaStream_t)stream);
}
