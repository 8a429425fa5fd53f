// Shared host/device helpers for the splitq kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "splitq.h"

namespace splitq {

// ---- error plumbing (thread-local message, negative status codes; no exceptions) ----------
void set_error(const char* fmt, ...);
int check_arch();  // SPLITQ_OK when the current device is compute capability 10.x
int validate_ladder(int S, const int* k, const int* bits, int n);

#define SPLITQ_REQUIRE(cond, code, ...)   \
    do {                                  \
        if (!(cond)) {                    \
            splitq::set_error(__VA_ARGS__); \
            return (code);                \
        }                                 \
    } while (0)

#define SPLITQ_CUDA(call)                                                                  \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            (void)cudaGetLastError(); /* reported here: do not leave it for the caller's next CUDA call */ \
            splitq::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                              __LINE__);                                                   \
            return SPLITQ_ERR_CUDA;                                                        \
        }                                                                                  \
    } while (0)

#define SPLITQ_LAUNCH_CHECK(name)                                                         \
    do {                                                                                  \
        cudaError_t e__ = cudaGetLastError();                                             \
        if (e__ != cudaSuccess) {                                                         \
            splitq::set_error("launch of %s failed: %s", name, cudaGetErrorString(e__));  \
            return SPLITQ_ERR_CUDA;                                                       \
        }                                                                                 \
    } while (0)

// ---- payload geometry (host + device) -----------------------------------------------------
__host__ __device__ inline uint32_t pad16(uint32_t n) { return (n + 15u) & ~15u; }
__host__ __device__ inline uint32_t row_bytes(uint32_t D, uint32_t bits) {
    return pad16((D * bits + 7u) >> 3);
}
__host__ __device__ inline int slot_of_bits(int bits) { return bits == 8 ? 0 : (bits == 4 ? 1 : 2); }

struct PayloadLayout {
    uint32_t bits_off, rowoff_off, scales_off, codes_off, n_scales;
};
// scale_mode SPLITQ_SCALE_GROUP / _TOKEN, or SPLITQ_SCALE_CHANNEL_BASE + channel method (one scale per
// channel: D entries; every token has the same width)
__host__ __device__ inline PayloadLayout payload_layout(uint32_t S, int scale_mode, uint32_t D = 0) {
    PayloadLayout l;
    l.bits_off = SPLITQ_HEADER_BYTES;
    l.rowoff_off = l.bits_off + pad16(S);
    l.scales_off = l.rowoff_off + pad16(4u * S);
    l.n_scales = scale_mode == SPLITQ_SCALE_GROUP ? 4u : (scale_mode == SPLITQ_SCALE_TOKEN ? S : D);
    l.codes_off = l.scales_off + pad16(4u * l.n_scales);
    return l;
}
__host__ __device__ inline bool is_channel_mode(int scale_mode) {
    return scale_mode >= SPLITQ_SCALE_CHANNEL_BASE && scale_mode <= SPLITQ_SCALE_CHANNEL_BASE + 3;
}
__host__ __device__ inline int channel_bits(int method) { return method == 0 ? 8 : (method == 1 ? 4 : 2); }

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- device helpers -------------------------------------------------------------------------
#ifdef __CUDACC__
// streaming 128-bit load that does not pollute L1 (inputs are read exactly once)
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// max of |x| over a warp on the raw bit pattern: non-negative floats order like unsigned ints,
// and sm_80+ has a single-instruction integer warp reduction (REDUX).
__device__ __forceinline__ float warp_absmax_bits(float v) {
    unsigned u = __float_as_uint(v) & 0x7fffffffu;
    return __uint_as_float(__reduce_max_sync(0xffffffffu, u));
}
#endif

}  // namespace splitq
