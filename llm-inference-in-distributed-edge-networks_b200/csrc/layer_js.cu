// Layer-to-layer Jensen-Shannon divergence of importance distributions.
//
// Reference: Notebooks/distributions_distance_across_layers.ipynb cells 12, 13, 16 --
//   KL(p||q) = sum(where(p != 0, p * log2(p / q), 0)),  JS(p, q) = (KL(p||m) + KL(q||m)) / 2, m = (p + q) / 2,
// evaluated there with numpy for every layer pair i < j and every sample in three nested Python
// loops.  Here one CTA takes one (pair, sample): it streams the two fp32 distributions once, does
// the arithmetic in fp64 like numpy, and reduces in a fixed tree, so the result does not depend on
// scheduling.  The (L, N, S) tensor is what the fused scoring kernel leaves on the device.
#include "common.cuh"

namespace splitq {

__global__ void __launch_bounds__(256)
layer_js_kernel(const float* __restrict__ dist, int L, int N, int S, double* __restrict__ js /* (L, L, N) */) {
    __shared__ double red[8];
    const int n = blockIdx.x, j = blockIdx.y, i = blockIdx.z;
    if (i >= j) return;
    const float* p = dist + ((size_t)i * N + n) * S;
    const float* q = dist + ((size_t)j * N + n) * S;
    double acc = 0.0;
    for (int s = threadIdx.x; s < S; s += 256) {
        const double a = (double)__ldg(p + s), b = (double)__ldg(q + s);
        const double m = 0.5 * (a + b);
        if (a != 0.0) acc += a * log2(a / m);
        if (b != 0.0) acc += b * log2(b / m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w];
        js[((size_t)i * L + j) * N + n] = 0.5 * t;
    }
}

}  // namespace splitq

using namespace splitq;

extern "C" int splitq_layer_js(const float* dist, int L, int N, int S, double* js_out, splitq_stream_t stream) {
    SPLITQ_REQUIRE(dist && js_out, SPLITQ_ERR_ARG, "layer_js: NULL buffer");
    SPLITQ_REQUIRE(L > 0 && N > 0 && S > 0 && L <= 65535 && N <= 2147483647 / 1, SPLITQ_ERR_ARG,
                   "layer_js: bad shape L=%d N=%d S=%d", L, N, S);
    if (int rc = check_arch()) return rc;
    if (L < 2) return SPLITQ_OK;
    layer_js_kernel<<<dim3(N, L, L), 256, 0, (cudaStream_t)stream>>>(dist, L, N, S, js_out);
    SPLITQ_LAUNCH_CHECK("layer_js_kernel");
    return SPLITQ_OK;
}
