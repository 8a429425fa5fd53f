// Status/error plumbing, size queries and the device-to-device hop of the C ABI.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace splitq {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_arch() {
    int dev = 0, major = 0;
    SPLITQ_CUDA(cudaGetDevice(&dev));
    SPLITQ_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    SPLITQ_REQUIRE(major == 10, SPLITQ_ERR_ARCH,
                   "device %d has compute capability %d.x; this library is built for sm_100a only", dev,
                   major);
    return SPLITQ_OK;
}

static bool valid_bits(int b) { return b == 2 || b == 4 || b == 8; }

const char* last_error() { return g_err; }

int validate_ladder(int S, const int* k, const int* bits, int n) {
    SPLITQ_REQUIRE(n >= 0 && n <= SPLITQ_MAX_LADDER, SPLITQ_ERR_ARG, "n_ladder=%d not in [0,%d]", n,
                   SPLITQ_MAX_LADDER);
    SPLITQ_REQUIRE(n == 0 || (k && bits), SPLITQ_ERR_ARG, "ladder arrays are NULL");
    long long used = 0;
    for (int i = 0; i < n; ++i) {
        SPLITQ_REQUIRE(valid_bits(bits[i]), SPLITQ_ERR_ARG, "ladder width %d not in {2,4,8}", bits[i]);
        SPLITQ_REQUIRE(k[i] >= 0, SPLITQ_ERR_ARG, "ladder count %d is negative", k[i]);
        for (int j = 0; j < i; ++j)
            SPLITQ_REQUIRE(bits[j] != bits[i], SPLITQ_ERR_ARG, "width %d appears twice in the ladder",
                           bits[i]);
        used += k[i];
    }
    SPLITQ_REQUIRE(used <= S, SPLITQ_ERR_ARG, "ladder selects %lld tokens but the window has %d", used, S);
    return SPLITQ_OK;
}

}  // namespace splitq

using namespace splitq;

extern "C" int splitq_version(void) { return 100; }

extern "C" const char* splitq_last_error(void) { return last_error(); }

extern "C" size_t splitq_payload_bytes(int S, int D, const int* ladder_k_host,
                                       const int* ladder_bits_host, int n_ladder, int scale_mode) {
    if (S <= 0 || D <= 0 || validate_ladder(S, ladder_k_host, ladder_bits_host, n_ladder) != SPLITQ_OK ||
        (scale_mode != SPLITQ_SCALE_GROUP && scale_mode != SPLITQ_SCALE_TOKEN))
        return 0;
    PayloadLayout l = payload_layout((uint32_t)S, scale_mode);
    size_t n = l.codes_off;
    long long used = 0;
    for (int i = 0; i < n_ladder; ++i) {
        n += (size_t)ladder_k_host[i] * row_bytes(D, ladder_bits_host[i]);
        used += ladder_k_host[i];
    }
    n += (size_t)(S - used) * row_bytes(D, SPLITQ_BITS_RAW);
    return n;
}

extern "C" int splitq_transfer_peer(void* dst, int dst_device, const void* src, int src_device,
                                    size_t bytes, splitq_stream_t stream) {
    SPLITQ_REQUIRE(dst && src, SPLITQ_ERR_ARG, "transfer: NULL buffer");
    if (bytes == 0) return SPLITQ_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dst_device == src_device)
        SPLITQ_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st));
    else
        SPLITQ_CUDA(cudaMemcpyPeerAsync(dst, dst_device, src, src_device, bytes, st));
    return SPLITQ_OK;
}
