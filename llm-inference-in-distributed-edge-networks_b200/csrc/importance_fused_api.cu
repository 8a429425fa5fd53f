// K2's exported entry points (include/splitq.h).  importance_fused.cu is built twice -- two and three
// epilogue warpgroups (512 / 640 threads per CTA) -- and each call is handed to the build that is faster
// for its shape.  Measured on B200, 32768 tokens per call (benchmarks/k2_shapes.py, profiles/r2_h_k2_shapes.txt):
// three groups win where items are long and every CTA has a few hundred tiles (S >= 4096 with several
// windows: -3 % at 8 x 4096, -4 ... -8 % at S = 8192, -3 ... -13 % at S = 16384), two groups win on short
// sequences and single windows (S <= 1024: 5-7 %), where the extra group's hand-offs per item cost more than
// its MUFU overlap gains.
#include <stdlib.h>

#include "common.cuh"

namespace splitq {
#define SPLITQ_K2_DECLARE(ns)                                                                                          \
    namespace ns {                                                                                                     \
    int tile_cols();                                                                                                   \
    size_t workspace_bytes(int B, int H, int KVH, int S, int hd);                                                      \
    int fused_impl(const float* Q, const long long* q_strides_host, const float* K, const long long* k_strides_host,  \
                   int B, int H, int KVH, int S, int hd, float scaling, int causal, int mode,                          \
                   const float* head_w_host, float* out_col, float* out_last, const float* head_w2_host,               \
                   float* out_col2, void* ws, size_t ws_bytes, splitq_stream_t stream);                                \
    }
SPLITQ_K2_DECLARE(k2_g2)
SPLITQ_K2_DECLARE(k2_g3)
#undef SPLITQ_K2_DECLARE

namespace {
// SPLITQ_K2_GROUPS=2|3 pins the build (the tests run every case on both, timing experiments compare them);
// read per call so that a test can change it; anything else: by shape
int pinned_groups() {
    const char* e = getenv("SPLITQ_K2_GROUPS");
    return e ? atoi(e) : 0;
}

bool three_groups(int B, int H, int S, int causal) {
    if (k2_g3::tile_cols() != 128) return false;   // the wide-tile build switch exists for two groups only
    const int pin = pinned_groups();
    if (pin == 2 || pin == 3) return pin == 3;
    const long long nb = cdiv(S, 128);
    const long long tiles = (long long)B * H * (causal ? nb * (nb + 1) / 2 : nb * nb);  // per pass
    return nb >= 32 && tiles >= 256ll * 148;  // items of up to >= 32 tiles, >= 256 tiles per CTA and pass
}

int dispatch(const float* Q, const long long* q_strides_host, const float* K, const long long* k_strides_host, int B,
             int H, int KVH, int S, int hd, float scaling, int causal, int mode, const float* head_w_host,
             float* out_col, float* out_last, const float* head_w2_host, float* out_col2, void* ws, size_t ws_bytes,
             splitq_stream_t stream) {
    // shapes are validated by the build that runs; a nonsensical one goes to the two-group build's checks
    const bool g3 = B > 0 && H > 0 && S > 0 && three_groups(B, H, S, causal);
    return (g3 ? k2_g3::fused_impl : k2_g2::fused_impl)(Q, q_strides_host, K, k_strides_host, B, H, KVH, S, hd, scaling,
                                                        causal, mode, head_w_host, out_col, out_last, head_w2_host,
                                                        out_col2, ws, ws_bytes, stream);
}
}  // namespace
}  // namespace splitq

using namespace splitq;

extern "C" int splitq_importance_fused_tile_cols(void) { return k2_g2::tile_cols(); }

// one size for both builds (the three-group build keeps more partial-result slots)
extern "C" size_t splitq_importance_fused_workspace_bytes(int B, int H, int KVH, int S, int hd) {
    const size_t a = k2_g2::workspace_bytes(B, H, KVH, S, hd), b = k2_g3::workspace_bytes(B, H, KVH, S, hd);
    return a > b ? a : b;
}

extern "C" int splitq_importance_fused(const float* Q, const long long* q_strides_host, const float* K,
                                       const long long* k_strides_host, int B, int H, int KVH, int S, int hd,
                                       float scaling, int causal, int mode, const float* head_w_host,
                                       float* out_col, float* out_last, void* ws, size_t ws_bytes,
                                       splitq_stream_t stream) {
    return dispatch(Q, q_strides_host, K, k_strides_host, B, H, KVH, S, hd, scaling, causal, mode, head_w_host, out_col,
                    out_last, nullptr, nullptr, ws, ws_bytes, stream);
}

extern "C" int splitq_importance_fused_dual(const float* Q, const long long* q_strides_host, const float* K,
                                            const long long* k_strides_host, int B, int H, int KVH, int S, int hd,
                                            float scaling, int causal, int mode, const float* head_w_host,
                                            float* out_col, float* out_last, const float* head_w2_host,
                                            float* out_col2, void* ws, size_t ws_bytes, splitq_stream_t stream) {
    return dispatch(Q, q_strides_host, K, k_strides_host, B, H, KVH, S, hd, scaling, causal, mode, head_w_host, out_col,
                    out_last, head_w2_host, out_col2, ws, ws_bytes, stream);
}
