/*
 * splitq.h -- C ABI of the B200 (sm_100a) split-inference boundary path.
 *
 * Drop-in boundary for sv-goat/LLM-Inference-in-Distributed-Edge-Networks.  The reference is
 * pure Python/ATen and has no FFI; each entry point below names the reference lines whose
 * work it replaces (paths relative to Experiments/ in the reference tree).  A maintainer
 * binds these with ctypes; the shipped binding is
 * llm-inference-in-distributed-edge-networks_b200/splitq/_native.py.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - the caller owns every buffer (inputs, outputs, payload, workspace); nothing here
 *     allocates, frees or synchronises; work is enqueued on `stream` (a cudaStream_t);
 *   - functions return SPLITQ_OK or a negative SPLITQ_ERR_* code and record a message that
 *     splitq_last_error() returns (thread-local); no C++ exception crosses this boundary;
 *   - tensors are dense, row-major, fp32; batch B = number of evaluation windows processed
 *     by one call (the reference always has B = 1, last_row_exp.py:20).
 *
 * Payload ("SPQ1", little-endian, one per window; windows of a batch are `payload_stride`
 * bytes apart; every section starts on a 16-byte boundary):
 *   [0,32)   u32 magic 'SPQ1', u32 scale_mode, u32 S, u32 D, u32 n_scales, u32 codes_offset,
 *            u32 payload_bytes, u32 flags (bit 0 = SPLITQ_HDR_DEQUANT_RECIP, else 0)
 *   bits     u8[S]   per-token width: 2, 4, 8 or 32 (= raw fp32 row)
 *   row_off  u32[S]  row start, in 16-byte units from codes_offset
 *   scales   f32[n_scales]  group mode: 4 slots (8-bit, 4-bit, 2-bit, unused); token mode: S;
 *            channel modes (scale_mode = SPLITQ_SCALE_CHANNEL_BASE + method): D, one per channel
 *   codes    rows in ascending token order; a b-bit row holds D two's-complement fields,
 *            element e at bit e*b (LSB first), padded to 16 bytes; a raw row is D floats.
 */
#ifndef SPLITQ_H_
#define SPLITQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPLITQ_OK 0
#define SPLITQ_ERR_ARG (-1)       /* bad shape / null pointer / unsupported size */
#define SPLITQ_ERR_ALIGN (-2)     /* pointer or stride not aligned as documented */
#define SPLITQ_ERR_WORKSPACE (-3) /* workspace or payload buffer too small */
#define SPLITQ_ERR_ARCH (-4)      /* current device is not compute capability 10.x */
#define SPLITQ_ERR_CUDA (-5)      /* a CUDA runtime call failed; see splitq_last_error() */

#define SPLITQ_MAGIC 0x31515053u
#define SPLITQ_HEADER_BYTES 32
#define SPLITQ_SCALE_GROUP 0 /* one absmax per width class (reference: one per selected set) */
#define SPLITQ_SCALE_TOKEN 1 /* one absmax per token (extension) */
#define SPLITQ_SCALE_CHANNEL_BASE 2 /* + SPLITQ_CHANNEL_* : one scale per channel (splitq_channel_quant_pack) */
/* OR into splitq_quant_pack's scale_mode: the receiver dequantises as (q * fl(1/L)) * m instead of
 * (q / L) * m.  The reference writes `quantized_values / 7 * max_val` (pythia_model.py:186); CPU
 * ATen divides (the default here, bit-identical), CUDA ATen -- the device every reference entry
 * point hard-codes (last_row_exp.py:57) -- multiplies a tensor by the reciprocal of a Python
 * scalar, which differs from the quotient by one ulp for some codes.  With this flag the
 * dequantised values are bit-identical to the reference run on a GPU.  Codes are the same either
 * way (`x / max_val` divides by a tensor on both devices).  Stored in header word 7. */
#define SPLITQ_FLAG_DEQUANT_RECIP 0x100
#define SPLITQ_HDR_DEQUANT_RECIP 1u
#define SPLITQ_BITS_RAW 32
#define SPLITQ_MAX_LADDER 3
#define SPLITQ_MAX_HEADS 128
#define SPLITQ_MAX_SELECT_S 32768
#define SPLITQ_NORM_LAYERNORM 0
#define SPLITQ_NORM_RMSNORM 1
#define SPLITQ_FLAG_CAUSAL_HINT 1u /* caller guarantees A[i][j] == 0 for j > i (HF causal eager) */

typedef void* splitq_stream_t; /* cudaStream_t */

int splitq_version(void);
const char* splitq_last_error(void);

/* ---- importance over materialised attention (K1) ---------------------------------------
 * Replaces get_importance_order's per-layer reductions: "regular_importance"
 * (Pythia-70M/last_row_exp.py:13-23, Qwen2-0.5B/main.py:46-56), "weighted_importance"
 * (Qwen2-0.5B/main.py:57-78) and "last_row" (last_row_exp.py:24-30, main.py:80-86).
 *   A        (B,H,Sq,Sk) attention probabilities
 *   head_w_host  H host floats or NULL (NULL = 1/H each, i.e. the plain head mean)
 *   out_col  (B,Sk)  (1/Sq) sum_i sum_h w_h A[b,h,i,j]   or NULL
 *   out_last (B,Sk)  sum_h w_h A[b,h,Sq-1,j]              or NULL
 *   ws       scratch of splitq_importance_colmean_workspace_bytes(); contents undefined after */
size_t splitq_importance_colmean_workspace_bytes(int B, int H, int Sq, int Sk);
int splitq_importance_colmean(const float* A, int B, int H, int Sq, int Sk, const float* head_w_host,
                              float* out_col, float* out_last, void* ws, size_t ws_bytes,
                              unsigned flags, splitq_stream_t stream);

/* ---- fused importance from Q and K (K2) ------------------------------------------------
 * Replaces the materialisation of attention probabilities by HF eager attention
 * (output_attentions=True, Pythia-70M/last_row_exp.py:92-95, Qwen2-0.5B/main.py:158-160) AND the
 * reductions above: column-mean and last-row importance straight from the post-RoPE query/key
 * states of one layer, two tcgen05 passes (row log-sum-exp, then column sums of exp(s - lse));
 * the (S,S) matrix is never written.  out_col / out_last as for splitq_importance_colmean.
 *   Q (B,H,S,hd), K (B,KVH,S,hd): fp32, head_dim contiguous, element strides of (b, head, s) in
 *   q_strides_host[3] / k_strides_host[3] (so HF's transposed views need no copy); hd must be 64.
 *   scaling = head_dim^-0.5; causal != 0 applies the lower-triangular mask.
 *   mode: SPLITQ_FUSED_F16X3 (operands split into h0 = fp16(x), h1 = fp16(x - h0); three fp16
 *   products h0*h0 + h1*h0 + h0*h1 accumulated in fp32: ~fp32 accuracy, meets 1e-5) or
 *   SPLITQ_FUSED_F16X1 (h0*h0 only, 11-bit mantissas).  Any finite magnitude works: a 128-row block
 *   whose largest |x| leaves [2^-10, 2^15) is multiplied by a power of two before the fp16 conversion
 *   and the factor is restored per tile; non-finite inputs give non-finite scores.
 *   Q must stay unchanged until the call's kernels retire: the first pass reads it in place.
 *   Two builds of the kernels exist (two / three epilogue warpgroups); each call picks one by shape.
 *   Scores are bit-reproducible for a given build; SPLITQ_K2_GROUPS=2|3 in the environment pins one.
 *   ws: 256-byte aligned scratch of splitq_importance_fused_workspace_bytes() (covers both builds). */
#define SPLITQ_FUSED_F16X3 0
#define SPLITQ_FUSED_F16X1 1
size_t splitq_importance_fused_workspace_bytes(int B, int H, int KVH, int S, int hd);
/* columns of one tensor-core tile of the passes (128 fixed rows x this many streamed rows); lets a
 * caller count the MMA work a call issues (bench.py's roofline) */
int splitq_importance_fused_tile_cols(void);
int splitq_importance_fused(const float* Q, const long long* q_strides_host, const float* K,
                            const long long* k_strides_host, int B, int H, int KVH, int S, int hd,
                            float scaling, int causal, int mode, const float* head_w_host,
                            float* out_col, float* out_last, void* ws, size_t ws_bytes,
                            splitq_stream_t stream);

/* Same call with a SECOND set of head weights over the same per-head column sums: one scoring pass
 * yields the plain column mean / last row (head_w_host, NULL = 1/H) and the "weighted_importance" of
 * Qwen2-0.5B/main.py:57-78 (head_w2_host -> out_col2) -- the reference's loop evaluates both methods
 * for every window (main.py:151-178). */
int splitq_importance_fused_dual(const float* Q, const long long* q_strides_host, const float* K,
                                 const long long* k_strides_host, int B, int H, int KVH, int S, int hd,
                                 float scaling, int causal, int mode, const float* head_w_host,
                                 float* out_col, float* out_last, const float* head_w2_host,
                                 float* out_col2, void* ws, size_t ws_bytes, splitq_stream_t stream);

/* ---- selection (K3a) -------------------------------------------------------------------
 * Replaces torch.argsort(importance)[:int(ratio*S)] (Pythia-70M/pythia_model.py:169-170,
 * Qwen2-0.5B/qwen_layer_wise.py:57), generalised to a ladder: the ladder_k_host[0] least
 * important tokens get ladder_bits_host[0] bits, the next ladder_k_host[1] get
 * ladder_bits_host[1], ...; all others stay fp32.  Ties go to the lower token index.
 *   imp (B,S); bits_out u8 (B,S); row_off_out u32 (B,S) in 16-byte units.  S <= 32768. */
int splitq_select_bits(const float* imp, int B, int S, int D, const int* ladder_k_host,
                       const int* ladder_bits_host, int n_ladder, uint8_t* bits_out,
                       uint32_t* row_off_out, splitq_stream_t stream);

/* Same selection with ONE rung of `bits` bits whose size differs per window.  Replaces the two
 * token-set rules of the reference's first experiment (Experiments/Pythia-70M/pythia_model.py):
 *   k_host != NULL              window b quantises its k_host[b] least important tokens -- the rule of
 *                               batched_quantization_helper (:141-142, `ordering[:int(0.1*ratio*len)]`);
 *   mass_threshold_host != NULL "top-rho", batched_top_rho_quantization (:93-109): walking the tokens
 *                               from the most important down, keep them while the importance mass
 *                               taken so far is below mass_threshold_host[b] (the reference uses
 *                               1 - 0.1*ratio) and quantise all the others.  (The reference then slices
 *                               with a token id where a rank is meant; the intended rule is built.)
 *                               Masses are summed as 2^-47 fixed-point integers of the scores clamped
 *                               to [0, 1], so the kept count is independent of summation order.
 * Exactly one of the two host arrays is given (B entries).  k_out (device, B ints, nullable) receives
 * the number of quantised tokens per window.  Payloads of such windows differ in size: size the
 * payload buffer for the all-raw case, splitq_payload_bytes(S, D, NULL, NULL, 0, mode). */
int splitq_select_bits_per_window(const float* imp, int B, int S, int D, int bits, const int* k_host,
                                  const double* mass_threshold_host, uint8_t* bits_out,
                                  uint32_t* row_off_out, int* k_out, splitq_stream_t stream);

/* ---- absmax + quantise + bit-pack (K3) -------------------------------------------------
 * Replaces the scale/clamp/round block (pythia_model.py:176-184, qwen_layer_wise.py:60-66)
 * and emits what the reference never materialises: integer codes packed into one payload.
 * Codes are round_half_even(clamp(h / m * L, -(L+1), L)), L = 2^(bits-1) - 1, with IEEE
 * division and no contraction, so they are bit-identical to the ATen sequence on CPU and on CUDA
 * (tests/test_cuda_aten_parity_gpu.py runs the reference's op sequence with torch on the GPU).
 * scale_mode = SPLITQ_SCALE_GROUP / _TOKEN, optionally | SPLITQ_FLAG_DEQUANT_RECIP. */
size_t splitq_payload_bytes(int S, int D, const int* ladder_k_host, const int* ladder_bits_host,
                            int n_ladder, int scale_mode);
size_t splitq_quant_pack_workspace_bytes(int B, int S, int D, int scale_mode);
int splitq_quant_pack(const float* h, int B, int S, int D, const uint8_t* bits,
                      const uint32_t* row_off, int scale_mode, uint8_t* payload,
                      size_t payload_stride, void* ws, size_t ws_bytes, splitq_stream_t stream);

/* Group-mode splitq_quant_pack for hosts that cannot allocate device memory.  It has no
 * counterpart in the reference.  Who needs it: a process whose serving runtime took all of HBM
 * into its own pool at start-up (cudaMalloc then fails or fragments, and it stalls every stream
 * of the device while it runs), or a plugin callback that is not allowed to allocate while it
 * enqueues work -- such callers have nowhere to get the class-maximum workspace from.  Here that
 * scratch is a fixed array inside the library (16 KiB on every GPU that loads it, held until
 * exit), so the call takes only inputs and outputs.  Results are identical to
 * splitq_quant_pack(..., SPLITQ_SCALE_GROUP, ...).
 *   - at most SPLITQ_NOALLOC_MAX_WINDOWS windows per call; larger B is refused (SPLITQ_ERR_ARG);
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *     with that device current); make that first call before any stream capture begins;
 *   - every call uses the same array and clears the part it needs first, so two calls whose GPU
 *     work overlaps (different streams, or different threads) race: serialise them;
 *   - no cudaMalloc / cudaFree / allocating library routine is called. */
#define SPLITQ_NOALLOC_MAX_WINDOWS 1024
int splitq_quant_pack_group_noalloc(const float* h, int B, int S, int D, const uint8_t* bits,
                                    const uint32_t* row_off, uint8_t* payload, size_t payload_stride,
                                    splitq_stream_t stream);

/* ---- dequantise (K4) -------------------------------------------------------------------
 * Replaces quantized/7*max + index_put (pythia_model.py:186-188, qwen_layer_wise.py:68-70):
 * h_out (B,S,D) = every token of the window, raw rows copied, coded rows as (q / L) * m with the
 * correctly rounded quotient (bit-identical to CPU ATen; within one ulp of CUDA ATen), or as
 * (q * fl(1/L)) * m when the payload header carries SPLITQ_HDR_DEQUANT_RECIP (bit-identical to CUDA
 * ATen).  Group mode with a zero class scale yields NaN rows, as the reference's 0/0 does. */
int splitq_dequant(const uint8_t* payload, size_t payload_stride, int B, int S, int D,
                   float* h_out, splitq_stream_t stream);

/* Same, fused with the first normalisation of the downstream half (the input_layernorm of
 * layer i+1, or the final norm when the split is after the last layer; reached through
 * pythia_model.py:165,193 / qwen_layer_wise.py:52,75).  Emits the fp32 residual stream and
 * the normalised activations; gamma2/beta2/normed2_out (nullable) give the second LayerNorm
 * GPT-NeoX's parallel-residual layer applies to the same input. */
int splitq_dequant_norm(const uint8_t* payload, size_t payload_stride, int B, int S, int D,
                        int norm_kind, const float* gamma, const float* beta, float eps,
                        const float* gamma2, const float* beta2, float* resid_out,
                        float* normed_out, float* normed2_out, splitq_stream_t stream);

/* ---- channel-wise quantisers ----------------------------------------------------------
 * Replaces QwenPointFiveBModel.channel_wise's per-channel Python loop
 * (Qwen2-0.5B/qwen_layer_wise.py:122-152): per-channel statistic over all B*S rows (absmax, or
 * mean + 1e-8 for SPLITQ_CHANNEL_1_MEAN) and the in-place quantise/dequantise of every element.
 *   h (B,S,D) -> h_out (B,S,D) (may alias h), stats_out f32[D] (the per-channel scale used),
 *   codes_out int8 (B,S,D) or NULL (the integer codes the reference never stores). */
#define SPLITQ_CHANNEL_8 0
#define SPLITQ_CHANNEL_4 1
#define SPLITQ_CHANNEL_1_MEAN 2
#define SPLITQ_CHANNEL_1_MAX 3
size_t splitq_channel_workspace_bytes(int B, int S, int D);
int splitq_channel_fake_quant(const float* h, int B, int S, int D, int method, float* stats_out,
                              float* h_out, int8_t* codes_out, void* ws, size_t ws_bytes,
                              splitq_stream_t stream);

/* ---- layer-distance analytics -----------------------------------------------------------
 * Replaces the triple Python loop of Notebooks/distributions_distance_across_layers.ipynb (cells 12,
 * 13, 16: KL in bits with a p != 0 guard, Jensen-Shannon divergence, every layer pair i < j, every
 * sample): dist (L,N,S) fp32 importance distributions of L layers over N samples ->
 * js_out (L,L,N) fp64, entry [i][j][n] = JS(dist[i][n], dist[j][n]) for i < j (other entries are
 * left untouched).  fp64 arithmetic, fixed reduction order. */
int splitq_layer_js(const float* dist, int L, int N, int S, double* js_out, splitq_stream_t stream);

/* The same quantisers emitting a payload instead of fake-quantised floats ("shares the
 * pack format, per-channel scales"): per window, the per-channel statistic (over that window's S rows;
 * the reference always has one window) goes into the scales section, every token row is packed at 8
 * (channel_8), 4 (channel_4) or 2 bits (the two ternary methods), header scale_mode =
 * SPLITQ_SCALE_CHANNEL_BASE + method.  splitq_dequant / splitq_dequant_norm read it back with the
 * reference's operation order (q * m / L for the absmax widths, q * u for the ternary ones), so
 * pack -> dequantise equals splitq_channel_fake_quant on one window bit for bit (a channel whose
 * statistic is 0 dequantises to NaN, as the reference's 0/0 does).  D % 4 == 0.
 *   ws: scratch of splitq_channel_workspace_bytes(B, S, D). */
size_t splitq_channel_payload_bytes(int S, int D, int method);
int splitq_channel_quant_pack(const float* h, int B, int S, int D, int method, uint8_t* payload,
                              size_t payload_stride, void* ws, size_t ws_bytes, splitq_stream_t stream);

/* ---- transfer --------------------------------------------------------------------------
 * The device hop the reference only simulates (in-place quant->dequant,
 * pythia_model.py:176-188).  cudaMemcpyPeerAsync over NVLink; src/dst may be one device. */
int splitq_transfer_peer(void* dst, int dst_device, const void* src, int src_device, size_t bytes,
                         splitq_stream_t stream);

/* ---- transfer between two processes (one per GPU) without any SM --------------------------
 * No reference counterpart (the reference only simulates the hop, pythia_model.py:176-188; its
 * distributed deployment is the paper's premise, not code).  The receiving rank's copy engine
 * pulls the partner's payload through a CUDA-IPC mapping (NVLink P2P between two GPUs, a plain
 * device copy when both ranks share one GPU), and the two sides order themselves with stream
 * memory operations on 32-bit flags, so nothing competes with the persistent scoring kernel for
 * SMs (splitq/transport.py IpcPairExchange shows the protocol).
 *   splitq_ipc_export: handle_out[SPLITQ_IPC_HANDLE_BYTES] + byte offset of dev_ptr inside its
 *     cudaMalloc allocation, to be sent to the partner process by any host channel.
 *   splitq_ipc_open / _close: map / unmap the partner's buffer (not in the exporting process).
 *   splitq_stream_write_u32: after all earlier work of the stream, store value at dev_addr.
 *   splitq_stream_wait_geq_u32: hold the stream until (int32)(*dev_addr - value) >= 0.
 *   splitq_copy_async: cudaMemcpyAsync over unified addressing (local, peer or IPC-mapped). */
#define SPLITQ_IPC_HANDLE_BYTES 64
int splitq_ipc_export(const void* dev_ptr, void* handle_out, size_t* offset_out);
int splitq_ipc_open(const void* handle, size_t offset, void** dev_ptr_out);
int splitq_ipc_close(void* dev_ptr, size_t offset);
int splitq_stream_write_u32(void* dev_addr, uint32_t value, splitq_stream_t stream);
int splitq_stream_wait_geq_u32(const void* dev_addr, uint32_t value, splitq_stream_t stream);
int splitq_copy_async(void* dst, const void* src, size_t bytes, splitq_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SPLITQ_H_ */
