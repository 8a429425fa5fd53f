// K2: fused attention-importance scoring -- the S x S probability matrix never reaches HBM.
//
// Reference: HF eager attention materialises softmax(Q K^T * scaling + causal mask) for every
// layer (third-party transformers eager_attention_forward; reference call sites
// Experiments/Pythia-70M/last_row_exp.py:92-95, Experiments/Qwen2-0.5B/main.py:158-160) only so
// that get_importance_order can reduce it to one score per token (last_row_exp.py:13-30).
// Here the same scores are computed from the post-RoPE Q and K directly, flash-style, in two
// tensor-core passes:
//   pass 1  row block I (128 queries) x key blocks J<=I : S = Q_I K_J^T  -> per-row log-sum-exp
//   pass 2  key block J (128 keys) x query blocks I>=J  : S^T = K_J Q_I^T -> p = exp(s - lse_i),
//           summed over queries per key (one TMEM lane = one key: no cross-thread reduction)
// QK^T runs on tcgen05 (accumulators in TMEM, the streamed operand staged by TMA with 128-byte
// swizzle).  The reference computes QK^T in true fp32; one 11-bit-mantissa product misses the 1e-5
// importance tolerance for trained-scale logits, so every operand x is pre-split into
// h0 = fp16(x) and h1 = fp16(x - h0) (22 mantissa bits together) and each tile accumulates
// h0*h0 + h1*h0 + h0*h1 in fp32 -- three kind::f16 products, K = 16 per instruction.  The error is
// that of the classic 3xTF32 scheme (same mantissa widths; the dropped h1*h1 term is ~2^-22) at 3/4
// of the tensor time of one tf32 product plus two fp16 corrections, and with half the operand bytes.
// Blocks whose values leave the fp16 range are multiplied by a power of two first (see the operand
// preparation below).  SPLITQ_FUSED_F16X1 keeps only h0*h0.  Warpgroups per CTA: control (warp 0 = TMA
// producer of the streamed operand, warps 1 and 2 = MMA issuers of alternate tiles), two or three
// epilogue groups (alternate tiles, three TMEM accumulators in rotation), and loaders that put the
// fixed operand of the next work item into tensor memory, so the MMA reads A from TMEM and only B
// from shared memory.  K is split by a small kernel before pass 1; Q is split by pass 1's own loaders
// on its way into tensor memory (kFuseQ), which also write the parts out for pass 2 to stream.
#include <cuda.h>
#include <cuda_fp16.h>
#include <math.h>

#include <algorithm>
#include <type_traits>
#include <mutex>

#include "common.cuh"

// This file is compiled once per epilogue layout (SPLITQ_K2_EPI_GROUPS = 2 and 3, see the Makefile): each
// build lives in its own namespace and importance_fused_api.cu picks one per call by the problem's shape.
#ifndef SPLITQ_K2_EPI_GROUPS
#define SPLITQ_K2_EPI_GROUPS 2
#endif
#define SPLITQ_K2_CAT_(a, b) a##b
#define SPLITQ_K2_CAT(a, b) SPLITQ_K2_CAT_(a, b)
#define SPLITQ_K2_NS SPLITQ_K2_CAT(k2_g, SPLITQ_K2_EPI_GROUPS)

namespace splitq {
namespace SPLITQ_K2_NS {

constexpr int kTile = 128;                           // rows of the fixed operand = M of every MMA = TMEM lanes
// Columns of one accumulator = rows of one streamed tile = N of every MMA: 128, or (build-time switch) 192 --
// the widest that leaves room for the fixed operand in tensor memory (2 accumulators x 192 + 2 operand
// buffers x 64 = all 512 columns).  192 is parity-green and slower at the bench shape (0.775 vs 0.725 ms on
// one box): 6-9 % more tile area is issued around the causal diagonal and only two accumulators are in
// flight; the MMA issue rate, which wider tiles amortise, is not what bounds the kernel.
#ifndef SPLITQ_K2_NTILE
#define SPLITQ_K2_NTILE 128
#endif
constexpr int kTileN = SPLITQ_K2_NTILE;
static_assert(kTileN == 128 || kTileN == 192, "streamed tiles of 128 or 192 rows");
constexpr int kHd = 64;
constexpr uint32_t kFixPartBytes = kTile * 128;      // fixed operand in shared memory (kASmem): 128 rows x 64 fp16 per part
constexpr uint32_t kPartBytes = kTileN * 128;        // kTileN rows x 64 fp16: one swizzle-128B slab (h0 or h1 part)
constexpr uint32_t kOperandBytes = 2 * kPartBytes;   // h0 + h1
#ifndef SPLITQ_K2_STAGES
#define SPLITQ_K2_STAGES (SPLITQ_K2_NTILE == 192 ? 3 : 4)
#endif
constexpr int kStages = SPLITQ_K2_STAGES;            // streamed tiles in flight; one barrier pair and one TMA issue per tile
// SPLITQ_K2_A_SMEM = 1: the fixed operand stays in shared memory (SS-mode MMA, brought in by TMA like
// the streamed one) instead of tensor memory; the 128 TMEM columns this frees hold a FOURTH
// accumulator, and the loader warps have nothing to do.
#ifndef SPLITQ_K2_A_SMEM
#define SPLITQ_K2_A_SMEM 0
#endif
constexpr bool kASmem = SPLITQ_K2_A_SMEM != 0;
constexpr int kAcc = kTileN == 192 ? 2 : (kASmem ? 4 : 3);  // TMEM accumulator buffers, used in rotation
static_assert(!kASmem || kTileN == 128, "the shared-memory fixed operand variant exists for 128-column tiles only");
constexpr int kEpiGroups = SPLITQ_K2_EPI_GROUPS;      // epilogue warpgroups; tile n goes to group n % kEpiGroups
constexpr int kChunk = kEpiGroups == 2 ? (kTileN == 192 ? 48 : 64) : 32;  // accumulator columns per register buffer (two buffers per thread)
constexpr int kChunks = kTileN / kChunk;
static_assert(kChunks % 2 == 0 && kChunks * kChunk == kTileN, "an even number of chunks per tile keeps the two register buffers in step");
// SPLITQ_K2_FUSE_QSPLIT = 1 (default): pass 1's loader warps read the fp32 queries themselves -- fold the softmax
// scale, split into fp16 h0 + h1, put the block into tensor memory AND write the parts out for pass 2 to stream
// -- so the operand-split launch before pass 1 only converts K (an eighth of the bytes at Qwen's 14:2 heads).
// The per-block facts the split used to provide for the whole launch (norm bound, pre-scale exponent) then
// exist per item only, and pass 1 picks the bounded-logit (FAST) or the general epilogue body item by item.
#ifndef SPLITQ_K2_FUSE_QSPLIT
#define SPLITQ_K2_FUSE_QSPLIT 1
#endif
constexpr bool kFuseQ = SPLITQ_K2_FUSE_QSPLIT != 0 && kTileN == kTile && !kASmem;
// Result slots nobody fills (a group without a tile in a short run, the second piece of an item that is not cut)
// must still hold the neutral element when the merge kernels read all of them.  The loader warps see every run
// of their CTA and know which slots the epilogue will write, so they store the neutral element into the others;
// only the build whose loaders idle (kASmem) leaves that to the split kernel.
constexpr bool kLoaderPresets = !kASmem;
constexpr int kEpiRegs = kEpiGroups == 2 ? 200 : 128; // setmaxnreg budgets; they add up to the launch-time pool
constexpr int kCtlRegs = 40;
constexpr int kLdrRegs = kEpiGroups == 2 ? 72 : 56;
#ifndef SPLITQ_K2_MMA_WARPS
#define SPLITQ_K2_MMA_WARPS 2
#endif
constexpr int kMmaWarps = SPLITQ_K2_MMA_WARPS;        // warps 1.. of the control group issue the MMAs of alternate tiles
static_assert(kMmaWarps == 1 || kMmaWarps == 2, "1 or 2 MMA-issuing warps");
constexpr int kThreads = 128 * (kEpiGroups + 2);      // warpgroups: control, epilogue x kEpiGroups, operand loaders
static_assert(kEpiGroups == 2 || kEpiGroups == 3, "2 or 3 epilogue groups");
static_assert(128 * (kCtlRegs + kLdrRegs + kEpiGroups * kEpiRegs) <= kThreads * (65536 / kThreads / 8 * 8),
              "setmaxnreg budgets exceed the registers the CTA is launched with");
constexpr uint32_t kTmemCols = 512;                   // 3 accumulators (128 each) + 2 fixed-operand buffers (h0 32 + h1 32 each), or 4 accumulators
constexpr uint32_t kTmemA = kAcc * kTileN;            // first column of the fixed-operand buffers
static_assert(kTmemA + (kASmem ? 0 : 2 * 64) <= 512, "tensor memory overflow");
constexpr uint32_t kTmemABuf = 64;                    // columns per fixed-operand buffer (two halfs per column)
constexpr uint32_t kASmemOff = kStages * kOperandBytes;              // two fixed-operand buffers (kASmem only)
constexpr uint32_t kBarOff = kASmemOff + (kASmem ? 2 * 2 * kFixPartBytes : 0);
constexpr uint32_t kLseOff = kBarOff + 512;            // 18 mbarriers + TMEM base slot
constexpr uint32_t kFlagOff = kLseOff + 4 * kEpiGroups * 2 * kTileN * 4;    // per-query values: per epilogue warp 2 buffers x kTileN floats
constexpr uint32_t kFactOff = kFlagOff + 256;                                // 3 floats per warp: launch-wide facts
// kFuseQ: per-item facts from the loaders -- a ring of 8 (largest squared row norm, exponent) pairs for the
// epilogue, and behind it two sets of 4 per-loader-warp partials (largest row norm, largest |x| bits)
constexpr int kFactRing = 8;
constexpr uint32_t kSmemBytes = kFactOff + 8 * kFactRing + 2 * 4 * 8;
// (the dynamic smem base must be 1024-aligned; checked in the kernel)

// ---- PTX wrappers ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Protocol guard: a wait that does not complete within ~2 s traps instead of hanging the GPU.  The
// debug instantiation of the kernel (any timing-experiment bit set in `mode`) also says which
// barrier; the production kernel only traps, because a dozen inlined printf call sequences cost
// ~10 % of a code size that sits at the edge of the instruction cache, and force registers that are
// live across the call (the loaders' prefetched operand row) through local memory.
template <bool VERBOSE>
__device__ __forceinline__ void mbar_wait_t(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    const long long t0 = clock64();
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return;
        if (clock64() - t0 > 4000000000LL) {
            if (VERBOSE)
                printf("splitq K2: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", blockIdx.x, threadIdx.x,
                       bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// both operands from shared memory (K-major, 128-byte swizzle)
__device__ __forceinline__ void tc_mma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// A operand (M = 128 rows on the 128 TMEM lanes, K along columns, two halfs per column) read from
// tensor memory; fp16 operands, K = 16 per instruction
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// 32 consecutive columns of this thread's TMEM lane <- registers
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint4 (&q)[8]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
        "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr),
          "r"(q[0].x), "r"(q[0].y), "r"(q[0].z), "r"(q[0].w), "r"(q[1].x), "r"(q[1].y), "r"(q[1].z), "r"(q[1].w),
          "r"(q[2].x), "r"(q[2].y), "r"(q[2].z), "r"(q[2].w), "r"(q[3].x), "r"(q[3].y), "r"(q[3].z), "r"(q[3].w),
          "r"(q[4].x), "r"(q[4].y), "r"(q[4].z), "r"(q[4].w), "r"(q[5].x), "r"(q[5].y), "r"(q[5].z), "r"(q[5].w),
          "r"(q[6].x), "r"(q[6].y), "r"(q[6].z), "r"(q[6].w), "r"(q[7].x), "r"(q[7].y), "r"(q[7].z), "r"(q[7].w)
        : "memory");
}
// 256-bit global accesses (sm_100): a lane moves one whole 32-byte sector, so the loaders' row-per-thread
// pattern does not depend on the L1 to put two 16-byte pieces of a sector together
__device__ __forceinline__ void ldg256(const float* p, float4& a, float4& b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}
__device__ __forceinline__ void stg256(uint4* p, const uint4& a, const uint4& b) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w),
                 "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
                 : "memory");
}
// 16 consecutive columns of this thread's TMEM lane <- registers
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint4 (&q)[4]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr),
          "r"(q[0].x), "r"(q[0].y), "r"(q[0].z), "r"(q[0].w), "r"(q[1].x), "r"(q[1].y), "r"(q[1].z), "r"(q[1].w),
          "r"(q[2].x), "r"(q[2].y), "r"(q[2].z), "r"(q[2].w), "r"(q[3].x), "r"(q[3].y), "r"(q[3].z), "r"(q[3].w)
        : "memory");
}
// one lane of a converged warp; the surrounding code stays warp-uniform so that descriptors and
// addresses live in uniform registers (a divergent `if (lane == 0)` costs ~6 R2UR moves per MMA)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
#ifdef SPLITQ_K2_NO_MUFU  // timing experiment only: how much of the epilogue is the MUFU pipe?
__device__ __forceinline__ float ex2(float x) { return x * 0.5f; }
__device__ __forceinline__ float ex2_ordered(float x) { return x * 0.5f; }
#else
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// volatile: the compiler keeps these in program order, which is how the epilogue loops force
// kExpBatch exp2 in flight before the first result is consumed.  Measured (benchmarks/probes/
// mufu_probe.cu): ptxas' own schedule consumes each result three MUFUs later, which caps a lone warp at
// 13.9 clk per warp-level EX2; batches of 16 reach 9.8 (the pipe's floor is 8).
__device__ __forceinline__ float ex2_ordered(float x) {
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
#endif
#ifndef SPLITQ_K2_GROUP_DELAY
#define SPLITQ_K2_GROUP_DELAY 0
#endif
#ifndef SPLITQ_K2_EXP_BATCH
#define SPLITQ_K2_EXP_BATCH 16
#endif
constexpr int kExpBatch = SPLITQ_K2_EXP_BATCH;
// K-major, 128-byte swizzle, 8-row groups 1024 B apart (what a {64 fp16, 128 rows} TMA box writes)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 with fp16 A and B (format 0), fp32 accumulate, K-major, M = N = 128
constexpr uint32_t kIdescF16 = (1u << 4) | (((uint32_t)kTileN >> 3) << 17) | ((128u >> 4) << 24);  // M = 128, N = kTileN

struct Params {
    int B, H, KVH, S, nb;  // nb = 128-row blocks per (b,h) = items
    int nbn;               // streamed tiles (kTileN rows) per (b,h)
    int per_bh;            // tiles of one (b,h) in one pass
    int causal, x3;    // x3: three products (h0*h0 + h1*h0 + h0*h1), else h0*h0 only
    int dbg;           // bit0: skip epilogue math, bit1: skip MMA issue, bit2: skip TMEM reads, bit4: loaders skip their global loads, bit5: never take the FAST epilogue
    int nq_blk, nk_blk;  // entries of bound2 / eblk that belong to Q blocks (B*H*nb) and to K blocks (B*KVH*nb)
    int tiles;         // tiles of one pass = B * H * tiles_per_bh
    float2* ml;        // (kSlots,B,H,S) pass-1 partial (max, sum) per result slot, pre-set to (-inf, 0)
    float* lse2;       // (B,H,S) log2-domain log-sum-exp (lse_combine_kernel)
    float* w2;         // (B,H,S) 2^-lse2 (FAST epilogue of pass 2)
    const float* bound2;  // per operand block: largest squared row norm (Q rows already carry scaling * log2 e)
    // exponent e of the power-of-two pre-scale 2^-e: with 128-column tiles every streamed tile is one block of
    // the split, so each block keeps its own exponent (eblk, per block); wider tiles straddle blocks and need one
    // exponent per head (ehead: B*H query heads, then B*KVH key heads; written by the fix-up launch)
    const int* eblk;
    const int* ehead;
    int n_heads;          // B*H + B*KVH
    float* part_col;   // (kSlots,B,H,S) partial sum_i p[i][j] per result slot, pre-set to 0
    float* part_last;  // (B,H,S) p[S-1][j]
    const uint4 *q16, *k16;  // dense (n,S,64) fp16 h0 part; the h1 part follows at +q_part / +k_part uint4
    size_t q_part, k_part;
    // kFuseQ, pass 1: the fp32 queries as the caller holds them (element strides of batch, head, row; head_dim
    // contiguous), scaling * log2(e), and writable views of what the loaders produce for pass 2
    const float* qsrc;
    long long qs0, qs1, qs2;
    int qvec;             // rows may be read as float4 (1) or as 32-byte pieces (2)
    float cs;
    uint4* q16_w;
    float* bound2_w;
    int* eblk_w;
};

// ---- schedule, shared by all warp roles --------------------------------------------------------
// The tiles of one pass form ONE sequence: (b,h) by (b,h), inside it item by item (an item = one
// fixed 128-row block: a query block in pass 1, a key block in pass 2), inside an item the streamed
// blocks in order.  CTA c takes the contiguous range [c*T/G, (c+1)*T/G): every CTA gets the same
// number of tiles (+-1) whatever the causal triangle looks like, neighbouring CTAs stream the same
// (b,h) operands (L2), and an item is cut at most once (G is chosen so that a range is at least as
// long as the longest item).  The two pieces of a cut item, and the epilogue groups inside a
// piece, put their partial results into kSlots = 2 x groups separate slots that are merged in a fixed order
// afterwards, so the result does not depend on the schedule's timing.
constexpr int kSlots = 2 * kEpiGroups;
struct Seg {
    int bh, blk, ta, tb;  // streamed blocks [ta, tb) of item (bh, blk)
    int part;             // 1 if an earlier CTA holds the first streamed blocks of this item
};
template <int PASS>
struct SegIter {
    int bh, blk, t, rem, nb, nbn, S, per_bh;
    int item;   // items entered so far on this CTA's range (0 for the first): index of the loaders' per-item facts
    bool causal;
    // streamed tiles [t0, t1) of item blk.  pass 1: fixed = query block blk, streamed = key tiles up to
    // its last query; pass 2: fixed = key block blk, streamed = query tiles from its first key on.
    __device__ __forceinline__ int t0(int b) const { return PASS == 1 ? 0 : (causal ? (b * kTile) / kTileN : 0); }
    __device__ __forceinline__ int t1(int b) const {
        return PASS == 1 ? (causal ? (min((b + 1) * kTile, S) + kTileN - 1) / kTileN : nbn) : nbn;
    }
    __device__ __forceinline__ void next_item() {
        ++item;
        if (++blk == nb) { blk = 0; ++bh; }
    }
    // position on the n-th tile (n >= 0) of this CTA's range
    __device__ __forceinline__ void init(const Params& p, int n) {
        nb = p.nb; nbn = p.nbn; S = p.S; per_bh = p.per_bh;
        causal = p.causal != 0;
        item = 0;   // counted from the range's first item: callers that need it start with n = 0
        const long long T = p.tiles;
        const int lo = (int)(T * blockIdx.x / gridDim.x), hi = (int)(T * (blockIdx.x + 1) / gridDim.x);
        rem = hi - lo - n;
        const int x = lo + n;
        bh = x / per_bh;
        int y = x - bh * per_bh;
        blk = 0;
        while (y >= t1(blk) - t0(blk)) {
            y -= t1(blk) - t0(blk);
            ++blk;
        }
        t = t0(blk) + y;
    }
    // the next run of tiles inside one item; false when the range is used up
    __device__ __forceinline__ bool next(Seg& s) {
        if (rem <= 0) return false;
        const int e = t1(blk);
        const int take = min(e - t, rem);
        s.bh = bh; s.blk = blk; s.ta = t; s.tb = t + take;
        s.part = t != t0(blk);
        rem -= take;
        if (s.tb == e) { next_item(); t = t0(blk); }
        else t = s.tb;
        return true;
    }
    // tile-wise use (epilogue): the current tile is (bh, blk, t); move n tiles on inside the range
    __device__ __forceinline__ bool valid() const { return rem > 0; }
    __device__ __forceinline__ void advance(int n) {
        rem -= n;
        t += n;
        if (rem <= 0) return;
        int e = t1(blk);
        while (t >= e) {
            const int carry = t - e;
            next_item();
            t = t0(blk) + carry;
            e = t1(blk);
        }
    }
};

// ---- epilogue building blocks -------------------------------------------------------------------
// NC (32 or 64) consecutive columns of this thread's TMEM lane; asynchronous until tc_ld_wait()
template <int NC>
__device__ __forceinline__ void tc_ld_issue(uint32_t taddr, float (&v)[NC]) {
    static_assert(NC % 32 == 0 || NC % 32 == 16, "chunks of 32 columns plus an optional 16");
    if (NC % 32 == 16) {
        float* q = v + (NC / 32) * 32;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=f"(q[0]), "=f"(q[1]), "=f"(q[2]), "=f"(q[3]), "=f"(q[4]), "=f"(q[5]), "=f"(q[6]), "=f"(q[7]),
              "=f"(q[8]), "=f"(q[9]), "=f"(q[10]), "=f"(q[11]), "=f"(q[12]), "=f"(q[13]), "=f"(q[14]), "=f"(q[15])
            : "r"(taddr + (NC / 32) * 32)
            : "memory");
    }
#pragma unroll
    for (int hlf = 0; hlf < NC / 32; ++hlf) {
        float* q = v + 32 * hlf;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
            "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=f"(q[0]), "=f"(q[1]), "=f"(q[2]), "=f"(q[3]), "=f"(q[4]), "=f"(q[5]), "=f"(q[6]), "=f"(q[7]),
              "=f"(q[8]), "=f"(q[9]), "=f"(q[10]), "=f"(q[11]), "=f"(q[12]), "=f"(q[13]), "=f"(q[14]), "=f"(q[15]),
              "=f"(q[16]), "=f"(q[17]), "=f"(q[18]), "=f"(q[19]), "=f"(q[20]), "=f"(q[21]), "=f"(q[22]), "=f"(q[23]),
              "=f"(q[24]), "=f"(q[25]), "=f"(q[26]), "=f"(q[27]), "=f"(q[28]), "=f"(q[29]), "=f"(q[30]), "=f"(q[31])
            : "r"(taddr + 32 * hlf)
            : "memory");
    }
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    return done != 0;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// per-query values (log2-domain lse, or w = 2^-lse) of queries q0..q0+3 of one (b,h); queries past S read `fill` (-> p = 0)
__device__ __forceinline__ float4 load_lse4(const float* __restrict__ row0, int q0, int S, bool vec, float fill) {
    if (vec && q0 + 3 < S) return __ldg(reinterpret_cast<const float4*>(row0 + q0));
    float4 r;
    r.x = q0 + 0 < S ? __ldg(row0 + q0 + 0) : fill;
    r.y = q0 + 1 < S ? __ldg(row0 + q0 + 1) : fill;
    r.z = q0 + 2 < S ? __ldg(row0 + q0 + 2) : fill;
    r.w = q0 + 3 < S ? __ldg(row0 + q0 + 3) : fill;
    return r;
}
// One column of this thread's TMEM lane (the last query's logit for pass 2's last-row importance)
__device__ __forceinline__ void tc_ld1_issue(uint32_t taddr, float& v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=f"(v) : "r"(taddr) : "memory");
}
// Masked tiles (the causal diagonal, the ragged last block) are handled per batch of kExpBatch
// columns with warp-uniform tests: a batch that is invalid for all 32 rows of the warp is skipped
// outright (no exp2 at all: the diagonal tile costs about half a normal one), a batch valid for
// all rows runs the plain code, and only the one or two batches the boundary crosses pay a
// compare + select per element.  r = this thread's row and w0 = the warp's first row, both
// relative to the tile; columns are tile-relative too.
//
// pass 1, NC columns (= keys cr0..cr0+NC-1) of one query row: online log-sum-exp in the log2
// domain.  EDGE: key c is valid iff c <= jmax (this row) ; jlo / jhi = the smallest / largest jmax in the warp.
// FAST variants: the operand split has proven |logit| < kFastBound for every (query, key) pair of the
// launch (largest query norm x largest key norm; the logits arrive in the log2 domain because the
// softmax scale is folded into Q), so exp2 can neither overflow nor lose the row's largest terms and
// needs no running maximum: pass 1 is a plain sum of exp2(v) -- MUFU + FADD per element -- and pass 2
// multiplies by w = 2^-lse of the query instead of subtracting lse first -- MUFU + FFMA.  With half the
// issue slots per element one warp alone keeps the MUFU pipe busy, so the two epilogue warps of a
// sub-partition no longer depend on being inside their exp2 code at the same time.
constexpr float kFastBound = 100.f;
// The FAST loops leave most issue slots free while the MUFU pipe (16 exp2 / clk / SM) is the busiest
// unit, so kPoly of every 16 exponentials are evaluated on the FMA pipe instead: round-to-nearest range
// reduction by the 1.5 * 2^23 trick (valid for |x| < 2^22; here |x| < 100), a degree-5 polynomial for
// 2^r on [-0.5, 0.5] (relative error 1.8e-7 in fp32, the same as ex2.approx) and an integer add of
// the exponent.
#ifndef SPLITQ_K2_POLY
#define SPLITQ_K2_POLY 0
#endif
constexpr int kPoly = SPLITQ_K2_POLY;
__device__ __forceinline__ float ex2_poly(float x) {
    const float t = x + 12582912.f;
    const float r = x - (t - 12582912.f);
    float p = 0.001326472731307149f;
    p = fmaf(p, r, 0.009671512991189957f);
    p = fmaf(p, r, 0.05550733581185341f);
    p = fmaf(p, r, 0.24022242426872253f);
    p = fmaf(p, r, 0.6931470036506653f);
    p = fmaf(p, r, 1.0f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// element j of a batch of 16 goes to the FMA pipe if it closes one of kPoly equal strides
__device__ __forceinline__ constexpr bool on_fma_pipe(int j) { return ((j & 15) + 1) * kPoly / 16 != (j & 15) * kPoly / 16; }
__device__ __forceinline__ float ex2_mixed(float x, int j) { return on_fma_pipe(j) ? ex2_poly(x) : ex2_ordered(x); }
template <int NC, bool EDGE>
__device__ __forceinline__ void pass1_fast(const float (&v)[NC], int cr0, int jmax, int jlo, int jhi, float& l) {
    constexpr int EB = NC < kExpBatch ? NC : kExpBatch;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int i0 = 0; i0 < NC; i0 += EB) {
        if (EDGE && cr0 + i0 > jhi) continue;            // nothing valid for any row of the warp
        float e[EB];
#pragma unroll
        for (int j = 0; j < EB; ++j) e[j] = ex2_mixed(v[i0 + j], j);
        if (EDGE && cr0 + i0 + EB - 1 > jlo) {           // the boundary crosses this batch
#pragma unroll
            for (int j = 0; j < EB; ++j) e[j] = (cr0 + i0 + j > jmax) ? 0.f : e[j];
        }
#pragma unroll
        for (int j = 0; j < EB; j += 4) {
            s0 += e[j + 0];
            s1 += e[j + 1];
            s2 += e[j + 2];
            s3 += e[j + 3];
        }
    }
    l += (s0 + s1) + (s2 + s3);
}
// ws = shared address of the NC per-query weights w = 2^-lse (queries >= S carry w = 0)
template <int NC, bool DIAG>
__device__ __forceinline__ void pass2_fast(const float (&v)[NC], uint32_t ws, int cr0, int r, int w0, float& colsum) {
    constexpr int EB = NC < kExpBatch ? NC : kExpBatch;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int i0 = 0; i0 < NC; i0 += EB) {
        if (DIAG && cr0 + i0 + EB <= w0) continue;       // all queries of the batch precede every key of the warp
        float e[EB];
#pragma unroll
        for (int j = 0; j < EB; ++j) e[j] = ex2_mixed(v[i0 + j], j);
        if (DIAG && cr0 + i0 < w0 + 31) {                // the diagonal crosses this batch
#pragma unroll
            for (int j = 0; j < EB; ++j) e[j] = (cr0 + i0 + j < r) ? 0.f : e[j];
        }
#pragma unroll
        for (int j = 0; j < EB; j += 4) {
            const float4 w4 = lds128(ws + 4u * (i0 + j));
            s0 = fmaf(e[j + 0], w4.x, s0);
            s1 = fmaf(e[j + 1], w4.y, s1);
            s2 = fmaf(e[j + 2], w4.z, s2);
            s3 = fmaf(e[j + 3], w4.w, s3);
        }
    }
    colsum += (s0 + s1) + (s2 + s3);
}
// General variants (any magnitude): cs = 2^(block exponents) rescales a tile whose operand blocks
// were pre-scaled into fp16 range (1 otherwise).
template <int NC, bool EDGE>
__device__ __forceinline__ void pass1_body(float (&v)[NC], int cr0, int jmax, int jlo, int jhi, float cs, float& m, float& l) {
    constexpr int EB = NC < kExpBatch ? NC : kExpBatch;
    float cm0 = -INFINITY, cm1 = -INFINITY;
#pragma unroll
    for (int i0 = 0; i0 < NC; i0 += EB) {
        if (EDGE && cr0 + i0 > jhi) continue;            // nothing valid for any row of the warp
        if (EDGE && cr0 + i0 + EB - 1 > jlo) {           // the boundary crosses this batch
#pragma unroll
            for (int j = 0; j < EB; ++j) v[i0 + j] = (cr0 + i0 + j > jmax) ? -INFINITY : v[i0 + j];
        }
#pragma unroll
        for (int j = 0; j < EB; j += 2) {
            cm0 = fmaxf(cm0, v[i0 + j]);
            cm1 = fmaxf(cm1, v[i0 + j + 1]);
        }
    }
    const float cm = fmaxf(cm0, cm1) * cs;
    if (cm > -INFINITY) {  // at least one valid column for this row in this chunk
        const float mn = fmaxf(m, cm);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int i0 = 0; i0 < NC; i0 += EB) {
            if (EDGE && cr0 + i0 > jhi) continue;
            float e[EB];
#pragma unroll
            for (int j = 0; j < EB; ++j) e[j] = ex2_ordered(fmaf(v[i0 + j], cs, -mn));
#pragma unroll
            for (int j = 0; j < EB; j += 4) {
                s0 += e[j + 0];
                s1 += e[j + 1];
                s2 += e[j + 2];
                s3 += e[j + 3];
            }
        }
        l = l * ex2(m - mn) + ((s0 + s1) + (s2 + s3));
        m = mn;
    }
}
// pass 2, NC columns (= queries cr0..cr0+NC-1) of one key row: sum of exp2(s*cs - lse2[query]); ls =
// shared address of the NC lse values (queries >= S carry lse = +inf -> 0, so the ragged last
// block needs no mask).  DIAG: the causal diagonal tile, query c is valid iff c >= r.
template <int NC, bool DIAG>
__device__ __forceinline__ void pass2_body(const float (&v)[NC], uint32_t ls, int cr0, int r, int w0, float cs, float& colsum) {
    constexpr int EB = NC < kExpBatch ? NC : kExpBatch;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int i0 = 0; i0 < NC; i0 += EB) {
        if (DIAG && cr0 + i0 + EB <= w0) continue;       // all queries of the batch precede every key of the warp
        float e[EB];
#pragma unroll
        for (int j = 0; j < EB; j += 4) {
            const float4 ls4 = lds128(ls + 4u * (i0 + j));
            e[j + 0] = ex2_ordered(fmaf(v[i0 + j + 0], cs, -ls4.x));
            e[j + 1] = ex2_ordered(fmaf(v[i0 + j + 1], cs, -ls4.y));
            e[j + 2] = ex2_ordered(fmaf(v[i0 + j + 2], cs, -ls4.z));
            e[j + 3] = ex2_ordered(fmaf(v[i0 + j + 3], cs, -ls4.w));
        }
        if (DIAG && cr0 + i0 < w0 + 31) {                // the diagonal crosses this batch
#pragma unroll
            for (int j = 0; j < EB; ++j) e[j] = (cr0 + i0 + j < r) ? 0.f : e[j];
        }
#pragma unroll
        for (int j = 0; j < EB; j += 4) {
            s0 += e[j + 0];
            s1 += e[j + 1];
            s2 += e[j + 2];
            s3 += e[j + 3];
        }
    }
    colsum += (s0 + s1) + (s2 + s3);
}

// PASS 1 = row log-sum-exp, 2 = column sums; DBG compiles the timing-experiment switches of Params::dbg in
template <int PASS, bool DBG>
__global__ void __launch_bounds__(kThreads, 1)
fused_importance_kernel(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapK,
                        const Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint32_t base;  // opaque copy: otherwise ptxas re-derives the shared-window address (S2UR + ULEA) at every use
    asm volatile("mov.u32 %0, %1;" : "=r"(base) : "r"(smem_u32(smem_raw)));
    if (base & 1023u) __trap();  // swizzle-128B slabs need 1024-byte alignment; no static smem precedes
    uint8_t* gen = smem_raw;
    const uint32_t ring_s = base;
    const uint32_t bar0 = base + kBarOff;
    const uint32_t bar_afull = bar0, bar_aempty = bar0 + 16;          // fixed operand in TMEM, 2 buffers
    const uint32_t bar_full = bar0 + 32, bar_empty = bar_full + 8 * kStages;
    const uint32_t bar_tfull = bar_empty + 8 * kStages, bar_tempty = bar_tfull + 8 * kAcc;
    const uint32_t bar_turn = bar_tempty + 8 * kAcc;                   // two MMA warps: "passed the waits of my tile"
    static_assert(32 + 16 * kStages + 16 * kAcc + 16 <= 504, "barrier region overflow");
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gen + kBarOff + 504);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_afull + 8 * a, kASmem ? 1 : 4);  // one arrive per loader warp, or the producer's expect_tx
            mbar_init(bar_aempty + 8 * a, kMmaWarps);  // tcgen05.commit after each issuing warp's last MMA of the item
        }
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, 1);
        }
        for (int w = 0; w < 2; ++w) mbar_init(bar_turn + 8 * w, 1);
        for (int a = 0; a < kAcc; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, 4);  // one arrive per warp of the epilogue group that drained it
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // launch-wide facts from the operand split (every CTA reduces the small per-block arrays itself):
    // the largest squared row norms of Q' = Q * scaling * log2(e) and of K bound every logit of the
    // launch by Cauchy-Schwarz; any non-zero block exponent means some block was pre-scaled
    // (kFuseQ, pass 1: the Q blocks' facts do not exist yet -- the loaders produce them item by item -- so only
    // the K blocks are reduced here)
    constexpr bool kItemFacts = kFuseQ && PASS == 1;
    {
        float q2 = 0.f, k2 = 0.f;
        int emax = 0;
        const int i0 = kItemFacts ? p.nq_blk : 0;
        for (int i = i0 + threadIdx.x; i < p.nq_blk + p.nk_blk; i += kThreads) {
            const float b2 = __ldg(p.bound2 + i);
            if (i < p.nq_blk) q2 = fmaxf(q2, b2 == b2 ? b2 : INFINITY);
            else k2 = fmaxf(k2, b2 == b2 ? b2 : INFINITY);
        }
        if (kTileN == kTile) {
            for (int i = i0 + threadIdx.x; i < p.nq_blk + p.nk_blk; i += kThreads) emax = max(emax, abs(__ldg(p.eblk + i)));
        } else {
            for (int i = threadIdx.x; i < p.n_heads; i += kThreads) emax = max(emax, abs(__ldg(p.ehead + i)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            q2 = fmaxf(q2, __shfl_xor_sync(0xffffffffu, q2, o));
            k2 = fmaxf(k2, __shfl_xor_sync(0xffffffffu, k2, o));
            emax = max(emax, __shfl_xor_sync(0xffffffffu, emax, o));
        }
        if (lane == 0) {
            float* f = reinterpret_cast<float*>(gen + kFlagOff) + 3 * warp;
            f[0] = q2; f[1] = k2; f[2] = __int_as_float(emax);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    bool fast, scaled;
    float k2_all;   // largest squared key row norm of the launch
    {
        float q2 = 0.f, k2 = 0.f;
        int emax = 0;
        const float* f = reinterpret_cast<const float*>(gen + kFlagOff);
        for (int w = 0; w < kThreads / 32; ++w) {
            q2 = fmaxf(q2, f[3 * w]); k2 = fmaxf(k2, f[3 * w + 1]); emax = max(emax, __float_as_int(f[3 * w + 2]));
        }
        scaled = emax != 0;
        // the norm caps keep the absolute fp16 floor of one operand (2^-25 per element) times the other
        // operand's norm far below the 1e-6 a logit may be off by
        fast = !scaled && q2 < 16384.f && k2 < 16384.f && q2 * k2 < kFastBound * kFastBound && !(DBG && (p.dbg & 32));
        k2_all = k2;
    }

    const CUtensorMap* str_map = PASS == 1 ? &mapK : &mapQ;  // streamed operand, h0 and h1 parts in dim 3
    const CUtensorMap* fix_map = PASS == 1 ? &mapQ : &mapK;  // fixed operand (kASmem: by TMA as well)
    const int group = p.H / p.KVH;
    const int wg = warp >> 2;  // warpgroup: 0 control, 1-2 epilogue, 3 fixed-operand loaders

    if (wg == 0) {
        reg_dec<kCtlRegs>();
        if (warp == 0) {
            // ===== TMA producer: one stage = one streamed 128-row tile (h0 + h1 parts); per tile ONE
            // barrier wait, one expect_tx and ONE 4-D TMA issue.  The per-transaction round trip
            // (try_wait + expect_tx + issue, ~300 cycles) was the limiter of a finer 16 KB-slab ring,
            // not bytes =====
            {
                uint32_t stage = 0, sphase = 0;
                const uint32_t tile_tx = p.x3 ? kOperandBytes : kPartBytes;
                SegIter<PASS> si;
                si.init(p, 0);
                Seg it;
                // kASmem: the same warp also brings in the fixed operand of each segment (one more tile of
                // the other tensor), up to two segments ahead of the streaming (two buffers)
                SegIter<PASS> sa;
                Seg ia;
                int ka = 0;                     // segments whose fixed operand has been requested
                bool a_more = false;
                if (kASmem) {
                    sa.init(p, 0);
                    a_more = sa.next(ia);
                }
                auto request_fixed = [&](bool block) {
                    const uint32_t ab = ka & 1, abphase = (ka >> 1) & 1;
                    if (block) mbar_wait_t<DBG>(bar_aempty + 8 * ab, abphase ^ 1);
                    else if (!mbar_test(bar_aempty + 8 * ab, abphase ^ 1)) return;
                    const int fb_ = ia.bh / p.H, fh_ = ia.bh % p.H;
                    const int fix_idx = PASS == 1 ? ia.bh : fb_ * p.KVH + fh_ / group;
                    if (elect_one()) {
                        mbar_expect_tx(bar_afull + 8 * ab, tile_tx);
                        tma_load_4d(ring_s + kASmemOff + ab * (2 * kFixPartBytes), fix_map, bar_afull + 8 * ab, 0, ia.blk * kTile,
                                    fix_idx, 0);
                    }
                    __syncwarp();
                    ++ka;
                    a_more = sa.next(ia);
                };
                for (int k = 0; si.next(it); ++k) {
                    const int b = it.bh / p.H, h = it.bh % p.H;
                    const int str_idx = PASS == 1 ? b * p.KVH + h / group : it.bh;
                    if (kASmem && ka == k) request_fixed(true);   // this segment's own operand (first segment only)
                    for (int t = it.ta; t < it.tb; ++t) {
                        if (kASmem && a_more && ka == k + 1) request_fixed(false);  // the next segment's, once its buffer is free
                        const uint32_t fb = bar_full + 8 * stage;
                        const uint32_t dst = ring_s + stage * kOperandBytes;
                        mbar_wait_t<DBG>(bar_empty + 8 * stage, sphase ^ 1);
                        if (elect_one()) {
                            mbar_expect_tx(fb, tile_tx);
                            tma_load_4d(dst, str_map, fb, 0, t * kTileN, str_idx, 0);
                        }
                        __syncwarp();
                        if (++stage == kStages) { stage = 0; sphase ^= 1; }
                    }
                    if (kASmem && a_more && ka == k + 1) request_fixed(true);  // must be on its way before its tiles are
                }
            }
        } else if (warp <= kMmaWarps) {
            // ===== MMA issuer(s): A (fixed operand) from TMEM, B (streamed slab) from shared memory.
            // With two issuing warps each takes every other tile: `tcgen05.mma` blocks while the
            // pipe's queue is full, so one warp's barrier waits run under the other's MMAs =====
            {
                const uint32_t mi = (uint32_t)warp - 1u;
                uint32_t n_tile = 0;
                uint32_t stage = 0, sphase = 0, acc = 0, aphase = 0;
                SegIter<PASS> si;
                si.init(p, 0);
                Seg it;
                for (int k = 0; si.next(it); ++k) {
                    const uint32_t ab = k & 1, abphase = (k >> 1) & 1;
                    const uint32_t a_h0 = tmem_base + kTmemA + ab * kTmemABuf, a_h1 = a_h0 + 32u;           // TMEM columns
                    const uint32_t as_h0 = ring_s + kASmemOff + ab * (2 * kFixPartBytes), as_h1 = as_h0 + kFixPartBytes;  // kASmem
                    mbar_wait_t<DBG>(bar_afull + 8 * ab, abphase);
                    tc_fence_after();
                    for (int t = it.ta; t < it.tb; ++t, ++n_tile) {
                        if (kMmaWarps > 1 && (n_tile & 1u) != mi) {  // the other warp's tile
                            if (++stage == kStages) { stage = 0; sphase ^= 1; }
                            if (++acc == kAcc) { acc = 0; aphase ^= 1; }
                            continue;
                        }
                        if (kMmaWarps > 1 && n_tile > 0) {
                            // The other warp must be past the waits of tile n - 1 first: parity waits are only
                            // sound one phase apart, so the two warps may not drift by more than a tile.
                            const uint32_t k_other = mi ? (n_tile >> 1) : (n_tile >> 1) - 1u;
                            mbar_wait_t<DBG>(bar_turn + 8 * (mi ^ 1u), k_other & 1u);
                        }
                        mbar_wait_t<DBG>(bar_tempty + 8 * acc, aphase ^ 1);
                        mbar_wait_t<DBG>(bar_full + 8 * stage, sphase);
                        if (kMmaWarps > 1) {
                            if (elect_one()) mbar_arrive(bar_turn + 8 * mi);
                            __syncwarp();
                        }
                        tc_fence_after();
                        const uint32_t d = tmem_base + acc * kTileN;
                        const uint32_t st0 = ring_s + stage * kOperandBytes;
                        if (elect_one()) {
                            if (!(DBG && (p.dbg & 2))) {
                                // F = fixed operand (A, TMEM), T = streamed operand (B, shared memory); K = 16 per
                                // instruction, 8 TMEM columns / 32 smem bytes per step
                                const uint32_t s_h0 = st0, s_h1 = st0 + kPartBytes;
#pragma unroll
                                for (int kk = 0; kk < 4; ++kk)   // h0(F) * h0(T)
                                    if (kASmem) tc_mma_f16_ss(d, umma_desc(as_h0 + kk * 32u), umma_desc(s_h0 + kk * 32u), kIdescF16, (uint32_t)kk);
                                    else tc_mma_f16_ts(d, a_h0 + kk * 8u, umma_desc(s_h0 + kk * 32u), kIdescF16, (uint32_t)kk);
                                if (p.x3) {
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk)  // h1(F) * h0(T)
                                        if (kASmem) tc_mma_f16_ss(d, umma_desc(as_h1 + kk * 32u), umma_desc(s_h0 + kk * 32u), kIdescF16, 1u);
                                        else tc_mma_f16_ts(d, a_h1 + kk * 8u, umma_desc(s_h0 + kk * 32u), kIdescF16, 1u);
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk)  // h0(F) * h1(T)
                                        if (kASmem) tc_mma_f16_ss(d, umma_desc(as_h0 + kk * 32u), umma_desc(s_h1 + kk * 32u), kIdescF16, 1u);
                                        else tc_mma_f16_ts(d, a_h0 + kk * 8u, umma_desc(s_h1 + kk * 32u), kIdescF16, 1u);
                                }
                            }
                            tc_commit(bar_empty + 8 * stage);   // stage reusable once these MMAs retire
                            tc_commit(bar_tfull + 8 * acc);     // accumulator ready for the epilogue
                        }
                        __syncwarp();
                        if (++stage == kStages) { stage = 0; sphase ^= 1; }
                        if (++acc == kAcc) { acc = 0; aphase ^= 1; }
                    }
                    if (elect_one()) tc_commit(bar_aempty + 8 * ab);      // fixed-operand buffer reusable
                    __syncwarp();
                }
            }
        }
    } else if (wg == kEpiGroups + 1) {
        // ===== fixed-operand loaders: one thread per row; fp16 h0 / h1 parts -> TMEM (A operand) =====
        reg_dec<kLdrRegs>();
        const int q4 = warp & 3;
        const int row = q4 * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q4 * 32) << 16;
        const uint4* f16 = PASS == 1 ? p.q16 : p.k16;
        const size_t fpart = PASS == 1 ? p.q_part : p.k_part;
        SegIter<PASS> si;
        si.init(p, 0);
        Seg it;
        // neutral elements for the slots of this run's rows that no epilogue group will write (kLoaderPresets):
        // the epilogue fills (piece, g) for every group g that owns a tile of the run -- tile n of the CTA's
        // range belongs to group n % groups -- and, if the run is not a whole item, the CTA holding the other
        // piece looks after that piece's slots
        int n_before = 0;   // tiles of this CTA's range before the current run
        auto preset_unwritten = [&](const Seg& r) {
            const int len = r.tb - r.ta;
            const int piece = r.ta != si.t0(r.blk) ? 1 : 0;
            const bool whole = piece == 0 && r.tb == si.t1(r.blk);
            const int g0 = n_before % kEpiGroups;
            n_before += len;
            const int gr = r.blk * kTile + row;
            if (!kLoaderPresets || gr >= p.S) return;
            const size_t rows = (size_t)p.B * p.H * p.S, o = (size_t)r.bh * p.S + gr;
#pragma unroll
            for (int pc = 0; pc < 2; ++pc) {
#pragma unroll
                for (int gq = 0; gq < kEpiGroups; ++gq) {
                    const int rel_g = gq - g0 + (gq < g0 ? kEpiGroups : 0);   // position of group gq in this run's rotation
                    const bool filled = pc == piece && rel_g < len;
                    const bool ours = pc == piece || whole;
                    if (ours && !filled) {
                        if (PASS == 1) p.ml[(pc * kEpiGroups + gq) * rows + o] = make_float2(-INFINITY, 0.f);
                        else p.part_col[(pc * kEpiGroups + gq) * rows + o] = 0.f;
                    }
                }
            }
        };
        if (kItemFacts) {
            // pass 1 with the fused operand split: this thread's row comes in as fp32, half a row (32 floats) at
            // a time; scaled, split into h0 = fp16(x), h1 = fp16(x - h0) exactly as split_f16_kernel does, stored
            // to tensor memory (16 columns per half row and part) and written out for pass 2's TMA.  The block's
            // largest squared row norm and largest |x| meet across the four warps in shared memory; a block
            // whose values leave the fp16 range (exponent e != 0, rare) is converted a second time with 2^-e.
            float2* ring = reinterpret_cast<float2*>(gen + kFactOff);
            float2* part = ring + kFactRing;
            float4 v[8];
            for (int k = 0; si.next(it); ++k) {
                const uint32_t ab = k & 1, abphase = (k >> 1) & 1;
                const int g = it.blk * kTile + row;
                const bool live = g < p.S && !(DBG && (p.dbg & 16));
                preset_unwritten(it);
                const float* src = p.qsrc + (it.bh / p.H) * p.qs0 + (it.bh % p.H) * p.qs1 + (long long)(live ? g : 0) * p.qs2;
                uint4* dst = p.q16_w + ((size_t)it.bh * p.S + (live ? g : 0)) * (kHd * 2 / 16);
                const uint32_t a0 = tmem_base + kTmemA + ab * kTmemABuf + lane_addr;
                auto load_half = [&](int half) {
                    if (live && p.qvec == 2) {
#pragma unroll
                        for (int i = 0; i < 8; i += 2) ldg256(src + half * 32 + i * 4, v[i], v[i + 1]);
                        return;
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float* s4 = src + half * 32 + i * 4;
                        if (!live) v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                        else if (p.qvec) v[i] = __ldg(reinterpret_cast<const float4*>(s4));
                        else v[i] = make_float4(__ldg(s4), __ldg(s4 + 1), __ldg(s4 + 2), __ldg(s4 + 3));
                    }
                };
                float ss = 0.f;
                uint32_t amax = 0u;
                auto convert_half = [&](int half, float sc, bool stats) {
                    uint4 h0[4], h1[4];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float4 x = v[i];
                        x.x *= p.cs; x.y *= p.cs; x.z *= p.cs; x.w *= p.cs;
                        if (stats) {
                            amax = max(amax, __float_as_uint(x.x) & 0x7fffffffu);
                            amax = max(amax, __float_as_uint(x.y) & 0x7fffffffu);
                            amax = max(amax, __float_as_uint(x.z) & 0x7fffffffu);
                            amax = max(amax, __float_as_uint(x.w) & 0x7fffffffu);
                            ss += fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, x.w * x.w)));
                        }
                        const float x0 = x.x * sc, x1 = x.y * sc, x2 = x.z * sc, x3 = x.w * sc;
                        const __half2 a01 = __floats2half2_rn(x0, x1), a23 = __floats2half2_rn(x2, x3);
                        const float2 f01 = __half22float2(a01), f23 = __half22float2(a23);
                        const __half2 l01 = __floats2half2_rn(x0 - f01.x, x1 - f01.y);
                        const __half2 l23 = __floats2half2_rn(x2 - f23.x, x3 - f23.y);
                        const uint32_t a01u = *reinterpret_cast<const uint32_t*>(&a01), a23u = *reinterpret_cast<const uint32_t*>(&a23);
                        const uint32_t l01u = *reinterpret_cast<const uint32_t*>(&l01), l23u = *reinterpret_cast<const uint32_t*>(&l23);
                        if (i & 1) { h0[i >> 1].z = a01u; h0[i >> 1].w = a23u; h1[i >> 1].z = l01u; h1[i >> 1].w = l23u; }
                        else { h0[i >> 1].x = a01u; h0[i >> 1].y = a23u; h1[i >> 1].x = l01u; h1[i >> 1].y = l23u; }
                    }
                    tc_st16(a0 + 16u * half, h0);
                    if (p.x3) tc_st16(a0 + 32u + 16u * half, h1);
                    if (live) {
#pragma unroll
                        for (int i = 0; i < 4; i += 2) {
                            stg256(dst + 4 * half + i, h0[i], h0[i + 1]);
                            if (p.x3) stg256(dst + p.q_part + 4 * half + i, h1[i], h1[i + 1]);
                        }
                    }
                };
                load_half(0);                                   // on its way before the buffer is free
                {   // the next item's row: into the L2 now, so that both of its half-row loads are short ones
                    SegIter<PASS> peek = si;
                    Seg nxt;
                    if (peek.next(nxt)) {
                        const int gn = nxt.blk * kTile + row;
                        if (gn < p.S && !(DBG && (p.dbg & 16))) {
                            const float* pn = p.qsrc + (nxt.bh / p.H) * p.qs0 + (nxt.bh % p.H) * p.qs1 + (long long)gn * p.qs2;
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(pn));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(pn + 32));
                        }
                    }
                }
                mbar_wait_t<false>(bar_aempty + 8 * ab, abphase ^ 1);
                tc_fence_after();
                convert_half(0, 1.0f, true);
                load_half(1);
                convert_half(1, 1.0f, true);
                // block facts: four warp partials, combined by every thread after a 128-thread barrier
                float rmax = ss == ss ? ss : INFINITY;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) rmax = fmaxf(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
                amax = __reduce_max_sync(0xffffffffu, amax);
                float2* mine = part + 4 * ab;
                if (lane == 0) mine[q4] = make_float2(rmax, __uint_as_float(amax));
                asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const float2 f = mine[w];
                    rmax = fmaxf(rmax, f.x);
                    amax = max(amax, __float_as_uint(f.y));
                }
                const int ex = (int)(amax >> 23) - 127;
                const int e = (ex >= 15 || (ex < -10 && amax != 0u)) ? ex - 14 : 0;
                if (e != 0) {
                    const float sc = __int_as_float((127 - min(max(e, -126), 126)) << 23);
                    load_half(0);
                    convert_half(0, sc, false);
                    load_half(1);
                    convert_half(1, sc, false);
                }
                if (q4 == 0 && lane == 0) {
                    ring[k & (kFactRing - 1)] = make_float2(rmax, __int_as_float(e));
                    p.bound2_w[(size_t)it.bh * p.nb + it.blk] = rmax;   // pass 2 reduces these over the launch
                    p.eblk_w[(size_t)it.bh * p.nb + it.blk] = e;
                }
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_afull + 8 * ab);
            }
        }
        for (int k = 0; !kItemFacts && !kASmem && si.next(it); ++k) {   // kASmem: the producer's TMA does this job
            const uint32_t ab = k & 1, abphase = (k >> 1) & 1;
            preset_unwritten(it);
            const int b = it.bh / p.H, h = it.bh % p.H;
            const int fix_idx = PASS == 1 ? it.bh : b * p.KVH + h / group;
            const int g = it.blk * kTile + row;
            const uint4* src = f16 + ((size_t)fix_idx * p.S + (g < p.S ? g : 0)) * (kHd * 2 / 16);
            uint4 q[8];  // this row: 64 halfs = 128 B = 32 TMEM columns per part; h0 is fetched ahead of the wait
#pragma unroll
            for (int i = 0; i < 8; ++i) q[i] = (g < p.S && !(DBG && (p.dbg & 16))) ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
            mbar_wait_t<false>(bar_aempty + 8 * ab, abphase ^ 1);
            tc_fence_after();
            const uint32_t a0 = tmem_base + kTmemA + ab * kTmemABuf + lane_addr;
            tc_st32(a0, q);
            if (p.x3) {
#pragma unroll
                for (int i = 0; i < 8; ++i) q[i] = (g < p.S && !(DBG && (p.dbg & 16))) ? __ldg(src + fpart + i) : make_uint4(0u, 0u, 0u, 0u);
                tc_st32(a0 + 32u, q);
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_afull + 8 * ab);
        }
    } else {
        reg_inc<kEpiRegs>();
        // ===== epilogue: two groups of 4 warps; group g takes the tiles of parity g of the CTA's
        // range, tile n lives in TMEM accumulator n % 3; a thread owns one TMEM lane (tile row).
        // Every warp runs on its own: no block or group barrier anywhere, only the accumulator
        // mbarriers.  The loop is software-pipelined across tiles AND items: while the exp2 loop
        // runs on one 64-column half, the tcgen05.ld of the next half (or of the next tile's first
        // half) is in flight, and the next tile's per-query values (pass 2; each warp stages all 128
        // of them for itself) are on their way from L2.  When the sequence moves to another item the
        // per-row partial results are stored to this (piece, group) slot.  The loop exists twice:
        // FAST (bounded logits, see pass1_fast) and the general form; the choice is launch-wide.
        // (kFuseQ, pass 1: a third form, MODE 2, takes the choice item by item from the loaders' facts.)
        auto epilogue = [&](auto mode_tag) {
        constexpr int MODE = decltype(mode_tag)::value;   // 0 general, 1 FAST, 2 per item (pass 1 only)
        constexpr bool FAST = MODE == 1;                  // launch-wide FAST: all of pass 2's uses
        bool item_fast = false;                           // MODE 2: the current item's choice ...
        int item_e = 0;                                   // ... and its query block's exponent
        auto fast_now = [&]() { return MODE == 2 ? item_fast : FAST; };
        const int q4 = warp & 3;                    // TMEM lane quarter this warp may read
        const int grp = wg - 1;                     // this group takes tiles grp, grp + kEpiGroups, ...
        const int row = q4 * 32 + lane;
        const uint32_t tmem_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const uint32_t lse_w = base + kLseOff + (uint32_t)(warp - 4) * (2u * kTileN * 4u);  // two kTileN-float buffers per warp
        const int S = p.S, dbg = DBG ? p.dbg : 0;
        const bool causal = p.causal != 0;
        // pass 2 stages one value per query of the tile: w = 2^-lse (FAST; 0 past S) or lse (+inf past S)
        const float* __restrict__ qval = FAST ? p.w2 : p.lse2;
        const float qfill = FAST ? 0.f : INFINITY;
        const bool lse_vec = (S & 3) == 0;
        const int last_tile = (S - 1) / kTileN, last_col = (S - 1) % kTileN;
        // one tile's per-query values: 32 lanes x float4 cover 128, lanes 0-15 fetch a second float4 for 192
        struct QV { float4 a, b; };
        auto load_qv = [&](int bh_, int t_) {
            QV q;
            const float* rowp = qval + (size_t)bh_ * S;
            q.a = load_lse4(rowp, t_ * kTileN + lane * 4, S, lse_vec, qfill);
            q.b = make_float4(qfill, qfill, qfill, qfill);
            if (kTileN > 128 && lane < (kTileN - 128) / 4) q.b = load_lse4(rowp, t_ * kTileN + 128 + lane * 4, S, lse_vec, qfill);
            return q;
        };
        auto stage_qv = [&](uint32_t buf, const QV& q) {
            sts128(lse_w + buf * (kTileN * 4u) + 16u * lane, q.a);
            if (kTileN > 128 && lane < (kTileN - 128) / 4) sts128(lse_w + buf * (kTileN * 4u) + 512u + 16u * lane, q.b);
        };
        const size_t rows = (size_t)p.B * p.H * S;
        SegIter<PASS> nx;                           // runs one tile ahead of the tile being processed
        nx.init(p, 0);
        const int first_bh = nx.bh, first_blk = nx.blk;
        const int first_part = nx.t != nx.t0(nx.blk);  // the range starts inside an item
        nx.advance(grp);
        int st_bh = -1, st_blk = 0;                 // item whose partial results are in registers
        float m = FAST ? 0.f : -INFINITY, l = 0.f;  // pass 1 (MODE 2: set when the first item is entered)
        float colsum = 0.f, lastp = 0.f;            // pass 2
        bool saw_last = false;

        auto flush = [&]() {
            const int g = st_blk * kTile + row;
            if (st_bh < 0 || g >= S) return;
            const int slot = ((st_bh == first_bh && st_blk == first_blk) ? first_part * kEpiGroups : 0) + grp;
            const size_t o = (size_t)st_bh * S + g;
            if (PASS == 1) {
                p.ml[slot * rows + o] = make_float2(m, l);
            } else {
                p.part_col[slot * rows + o] = colsum;
                if (saw_last) p.part_last[o] = lastp;
            }
        };
        // general form only: 2^(exponents) of the query head and the key head of this item, 1 unless a head
        // was pre-scaled into fp16 range (ehead: query heads first, then key heads)
        auto tile_scale = [&](int bh, int blk, int t) -> float {
            if (fast_now() || !(scaled || (MODE == 2 && item_e != 0))) return 1.0f;
            const int kv = (bh / p.H) * p.KVH + (bh % p.H) / group;
            int e;
            if (MODE == 2) {        // the query block's exponent came with the item, the key blocks' are final
                e = item_e + (scaled ? __ldg(p.eblk + (size_t)p.B * p.H * p.nb + (size_t)kv * p.nb + t) : 0);
            } else if (kTileN == kTile) {  // the streamed tile t is block t of the split
                const int qb = PASS == 1 ? blk : t, kb = PASS == 1 ? t : blk;
                e = __ldg(p.eblk + (size_t)bh * p.nb + qb) + __ldg(p.eblk + (size_t)p.B * p.H * p.nb + (size_t)kv * p.nb + kb);
            } else {
                e = __ldg(p.ehead + bh) + __ldg(p.ehead + p.B * p.H + kv);
            }
            return __int_as_float((127 + min(max(e, -126), 127)) << 23);
        };

        float va[kChunk], vb[kChunk];               // chunk c of a tile sits in va (c even) or vb (c odd)
        if (dbg & 4) {
#pragma unroll
            for (int i = 0; i < kChunk; ++i) va[i] = vb[i] = (float)(row + i) * 0.01f;
        }
        uint32_t lb = 0;                            // per-query value buffer of the current tile
        uint32_t acc_i = (uint32_t)grp % kAcc, acc_ph = 0;  // accumulator (tile number % kAcc) and barrier phase of the current tile
        bool have = nx.valid();
#if SPLITQ_K2_GROUP_DELAY > 0
        // phase offset between the epilogue groups: two warps that share a sub-partition's MUFU pipe keep
        // whatever offset they start with (see benchmarks/experiments/README.md)
        if (grp > 0) {
            const long long t0 = clock64();
            while (clock64() - t0 < (long long)SPLITQ_K2_GROUP_DELAY * grp) {}
        }
#endif
        if (have) {
            if (PASS == 2) {
                stage_qv(0u, load_qv(nx.bh, nx.t));
                __syncwarp();
            }
            mbar_wait_t<DBG>(bar_tfull + 8 * acc_i, acc_ph);
            tc_fence_after();
            if (!(dbg & 4)) tc_ld_issue<kChunk>(tmem_lane + acc_i * kTileN, va);
            tc_ld_wait();
        }
        while (have) {
            // ---- the tile being processed: (bh, blk, t), where nx stands before it moves on
            const int bh = nx.bh, blk = nx.blk, t = nx.t, item = nx.item;
            const uint32_t a = acc_i;
            const uint32_t cur_tmem = tmem_lane + a * kTileN;
            // ---- look one tile ahead: is its accumulator ready (asked early, the answer takes ~100
            // cycles), and start fetching its per-query values
            uint32_t na = a + (uint32_t)(kEpiGroups % kAcc), nph = acc_ph;
            if (na >= (uint32_t)kAcc) { na -= (uint32_t)kAcc; nph ^= 1u; }
            if (kEpiGroups >= kAcc) nph ^= 1u;
            nx.advance(kEpiGroups);
            const bool has_next = nx.valid();
            bool ready = has_next && mbar_test(bar_tfull + 8 * na, nph);
            QV lse_nx = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
            if (PASS == 2 && has_next) lse_nx = load_qv(nx.bh, nx.t);
            // ---- entering a new item: store what the registers hold
            if (bh != st_bh || blk != st_blk) {
                flush();
                if (MODE == 2) {
                    // the loaders' facts of this item (written before the block went to the MMAs whose result
                    // this warp has just waited for): FAST iff nothing was pre-scaled and Cauchy-Schwarz bounds
                    // every logit of the item, with the same norm caps as the launch-wide test
                    const float2 f = reinterpret_cast<const float2*>(gen + kFactOff)[item & (kFactRing - 1)];
                    item_e = __float_as_int(f.y);
                    item_fast = !scaled && item_e == 0 && f.x < 16384.f && k2_all < 16384.f &&
                                f.x * k2_all < kFastBound * kFastBound && !(DBG && (p.dbg & 32));
                }
                m = fast_now() ? 0.f : -INFINITY; l = 0.f; colsum = 0.f; lastp = 0.f; saw_last = false;
                st_bh = bh;
                st_blk = blk;
            }
            const float cs = tile_scale(bh, blk, t);
            const bool has_last = PASS == 2 && last_tile == t;
            saw_last = saw_last || has_last;
            const uint32_t lse_t = lse_w + lb * (kTileN * 4u);
            // Tile-relative position of the causal boundary.  The fixed block's row r (absolute index
            // blk * 128 + r) meets the streamed tile's column c (absolute t * kTileN + c):
            //   pass 1 (rows = queries, columns = keys): key c is valid iff c <= rel + r, and c < S - t * kTileN;
            //   pass 2 (rows = keys, columns = queries): query c is valid iff c >= rel + r (queries past S carry
            //   a per-query value that yields 0).
            const int rel = causal ? blk * kTile - t * kTileN : (PASS == 1 ? kTileN : -kTile);
            const int s_hi = S - 1 - t * kTileN;     // last key of the sequence, tile-relative
            const int jmax = min(min(rel + row, kTileN - 1), s_hi);   // may be negative: no valid key for this row
            const int jlo = min(min(rel + q4 * 32, kTileN - 1), s_hi), jhi = min(min(rel + q4 * 32 + 31, kTileN - 1), s_hi);
            const bool edge1 = jlo < kTileN - 1;      // some (row, key) of this warp's part of the tile is masked
            const bool diag = rel + q4 * 32 + 31 > 0; // pass 2: some (key, query) of this warp's part is masked
            float lastv = 0.f;
            if (has_last && !(dbg & 4)) tc_ld1_issue(cur_tmem + (uint32_t)last_col, lastv);  // lands with the next wait
            // one tile's chunks, once per epilogue body: the choice between the FAST and the general body is made
            // here, outside the chunk loop, so that each copy is as tight as a loop that only knows one of them
            auto run_tile = [&](auto fast_c) {
                constexpr bool FASTC = decltype(fast_c)::value;
    #pragma unroll
                for (int c = 0; c < kChunks; ++c) {
                    float (&cur)[kChunk] = (c & 1) ? vb : va;
                    float (&oth)[kChunk] = (c & 1) ? va : vb;
                    if (c + 1 < kChunks) {                // next chunk of this tile in flight ...
                        if (!(dbg & 4)) tc_ld_issue<kChunk>(cur_tmem + (c + 1) * kChunk, oth);
                    } else if (has_next) {                // ... or the next tile's first chunk, if its MMAs are done
                        if (PASS == 2) stage_qv(lb ^ 1u, lse_nx);
                        if (!ready) ready = mbar_test(bar_tfull + 8 * na, nph);
                        ready = __all_sync(0xffffffffu, ready);
                        if (ready) {
                            tc_fence_after();
                            if (!(dbg & 4)) tc_ld_issue<kChunk>(tmem_lane + na * kTileN, oth);
                        }
                    }
                    if (dbg & 1) { colsum += cur[0] + cur[kChunk - 1]; l += cur[1]; }
                    else if (PASS == 1) {
                        if (FASTC) {
                            if (edge1) pass1_fast<kChunk, true>(cur, c * kChunk, jmax, jlo, jhi, l);
                            else pass1_fast<kChunk, false>(cur, c * kChunk, jmax, jlo, jhi, l);
                        } else {
                            if (edge1) pass1_body<kChunk, true>(cur, c * kChunk, jmax, jlo, jhi, cs, m, l);
                            else pass1_body<kChunk, false>(cur, c * kChunk, jmax, jlo, jhi, cs, m, l);
                        }
                    } else {
                        if (FASTC) {
                            if (diag) pass2_fast<kChunk, true>(cur, lse_t + c * (kChunk * 4u), c * kChunk, rel + row, rel + q4 * 32, colsum);
                            else pass2_fast<kChunk, false>(cur, lse_t + c * (kChunk * 4u), c * kChunk, rel + row, rel + q4 * 32, colsum);
                        } else {
                            if (diag) pass2_body<kChunk, true>(cur, lse_t + c * (kChunk * 4u), c * kChunk, rel + row, rel + q4 * 32, cs, colsum);
                            else pass2_body<kChunk, false>(cur, lse_t + c * (kChunk * 4u), c * kChunk, rel + row, rel + q4 * 32, cs, colsum);
                        }
                    }
                    if (c + 1 < kChunks) {
                        tc_ld_wait();
                        if (c == 0 && has_last) {         // p[S-1][key]: the last query is never masked
                            const float qv = lds32(lse_t + 4u * (uint32_t)last_col);
                            lastp = FASTC ? ex2(lastv) * qv : ex2(fmaf(lastv, cs, -qv));
                        }
                        if (c + 2 == kChunks) {           // the last chunk is in registers: accumulator drained
                            tc_fence_before();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(bar_tempty + 8 * a);
                        }
                    } else if (has_next) {
                        if (!ready) {
                            mbar_wait_t<DBG>(bar_tfull + 8 * na, nph);
                            tc_fence_after();
                            if (!(dbg & 4)) tc_ld_issue<kChunk>(tmem_lane + na * kTileN, oth);
                        }
                        tc_ld_wait();                     // next tile's first chunk landed
                        __syncwarp();                     // the per-query values staged by the other lanes are visible
                        lb ^= 1u;
                        acc_i = na;
                        acc_ph = nph;
                    }
                }
            };
            if (MODE == 2) {
                if (item_fast) run_tile(std::true_type{});
                else run_tile(std::false_type{});
            } else {
                run_tile(std::integral_constant<bool, FAST>{});
            }
            have = has_next;
        }
        flush();
        };
        if constexpr (kItemFacts) {
#ifdef SPLITQ_K2_FORCE_FAST1   // timing experiment: the launch-wide FAST loop on top of the fused split (valid for tame inputs only)
            epilogue(std::integral_constant<int, 1>{});
#else
            epilogue(std::integral_constant<int, 2>{});
#endif
        } else {
            if (fast) epilogue(std::integral_constant<int, 1>{});
            else epilogue(std::integral_constant<int, 0>{});
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// ---- operand preparation: fp32 -> fp16 h0 = fp16(x) and h1 = fp16(x - h0), gathered from arbitrary
// (b,h,s) strides into dense (n, S, 64) fp16 arrays (one 128-byte row per token).  One CTA = one
// 128-row block of one head of Q (blockIdx.y < BH) or of K.  Q is multiplied by scaling * log2(e)
// first, so the tensor-core products are log2-domain logits.  With kFuseQ (the default) the Q blocks are
// converted by pass 1's loader warps instead -- same arithmetic, same facts, written to the same arrays --
// and this launch covers the K blocks only: 34 MB instead of 300 MB at the bench shape (52 -> 9 us; the
// result slots are looked after by the passes' loader warps, see kLoaderPresets).  Per block the kernel (or
// the loader) also records
//   bound2: the largest squared row norm (of the values as they enter the products' true magnitude),
//           from which the passes bound every logit of the launch (FAST epilogue), and
//   eblk:   e such that the block was multiplied by 2^-e before the fp16 conversion: operands beyond
//           the fp16 range (or far below it) lose nothing but a power of two, which the epilogue
//           multiplies back per tile (e = 0 for any ordinary activation).  Inside a block that is not
//           pre-scaled, elements below 2^-3 carry an absolute error of up to 2^-25 instead of a
//           relative one of 2^-23 (the low fp16 part is subnormal there).
struct SplitArgs {
    const float *q, *k;
    long long qs[3], ks[3];   // element strides of (b, head, s)
    int H, KVH, S, nb, BH;
    int vecq, veck;
    float cs;                 // scaling * log2(e), folded into Q
    uint2 *q0, *q1, *k0, *k1; // h0 / h1 parts
    float2* ml;               // result slots, preset here only if the loaders do not (kLoaderPresets): (kSlots, rows) (max, sum) = (-inf, 0)
    float* part_col;          // (kSlots, rows) = 0
    size_t rows;
    float* bound2;            // (BH + BKV) * nb: largest squared row norm per block
    int* eblk;                // (BH + BKV) * nb: exponent each block would like for itself
    int* ehead;               // (BH + BKV): exponent every block of the head ends up with
    int fuse_q;               // kFuseQ: pass 1 converts the Q blocks itself
    int y0;                   // first (b,head) row of the launch: B*H when only the K blocks are launched
};
// exponent all blocks of one head share: a tile of the passes spans more than one 128-row block of the
// streamed operand, so the power of two must be uniform along a head.  If any block is beyond the fp16
// range, the largest request wins; if EVERY block is far below it, the least negative one; otherwise 0.
__device__ __forceinline__ int head_exponent(const int* __restrict__ e, int nb) {
    int emax = -1000;
    for (int i = 0; i < nb; ++i) emax = max(emax, __ldg(e + i));
    return emax;  // > 0: some block is too large; < 0: every block is tiny; 0 otherwise
}
// FIXUP = false: convert every block with the exponent it asks for itself and record the request.
// FIXUP = true (second launch, a few microseconds when nothing is to do): blocks whose own exponent
// differs from the head's are converted again with the head's; records ehead.
template <bool FIXUP>
__global__ void __launch_bounds__(256)
split_f16_kernel(const SplitArgs a) {
    __shared__ float s_red[2][8];
    const int y = blockIdx.y + a.y0;
    const bool isq = y < a.BH;
    const int bh = isq ? y : y - a.BH;
    const int heads = isq ? a.H : a.KVH;
    int e_forced = 0;
    if (FIXUP) {
        e_forced = head_exponent(a.eblk + (size_t)y * a.nb, a.nb);
        if (blockIdx.x == 0 && threadIdx.x == 0) a.ehead[y] = e_forced;
        if (e_forced == __ldg(a.eblk + (size_t)y * a.nb + blockIdx.x)) return;  // block-uniform
    }
    if (!FIXUP && isq && a.fuse_q) return;   // (not launched at all in the default build, see y0)
    const int b = bh / heads, h = bh - b * heads;
    const long long* st = isq ? a.qs : a.ks;
    const float* src0 = (isq ? a.q : a.k) + b * st[0] + h * st[1];
    const bool vec = isq ? a.vecq : a.veck;
    const float pre = isq ? a.cs : 1.0f;
    const int row0 = blockIdx.x * kTile;
    constexpr int UN = kTile * (kHd / 4) / 256;  // 8 float4 per thread; 16 consecutive threads share a row
    float4 v[UN];
#pragma unroll
    for (int u = 0; u < UN; ++u) {
        const int i = threadIdx.x + 256 * u;
        const int r = row0 + (i >> 4);
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < a.S) {
            const float* src = src0 + (long long)r * st[2] + (i & 15) * 4;
            if (vec) v[u] = __ldg(reinterpret_cast<const float4*>(src));
            else v[u] = make_float4(__ldg(src), __ldg(src + 1), __ldg(src + 2), __ldg(src + 3));
        }
        v[u].x *= pre; v[u].y *= pre; v[u].z *= pre; v[u].w *= pre;
    }
    int e = e_forced;
    if (!FIXUP) {
        uint32_t amax = 0u;
        float rmax = 0.f;
#pragma unroll
        for (int u = 0; u < UN; ++u) {
            amax = max(amax, __float_as_uint(v[u].x) & 0x7fffffffu);
            amax = max(amax, __float_as_uint(v[u].y) & 0x7fffffffu);
            amax = max(amax, __float_as_uint(v[u].z) & 0x7fffffffu);
            amax = max(amax, __float_as_uint(v[u].w) & 0x7fffffffu);
            float ss = fmaf(v[u].x, v[u].x, fmaf(v[u].y, v[u].y, fmaf(v[u].z, v[u].z, v[u].w * v[u].w)));
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);  // the 16 lanes of one row
            rmax = fmaxf(rmax, ss == ss ? ss : INFINITY);
        }
        amax = __reduce_max_sync(0xffffffffu, amax);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) rmax = fmaxf(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) {
            s_red[0][warp] = __uint_as_float(amax);
            s_red[1][warp] = rmax;
        }
        __syncthreads();
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            amax = max(amax, __float_as_uint(s_red[0][w]));
            rmax = fmaxf(rmax, s_red[1][w]);
        }
        // a block whose maximum lies in [2^-10, 2^15) asks for nothing (every ordinary activation); otherwise
        // 2^-e brings the block maximum into [2^14, 2^15): down from beyond the fp16 range, or up from so far
        // below it that the fp16 parts would be subnormal (e < 0)
        const int ex = (int)(amax >> 23) - 127;
        e = (ex >= 15 || (ex < -10 && amax != 0u)) ? ex - 14 : 0;
        if (threadIdx.x == 0) {
            a.bound2[(size_t)y * a.nb + blockIdx.x] = rmax;
            a.eblk[(size_t)y * a.nb + blockIdx.x] = e;
        }
    }
    const float sc = __int_as_float((127 - min(max(e, -126), 126)) << 23);
    const size_t n4 = (size_t)a.S * (kHd / 4);
    uint2* o0 = (isq ? a.q0 : a.k0) + (size_t)bh * n4 + (size_t)row0 * (kHd / 4);
    uint2* o1 = (isq ? a.q1 : a.k1) + (size_t)bh * n4 + (size_t)row0 * (kHd / 4);
#pragma unroll
    for (int u = 0; u < UN; ++u) {
        const int i = threadIdx.x + 256 * u;
        const int r = row0 + (i >> 4);
        if (r < a.S) {
            const float x0 = v[u].x * sc, x1 = v[u].y * sc, x2 = v[u].z * sc, x3 = v[u].w * sc;
            const __half2 a01 = __floats2half2_rn(x0, x1), a23 = __floats2half2_rn(x2, x3);
            const float2 f01 = __half22float2(a01), f23 = __half22float2(a23);
            const __half2 l01 = __floats2half2_rn(x0 - f01.x, x1 - f01.y);
            const __half2 l23 = __floats2half2_rn(x2 - f23.x, x3 - f23.y);
            o0[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&a01), *reinterpret_cast<const uint32_t*>(&a23));
            o1[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&l01), *reinterpret_cast<const uint32_t*>(&l23));
            if (!kLoaderPresets && !FIXUP && isq && (i & 15) == 0) {  // the Q blocks also preset the result slots of their (b,h,s) rows
                const size_t rr = (size_t)bh * a.S + r;
#pragma unroll
                for (int sl = 0; sl < kSlots; ++sl) {
                    a.ml[sl * a.rows + rr] = make_float2(-INFINITY, 0.f);
                    a.part_col[sl * a.rows + rr] = 0.f;
                }
            }
        }
    }
}

// pass 1 leaves up to kSlots (max, sum) pairs per row (two pieces of a cut item x two epilogue
// groups; unused slots hold (-inf, 0)); merge them in slot order into the log2-domain lse
__global__ void __launch_bounds__(256)
lse_combine_kernel(const float2* __restrict__ ml, size_t n, float* __restrict__ lse2, float* __restrict__ w2) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2 v[kSlots];
    float mm = -INFINITY;
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl) {
        v[sl] = ml[sl * n + i];
        mm = fmaxf(mm, v[sl].x);
    }
    // every row has at least one valid column (key 0), so the largest max is finite
    float sum = 0.f;
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl) sum += v[sl].x > -INFINITY ? v[sl].y * exp2f(v[sl].x - mm) : 0.f;
    const float lse = mm + log2f(sum);
    lse2[i] = lse;
    w2[i] = exp2f(-lse);  // only read by the FAST epilogue, whose bound keeps it a normal number
}

struct HeadW {
    float w[SPLITQ_MAX_HEADS];
};
// 32 columns per CTA, one warp per head chain: warp g sums the heads h = g, g + 8, ... of its 32 columns (the loop
// over heads is pure load latency -- 2 x groups + 1 loads per head and column -- so eight short chains instead of
// one long one: 14.9 -> 8.1 us under ncu, with four chains as with eight), then the partial sums meet in shared
// memory in fixed order
__global__ void __launch_bounds__(256)
fused_finalize_kernel(const float* __restrict__ part_col, const float* __restrict__ part_last, size_t rows, int H, int S,
                      HeadW hw, float* __restrict__ out_col, float* __restrict__ out_last, HeadW hw2,
                      float* __restrict__ out_col2) {
    __shared__ float red[3][8][32];
    const int jj = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + jj;
    const int b = blockIdx.y;
    float c = 0.f, l = 0.f, c2 = 0.f;
    if (j < S) {
        for (int h = g; h < H; h += 8) {
            const size_t o = ((size_t)b * H + h) * S + j;
            float pc = 0.f;
#pragma unroll
            for (int sl = 0; sl < kSlots; ++sl) pc += part_col[sl * rows + o];  // fixed order: schedule-independent
            c += hw.w[h] * pc;
            c2 += hw2.w[h] * pc;   // a second set of head weights over the same per-head sums (weighted_importance)
            l += hw.w[h] * part_last[o];
        }
    }
    red[0][g][jj] = c; red[1][g][jj] = l; red[2][g][jj] = c2;
    __syncthreads();
    if (g != 0 || j >= S) return;
    c = l = c2 = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { c += red[0][k][jj]; l += red[1][k][jj]; c2 += red[2][k][jj]; }
    if (out_col) out_col[(size_t)b * S + j] = c / (float)S;
    if (out_last) out_last[(size_t)b * S + j] = l;
    if (out_col2) out_col2[(size_t)b * S + j] = c2 / (float)S;
}

// ---- last-row importance only: a single-query GEMV per head, no S x S work at all ---------------
// logit2[b][h][j] = (q[b,h,S-1,:] . k[b,kvh,j,:]) * scaling * log2(e), exact fp32 FMAs on CUDA cores
__global__ void __launch_bounds__(256)
last_row_logits_kernel(const float* __restrict__ Q, long long qsb, long long qsh, long long qss,
                       const float* __restrict__ K, long long ksb, long long ksh, long long kss, int H, int KVH,
                       int S, float cs, float* __restrict__ logit2) {
    __shared__ float qs[kHd];
    const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
    const float* q = Q + b * qsb + h * qsh + (long long)(S - 1) * qss;
    if (threadIdx.x < kHd) qs[threadIdx.x] = q[threadIdx.x];
    __syncthreads();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= S) return;
    const float* kr = K + b * ksb + (h / (H / KVH)) * ksh + (long long)j * kss;
    float acc = 0.f;
#pragma unroll
    for (int d = 0; d < kHd; ++d) acc = fmaf(qs[d], __ldg(kr + d), acc);
    logit2[(size_t)bh * S + j] = acc * cs;
}
// out_last[b][j] = sum_h w_h * softmax_j(logits of head h); one CTA per window
__global__ void __launch_bounds__(1024)
last_row_softmax_kernel(const float* __restrict__ logit2, int H, int S, HeadW hw, float* __restrict__ out_last) {
    __shared__ float red[32];
    __shared__ float s_m, s_l;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = tid; j < S; j += blockDim.x) out_last[(size_t)b * S + j] = 0.f;
    for (int h = 0; h < H; ++h) {
        const float* x = logit2 + ((size_t)b * H + h) * S;
        float m = -INFINITY;
        for (int j = tid; j < S; j += blockDim.x) m = fmaxf(m, x[j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) red[warp] = m;
        __syncthreads();
        if (warp == 0) {
            float v = lane < (blockDim.x >> 5) ? red[lane] : -INFINITY;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
            if (lane == 0) s_m = v;
        }
        __syncthreads();
        m = s_m;
        float l = 0.f;
        for (int j = tid; j < S; j += blockDim.x) l += exp2f(x[j] - m);
        l = warp_sum(l);
        if (lane == 0) red[warp] = l;
        __syncthreads();
        if (warp == 0) {
            float v = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
            v = warp_sum(v);
            if (lane == 0) s_l = v;
        }
        __syncthreads();
        const float scale = hw.w[h] / s_l;
        for (int j = tid; j < S; j += blockDim.x) out_last[(size_t)b * S + j] += scale * exp2f(x[j] - m);
        __syncthreads();
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// the (n, S, 64) fp16 h0 array and the h1 array behind it, viewed as a 4-D tensor
// (64 halfs, S rows, n, 2 parts); box = 64 x 128 x 1 x parts, 128-byte swizzle, rows past S read as zero
static int make_map(CUtensorMap* map, void* base, int S, long long n, size_t part_stride_bytes, int parts) {
    EncodeTiledFn enc = encode_fn();
    SPLITQ_REQUIRE(enc != nullptr, SPLITQ_ERR_CUDA, "importance_fused: cuTensorMapEncodeTiled is unavailable");
    cuuint64_t dims[4] = {(cuuint64_t)kHd, (cuuint64_t)S, (cuuint64_t)n, 2};
    cuuint64_t strides[3] = {(cuuint64_t)kHd * 2, (cuuint64_t)S * kHd * 2, (cuuint64_t)part_stride_bytes};
    cuuint32_t box[4] = {(cuuint32_t)kHd, (cuuint32_t)kTileN, 1, (cuuint32_t)parts};  // one streamed tile
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, base, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    SPLITQ_REQUIRE(r == CUDA_SUCCESS, SPLITQ_ERR_CUDA, "importance_fused: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return SPLITQ_OK;
}

static size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

struct WsLayout {
    size_t q0, q1, k0, k1, lse, pcol, plast, ml, w2, bound2, eblk, ehead, total;  // fp16 h0 / h1 parts of Q and K, per-(b,h,s) floats, per-block / per-head facts
};
static WsLayout ws_layout(int B, int H, int KVH, int S) {
    WsLayout w;
    const size_t q = align256((size_t)B * H * S * kHd * 2), k = align256((size_t)B * KVH * S * kHd * 2);
    const size_t v = align256((size_t)B * H * S * 4);
    const size_t blk = align256((size_t)B * (H + KVH) * cdiv(S, kTile) * 4);
    w.q0 = 0; w.q1 = q; w.k0 = 2 * q; w.k1 = 2 * q + k;
    w.lse = 2 * q + 2 * k; w.pcol = w.lse + v; w.plast = w.pcol + kSlots * v; w.ml = w.plast + v;
    w.w2 = w.ml + 2 * kSlots * v; w.bound2 = w.w2 + v; w.eblk = w.bound2 + blk;
    w.ehead = w.eblk + blk;
    w.total = w.ehead + align256((size_t)B * (H + KVH) * 4);
    return w;
}

// ---- this build's entry points (C++ linkage; the exported C symbols are in importance_fused_api.cu) ----
int tile_cols() { return kTileN; }

size_t workspace_bytes(int B, int H, int KVH, int S, int hd) {
    if (B <= 0 || H <= 0 || KVH <= 0 || S <= 0 || hd != kHd) return 0;
    return ws_layout(B, H, KVH, S).total;
}

int fused_impl(const float* Q, const long long* q_strides_host, const float* K, const long long* k_strides_host,
                      int B, int H, int KVH, int S, int hd, float scaling, int causal, int mode,
                      const float* head_w_host, float* out_col, float* out_last, const float* head_w2_host,
                      float* out_col2, void* ws, size_t ws_bytes, splitq_stream_t stream) {
    SPLITQ_REQUIRE(Q && K && (out_col || out_last || out_col2) && q_strides_host && k_strides_host, SPLITQ_ERR_ARG,
                   "importance_fused: NULL argument");
    SPLITQ_REQUIRE(!out_col2 || head_w2_host, SPLITQ_ERR_ARG, "importance_fused: out_col2 given without head_w2_host");
    SPLITQ_REQUIRE(B > 0 && H > 0 && KVH > 0 && S > 0 && H % KVH == 0 && H <= SPLITQ_MAX_HEADS, SPLITQ_ERR_ARG,
                   "importance_fused: bad shape B=%d H=%d KVH=%d S=%d", B, H, KVH, S);
    SPLITQ_REQUIRE(hd == kHd, SPLITQ_ERR_ARG, "importance_fused: head_dim %d unsupported (only 64); use the "
                   "materialised path splitq_importance_colmean", hd);
    const int dbg = (mode >> 4) & 63;  // undocumented timing-experiment bits
    mode &= 15;
    SPLITQ_REQUIRE(mode == SPLITQ_FUSED_F16X3 || mode == SPLITQ_FUSED_F16X1, SPLITQ_ERR_ARG,
                   "importance_fused: unknown mode %d", mode);
    SPLITQ_REQUIRE(scaling > 0.f, SPLITQ_ERR_ARG, "importance_fused: scaling must be positive");
    const WsLayout wl = ws_layout(B, H, KVH, S);
    SPLITQ_REQUIRE(ws && ws_bytes >= wl.total, SPLITQ_ERR_WORKSPACE, "importance_fused: workspace %zu < %zu bytes",
                   ws_bytes, wl.total);
    SPLITQ_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255u) == 0, SPLITQ_ERR_ALIGN,
                   "importance_fused: workspace must be 256-byte aligned");
    if (int rc = check_arch()) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* w8 = (uint8_t*)ws;
    if (!out_col && !out_col2 && causal) {
        // "last_row" alone needs one query row per head: GEMV + softmax, exact fp32, no tensor cores
        HeadW hw1;
        for (int h = 0; h < H; ++h) hw1.w[h] = head_w_host ? head_w_host[h] : 1.0f / (float)H;
        float* logit2 = (float*)(w8 + wl.lse);
        SPLITQ_REQUIRE((long long)B * H <= 65535, SPLITQ_ERR_ARG, "importance_fused: B*H=%lld > 65535", (long long)B * H);
        last_row_logits_kernel<<<dim3(cdiv(S, 256), B * H), 256, 0, st>>>(
            Q, q_strides_host[0], q_strides_host[1], q_strides_host[2], K, k_strides_host[0], k_strides_host[1],
            k_strides_host[2], H, KVH, S, scaling * 1.4426950408889634f, logit2);
        last_row_softmax_kernel<<<B, 1024, 0, st>>>(logit2, H, S, hw1, out_last);
        SPLITQ_LAUNCH_CHECK("last_row kernels");
        return SPLITQ_OK;
    }

    const int nb = cdiv(S, kTile);
    {   // operand split: one launch for Q and K
        auto vec_ok = [](const float* p, const long long* s) {
            return aligned16(p) && s[0] % 4 == 0 && s[1] % 4 == 0 && s[2] % 4 == 0;
        };
        SPLITQ_REQUIRE((long long)B * (H + KVH) <= 65535, SPLITQ_ERR_ARG, "importance_fused: B*(H+KVH)=%lld > 65535",
                       (long long)B * (H + KVH));
        SplitArgs a;
        a.q = Q; a.k = K;
        for (int i = 0; i < 3; ++i) { a.qs[i] = q_strides_host[i]; a.ks[i] = k_strides_host[i]; }
        a.H = H; a.KVH = KVH; a.S = S; a.nb = nb; a.BH = B * H;
        a.vecq = vec_ok(Q, q_strides_host); a.veck = vec_ok(K, k_strides_host);
        a.cs = scaling * 1.4426950408889634f;
        a.q0 = (uint2*)(w8 + wl.q0); a.q1 = (uint2*)(w8 + wl.q1); a.k0 = (uint2*)(w8 + wl.k0); a.k1 = (uint2*)(w8 + wl.k1);
        a.ml = (float2*)(w8 + wl.ml); a.part_col = (float*)(w8 + wl.pcol); a.rows = (size_t)B * H * S;
        a.bound2 = (float*)(w8 + wl.bound2); a.eblk = (int*)(w8 + wl.eblk); a.ehead = (int*)(w8 + wl.ehead);
        a.fuse_q = kFuseQ ? 1 : 0;
        a.y0 = (kFuseQ && kLoaderPresets) ? B * H : 0;   // nothing left to do for the Q blocks' CTAs
        split_f16_kernel<false><<<dim3(nb, B * (H + KVH) - a.y0), 256, 0, st>>>(a);
        a.y0 = 0;
        if (kTileN != kTile)  // tiles that straddle blocks need head-uniform exponents; exits at once normally
            split_f16_kernel<true><<<dim3(nb, B * (H + KVH)), 256, 0, st>>>(a);
        SPLITQ_LAUNCH_CHECK("split_f16_kernel");
    }

    const int parts = mode == SPLITQ_FUSED_F16X3 ? 2 : 1;
    CUtensorMap mq, mk;
    if (int rc = make_map(&mq, w8 + wl.q0, S, (long long)B * H, wl.q1 - wl.q0, parts)) return rc;
    if (int rc = make_map(&mk, w8 + wl.k0, S, (long long)B * KVH, wl.k1 - wl.k0, parts)) return rc;

    Params p;
    p.B = B; p.H = H; p.KVH = KVH; p.S = S; p.nb = nb;
    p.nq_blk = B * H * nb; p.nk_blk = B * KVH * nb;
    p.w2 = (float*)(w8 + wl.w2); p.bound2 = (const float*)(w8 + wl.bound2);
    p.eblk = (const int*)(w8 + wl.eblk); p.ehead = (const int*)(w8 + wl.ehead);
    p.n_heads = B * (H + KVH);
    p.nbn = cdiv(S, kTileN);
    p.causal = causal ? 1 : 0;
    p.x3 = mode == SPLITQ_FUSED_F16X3;
    p.dbg = dbg;
    p.lse2 = (float*)(w8 + wl.lse);
    p.ml = (float2*)(w8 + wl.ml);
    p.part_col = (float*)(w8 + wl.pcol);
    p.part_last = (float*)(w8 + wl.plast);
    p.q16 = (const uint4*)(w8 + wl.q0); p.k16 = (const uint4*)(w8 + wl.k0);
    p.q_part = (wl.q1 - wl.q0) / 16; p.k_part = (wl.k1 - wl.k0) / 16;
    p.qsrc = Q; p.qs0 = q_strides_host[0]; p.qs1 = q_strides_host[1]; p.qs2 = q_strides_host[2];
    p.qvec = aligned16(Q) && p.qs0 % 4 == 0 && p.qs1 % 4 == 0 && p.qs2 % 4 == 0;
    if (p.qvec && (reinterpret_cast<uintptr_t>(Q) & 31u) == 0 && p.qs0 % 8 == 0 && p.qs1 % 8 == 0 && p.qs2 % 8 == 0) p.qvec = 2;
    p.cs = scaling * 1.4426950408889634f;
    p.q16_w = (uint4*)(w8 + wl.q0); p.bound2_w = (float*)(w8 + wl.bound2); p.eblk_w = (int*)(w8 + wl.eblk);

    // one-time per-device set-up; the entry point is re-entrant (any thread, any stream), so the
    // bookkeeping is under a lock -- the attribute calls themselves are idempotent
    static std::mutex attr_mu;
    static bool attr_set[64] = {};
    int dev = 0, sms = 148;
    SPLITQ_CUDA(cudaGetDevice(&dev));
    SPLITQ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    {
    std::lock_guard<std::mutex> attr_lock(attr_mu);
    if (dev >= 64 || !attr_set[dev]) {
        SPLITQ_CUDA(cudaFuncSetAttribute(fused_importance_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        SPLITQ_CUDA(cudaFuncSetAttribute(fused_importance_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        SPLITQ_CUDA(cudaFuncSetAttribute(fused_importance_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        SPLITQ_CUDA(cudaFuncSetAttribute(fused_importance_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        if (dev < 64) attr_set[dev] = true;
    }
    }
    // Tiles of one (b,h): item blk (a 128-row block of the fixed operand) meets the streamed kTileN-row tiles
    // [0, ceil(min((blk+1)*128, S) / kTileN)) in pass 1 (keys up to its last query) and
    // [floor(blk*128 / kTileN), nbn) in pass 2 (queries from its first key on); the two counts differ
    // unless the tile is 128 wide.  SegIter<PASS>::t0/t1 are the device-side twins of these formulas.
    long long per_bh[2] = {0, 0};
    for (int blk = 0; blk < p.nb; ++blk) {
        per_bh[0] += p.causal ? cdiv(std::min((blk + 1) * kTile, S), kTileN) : p.nbn;
        per_bh[1] += p.causal ? p.nbn - (blk * kTile) / kTileN : p.nbn;
    }
    const size_t rows = (size_t)B * H * S;
    for (int pass = 0; pass < 2; ++pass) {
        const long long tiles = (long long)B * H * per_bh[pass];
        SPLITQ_REQUIRE(tiles < (1ll << 31), SPLITQ_ERR_ARG, "importance_fused: %lld tiles per pass do not fit 31 bits", tiles);
        p.per_bh = (int)per_bh[pass];
        p.tiles = (int)tiles;
        // every CTA's range is at least as long as the longest item (nbn tiles), so an item is cut at most once
        const int grid = (int)std::max<long long>(1, std::min<long long>(sms, tiles / p.nbn));
        if (pass == 0) {
            if (dbg) fused_importance_kernel<1, true><<<grid, kThreads, kSmemBytes, st>>>(mq, mk, p);
            else fused_importance_kernel<1, false><<<grid, kThreads, kSmemBytes, st>>>(mq, mk, p);
            SPLITQ_LAUNCH_CHECK("fused_importance_kernel (pass 1)");
            lse_combine_kernel<<<(unsigned)cdiv((long long)rows, 256), 256, 0, st>>>(p.ml, rows, p.lse2, p.w2);
        } else {
            if (dbg) fused_importance_kernel<2, true><<<grid, kThreads, kSmemBytes, st>>>(mq, mk, p);
            else fused_importance_kernel<2, false><<<grid, kThreads, kSmemBytes, st>>>(mq, mk, p);
            SPLITQ_LAUNCH_CHECK("fused_importance_kernel (pass 2)");
        }
    }
    HeadW hw;
    for (int h = 0; h < H; ++h) hw.w[h] = head_w_host ? head_w_host[h] : 1.0f / (float)H;
    dim3 fg(cdiv(S, 32), B);
    HeadW hw2;
    for (int h = 0; h < H; ++h) hw2.w[h] = (out_col2 && head_w2_host) ? head_w2_host[h] : 0.f;
    fused_finalize_kernel<<<fg, 256, 0, st>>>(p.part_col, p.part_last, rows, H, S, hw, out_col, out_last, hw2, out_col2);
    SPLITQ_LAUNCH_CHECK("fused_finalize_kernel");
    return SPLITQ_OK;
}

}  // namespace SPLITQ_K2_NS
}  // namespace splitq
