// Microbenchmark: tensor-memory read bandwidth (tcgen05.ld) and tcgen05.mma issue rate on one SM,
// alone and together -- the two things the "MMA + TMEM-read" leg of K2 is made of.
//   ld  : W warps (each on its own lane quarter, W/4 warps per quarter) stream 32x32b.x32 / .x64 loads
//   mma : one thread issues kind::f16 M=128 MMAs (N = 128 or 256; A from TMEM or from shared memory)
// Prints clocks per instruction and the implied bytes/clk/SM and flop/clk/SM.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tmem_probe tmem_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (!done)
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                 ::"r"(d), "r"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
template <int NC>
__device__ __forceinline__ void ld_cols(uint32_t taddr, float (&v)[NC]) {
#pragma unroll
    for (int h = 0; h < NC / 32; ++h) {
        float* q = v + 32 * h;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
            "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=f"(q[0]), "=f"(q[1]), "=f"(q[2]), "=f"(q[3]), "=f"(q[4]), "=f"(q[5]), "=f"(q[6]), "=f"(q[7]),
              "=f"(q[8]), "=f"(q[9]), "=f"(q[10]), "=f"(q[11]), "=f"(q[12]), "=f"(q[13]), "=f"(q[14]), "=f"(q[15]),
              "=f"(q[16]), "=f"(q[17]), "=f"(q[18]), "=f"(q[19]), "=f"(q[20]), "=f"(q[21]), "=f"(q[22]), "=f"(q[23]),
              "=f"(q[24]), "=f"(q[25]), "=f"(q[26]), "=f"(q[27]), "=f"(q[28]), "=f"(q[29]), "=f"(q[30]), "=f"(q[31])
            : "r"(taddr + 32 * h) : "memory");
    }
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

struct Out {
    long long ld_clk, mma_clk;
    float sink;
};

// ld_warps: 0, 4, 8 or 12 loading warps (warps 4.. of the CTA); ld_cols_per: 32 or 64 columns per wait
// mma: 0 none, 1 TS N=128, 2 SS N=128, 3 TS N=256, 4 SS N=256, 5 TS N=192, 6 SS N=192, 7 TS N=64
__global__ void __launch_bounds__(512, 1) probe(int ld_warps, int ld_per, int mma, int iters, int chains, Out* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t base = smem_u32(smem);
    base = (base + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&bars[0]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    float sink = 0.f;
    long long t_ld = 0, t_mma = 0;
    if (warp == 1 && mma) {
        if (lane == 0) {
            const int n = mma >= 7 ? 64 : (mma >= 5 ? 192 : ((mma >= 3) ? 256 : 128));
            const uint32_t idesc = (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
            const uint64_t bd = umma_desc(base), ad = umma_desc(base + 32 * 1024);
            uint32_t acc_off = 0;
            const long long t0 = clock64();
            for (int i = 0; i < iters; ++i) {
                // `chains` accumulators in rotation: consecutive MMAs on one accumulator depend on each other
                // (kept to an add and a select: a runtime modulo here made the issuing thread the limit, 157 clk
                // per MMA whatever N)
                const uint32_t d = tmem + acc_off;
                acc_off = (acc_off + (uint32_t)n >= (uint32_t)(chains * n)) ? 0u : acc_off + (uint32_t)n;
                if (mma == 1 || mma == 3 || mma == 5 || mma == 7) mma_ts(d, tmem + 384u + (uint32_t)(i & 3) * 8u, bd + (uint64_t)((i & 3) * 2), idesc);
                else mma_ss(d, ad + (uint64_t)((i & 3) * 2), bd + (uint64_t)((i & 3) * 2), idesc);
            }
            tc_commit(smem_u32(&bars[0]));
            mbar_wait(smem_u32(&bars[0]), 0);
            t_mma = clock64() - t0;
        }
        __syncwarp();
    } else if (warp >= 4 && warp < 4 + ld_warps) {
        const uint32_t lane_addr = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        float v[64];
        const long long t0 = clock64();
        if (ld_per == 32) {
            for (int i = 0; i < iters; ++i) {
                ld_cols<32>(lane_addr + (uint32_t)((i * 32) & 255), reinterpret_cast<float(&)[32]>(v));
                ld_wait();
                sink += v[0] + v[31];
            }
        } else {
            for (int i = 0; i < iters; ++i) {
                ld_cols<64>(lane_addr + (uint32_t)((i * 64) & 255), v);
                ld_wait();
                sink += v[0] + v[63];
            }
        }
        t_ld = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (blockIdx.x == 0) {
        if (threadIdx.x == 32) out->mma_clk = t_mma;
        if (threadIdx.x == 128) { out->ld_clk = t_ld; out->sink = sink; }
    }
    if (sink == 123.456f) out->sink = sink;
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

int main() {
    Out* out;
    cudaMallocManaged(&out, sizeof(Out));
    const int smem = 64 * 1024 + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 4096;
    const char* mma_names[] = {"none", "TS N=128", "SS N=128", "TS N=256", "SS N=256", "TS N=192", "SS N=192", "TS N=64"};
    for (int grid : {148}) {
        for (int mma = 0; mma <= 7; ++mma) {
            for (int lw : {0, 8}) {
                for (int per : {32, 64}) {
                  for (int chains : {1, 2, 3}) {
                    if (lw == 0 && per == 64) continue;
                    if (lw == 0 && mma == 0) continue;
                    if (mma == 0 && chains > 1) continue;
                    if (lw != 0 && chains != 2) continue;
                    const int nn = mma >= 7 ? 64 : (mma >= 5 ? 192 : (mma >= 3 ? 256 : 128));
                    if (chains * nn > 384) continue;   // accumulators below the A operand's columns
                    *out = Out{0, 0, 0.f};
                    probe<<<grid, 512, smem>>>(lw, per, mma, iters, chains, out);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
                    const double ld_clk = (double)out->ld_clk / iters, mma_clk = (double)out->mma_clk / iters;
                    const int n = mma >= 7 ? 64 : (mma >= 5 ? 192 : (mma >= 3 ? 256 : 128));
                    printf("grid %3d mma %-9s chains %d ld warps %2d x%-2d : ", grid, mma_names[mma], chains, lw, per);
                    if (lw) printf("%.1f clk per ld+wait per warp -> %.1f B/clk/SM  ", ld_clk, lw * per * 128.0 / ld_clk);
                    if (mma) printf("%.1f clk per MMA -> %.0f flop/clk/SM", mma_clk, 2.0 * 128 * n * 16 / mma_clk);
                    printf("\n");
                  }
                }
            }
        }
    }
    return 0;
}
