// Microbenchmark: what is the ex2 (MUFU) floor of the K2 epilogue on this GPU?
// Measures clocks per warp-level MUFU.EX2 per SM sub-partition, alone and inside the
// FFMA -> EX2 -> FADD mix of the pass-2 body, for 1, 2 and 4 warps per sub-partition.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mufu_probe mufu_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int MIX>
__global__ void probe(float* out, long long* clk, int iters, float cs, float off) {
    float v[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = (float)(threadIdx.x + i) * 1e-3f;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
            if (MIX >= 2) {
                // batched form: MIX exp2 results land in temporaries before any of them is consumed
                constexpr int BM = MIX >= 2 ? MIX : 4;
                if (i % BM == 0) {
                    float t[BM];
#pragma unroll
                    for (int j = 0; j < BM; ++j) t[j] = ex2(fmaf(v[i + j], cs, -off));
#pragma unroll
                    for (int j = 0; j < BM; j += 4) {
                        s0 += t[j + 0];
                        s1 += t[j + 1];
                        s2 += t[j + 2];
                        s3 += t[j + 3];
                    }
                }
            } else if (MIX) {
                s0 += ex2(fmaf(v[i + 0], cs, -off));
                s1 += ex2(fmaf(v[i + 1], cs, -off));
                s2 += ex2(fmaf(v[i + 2], cs, -off));
                s3 += ex2(fmaf(v[i + 3], cs, -off));
            } else {
                v[i + 0] = ex2(v[i + 0]);
                v[i + 1] = ex2(v[i + 1]);
                v[i + 2] = ex2(v[i + 2]);
                v[i + 3] = ex2(v[i + 3]);
            }
        }
        off += 1e-6f;
    }
    const long long t1 = clock64();
    float acc = (s0 + s1) + (s2 + s3);
#pragma unroll
    for (int i = 0; i < 64; ++i) acc += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
    float* out;
    long long* clk;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&clk, 148 * sizeof(long long));
    const int iters = 2000;
    for (int mix = 0; mix < 5; ++mix)
        for (int warps_per_smsp = 1; warps_per_smsp <= 4; warps_per_smsp *= 2) {
            const int threads = 128 * warps_per_smsp;
            for (int rep = 0; rep < 2; ++rep) {
                if (mix == 1) probe<1><<<148, threads>>>(out, clk, iters, 0.18f, 3.f);
                else if (mix == 2) probe<8><<<148, threads>>>(out, clk, iters, 0.18f, 3.f);
                else if (mix == 3) probe<16><<<148, threads>>>(out, clk, iters, 0.18f, 3.f);
                else if (mix == 4) probe<32><<<148, threads>>>(out, clk, iters, 0.18f, 3.f);
                else probe<0><<<148, threads>>>(out, clk, iters, 0.18f, 3.f);
            }
            cudaDeviceSynchronize();
            long long h[148];
            cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
            double c = 0;
            for (int i = 0; i < 148; ++i) c += (double)h[i];
            c /= 148;
            const double warp_instr_per_smsp = (double)iters * 64 * warps_per_smsp;
            printf("%s warps/SMSP=%d: %.2f clk per warp-level EX2 per sub-partition (%.1f ex2/clk/SM)\n",
                   mix == 0 ? "ex2 only        " : mix == 1 ? "ffma+ex2+fadd   " : mix == 2 ? "batched by 8    " : mix == 3 ? "batched by 16   " : "batched by 32   ", warps_per_smsp, c / warp_instr_per_smsp,
                   32.0 * 4 * warp_instr_per_smsp / c);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
