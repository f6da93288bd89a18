// microbench.cu -- pipe-rate probes for sm_100a (B200): what the FMA pipes, the
// XU (MUFU) pipe and the FP64 pipe sustain per clock per SM, alone and mixed in
// the ratio of the direct-sum inner loop.  Cycle counts come from clock64()
// inside the kernel (per CTA, max over CTAs), so the rates do not depend on the
// SM clock at the time of the run.  Output: one JSON object per line.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CHAINS 8

__device__ __forceinline__ float rsq(float x)
{
    float y;
    asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ double rsq64(double x)
{
    double y;
    asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ float ffma(float a, float b, float c)
{
    float d;
    asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long d, ua, ub, uc;
    ua = *reinterpret_cast<unsigned long long *>(&a);
    ub = *reinterpret_cast<unsigned long long *>(&b);
    uc = *reinterpret_cast<unsigned long long *>(&c);
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(ua), "l"(ub), "l"(uc));
    return *reinterpret_cast<float2 *>(&d);
}
__device__ __forceinline__ double dfma(double a, double b, double c)
{
    double d;
    asm volatile("fma.rn.f64 %0, %1, %2, %3;" : "=d"(d) : "d"(a), "d"(b), "d"(c));
    return d;
}

// MODE 0: NF scalar FFMA + NM MUFU.RSQ per chain per iteration
// MODE 1: NF packed FFMA2 + NM MUFU.RSQ
// MODE 2: NF DFMA + NM MUFU.RSQ64H (rsqrt.approx.f64)
template <int MODE, int NF, int NM>
__global__ void __launch_bounds__(256) probe(float *out, long long *cycles, int iters, float seed)
{
    float a[CHAINS], m[CHAINS];
    float2 p[CHAINS];
    double d[CHAINS], dm[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        a[c] = seed + c + threadIdx.x * 1e-3f;
        m[c] = 1.5f + c;
        p[c] = make_float2(a[c], a[c] + 0.5f);
        d[c] = a[c];
        dm[c] = 1.5 + c;
    }
    const float k1 = 0.9999f, k2 = 1e-6f;
    const float2 q1 = make_float2(k1, k1), q2 = make_float2(k2, k2);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int f = 0; f < (NF > NM ? NF : NM); ++f) {
#pragma unroll
            for (int c = 0; c < CHAINS; ++c) {
                if (f < NF) {
                    if (MODE == 0) a[c] = ffma(a[c], k1, k2);
                    if (MODE == 1) p[c] = ffma2(p[c], q1, q2);
                    if (MODE == 2) d[c] = dfma(d[c], 0.9999, 1e-6);
                }
                if (f < NM) {
                    if (MODE == 2) dm[c] = rsq64(dm[c]);
                    else m[c] = rsq(m[c]);
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) s += a[c] + m[c] + p[c].x + p[c].y + (float)d[c] + (float)dm[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE, int NF, int NM>
void run(const char *name, int ctas_per_sm, int sms)
{
    const int grid = sms * ctas_per_sm, block = 256, iters = 4096;
    float *out;
    long long *cyc;
    cudaMalloc(&out, sizeof(float) * grid * block);
    cudaMalloc(&cyc, sizeof(long long) * grid);
    probe<MODE, NF, NM><<<grid, block>>>(out, cyc, 16, 1.f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<MODE, NF, NM><<<grid, block>>>(out, cyc, iters, 1.f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long *h = (long long *)malloc(sizeof(long long) * grid);
    cudaMemcpy(h, cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    // per SM: ctas_per_sm CTAs x 8 warps x CHAINS x iters x N ops (warp instructions)
    const double warp_instr_f = (double)ctas_per_sm * 8 * CHAINS * iters * NF;
    const double warp_instr_m = (double)ctas_per_sm * 8 * CHAINS * iters * NM;
    const double lanes_f = warp_instr_f * 32 * (MODE == 1 ? 2 : 1) / (double)mx;
    const double lanes_m = warp_instr_m * 32 / (double)mx;
    printf("{\"probe\": \"%s\", \"mode\": %d, \"fma_per_iter\": %d, \"mufu_per_iter\": %d, "
           "\"ctas_per_sm\": %d, \"cycles\": %lld, \"ms\": %.4f, \"mhz_implied\": %.0f, "
           "\"fma_lane_ops_per_clk_per_sm\": %.2f, \"mufu_lane_ops_per_clk_per_sm\": %.2f, "
           "\"cycles_per_smsp_per_chain_iter\": %.3f}\n",
           name, MODE, NF, NM, ctas_per_sm, mx, ms, (double)mx / (ms * 1e3), lanes_f, lanes_m,
           (double)mx / ((double)ctas_per_sm * 2 * CHAINS * iters));
    free(h);
    cudaFree(out);
    cudaFree(cyc);
}

int main()
{
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) {
        fprintf(stderr, "no CUDA device\n");
        return 1;
    }
    const int sms = prop.multiProcessorCount;
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", prop.name, sms, prop.clockRate);
    run<0, 1, 0>("ffma", 2, sms);
    run<1, 1, 0>("ffma2", 2, sms);
    run<0, 0, 1>("mufu_rsq", 2, sms);
    run<0, 12, 1>("ffma12+mufu1 (scalar loop ratio)", 2, sms);
    run<1, 6, 1>("ffma2x6+mufu1 (packed loop ratio)", 2, sms);
    run<1, 4, 1>("ffma2x4+mufu1", 2, sms);
    run<1, 3, 1>("ffma2x3+mufu1", 2, sms);
    run<1, 8, 1>("ffma2x8+mufu1", 2, sms);
    run<0, 6, 1>("ffma6+mufu1", 2, sms);
    run<2, 1, 0>("dfma", 2, sms);
    run<2, 0, 1>("mufu_rsq64h", 2, sms);
    run<2, 17, 1>("dfma17+mufu64 (fp64 loop ratio)", 2, sms);
    run<0, 1, 0>("ffma occ1", 1, sms);
    run<1, 1, 0>("ffma2 occ4", 4, sms);
    return 0;
}
