// innerloop_bench.cu -- A/B harness for the inner loop of the FP32 direct-sum kernel
// (n-body-simulation_b200/csrc/nb_direct_f32.cu) on sm_100a.  Every variant runs the same
// arithmetic -- all (i, j) interactions of a thread's i-bodies against a 256-body tile in shared
// memory, FP32 partial sums flushed into FP64 running sums once per tile (or per FLUSH tiles) --
// with a different register blocking / operand packing / j-broadcast scheme, on a full machine
// (148 SMs x CTAS CTAs x NW warps).  Reported: clk per interaction-warp per SM sub-partition
// (cycles from clock64(), max over CTAs), the number the production kernel is judged by
// (13.6 in round 1; 11.46 = the bare 5.5 FFMA2 + 1 MUFU mix).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o innerloop_bench innerloop_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define TQ 64 // quads per tile

__device__ __forceinline__ float rsq(float x)
{
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- variant A: j-packed (production, round 1): 4 i-bodies per thread, float2 = two j-bodies
template <bool UNI>
__device__ __forceinline__ void pair_jpacked(float2 xj, float2 yj, float2 zj, float2 mj, const float (&xi)[4],
                                             const float (&yi)[4], const float (&zi)[4], float2 e2,
                                             float2 (&ax)[4], float2 (&ay)[4], float2 (&az)[4])
{
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float2 dx = __fadd2_rn(xj, make_float2(-xi[r], -xi[r]));
        float2 dy = __fadd2_rn(yj, make_float2(-yi[r], -yi[r]));
        float2 dz = __fadd2_rn(zj, make_float2(-zi[r], -zi[r]));
        float2 r2 = __ffma2_rn(dx, dx, e2);
        r2 = __ffma2_rn(dy, dy, r2);
        r2 = __ffma2_rn(dz, dz, r2);
        float2 ri = make_float2(rsq(r2.x), rsq(r2.y));
        float2 ri2 = __fmul2_rn(ri, ri);
        float2 s = UNI ? __fmul2_rn(ri2, ri) : __fmul2_rn(ri2, __fmul2_rn(mj, ri));
        ax[r] = __ffma2_rn(dx, s, ax[r]);
        ay[r] = __ffma2_rn(dy, s, ay[r]);
        az[r] = __ffma2_rn(dz, s, az[r]);
    }
}

// ---- variant B/C: i-packed: float2 = two i-bodies, the j-body is a broadcast scalar.  NP i-pairs
// per thread (NP = 2: 4 i-bodies, NP = 4: 8 i-bodies); accumulators are per i only (half of A's).
template <bool UNI, int NP>
__device__ __forceinline__ void one_j_ipacked(float xj, float yj, float zj, float mj, const float2 (&nxi)[NP],
                                              const float2 (&nyi)[NP], const float2 (&nzi)[NP], float e2,
                                              float2 (&ax)[NP], float2 (&ay)[NP], float2 (&az)[NP])
{
#pragma unroll
    for (int p = 0; p < NP; ++p) {
        float2 dx = __fadd2_rn(nxi[p], make_float2(xj, xj));
        float2 dy = __fadd2_rn(nyi[p], make_float2(yj, yj));
        float2 dz = __fadd2_rn(nzi[p], make_float2(zj, zj));
        float2 r2 = __ffma2_rn(dx, dx, make_float2(e2, e2));
        r2 = __ffma2_rn(dy, dy, r2);
        r2 = __ffma2_rn(dz, dz, r2);
        float2 ri = make_float2(rsq(r2.x), rsq(r2.y));
        float2 ri2 = __fmul2_rn(ri, ri);
        float2 s = UNI ? __fmul2_rn(ri2, ri) : __fmul2_rn(ri2, __fmul2_rn(make_float2(mj, mj), ri));
        ax[p] = __ffma2_rn(dx, s, ax[p]);
        ay[p] = __ffma2_rn(dy, s, ay[p]);
        az[p] = __ffma2_rn(dz, s, az[p]);
    }
}

struct Args {
    const float4 *src; // one tile: TQ quads x {X4,Y4,Z4,M4}
    float *out;
    long long *cycles;
    int ntiles;
    int flush; // flush FP32 partials into the FP64 sums every `flush` tiles
};

// VAR 0: j-packed 4i (production)   VAR 1: i-packed NP pairs   VAR 2: i-packed, j via warp shuffle
template <int VAR, bool UNI, int NP, int NW, int MINB, int UNROLL>
__global__ void __launch_bounds__(NW * 32, MINB) bench_kernel(const Args a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4(*tile)[TQ * 4] = reinterpret_cast<float4(*)[TQ * 4]>(smem_raw);
    double(*dacc)[NW * 32] = reinterpret_cast<double(*)[NW * 32]>(smem_raw + 2 * TQ * 4 * sizeof(float4));
    for (int i = threadIdx.x; i < TQ * 4; i += NW * 32) {
        tile[0][i] = a.src[i];
        tile[1][i] = a.src[i];
    }
    const int t = threadIdx.x;
    const float e2 = 0.01f;
    constexpr int NACC = (VAR == 0 ? 12 : 6 * NP);
#pragma unroll
    for (int c = 0; c < NACC; ++c) dacc[c][t] = 0.0;
    __syncthreads();
    const float base = 0.001f * (float)(blockIdx.x * blockDim.x + t);
    long long t0 = 0, t1 = 0;

    if (VAR == 0) {
        float xi[4], yi[4], zi[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) { xi[r] = base + r; yi[r] = base - r; zi[r] = 0.5f * base + r; }
        float2 ax[4], ay[4], az[4];
        const float2 e22 = make_float2(e2, e2);
        t0 = clock64();
        for (int k = 0; k < a.ntiles; ++k) {
            if (k % a.flush == 0) {
#pragma unroll
                for (int r = 0; r < 4; ++r) ax[r] = ay[r] = az[r] = make_float2(0.f, 0.f);
            }
            const float4 *__restrict__ tl = tile[k & 1];
#pragma unroll UNROLL
            for (int jq = 0; jq < TQ; ++jq) {
                const float4 JX = tl[jq * 4 + 0], JY = tl[jq * 4 + 1], JZ = tl[jq * 4 + 2];
                float4 JM = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!UNI) JM = tl[jq * 4 + 3];
                pair_jpacked<UNI>(make_float2(JX.x, JX.y), make_float2(JY.x, JY.y), make_float2(JZ.x, JZ.y),
                                  make_float2(JM.x, JM.y), xi, yi, zi, e22, ax, ay, az);
                pair_jpacked<UNI>(make_float2(JX.z, JX.w), make_float2(JY.z, JY.w), make_float2(JZ.z, JZ.w),
                                  make_float2(JM.z, JM.w), xi, yi, zi, e22, ax, ay, az);
            }
            __syncwarp();
            if ((k + 1) % a.flush == 0) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    dacc[0 + r][t] += (double)(ax[r].x + ax[r].y);
                    dacc[4 + r][t] += (double)(ay[r].x + ay[r].y);
                    dacc[8 + r][t] += (double)(az[r].x + az[r].y);
                }
            }
        }
        t1 = clock64();
    } else {
        float2 nxi[NP], nyi[NP], nzi[NP], ax[NP], ay[NP], az[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            nxi[p] = make_float2(-(base + 2 * p), -(base + 2 * p + 1));
            nyi[p] = make_float2(-(base - 2 * p), -(base - 2 * p - 1));
            nzi[p] = make_float2(-(0.5f * base + 2 * p), -(0.5f * base + 2 * p + 1));
        }
        const int lane = t & 31;
        t0 = clock64();
        for (int k = 0; k < a.ntiles; ++k) {
            if (k % a.flush == 0) {
#pragma unroll
                for (int p = 0; p < NP; ++p) ax[p] = ay[p] = az[p] = make_float2(0.f, 0.f);
            }
            const float4 *__restrict__ tl = tile[k & 1];
            if (VAR == 1) {
#pragma unroll UNROLL
                for (int jq = 0; jq < TQ; ++jq) {
                    const float4 JX = tl[jq * 4 + 0], JY = tl[jq * 4 + 1], JZ = tl[jq * 4 + 2];
                    float4 JM = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!UNI) JM = tl[jq * 4 + 3];
                    one_j_ipacked<UNI, NP>(JX.x, JY.x, JZ.x, JM.x, nxi, nyi, nzi, e2, ax, ay, az);
                    one_j_ipacked<UNI, NP>(JX.y, JY.y, JZ.y, JM.y, nxi, nyi, nzi, e2, ax, ay, az);
                    one_j_ipacked<UNI, NP>(JX.z, JY.z, JZ.z, JM.z, nxi, nyi, nzi, e2, ax, ay, az);
                    one_j_ipacked<UNI, NP>(JX.w, JY.w, JZ.w, JM.w, nxi, nyi, nzi, e2, ax, ay, az);
                }
            } else {
                // the north star's "warp-shuffle broadcast": every lane holds one j-body of a
                // 32-body group in registers (one LDS per 32 j-bodies per component) and the
                // group is fanned out with SHFL.IDX, one j per trip
                const float *__restrict__ tf = reinterpret_cast<const float *>(tl);
                for (int g = 0; g < TQ * 4 / 32; ++g) {
                    const int q = g * 8 + (lane >> 2), c = lane & 3; // quad, slot inside the quad
                    const float mx = tf[q * 16 + 0 + c], my = tf[q * 16 + 4 + c], mz = tf[q * 16 + 8 + c];
                    float mm = 0.f;
                    if (!UNI) mm = tf[q * 16 + 12 + c];
#pragma unroll UNROLL
                    for (int l = 0; l < 32; ++l) {
                        const float xj = __shfl_sync(0xffffffffu, mx, l), yj = __shfl_sync(0xffffffffu, my, l),
                                    zj = __shfl_sync(0xffffffffu, mz, l);
                        float mj = 0.f;
                        if (!UNI) mj = __shfl_sync(0xffffffffu, mm, l);
                        one_j_ipacked<UNI, NP>(xj, yj, zj, mj, nxi, nyi, nzi, e2, ax, ay, az);
                    }
                }
            }
            __syncwarp();
            if ((k + 1) % a.flush == 0) {
#pragma unroll
                for (int p = 0; p < NP; ++p) {
                    dacc[6 * p + 0][t] += (double)ax[p].x;
                    dacc[6 * p + 1][t] += (double)ax[p].y;
                    dacc[6 * p + 2][t] += (double)ay[p].x;
                    dacc[6 * p + 3][t] += (double)ay[p].y;
                    dacc[6 * p + 4][t] += (double)az[p].x;
                    dacc[6 * p + 5][t] += (double)az[p].y;
                }
            }
        }
        t1 = clock64();
    }
    double s = 0.0;
#pragma unroll
    for (int c = 0; c < NACC; ++c) s += dacc[c][t];
    a.out[blockIdx.x * blockDim.x + t] = (float)s;
    if (t == 0) a.cycles[blockIdx.x] = t1 - t0;
}

// ---- variant W: both FMA pipes.  B200 has 64 FP64 lanes per SM next to the 128 FP32 lanes, idle in
// an FP32 kernel.  N32 warps run the i-packed FP32 loop (NP pairs per thread), N64 warps of the same
// CTA run the SAME interactions for other i-bodies in FP64 (R64 i-bodies per thread, rsqrt seed from
// MUFU.RSQ64H + one correction: 14 FP64 operations per interaction), so both pipes are fed from the
// one SM sub-partition scheduler.  The FP64 warps convert the tile to doubles once per tile.
__device__ __forceinline__ double rsq64(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}

template <int NP, int N32, int N64, int R64, bool CORRECT>
__global__ void __launch_bounds__((N32 + N64) * 32, 1) hetero_kernel(const Args a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int NT = (N32 + N64) * 32;
    float4(*tile)[TQ * 4] = reinterpret_cast<float4(*)[TQ * 4]>(smem_raw);                   // 2 x 4 KB
    double4(*dtile)[TQ * 4] = reinterpret_cast<double4(*)[TQ * 4]>(smem_raw + 2 * TQ * 4 * 16); // 2 x 8 KB
    double(*dacc)[N32 * 32] = reinterpret_cast<double(*)[N32 * 32]>(smem_raw + 2 * TQ * 4 * 16 + 2 * TQ * 4 * 32);
    for (int i = threadIdx.x; i < TQ * 4; i += NT) {
        tile[0][i] = a.src[i];
        tile[1][i] = a.src[i];
    }
    const int t = threadIdx.x;
    const int warp = t >> 5;
    const float e2 = 0.01f;
    const float base = 0.001f * (float)(blockIdx.x * blockDim.x + t);
    long long t0 = 0, t1 = 0;
    double sum = 0.0;
    __syncthreads();
    if (warp < N32) {
#pragma unroll
        for (int c = 0; c < 6 * NP; ++c) dacc[c][t] = 0.0;
        float2 nxi[NP], nyi[NP], nzi[NP], ax[NP], ay[NP], az[NP];
#pragma unroll
        for (int p = 0; p < NP; ++p) {
            nxi[p] = make_float2(-(base + 2 * p), -(base + 2 * p + 1));
            nyi[p] = make_float2(-(base - 2 * p), -(base - 2 * p - 1));
            nzi[p] = make_float2(-(0.5f * base + 2 * p), -(0.5f * base + 2 * p + 1));
        }
        t0 = clock64();
        for (int k = 0; k < a.ntiles; ++k) {
#pragma unroll
            for (int p = 0; p < NP; ++p) ax[p] = ay[p] = az[p] = make_float2(0.f, 0.f);
            const float4 *__restrict__ tl = tile[k & 1];
#pragma unroll 1
            for (int jq = 0; jq < TQ; ++jq) {
                const float4 JX = tl[jq * 4 + 0], JY = tl[jq * 4 + 1], JZ = tl[jq * 4 + 2];
                one_j_ipacked<true, NP>(JX.x, JY.x, JZ.x, 0.f, nxi, nyi, nzi, e2, ax, ay, az);
                one_j_ipacked<true, NP>(JX.y, JY.y, JZ.y, 0.f, nxi, nyi, nzi, e2, ax, ay, az);
                one_j_ipacked<true, NP>(JX.z, JY.z, JZ.z, 0.f, nxi, nyi, nzi, e2, ax, ay, az);
                one_j_ipacked<true, NP>(JX.w, JY.w, JZ.w, 0.f, nxi, nyi, nzi, e2, ax, ay, az);
            }
            __syncwarp();
#pragma unroll
            for (int p = 0; p < NP; ++p) {
                dacc[6 * p + 0][t] += (double)ax[p].x;
                dacc[6 * p + 1][t] += (double)ax[p].y;
                dacc[6 * p + 2][t] += (double)ay[p].x;
                dacc[6 * p + 3][t] += (double)ay[p].y;
                dacc[6 * p + 4][t] += (double)az[p].x;
                dacc[6 * p + 5][t] += (double)az[p].y;
            }
        }
        t1 = clock64();
#pragma unroll
        for (int c = 0; c < 6 * NP; ++c) sum += dacc[c][t];
    } else if (N64 > 0) {
        const int t64 = t - N32 * 32; // 0 .. N64*32-1
        double xi[R64 > 0 ? R64 : 1], yi[R64 > 0 ? R64 : 1], zi[R64 > 0 ? R64 : 1];
        double ax[R64 > 0 ? R64 : 1], ay[R64 > 0 ? R64 : 1], az[R64 > 0 ? R64 : 1];
#pragma unroll
        for (int r = 0; r < R64; ++r) {
            xi[r] = base + r; yi[r] = base - r; zi[r] = 0.5 * base + r;
            ax[r] = ay[r] = az[r] = 0.0;
        }
        const double de2 = 0.01;
        t0 = clock64();
        for (int k = 0; k < a.ntiles; ++k) {
            // the FP64 warps turn the float tile into doubles (their own copy), once per tile
            const float *__restrict__ tf = reinterpret_cast<const float *>(tile[k & 1]);
            double4 *__restrict__ dt = dtile[k & 1];
            for (int j = t64; j < TQ * 4; j += N64 * 32) {
                const int q = j >> 2, c = j & 3;
                dt[j] = make_double4((double)tf[q * 16 + c], (double)tf[q * 16 + 4 + c], (double)tf[q * 16 + 8 + c], 0.0);
            }
            asm volatile("bar.sync 1, %0;" ::"r"(N64 * 32) : "memory");
#pragma unroll 2
            for (int j = 0; j < TQ * 4; ++j) {
                const double4 b = dt[j];
#pragma unroll
                for (int r = 0; r < R64; ++r) {
                    const double dx = b.x - xi[r], dy = b.y - yi[r], dz = b.z - zi[r];
                    const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, de2)));
                    const double y0 = rsq64(r2);
                    const double y02 = y0 * y0;
                    double s = y02 * y0;
                    if (CORRECT) { // one correction of the cube: y0^3 (1 + 3/2 e), e = 1 - r2 y0^2
                        const double e = fma(-r2, y02, 1.0);
                        s = fma(s * e, 1.5, s);
                    }
                    ax[r] = fma(dx, s, ax[r]);
                    ay[r] = fma(dy, s, ay[r]);
                    az[r] = fma(dz, s, az[r]);
                }
            }
            asm volatile("bar.sync 1, %0;" ::"r"(N64 * 32) : "memory");
        }
        t1 = clock64();
#pragma unroll
        for (int r = 0; r < R64; ++r) sum += ax[r] + ay[r] + az[r];
    }
    a.out[blockIdx.x * blockDim.x + t] = (float)sum;
    // per CTA: [0] cycles of an FP32 warp, [1] cycles of an FP64 warp
    if (t == 0) a.cycles[2 * blockIdx.x] = t1 - t0;
    if (t == N32 * 32) a.cycles[2 * blockIdx.x + 1] = t1 - t0;
}

template <int NP, int N32, int N64, int R64, bool CORRECT>
static void run_hetero(const char *name, const Args &base, int sms)
{
    Args a = base;
    const int grid = sms;
    constexpr int NT = (N32 + N64) * 32;
    const size_t smem = 2 * TQ * 4 * 16 + 2 * TQ * 4 * 32 + (size_t)6 * NP * N32 * 32 * sizeof(double);
    cudaFuncSetAttribute(hetero_kernel<NP, N32, N64, R64, CORRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, hetero_kernel<NP, N32, N64, R64, CORRECT>);
    cudaMemset(a.cycles, 0, sizeof(long long) * 2 * grid);
    Args w = a;
    w.ntiles = 8;
    hetero_kernel<NP, N32, N64, R64, CORRECT><<<grid, NT, smem>>>(w);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    hetero_kernel<NP, N32, N64, R64, CORRECT><<<grid, NT, smem>>>(a);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        printf("{\"variant\": \"%s\", \"error\": \"%s\"}\n", name, cudaGetErrorString(err));
        return;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long *h = (long long *)malloc(sizeof(long long) * 2 * grid);
    cudaMemcpy(h, a.cycles, sizeof(long long) * 2 * grid, cudaMemcpyDeviceToHost);
    long long m32 = 0, m64 = 0;
    for (int i = 0; i < grid; ++i) {
        m32 = h[2 * i] > m32 ? h[2 * i] : m32;
        m64 = h[2 * i + 1] > m64 ? h[2 * i + 1] : m64;
    }
    free(h);
    const long long mx = m32 > m64 ? m32 : m64;
    // interactions per SM, in warp units (32 lanes)
    const double iw = (double)a.ntiles * TQ * 4 * ((double)N32 * 2 * NP + (double)N64 * R64);
    const double clk = (double)mx * 4.0 / iw; // per sub-partition
    const double total = iw * 32.0 * grid;
    printf("{\"variant\": \"%s\", \"fp32_warps\": %d, \"fp32_i_per_thread\": %d, \"fp64_warps\": %d, "
           "\"fp64_i_per_thread\": %d, \"fp64_share\": %.3f, \"regs\": %d, \"cycles_fp32_warp\": %lld, "
           "\"cycles_fp64_warp\": %lld, \"ms\": %.3f, \"mhz_implied\": %.0f, \"clk_per_interaction_warp\": %.3f, "
           "\"interactions_per_s_at_event_time\": %.4e, \"frac_of_nominal_fp32_roofline\": %.4f}\n",
           name, N32, 2 * NP, N64, R64, (double)N64 * R64 / ((double)N32 * 2 * NP + (double)N64 * R64), fa.numRegs,
           m32, m64, ms, (double)mx / (ms * 1e3), clk, total / (ms * 1e-3), total / (ms * 1e-3) * 20.0 / 74.45e12);
    fflush(stdout);
}

template <int VAR, bool UNI, int NP, int NW, int MINB, int UNROLL>
static void run(const char *name, int ctas_per_sm, int flush, const Args &base, int sms)
{
    Args a = base;
    a.flush = flush;
    const int grid = sms * ctas_per_sm;
    int occ = 0;
    const size_t smem = 2 * TQ * 4 * sizeof(float4) + (size_t)(VAR == 0 ? 12 : 6 * NP) * NW * 32 * sizeof(double);
    cudaFuncSetAttribute(bench_kernel<VAR, UNI, NP, NW, MINB, UNROLL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bench_kernel<VAR, UNI, NP, NW, MINB, UNROLL>, NW * 32, smem);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, bench_kernel<VAR, UNI, NP, NW, MINB, UNROLL>);
    Args w = a;
    w.ntiles = 8;
    bench_kernel<VAR, UNI, NP, NW, MINB, UNROLL><<<grid, NW * 32, smem>>>(w);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench_kernel<VAR, UNI, NP, NW, MINB, UNROLL><<<grid, NW * 32, smem>>>(a);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        printf("{\"variant\": \"%s\", \"error\": \"%s\"}\n", name, cudaGetErrorString(err));
        return;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long *h = (long long *)malloc(sizeof(long long) * grid);
    cudaMemcpy(h, a.cycles, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    free(h);
    const int ibodies = VAR == 0 ? 4 : 2 * NP;
    const double inter_per_thread = (double)a.ntiles * TQ * 4 * ibodies;
    const double warps_per_smsp = (double)ctas_per_sm * NW / 4.0;
    const double clk = (double)mx / (inter_per_thread * warps_per_smsp);
    const double total = inter_per_thread * 32.0 * NW * grid;
    printf("{\"variant\": \"%s\", \"uniform_mass\": %d, \"i_bodies_per_thread\": %d, \"warps_per_cta\": %d, "
           "\"ctas_per_sm\": %d, \"occupancy_limit\": %d, \"regs\": %d, \"unroll\": %d, \"flush_every\": %d, "
           "\"cycles\": %lld, \"ms\": %.3f, \"clk_per_interaction_warp\": %.3f, "
           "\"interactions_per_s_at_event_time\": %.4e, \"frac_of_nominal_roofline\": %.4f}\n",
           name, (int)UNI, ibodies, NW, ctas_per_sm, occ, fa.numRegs, UNROLL, flush, mx, ms, clk,
           total / (ms * 1e-3), total / (ms * 1e-3) * 20.0 / 74.45e12);
    fflush(stdout);
}

int main(int argc, char **argv)
{
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) {
        fprintf(stderr, "no CUDA device\n");
        return 1;
    }
    const int sms = prop.multiProcessorCount;
    const int ntiles = argc > 1 ? atoi(argv[1]) : 2000;
    float4 *src;
    float *out;
    long long *cyc;
    cudaMalloc(&src, sizeof(float4) * TQ * 4);
    cudaMalloc(&out, sizeof(float) * sms * 8 * 1024);
    cudaMalloc(&cyc, sizeof(long long) * sms * 16);
    float4 h[TQ * 4];
    for (int i = 0; i < TQ * 4; ++i) {
        const float b = (float)(i % 97) * 0.37f;
        h[i] = make_float4(b, b + 0.1f, b - 0.2f, b * 0.5f + 1.f);
    }
    cudaMemcpy(src, h, sizeof(h), cudaMemcpyHostToDevice);
    Args a{src, out, cyc, ntiles, 1};
    printf("{\"device\": \"%s\", \"sms\": %d, \"ntiles\": %d}\n", prop.name, sms, ntiles);
    if (argc > 2 && !strcmp(argv[2], "hetero")) {
        //         NP N32 N64 R64 CORRECT
        run_hetero<4, 16, 0, 0, true>("W0 16 FP32 warps x 8i (control = E)", a, sms);
        run_hetero<4, 12, 0, 0, true>("W0 12 FP32 warps x 8i alone", a, sms);
        run_hetero<4, 12, 4, 2, true>("W  12 FP32 x 8i + 4 FP64 x 2i", a, sms);
        run_hetero<4, 12, 4, 4, true>("W  12 FP32 x 8i + 4 FP64 x 4i", a, sms);
        run_hetero<4, 12, 4, 6, true>("W  12 FP32 x 8i + 4 FP64 x 6i", a, sms);
        run_hetero<4, 12, 4, 8, true>("W  12 FP32 x 8i + 4 FP64 x 8i", a, sms);
        run_hetero<4, 12, 4, 6, false>("W  12 FP32 x 8i + 4 FP64 x 6i, seed only (12 FP64 ops)", a, sms);
        run_hetero<3, 12, 4, 4, true>("W  12 FP32 x 6i + 4 FP64 x 4i", a, sms);
        run_hetero<3, 12, 4, 6, true>("W  12 FP32 x 6i + 4 FP64 x 6i", a, sms);
        run_hetero<4, 8, 8, 4, true>("W  8 FP32 x 8i + 8 FP64 x 4i", a, sms);
        run_hetero<4, 8, 4, 6, true>("W  8 FP32 x 8i + 4 FP64 x 6i", a, sms);
        run_hetero<4, 0, 16, 4, true>("W64 16 FP64 warps x 4i alone", a, sms);
        run_hetero<4, 0, 8, 4, true>("W64 8 FP64 warps x 4i alone", a, sms);
        return 0;
    }
    //          VAR UNI  NP NW MINB UNROLL
    run<0, true, 2, 8, 2, 1>("A  j-packed 4i, 2x8 warps (production, equal masses)", 2, 1, a, sms);
    run<0, false, 2, 8, 2, 1>("A  j-packed 4i, 2x8 warps (production, general masses)", 2, 1, a, sms);
    run<0, true, 2, 8, 2, 2>("A2 j-packed 4i, unroll 2", 2, 1, a, sms);
    run<0, true, 2, 8, 2, 1>("A  j-packed 4i, flush every 4 tiles", 2, 4, a, sms);
    run<1, true, 2, 8, 2, 1>("B  i-packed 4i, 2x8 warps, 128 regs", 2, 1, a, sms);
    run<1, true, 2, 8, 4, 1>("B4 i-packed 4i, 4x8 warps, 64 regs", 4, 1, a, sms);
    run<1, true, 2, 8, 3, 1>("B3 i-packed 4i, 3x8 warps, 80 regs", 3, 1, a, sms);
    run<1, true, 2, 8, 2, 2>("B  i-packed 4i, unroll 2", 2, 1, a, sms);
    run<1, true, 4, 8, 2, 1>("C  i-packed 8i, 2x8 warps, 128 regs", 2, 1, a, sms);
    run<1, false, 4, 8, 2, 1>("C  i-packed 8i, 2x8 warps, general masses", 2, 1, a, sms);
    run<1, true, 4, 8, 2, 2>("C2 i-packed 8i, unroll 2", 2, 1, a, sms);
    run<1, true, 4, 8, 2, 1>("C  i-packed 8i, flush every 4 tiles", 2, 4, a, sms);
    run<1, true, 4, 8, 1, 1>("D  i-packed 8i, 1x8 warps, 255 regs", 1, 1, a, sms);
    run<1, true, 4, 8, 1, 2>("D2 i-packed 8i, 1x8 warps, unroll 2", 1, 1, a, sms);
    run<1, true, 4, 16, 1, 1>("E  i-packed 8i, 1x16 warps, 128 regs", 1, 1, a, sms);
    run<1, true, 6, 8, 1, 1>("F  i-packed 12i, 1x8 warps", 1, 1, a, sms);
    run<1, true, 8, 8, 1, 1>("G  i-packed 16i, 1x8 warps", 1, 1, a, sms);
    run<1, true, 3, 8, 2, 1>("H  i-packed 6i, 2x8 warps", 2, 1, a, sms);
    run<2, true, 2, 8, 2, 1>("S  i-packed 4i, j by warp shuffle (north star)", 2, 1, a, sms);
    run<2, true, 4, 8, 2, 1>("S8 i-packed 8i, j by warp shuffle", 2, 1, a, sms);
    run<2, true, 4, 8, 2, 4>("S8 i-packed 8i, j by warp shuffle, unroll 4", 2, 1, a, sms);
    return 0;
}
