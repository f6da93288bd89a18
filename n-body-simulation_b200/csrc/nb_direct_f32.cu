// nb_direct_f32.cu -- the FP32 O(N^2) force + leapfrog kernel for sm_100a.
//
// Replaces acceleration<<<>>> + update_velocities<<<>>> + update_positions<<<>>>
// of the reference (cuda/exhaustive.cu:273-388; math of serial/exhaustive.c:155-181
// and :218-250) with ONE launch per step:
//   * bodies live in HBM as 64-byte quads {X4,Y4,Z4,M4}; a thread owns one quad
//     = a register block of 4 i-bodies (3 x 128-bit loads);
//   * j-bodies stream through a 4-stage shared-memory ring filled by 1-D bulk
//     async copies (TMA, SASS UBLKCP) and handed over with mbarriers; the copy
//     for tile k+2 is issued by one thread at the top of tile k (a dedicated
//     producer warp would leave one of the four SM sub-partitions with fewer
//     compute warps: measured 12 % loss, profiles/r1_notes.md); every warp
//     reads a j-quad with four broadcast LDS.128;
//   * the inner loop works on PAIRS of j-bodies with the sm_100 packed FP32
//     instructions (FADD2/FFMA2/FMUL2): per 2 interactions 3 FADD2 + 6 FFMA2 +
//     3 FMUL2 + 2 MUFU.RSQ, i.e. 12 FMA-pipe lane-operations and one MUFU per
//     interaction in 7 issue slots;
//   * FP32 partial sums are flushed into FP64 running sums once per tile
//     (256 j-bodies), so the 1e6-term sums keep ~1e-7 relative accuracy;
//   * the (i-block x j-tile) work list is cut evenly over a persistent grid
//     (nb_common.cuh WorkSplit), the last CTA to finish an i-block adds the
//     partial slots in fixed order and runs the kick/drift epilogue, so
//     positions and velocities make one HBM round trip per step.
#include "nb_direct.cuh"

namespace nb {
namespace {

constexpr int STAGES = 4;
#ifndef NB_F32_UNROLL
#define NB_F32_UNROLL 1 // j-quads per inner-loop trip; 1 measured fastest (63.97 vs 65.63 ms at 400k bodies for 2, 65.76 for 4)
#endif
constexpr int kUnrollJ = NB_F32_UNROLL;

__device__ __forceinline__ void named_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// One j-pair against the thread's four i-bodies, packed f32x2 arithmetic.
// nxi/nyi/nzi hold -x_i etc. (ptxas folds the negation and the scalar
// broadcast into the FADD2 operand: "FADD2 R, Rj.F32x2.HI_LO, -Ri.F32").
// UNI (all masses equal): the mass multiply leaves the loop -- a_i = m * sum_j d r^-3 -- and is
// applied once in FP64 in the epilogue: 11 instead of 12 FMA-pipe operations per interaction.
template <bool UNI>
__device__ __forceinline__ void pair_packed(float2 xj, float2 yj, float2 zj, float2 mj,
                                            const float4 &XI, const float4 &YI, const float4 &ZI,
                                            float2 e2, float2 (&ax)[4], float2 (&ay)[4],
                                            float2 (&az)[4])
{
    const float xi[4] = {XI.x, XI.y, XI.z, XI.w};
    const float yi[4] = {YI.x, YI.y, YI.z, YI.w};
    const float zi[4] = {ZI.x, ZI.y, ZI.z, ZI.w};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float2 dx = __fadd2_rn(xj, make_float2(-xi[r], -xi[r]));
        float2 dy = __fadd2_rn(yj, make_float2(-yi[r], -yi[r]));
        float2 dz = __fadd2_rn(zj, make_float2(-zi[r], -zi[r]));
        float2 r2 = __ffma2_rn(dx, dx, e2);
        r2 = __ffma2_rn(dy, dy, r2);
        r2 = __ffma2_rn(dz, dz, r2);
        float2 ri = make_float2(rsqrt_approx(r2.x), rsqrt_approx(r2.y));
        float2 ri2 = __fmul2_rn(ri, ri);
        float2 s = UNI ? __fmul2_rn(ri2, ri) : __fmul2_rn(ri2, __fmul2_rn(mj, ri));
        ax[r] = __ffma2_rn(dx, s, ax[r]);
        ay[r] = __ffma2_rn(dy, s, ay[r]);
        az[r] = __ffma2_rn(dz, s, az[r]);
    }
}

// Scalar twin (NB_TUNE_SCALAR_FP32), for A/B measurements of the packed path.
__device__ __forceinline__ void quad_scalar(const float4 &JX, const float4 &JY, const float4 &JZ,
                                            const float4 &JM, const float4 &XI, const float4 &YI,
                                            const float4 &ZI, float e2, float2 (&ax)[4],
                                            float2 (&ay)[4], float2 (&az)[4])
{
    const float xi[4] = {XI.x, XI.y, XI.z, XI.w};
    const float yi[4] = {YI.x, YI.y, YI.z, YI.w};
    const float zi[4] = {ZI.x, ZI.y, ZI.z, ZI.w};
    const float xj[4] = {JX.x, JX.y, JX.z, JX.w};
    const float yj[4] = {JY.x, JY.y, JY.z, JY.w};
    const float zj[4] = {JZ.x, JZ.y, JZ.z, JZ.w};
    const float mj[4] = {JM.x, JM.y, JM.z, JM.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            float dx = xj[j] - xi[r], dy = yj[j] - yi[r], dz = zj[j] - zi[r];
            float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, e2)));
            float ri = rsqrt_approx(r2);
            float s = (ri * ri) * (mj[j] * ri);
            ax[r].x = fmaf(dx, s, ax[r].x);
            ay[r].x = fmaf(dy, s, ay[r].x);
            az[r].x = fmaf(dz, s, az[r].x);
        }
    }
}

// kick / drift on one quad component, written with explicit roundings so the
// operation order is the reference's: v += (a*dt)/2 ; x += v*dt
// (serial/exhaustive.c:221-232).
__device__ __forceinline__ float kick(float v, float a, float dt)
{
    return __fadd_rn(v, __fmul_rn(__fmul_rn(a, dt), 0.5f));
}
__device__ __forceinline__ float drift(float x, float v, float dt)
{
    return __fadd_rn(x, __fmul_rn(v, dt));
}

// MODE: 0 scalar FP32 (A/B only), 1 packed f32x2, 2 packed + uniform mass
template <int NW, int TQ, int MODE, int MINB>
__global__ void __launch_bounds__(NW * 32, MINB)
    direct_f32_kernel(const StepParams p)
{
    constexpr int QPB = NW * 32; // quads (threads) per i-block
    __shared__ __align__(128) float4 tiles[STAGES][TQ * 4];
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ unsigned s_last;
    // FP64 running sums live in shared memory (12 per thread, touched once per
    // tile) so the inner loop keeps its registers for the 4x2 interaction block.
    __shared__ double dacc[12][QPB];

    const int lane = threadIdx.x & 31;
    const float4 *__restrict__ q = static_cast<const float4 *>(p.q_cur);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], NW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const unsigned long long w0 = p.ws.bound(blockIdx.x);
    const unsigned long long w1 = p.ws.bound(blockIdx.x + 1);
    const unsigned n_jtiles = p.ws.n_jtiles;

    // -------- j-tile ring: tile k of this CTA's list lives in stage k % STAGES ----
    // Thread 0 issues the copy of tile k + AHEAD at the top of tile k, into the
    // stage all warps released two tiles ago, so it never waits on a laggard.
    constexpr unsigned AHEAD = STAGES - 2;
    const unsigned n_tiles = (unsigned)(w1 - w0);
    const unsigned tile0 = (unsigned)(w0 % n_jtiles);
    auto global_tile = [&](unsigned t) {
        const unsigned g = p.jt_begin + t;
        return g >= p.jt_total ? g - p.jt_total : g;
    };
    auto issue_tile = [&](unsigned k) {
        const unsigned s = k % STAGES;
        if (k >= STAGES) mbar_wait(&empty_bar[s], ((k / STAGES) - 1u) & 1u);
        const unsigned first = global_tile((tile0 + k) % n_jtiles) * TQ;
        const unsigned nq = min((unsigned)TQ, p.n_units - first);
        const unsigned bytes = nq * 64u;
        mbar_arrive_expect_tx(&full_bar[s], bytes);
        tma_load_1d(&tiles[s][0], q + (size_t)first * 4, bytes, &full_bar[s]);
    };
    if (threadIdx.x == 0)
        for (unsigned k = 0; k < AHEAD && k < n_tiles; ++k) issue_tile(k);

    const int tidc = threadIdx.x; // 0 .. QPB-1
    const float dtf = (float)p.dt;
    const float e2s = (float)p.eps2;
    const float2 e2 = make_float2(e2s, e2s);
    unsigned stage = 0, phase = 0, kt = 0;

    unsigned long long w = w0;
    while (w < w1) {
        const unsigned ib = (unsigned)(w / n_jtiles);
        const unsigned t0 = (unsigned)(w - (unsigned long long)ib * n_jtiles);
        const unsigned cnt = (unsigned)min((unsigned long long)(n_jtiles - t0), w1 - w);
        const unsigned iq = p.i_begin + ib * QPB + tidc;
        const bool valid = iq < p.i_end;
        const unsigned iqc = valid ? iq : p.i_begin;
        const float4 XI = q[(size_t)iqc * 4 + 0];
        const float4 YI = q[(size_t)iqc * 4 + 1];
        const float4 ZI = q[(size_t)iqc * 4 + 2];

#pragma unroll
        for (int c = 0; c < 12; ++c) dacc[c][tidc] = 0.0;

        for (unsigned k = 0; k < cnt; ++k) {
            const unsigned first = global_tile(t0 + k) * TQ;
            const int nq = (int)min((unsigned)TQ, p.n_units - first);
            float2 ax[4], ay[4], az[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) ax[r] = ay[r] = az[r] = make_float2(0.f, 0.f);

            if (threadIdx.x == 0 && kt + AHEAD < n_tiles) issue_tile(kt + AHEAD);
            ++kt;
            mbar_wait(&full_bar[stage], phase);
            const float4 *__restrict__ tile = &tiles[stage][0];
#pragma unroll kUnrollJ
            for (int jq = 0; jq < nq; ++jq) {
                const float4 JX = tile[jq * 4 + 0];
                const float4 JY = tile[jq * 4 + 1];
                const float4 JZ = tile[jq * 4 + 2];
                const float4 JM = tile[jq * 4 + 3];
                if (MODE != 0) {
                    pair_packed<MODE == 2>(make_float2(JX.x, JX.y), make_float2(JY.x, JY.y),
                                           make_float2(JZ.x, JZ.y), make_float2(JM.x, JM.y), XI, YI,
                                           ZI, e2, ax, ay, az);
                    pair_packed<MODE == 2>(make_float2(JX.z, JX.w), make_float2(JY.z, JY.w),
                                           make_float2(JZ.z, JZ.w), make_float2(JM.z, JM.w), XI, YI,
                                           ZI, e2, ax, ay, az);
                } else {
                    quad_scalar(JX, JY, JZ, JM, XI, YI, ZI, e2s, ax, ay, az);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            // flush the tile's FP32 partials into the FP64 running sums
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                dacc[0 + r][tidc] += (double)(ax[r].x + ax[r].y);
                dacc[4 + r][tidc] += (double)(ay[r].x + ay[r].y);
                dacc[8 + r][tidc] += (double)(az[r].x + az[r].y);
            }
        }
        w += cnt;
        double dax[4], day[4], daz[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            dax[r] = dacc[0 + r][tidc];
            day[r] = dacc[4 + r][tidc];
            daz[r] = dacc[8 + r][tidc];
        }

        // ---------------- combine partial j-ranges of this i-block ----------------
        bool do_epilogue = true;
        if (cnt != n_jtiles) {
            const unsigned long long x0 = (unsigned long long)ib * n_jtiles;
            const unsigned c_first = p.ws.owner(x0);
            const unsigned c_last = p.ws.owner(x0 + n_jtiles - 1);
            const unsigned nslots = c_last - c_first + 1;
            const unsigned slot = blockIdx.x - c_first;
            double2 *ws = reinterpret_cast<double2 *>(p.workspace) +
                          ((size_t)ib * p.max_slots) * (size_t)(QPB * 6);
            double2 *mine = ws + (size_t)slot * (QPB * 6);
            mine[0 * QPB + tidc] = make_double2(dax[0], dax[1]);
            mine[1 * QPB + tidc] = make_double2(dax[2], dax[3]);
            mine[2 * QPB + tidc] = make_double2(day[0], day[1]);
            mine[3 * QPB + tidc] = make_double2(day[2], day[3]);
            mine[4 * QPB + tidc] = make_double2(daz[0], daz[1]);
            mine[5 * QPB + tidc] = make_double2(daz[2], daz[3]);
            __threadfence();
            named_sync(1, QPB);
            if (tidc == 0) {
                const unsigned old = atomicAdd(&p.counters[ib], 1u);
                const unsigned last = (old == nslots - 1);
                if (last) p.counters[ib] = 0; // ready for the next launch
                s_last = last;
            }
            named_sync(1, QPB);
            do_epilogue = (s_last != 0);
            if (do_epilogue) {
                __threadfence();
#pragma unroll
                for (int r = 0; r < 4; ++r) dax[r] = day[r] = daz[r] = 0.0;
                for (unsigned s = 0; s < nslots; ++s) {
                    const double2 *src = ws + (size_t)s * (QPB * 6);
                    double2 v;
                    v = __ldcg(&src[0 * QPB + tidc]); dax[0] += v.x; dax[1] += v.y;
                    v = __ldcg(&src[1 * QPB + tidc]); dax[2] += v.x; dax[3] += v.y;
                    v = __ldcg(&src[2 * QPB + tidc]); day[0] += v.x; day[1] += v.y;
                    v = __ldcg(&src[3 * QPB + tidc]); day[2] += v.x; day[3] += v.y;
                    v = __ldcg(&src[4 * QPB + tidc]); daz[0] += v.x; daz[1] += v.y;
                    v = __ldcg(&src[5 * QPB + tidc]); daz[2] += v.x; daz[3] += v.y;
                }
            }
            named_sync(1, QPB); // s_last may be rewritten by the next i-block
        }

        // ---------------- multi-rank: sums parked between the two j-ranges --------
        if (do_epilogue && (p.flags & (NB_F_PARTIAL_IN | NB_F_PARTIAL_OUT))) {
            double2 *park = reinterpret_cast<double2 *>(p.partial) + (size_t)ib * (QPB * 6);
            if (p.flags & NB_F_PARTIAL_IN) {
                double2 v;
                v = park[0 * QPB + tidc]; dax[0] += v.x; dax[1] += v.y;
                v = park[1 * QPB + tidc]; dax[2] += v.x; dax[3] += v.y;
                v = park[2 * QPB + tidc]; day[0] += v.x; day[1] += v.y;
                v = park[3 * QPB + tidc]; day[2] += v.x; day[3] += v.y;
                v = park[4 * QPB + tidc]; daz[0] += v.x; daz[1] += v.y;
                v = park[5 * QPB + tidc]; daz[2] += v.x; daz[3] += v.y;
            } else {
                park[0 * QPB + tidc] = make_double2(dax[0], dax[1]);
                park[1 * QPB + tidc] = make_double2(dax[2], dax[3]);
                park[2 * QPB + tidc] = make_double2(day[0], day[1]);
                park[3 * QPB + tidc] = make_double2(day[2], day[3]);
                park[4 * QPB + tidc] = make_double2(daz[0], daz[1]);
                park[5 * QPB + tidc] = make_double2(daz[2], daz[3]);
                do_epilogue = false;
            }
        }

        // ---------------- epilogue: kick2 / kick1 / drift, one HBM round trip ------
        if (do_epilogue && valid) {
            float a[3][4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                a[0][r] = (float)(p.G * dax[r]);
                a[1][r] = (float)(p.G * day[r]);
                a[2][r] = (float)(p.G * daz[r]);
            }
            if (p.flags & NB_F_STORE_ACC) {
                float4 *acc = static_cast<float4 *>(p.acc);
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    acc[(size_t)iq * 3 + c] = make_float4(a[c][0], a[c][1], a[c][2], a[c][3]);
            }
            if (p.flags & (NB_F_KICK2 | NB_F_KICK1_DRIFT)) {
                float4 *vel = static_cast<float4 *>(p.vel);
                float4 *qn = static_cast<float4 *>(p.q_next);
                const float xin[3][4] = {{XI.x, XI.y, XI.z, XI.w},
                                         {YI.x, YI.y, YI.z, YI.w},
                                         {ZI.x, ZI.y, ZI.z, ZI.w}};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float4 V = vel[(size_t)iq * 3 + c];
                    float v[4] = {V.x, V.y, V.z, V.w};
                    float x[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const bool real = (iq * 4u + r) < p.n_real;
                        if (p.flags & NB_F_KICK2) v[r] = kick(v[r], a[c][r], dtf);
                        if (p.flags & NB_F_KICK1_DRIFT) v[r] = kick(v[r], a[c][r], dtf);
                        x[r] = (p.flags & NB_F_KICK1_DRIFT) ? drift(xin[c][r], v[r], dtf)
                                                            : xin[c][r];
                        if (!real) { v[r] = 0.f; x[r] = xin[c][r]; }
                    }
                    vel[(size_t)iq * 3 + c] = make_float4(v[0], v[1], v[2], v[3]);
                    if (p.flags & NB_F_KICK1_DRIFT)
                        qn[(size_t)iq * 4 + c] = make_float4(x[0], x[1], x[2], x[3]);
                }
            }
        }
    }
}


// ======================= i-packed kernel (round 2, the large size class) =======================
// The packed lanes hold two I-bodies instead of two j-bodies; the j-body is a broadcast scalar
// operand (SASS: "FADD2 R, -Ri.F32x2.HI_LO, Rj.F32", eps2 as an immediate of the first FFMA2).
// Accumulators are then per i-body only -- half as many registers as the j-packed form -- so a
// thread carries QPT = 2 quads = 8 i-bodies in 128 registers: half the LDS and loop overhead per
// interaction, four independent dependency chains per j.  Measured with tools/innerloop_bench.cu
// (profiles/r2_innerloop_variants_1.jsonl): 12.6-12.9 clk per interaction-warp against 13.2 for
// the j-packed loop; the same file shows the north star's warp-shuffle broadcast of j-bodies at
// 13.4-15.8 clk (SHFL fan-out costs issue slots the loop does not have).
template <bool UNI, int NP>
__device__ __forceinline__ void one_j_ipacked(float xj, float yj, float zj, float mj,
                                              const float2 (&xi)[NP], const float2 (&yi)[NP],
                                              const float2 (&zi)[NP], float e2s, float2 (&ax)[NP],
                                              float2 (&ay)[NP], float2 (&az)[NP])
{
#pragma unroll
    for (int p = 0; p < NP; ++p) {
        const float2 dx = __fadd2_rn(make_float2(xj, xj), make_float2(-xi[p].x, -xi[p].y));
        const float2 dy = __fadd2_rn(make_float2(yj, yj), make_float2(-yi[p].x, -yi[p].y));
        const float2 dz = __fadd2_rn(make_float2(zj, zj), make_float2(-zi[p].x, -zi[p].y));
        float2 r2 = __ffma2_rn(dx, dx, make_float2(e2s, e2s));
        r2 = __ffma2_rn(dy, dy, r2);
        r2 = __ffma2_rn(dz, dz, r2);
        const float2 ri = make_float2(rsqrt_approx(r2.x), rsqrt_approx(r2.y));
        const float2 ri2 = __fmul2_rn(ri, ri);
        const float2 s = UNI ? __fmul2_rn(ri2, ri) : __fmul2_rn(ri2, __fmul2_rn(make_float2(mj, mj), ri));
        ax[p] = __ffma2_rn(dx, s, ax[p]);
        ay[p] = __ffma2_rn(dy, s, ay[p]);
        az[p] = __ffma2_rn(dz, s, az[p]);
    }
}

// MODE: 1 general masses, 2 uniform mass.  QPT quads (4 i-bodies each) per thread.
template <int NW, int TQ, int MODE, int MINB, int QPT>
__global__ void __launch_bounds__(NW * 32, MINB)
    direct_f32_ip_kernel(const StepParams p)
{
    constexpr int NT = NW * 32;      // threads
    constexpr int NP = 2 * QPT;      // i-pairs per thread
    constexpr int QPB = NT * QPT;    // quads per i-block
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4(*tiles)[TQ * 4] = reinterpret_cast<float4(*)[TQ * 4]>(smem_raw);
    // FP64 running sums, 6 per i-pair, touched once per tile
    double(*dacc)[NT] = reinterpret_cast<double(*)[NT]>(smem_raw + (size_t)STAGES * TQ * 4 * sizeof(float4));
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ unsigned s_last;

    const int lane = threadIdx.x & 31;
    const float4 *__restrict__ q = static_cast<const float4 *>(p.q_cur);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], NW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const unsigned long long w0 = p.ws.bound(blockIdx.x);
    const unsigned long long w1 = p.ws.bound(blockIdx.x + 1);
    const unsigned n_jtiles = p.ws.n_jtiles;

    // j-tile ring fed by thread 0, two tiles ahead (as in direct_f32_kernel above)
    constexpr unsigned AHEAD = STAGES - 2;
    const unsigned n_tiles = (unsigned)(w1 - w0);
    const unsigned tile0 = (unsigned)(w0 % n_jtiles);
    auto global_tile = [&](unsigned t) {
        const unsigned g = p.jt_begin + t;
        return g >= p.jt_total ? g - p.jt_total : g;
    };
    auto issue_tile = [&](unsigned k) {
        const unsigned s = k % STAGES;
        if (k >= STAGES) mbar_wait(&empty_bar[s], ((k / STAGES) - 1u) & 1u);
        const unsigned first = global_tile((tile0 + k) % n_jtiles) * TQ;
        const unsigned nq = min((unsigned)TQ, p.n_units - first);
        const unsigned bytes = nq * 64u;
        mbar_arrive_expect_tx(&full_bar[s], bytes);
        tma_load_1d(&tiles[s][0], q + (size_t)first * 4, bytes, &full_bar[s]);
    };
    if (threadIdx.x == 0)
        for (unsigned k = 0; k < AHEAD && k < n_tiles; ++k) issue_tile(k);

    const int tidc = threadIdx.x;
    const float dtf = (float)p.dt;
    const float e2s = (float)p.eps2;
    unsigned stage = 0, phase = 0, kt = 0;

    unsigned long long w = w0;
    while (w < w1) {
        const unsigned ib = (unsigned)(w / n_jtiles);
        const unsigned t0 = (unsigned)(w - (unsigned long long)ib * n_jtiles);
        const unsigned cnt = (unsigned)min((unsigned long long)(n_jtiles - t0), w1 - w);
        // the thread's QPT quads; pairs 2u, 2u+1 are (x0,x1), (x2,x3) of quad u
        unsigned iq[QPT];
        bool valid[QPT];
        float2 xi[NP], yi[NP], zi[NP];
#pragma unroll
        for (int u = 0; u < QPT; ++u) {
            iq[u] = p.i_begin + ib * QPB + u * NT + tidc; // coalesced: quad u of thread t
            valid[u] = iq[u] < p.i_end;
            const unsigned iqc = valid[u] ? iq[u] : p.i_begin;
            const float4 XI = q[(size_t)iqc * 4 + 0];
            const float4 YI = q[(size_t)iqc * 4 + 1];
            const float4 ZI = q[(size_t)iqc * 4 + 2];
            xi[2 * u] = make_float2(XI.x, XI.y); xi[2 * u + 1] = make_float2(XI.z, XI.w);
            yi[2 * u] = make_float2(YI.x, YI.y); yi[2 * u + 1] = make_float2(YI.z, YI.w);
            zi[2 * u] = make_float2(ZI.x, ZI.y); zi[2 * u + 1] = make_float2(ZI.z, ZI.w);
        }
#pragma unroll
        for (int c = 0; c < 6 * NP; ++c) dacc[c][tidc] = 0.0;

        for (unsigned k = 0; k < cnt; ++k) {
            const unsigned first = global_tile(t0 + k) * TQ;
            const int nq = (int)min((unsigned)TQ, p.n_units - first);
            float2 ax[NP], ay[NP], az[NP];
#pragma unroll
            for (int r = 0; r < NP; ++r) ax[r] = ay[r] = az[r] = make_float2(0.f, 0.f);

            if (threadIdx.x == 0 && kt + AHEAD < n_tiles) issue_tile(kt + AHEAD);
            ++kt;
            mbar_wait(&full_bar[stage], phase);
            const float4 *__restrict__ tile = &tiles[stage][0];
#pragma unroll 1
            for (int jq = 0; jq < nq; ++jq) {
                const float4 JX = tile[jq * 4 + 0];
                const float4 JY = tile[jq * 4 + 1];
                const float4 JZ = tile[jq * 4 + 2];
                float4 JM = make_float4(0.f, 0.f, 0.f, 0.f);
                if (MODE != 2) JM = tile[jq * 4 + 3];
                one_j_ipacked<MODE == 2, NP>(JX.x, JY.x, JZ.x, JM.x, xi, yi, zi, e2s, ax, ay, az);
                one_j_ipacked<MODE == 2, NP>(JX.y, JY.y, JZ.y, JM.y, xi, yi, zi, e2s, ax, ay, az);
                one_j_ipacked<MODE == 2, NP>(JX.z, JY.z, JZ.z, JM.z, xi, yi, zi, e2s, ax, ay, az);
                one_j_ipacked<MODE == 2, NP>(JX.w, JY.w, JZ.w, JM.w, xi, yi, zi, e2s, ax, ay, az);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            // flush the tile's FP32 partials into the FP64 running sums
#pragma unroll
            for (int r = 0; r < NP; ++r) {
                dacc[6 * r + 0][tidc] += (double)ax[r].x;
                dacc[6 * r + 1][tidc] += (double)ax[r].y;
                dacc[6 * r + 2][tidc] += (double)ay[r].x;
                dacc[6 * r + 3][tidc] += (double)ay[r].y;
                dacc[6 * r + 4][tidc] += (double)az[r].x;
                dacc[6 * r + 5][tidc] += (double)az[r].y;
            }
        }
        w += cnt;

        // From here on one quad at a time (u): sums of quad u = pairs 2u, 2u+1.
        // Workspace / parking layout per i-block: [slot][u][6][NT] double2 -- plane c of quad u
        // holds (dax0,dax1),(dax2,dax3),(day0,day1),... like the j-packed kernel's per-quad planes.
        bool do_epilogue = true;
        unsigned nslots = 1;
        double2 *ws = nullptr;
        if (cnt != n_jtiles) {
            const unsigned long long x0 = (unsigned long long)ib * n_jtiles;
            const unsigned c_first = p.ws.owner(x0);
            const unsigned c_last = p.ws.owner(x0 + n_jtiles - 1);
            nslots = c_last - c_first + 1;
            const unsigned slot = blockIdx.x - c_first;
            ws = reinterpret_cast<double2 *>(p.workspace) + ((size_t)ib * p.max_slots) * (size_t)(QPB * 6);
            double2 *mine = ws + (size_t)slot * (QPB * 6);
#pragma unroll
            for (int u = 0; u < QPT; ++u)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    // component c of pairs 2u (bodies 0,1) and 2u+1 (bodies 2,3)
                    mine[(u * 6 + 2 * c + 0) * NT + tidc] =
                        make_double2(dacc[12 * u + 2 * c][tidc], dacc[12 * u + 2 * c + 1][tidc]);
                    mine[(u * 6 + 2 * c + 1) * NT + tidc] =
                        make_double2(dacc[12 * u + 6 + 2 * c][tidc], dacc[12 * u + 6 + 2 * c + 1][tidc]);
                }
            __threadfence();
            named_sync(1, NT);
            if (tidc == 0) {
                const unsigned old = atomicAdd(&p.counters[ib], 1u);
                const unsigned last = (old == nslots - 1);
                if (last) p.counters[ib] = 0; // ready for the next launch
                s_last = last;
            }
            named_sync(1, NT);
            do_epilogue = (s_last != 0);
            if (do_epilogue) __threadfence();
            named_sync(1, NT); // s_last may be rewritten by the next i-block
        }
        if (!do_epilogue) continue;

        double2 *park = (p.flags & (NB_F_PARTIAL_IN | NB_F_PARTIAL_OUT))
                            ? reinterpret_cast<double2 *>(p.partial) + (size_t)ib * (QPB * 6)
                            : nullptr;
#pragma unroll
        for (int u = 0; u < QPT; ++u) {
            // d[c][r]: component c (x,y,z) of body r of quad u
            double d[3][4];
            if (cnt != n_jtiles) { // add the partial j-ranges in slot order (deterministic)
#pragma unroll
                for (int c = 0; c < 3; ++c) d[c][0] = d[c][1] = d[c][2] = d[c][3] = 0.0;
                for (unsigned s = 0; s < nslots; ++s) {
                    const double2 *src = ws + (size_t)s * (QPB * 6);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const double2 v0 = __ldcg(&src[(u * 6 + 2 * c + 0) * NT + tidc]);
                        const double2 v1 = __ldcg(&src[(u * 6 + 2 * c + 1) * NT + tidc]);
                        d[c][0] += v0.x; d[c][1] += v0.y; d[c][2] += v1.x; d[c][3] += v1.y;
                    }
                }
            } else {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    d[c][0] = dacc[12 * u + 2 * c][tidc];
                    d[c][1] = dacc[12 * u + 2 * c + 1][tidc];
                    d[c][2] = dacc[12 * u + 6 + 2 * c][tidc];
                    d[c][3] = dacc[12 * u + 6 + 2 * c + 1][tidc];
                }
            }
            // multi-rank: sums parked between the own-slice and the gathered j-ranges
            if (park) {
                if (p.flags & NB_F_PARTIAL_IN) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const double2 v0 = park[(u * 6 + 2 * c + 0) * NT + tidc];
                        const double2 v1 = park[(u * 6 + 2 * c + 1) * NT + tidc];
                        d[c][0] += v0.x; d[c][1] += v0.y; d[c][2] += v1.x; d[c][3] += v1.y;
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        park[(u * 6 + 2 * c + 0) * NT + tidc] = make_double2(d[c][0], d[c][1]);
                        park[(u * 6 + 2 * c + 1) * NT + tidc] = make_double2(d[c][2], d[c][3]);
                    }
                    continue; // no epilogue in a PARTIAL_OUT launch
                }
            }
            if (!valid[u]) continue;
            // ---------------- epilogue: kick2 / kick1 / drift, one HBM round trip ------
            float a[3][4];
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int r = 0; r < 4; ++r) a[c][r] = (float)(p.G * d[c][r]);
            const unsigned iqu = iq[u];
            if (p.flags & NB_F_STORE_ACC) {
                float4 *acc = static_cast<float4 *>(p.acc);
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    acc[(size_t)iqu * 3 + c] = make_float4(a[c][0], a[c][1], a[c][2], a[c][3]);
            }
            if (p.flags & (NB_F_KICK2 | NB_F_KICK1_DRIFT)) {
                float4 *vel = static_cast<float4 *>(p.vel);
                float4 *qn = static_cast<float4 *>(p.q_next);
                const float xin[3][4] = {{xi[2 * u].x, xi[2 * u].y, xi[2 * u + 1].x, xi[2 * u + 1].y},
                                         {yi[2 * u].x, yi[2 * u].y, yi[2 * u + 1].x, yi[2 * u + 1].y},
                                         {zi[2 * u].x, zi[2 * u].y, zi[2 * u + 1].x, zi[2 * u + 1].y}};
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float4 V = vel[(size_t)iqu * 3 + c];
                    float v[4] = {V.x, V.y, V.z, V.w};
                    float x[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const bool real = (iqu * 4u + r) < p.n_real;
                        if (p.flags & NB_F_KICK2) v[r] = kick(v[r], a[c][r], dtf);
                        if (p.flags & NB_F_KICK1_DRIFT) v[r] = kick(v[r], a[c][r], dtf);
                        x[r] = (p.flags & NB_F_KICK1_DRIFT) ? drift(xin[c][r], v[r], dtf) : xin[c][r];
                        if (!real) { v[r] = 0.f; x[r] = xin[c][r]; }
                    }
                    vel[(size_t)iqu * 3 + c] = make_float4(v[0], v[1], v[2], v[3]);
                    if (p.flags & NB_F_KICK1_DRIFT)
                        qn[(size_t)iqu * 4 + c] = make_float4(x[0], x[1], x[2], x[3]);
                }
            }
        }
    }
}

} // namespace

// ------------------------------------------------------------------ host ----
static void fill_geom(const DirectPlan &plan, LaunchGeom *g, int tune_flags)
{
    const unsigned long long total = (unsigned long long)plan.n_iblocks * g->n_jtiles;
    g->grid = (tune_flags & 2 /* NB_TUNE_NO_SPLIT */) ? plan.n_iblocks : plan.grid;
    if ((unsigned long long)g->grid > total) g->grid = (unsigned)total;
    if (g->grid == 0) g->grid = 1;
    WorkSplit ws{total, plan.n_iblocks, g->n_jtiles, g->grid};
    g->max_slots = 1;
    for (unsigned ib = 0; ib < plan.n_iblocks && total; ++ib) {
        const unsigned long long x0 = (unsigned long long)ib * g->n_jtiles;
        const unsigned s = ws.owner(x0 + g->n_jtiles - 1) - ws.owner(x0) + 1;
        if (s > g->max_slots) g->max_slots = s;
    }
}

void finish_plan(DirectPlan *plan, const SliceGeom &sg, int tune_flags)
{
    const unsigned T = plan->units_per_tile;
    plan->geom[NB_GEOM_FULL] = LaunchGeom{plan->n_jtiles, 0, 0, 1};
    plan->can_split = sg.own_units < sg.n_units && sg.own_units % T == 0 && sg.own_begin % T == 0;
    if (plan->can_split) {
        const unsigned own_tiles = sg.own_units / T, first = sg.own_begin / T;
        plan->geom[NB_GEOM_OWN] = LaunchGeom{own_tiles, first, 0, 1};
        plan->geom[NB_GEOM_REST] =
            LaunchGeom{plan->n_jtiles - own_tiles, (first + own_tiles) % plan->n_jtiles, 0, 1};
    } else {
        plan->geom[NB_GEOM_OWN] = plan->geom[NB_GEOM_REST] = plan->geom[NB_GEOM_FULL];
    }
    unsigned max_slots = 1;
    for (int k = 0; k < 3; ++k) {
        fill_geom(*plan, &plan->geom[k], tune_flags);
        if (plan->geom[k].max_slots > max_slots) max_slots = plan->geom[k].max_slots;
    }
    for (int k = 0; k < 3; ++k) plan->geom[k].max_slots = max_slots; // one workspace stride
    plan->workspace_bytes =
        (size_t)plan->n_iblocks * max_slots * plan->slot_doubles * sizeof(double);
    plan->partial_bytes =
        plan->can_split ? (size_t)plan->n_iblocks * plan->slot_doubles * sizeof(double) : 0;
    plan->counter_bytes = (size_t)plan->n_iblocks * sizeof(unsigned);
}

void apply_geom(const DirectPlan &plan, int geom, StepParams *p)
{
    const LaunchGeom &g = plan.geom[geom];
    p->max_slots = g.max_slots;
    p->jt_begin = g.jt_begin;
    p->jt_total = plan.n_jtiles;
    p->ws = WorkSplit{(unsigned long long)plan.n_iblocks * g.n_jtiles, plan.n_iblocks, g.n_jtiles,
                      g.grid};
}

namespace {
// Size classes.  i-block = NW*32*QPT quads, j-tile = TQ*4 bodies, MINB resident CTAs per SM asked
// of ptxas (register cap).  QPT = 0: the j-packed kernel (one quad per thread); QPT = 2: the
// i-packed kernel (two quads = 8 i-bodies per thread, dynamic shared memory for the FP64 sums).
// tune_flags bits 8..11 = variant + 1 forces a class (A/B runs).
struct VariantF32 {
    int nw, tq, minb, qpt;
};
constexpr int kNumVariants = 13;
constexpr VariantF32 kVariants[kNumVariants] = {
    {8, 64, 2, 0},  // 0 j-packed, 2 x 8 warps per SM, <=128 regs (round 1's large class)
    {4, 64, 4, 0},  // 1 medium: 4 x 4 warps per SM
    {1, 16, 8, 0},  // 2 small : latency regime
    {16, 64, 1, 2}, // 3 i-packed 8i, 1 x 16 warps, <=128 regs
    {12, 64, 1, 2}, // 4 i-packed 8i, 1 x 12 warps, <=168 regs
    {8, 64, 2, 2},  // 5 i-packed 8i, 2 x 8 warps, <=128 regs
    {8, 64, 1, 2},  // 6 i-packed 8i, 1 x 8 warps, <=255 regs
    {8, 64, 1, 3},  // 7 i-packed 12i, 1 x 8 warps, <=255 regs
    {4, 64, 2, 2},  // 8 i-packed 8i, 2 x 4 warps, <=255 regs
    {8, 128, 1, 2}, // 9 variant 6 with j-tiles of 128 quads (half the flushes into the FP64 sums)
    {8, 32, 1, 2},  // 10 ... of 32 quads
    {8, 256, 1, 2}, // 11 ... of 256 quads
    {8, 192, 1, 2}, // 12 ... of 192 quads
};
#ifndef NB_F32_LARGE_VARIANT
#define NB_F32_LARGE_VARIANT 6
#endif

constexpr size_t ip_smem_bytes(int nw, int tq, int qpt)
{
    return (size_t)STAGES * tq * 64 + (size_t)12 * qpt * nw * 32 * sizeof(double);
}

template <int NW, int TQ, int MINB>
cudaError_t launch_variant(const DirectPlan &plan, const StepParams &p, int mode,
                           cudaStream_t stream)
{
    if (mode == 2)
        direct_f32_kernel<NW, TQ, 2, MINB><<<p.ws.n_ctas, plan.block, 0, stream>>>(p);
    else if (mode == 1)
        direct_f32_kernel<NW, TQ, 1, MINB><<<p.ws.n_ctas, plan.block, 0, stream>>>(p);
    else
        direct_f32_kernel<NW, TQ, 0, MINB><<<p.ws.n_ctas, plan.block, 0, stream>>>(p);
    return cudaGetLastError();
}

template <int NW, int TQ, int MINB>
int occupancy_variant(bool packed)
{
    int n = 0;
    if (packed)
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &n, direct_f32_kernel<NW, TQ, 1, MINB>, NW * 32, 0);
    else
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &n, direct_f32_kernel<NW, TQ, 0, MINB>, NW * 32, 0);
    return n;
}

// i-packed kernels: more than 48 KB of dynamic shared memory, opted into once per device
template <int NW, int TQ, int MINB, int QPT>
cudaError_t prepare_ip()
{
    const int bytes = (int)ip_smem_bytes(NW, TQ, QPT);
    cudaError_t e = cudaFuncSetAttribute(direct_f32_ip_kernel<NW, TQ, 1, MINB, QPT>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(direct_f32_ip_kernel<NW, TQ, 2, MINB, QPT>,
                                cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
template <int NW, int TQ, int MINB, int QPT>
int occupancy_ip()
{
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, direct_f32_ip_kernel<NW, TQ, 1, MINB, QPT>, NW * 32,
                                                  ip_smem_bytes(NW, TQ, QPT));
    return n;
}
template <int NW, int TQ, int MINB, int QPT>
cudaError_t launch_ip(const DirectPlan &plan, const StepParams &p, int mode, cudaStream_t stream)
{
    const size_t smem = ip_smem_bytes(NW, TQ, QPT);
    if (mode == 2)
        direct_f32_ip_kernel<NW, TQ, 2, MINB, QPT><<<p.ws.n_ctas, plan.block, smem, stream>>>(p);
    else
        direct_f32_ip_kernel<NW, TQ, 1, MINB, QPT><<<p.ws.n_ctas, plan.block, smem, stream>>>(p);
    return cudaGetLastError();
}
} // namespace

cudaError_t plan_direct_f32(DirectPlan *plan, const SliceGeom &sg, int sm_count, int ctas_per_sm,
                            int tune_flags)
{
    const unsigned slice_units = sg.slice_units, n_units = sg.n_units;
    const unsigned slice_bodies = slice_units * 4u;
    // size class by work (i x j), not by slice length: a rank of an 8-GPU run owns 125k of 1M
    // bodies and still wants the large class.  Measured on B200 (tools/probe_perf.py, round 2,
    // profiles/r2_notes.md): the i-packed large class wins from 20k^2 up (181 vs 187 us; 0.71 of
    // the roofline at 30k^2, 0.77 at 1M^2), the small class below (10k^2: 60 vs 77 us).
    const double work = (double)slice_bodies * (double)n_units * 4.0;
    const bool packed = !(tune_flags & 1);
    int v = work >= 3.5e8 ? (packed ? NB_F32_LARGE_VARIANT : 0) : 2;
    // in between (15k^2: 104 us against 113 for the small and 127 for the large class): the
    // i-packed kernel with two CTAs of 4 warps per SM
    if (packed && work >= 2.0e8 && work < 3.5e8) v = 8;
    // j-tiles of 128 quads instead of 64 halve the flushes of the FP32 partial sums into the FP64
    // ones: 1M bodies 348.1 -> 343.5 ms; 256 quads: 341.7 ms, but too coarse a work unit below
    // ~500k bodies (30k: 0.34 -> 0.44 ms) and, on several ranks, more padding (slices are whole tiles)
    if (v == NB_F32_LARGE_VARIANT && v == 6)
        v = (sg.own_units == sg.n_units && n_units >= 200000u) ? 11 : 9;
    const int forced = (tune_flags >> 8) & 0xf;
    if (forced >= 1 && forced <= kNumVariants) v = forced - 1;
    if (!packed && kVariants[v].qpt) v = 0; // the scalar A/B twin exists in the j-packed kernel only
    const VariantF32 &k = kVariants[v];
    const unsigned qpt = k.qpt ? (unsigned)k.qpt : 1u;
    plan->precision = 1;
    plan->variant = v;
    plan->units_per_iblock = (unsigned)k.nw * 32u * qpt;
    plan->units_per_tile = (unsigned)k.tq;
    plan->n_iblocks = (slice_units + plan->units_per_iblock - 1) / plan->units_per_iblock;
    plan->n_jtiles = (n_units + plan->units_per_tile - 1) / plan->units_per_tile;
    plan->block = (unsigned)k.nw * 32u;
    plan->slot_doubles = (size_t)plan->units_per_iblock * 12;
    int occ = 0;
    cudaError_t err = cudaSuccess;
    switch (v) {
    case 0: occ = occupancy_variant<8, 64, 2>(packed); break;
    case 1: occ = occupancy_variant<4, 64, 4>(packed); break;
    case 2: occ = occupancy_variant<1, 16, 8>(packed); break;
    case 3: err = prepare_ip<16, 64, 1, 2>(); occ = occupancy_ip<16, 64, 1, 2>(); break;
    case 4: err = prepare_ip<12, 64, 1, 2>(); occ = occupancy_ip<12, 64, 1, 2>(); break;
    case 5: err = prepare_ip<8, 64, 2, 2>(); occ = occupancy_ip<8, 64, 2, 2>(); break;
    case 6: err = prepare_ip<8, 64, 1, 2>(); occ = occupancy_ip<8, 64, 1, 2>(); break;
    case 7: err = prepare_ip<8, 64, 1, 3>(); occ = occupancy_ip<8, 64, 1, 3>(); break;
    case 9: err = prepare_ip<8, 128, 1, 2>(); occ = occupancy_ip<8, 128, 1, 2>(); break;
    case 10: err = prepare_ip<8, 32, 1, 2>(); occ = occupancy_ip<8, 32, 1, 2>(); break;
    case 11: err = prepare_ip<8, 256, 1, 2>(); occ = occupancy_ip<8, 256, 1, 2>(); break;
    case 12: err = prepare_ip<8, 192, 1, 2>(); occ = occupancy_ip<8, 192, 1, 2>(); break;
    default: err = prepare_ip<4, 64, 2, 2>(); occ = occupancy_ip<4, 64, 2, 2>(); break;
    }
    if (err != cudaSuccess) return err;
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (occ < 1) occ = 1;
    int per_sm = ctas_per_sm > 0 ? ctas_per_sm : k.minb;
    if (per_sm > occ) per_sm = occ;
    plan->grid = (unsigned)(sm_count * per_sm);
    finish_plan(plan, sg, tune_flags);
    return cudaSuccess;
}

cudaError_t launch_direct_f32(const DirectPlan &plan, int geom, StepParams p, int tune_flags,
                              cudaStream_t stream)
{
    apply_geom(plan, geom, &p);
    // scalar (A/B) : packed : packed with the mass multiply hoisted (p.uniform_mass set by the engine)
    const int mode = (tune_flags & 1) ? 0 : (p.uniform_mass ? 2 : 1);
    switch (plan.variant) {
    case 0: return launch_variant<8, 64, 2>(plan, p, mode, stream);
    case 1: return launch_variant<4, 64, 4>(plan, p, mode, stream);
    case 2: return launch_variant<1, 16, 8>(plan, p, mode, stream);
    case 3: return launch_ip<16, 64, 1, 2>(plan, p, mode, stream);
    case 4: return launch_ip<12, 64, 1, 2>(plan, p, mode, stream);
    case 5: return launch_ip<8, 64, 2, 2>(plan, p, mode, stream);
    case 6: return launch_ip<8, 64, 1, 2>(plan, p, mode, stream);
    case 7: return launch_ip<8, 64, 1, 3>(plan, p, mode, stream);
    case 9: return launch_ip<8, 128, 1, 2>(plan, p, mode, stream);
    case 10: return launch_ip<8, 32, 1, 2>(plan, p, mode, stream);
    case 11: return launch_ip<8, 256, 1, 2>(plan, p, mode, stream);
    case 12: return launch_ip<8, 192, 1, 2>(plan, p, mode, stream);
    default: return launch_ip<4, 64, 2, 2>(plan, p, mode, stream);
    }
}

} // namespace nb
