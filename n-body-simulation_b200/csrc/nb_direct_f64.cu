// nb_direct_f64.cu -- the FP64 O(N^2) force + leapfrog kernel for sm_100a.
//
// Same structure as the FP32 kernel (nb_direct_f32.cu): TMA-fed mbarrier ring
// of j-tiles, persistent grid over an evenly cut (i-block x j-tile) list,
// last-arriver epilogue with the kick/drift fused in.  Differences:
//   * the reference's precision (serial/exhaustive.c is FP64 throughout), so
//     this is the kernel the parity tests hold to 1e-10 / 1e-12;
//   * bodies are double4 {x,y,z,m} (32 B), two i-bodies per thread;
//   * pow(r2, -1.5) (serial/exhaustive.c:173) starts from y0 = r2^-1/2 of
//     MUFU.RSQ64H (rsqrt.approx.ftz.f64, relative error e ~ 2^-20) and corrects
//     the CUBE directly: with e = 1 - r2*y0^2,
//         r2^-3/2 = y0^3 (1 - e)^-3/2 = y0^3 (1 + e (3/2 + 15/8 e) + O(e^3)),
//     O(e^3) < 2^-58.  y0^2 is shared by e and the cube: 7 FP64 operations for
//     m * r2^-3/2;
//   * accumulation is plain FP64 in j order inside a CTA's j-range; partial
//     ranges are added in fixed slot order.
// Per interaction: 3 DADD + 3 DFMA (r2) + 7 + 3 DFMA = 16 FP64-pipe operations
// + 1 MUFU.  Equal masses (UNI, detected at upload like in the FP32 engine): m_j leaves the loop
// (r2^-3/2 alone is 6 operations: 15 per interaction) and the epilogue's G carries G*m; padding
// bodies, which cannot be massless then, are parked at 1e150 where y0^3 underflows to exactly 0.
// IPT i-bodies per thread (2 or 4): every j-body read from shared memory (two LDS.128, or
// LDS.128 + LDS.64 without the mass) serves IPT interactions -- an FP64 instruction occupies the
// dispatch port for two cycles and everything else for one, so the loads and the
// loop counter are not free.
#include "nb_direct.cuh"

namespace nb {
namespace {

constexpr int STAGES64 = 4;

__device__ __forceinline__ void named_sync64(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// (r2)^(-3/2), six operations
__device__ __forceinline__ double inv_r3(double r2)
{
    const double y0 = rsqrt_approx_f64(r2);
    const double y02 = y0 * y0;
    const double e = fma(-r2, y02, 1.0);
    const double pe = fma(1.875, e, 1.5);
    const double q = e * pe;
    const double u = y02 * y0;
    return fma(u, q, u);
}

// m * (r2)^(-3/2)
__device__ __forceinline__ double inv_r3_times_m(double r2, double m)
{
    const double y0 = rsqrt_approx_f64(r2);
    const double y02 = y0 * y0;
    const double e = fma(-r2, y02, 1.0);
    const double pe = fma(1.875, e, 1.5);
    const double q = e * pe;
    const double my = m * y0;
    const double u = y02 * my;
    return fma(u, q, u);
}

// v += (a*dt)/2.0 and x += v*dt with the reference's roundings
// (serial/exhaustive.c:221-232): no FMA contraction.
__device__ __forceinline__ double kick64(double v, double a, double dt)
{
    return __dadd_rn(v, __dmul_rn(__dmul_rn(a, dt), 0.5));
}
__device__ __forceinline__ double drift64(double x, double v, double dt)
{
    return __dadd_rn(x, __dmul_rn(v, dt));
}

template <int NW, int TB, int MINB, bool UNI, int IPT>
__global__ void __launch_bounds__(NW * 32, MINB)
    direct_f64_kernel(const StepParams p)
{
    static_assert(IPT == 2 || IPT == 4, "i-bodies per thread");
    constexpr int TPB = NW * 32;    // compute threads
    constexpr int BPB = TPB * IPT;  // bodies per i-block
    constexpr int HALF = IPT / 2;   // double2 words per thread and component
    __shared__ __align__(128) double4 tiles[STAGES64][TB];
    __shared__ __align__(8) uint64_t full_bar[STAGES64];
    __shared__ __align__(8) uint64_t empty_bar[STAGES64];
    __shared__ unsigned s_last;

    const int lane = threadIdx.x & 31;
    const double4 *__restrict__ q = static_cast<const double4 *>(p.q_cur);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES64; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], NW);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const unsigned long long w0 = p.ws.bound(blockIdx.x);
    const unsigned long long w1 = p.ws.bound(blockIdx.x + 1);
    const unsigned n_jtiles = p.ws.n_jtiles;

    // j-tile ring fed by thread 0, two tiles ahead (see nb_direct_f32.cu)
    constexpr unsigned AHEAD = STAGES64 - 2;
    const unsigned n_tiles = (unsigned)(w1 - w0);
    const unsigned tile0 = (unsigned)(w0 % n_jtiles);
    auto global_tile = [&](unsigned t) {
        const unsigned g = p.jt_begin + t;
        return g >= p.jt_total ? g - p.jt_total : g;
    };
    auto issue_tile = [&](unsigned k) {
        const unsigned s = k % STAGES64;
        if (k >= STAGES64) mbar_wait(&empty_bar[s], ((k / STAGES64) - 1u) & 1u);
        const unsigned first = global_tile((tile0 + k) % n_jtiles) * TB;
        const unsigned nb = min((unsigned)TB, p.n_units - first);
        const unsigned bytes = nb * 32u;
        mbar_arrive_expect_tx(&full_bar[s], bytes);
        tma_load_1d(&tiles[s][0], q + first, bytes, &full_bar[s]);
    };
    if (threadIdx.x == 0)
        for (unsigned k = 0; k < AHEAD && k < n_tiles; ++k) issue_tile(k);
    unsigned kt = 0;

    const int tidc = threadIdx.x;
    unsigned stage = 0, phase = 0;
    unsigned long long w = w0;
    while (w < w1) {
        const unsigned ib = (unsigned)(w / n_jtiles);
        const unsigned t0 = (unsigned)(w - (unsigned long long)ib * n_jtiles);
        const unsigned cnt = (unsigned)min((unsigned long long)(n_jtiles - t0), w1 - w);
        unsigned bi[IPT];
        bool valid[IPT];
        double xi[IPT], yi[IPT], zi[IPT];
#pragma unroll
        for (int r = 0; r < IPT; ++r) {
            bi[r] = p.i_begin + ib * BPB + r * TPB + tidc;
            valid[r] = bi[r] < p.i_end;
            const double4 b = q[valid[r] ? bi[r] : p.i_begin];
            xi[r] = b.x; yi[r] = b.y; zi[r] = b.z;
        }
        double ax[IPT], ay[IPT], az[IPT];
#pragma unroll
        for (int r = 0; r < IPT; ++r) ax[r] = ay[r] = az[r] = 0.0;

        for (unsigned k = 0; k < cnt; ++k) {
            const unsigned first = global_tile(t0 + k) * TB;
            const int nb = (int)min((unsigned)TB, p.n_units - first);
            if (threadIdx.x == 0 && kt + AHEAD < n_tiles) issue_tile(kt + AHEAD);
            ++kt;
            mbar_wait(&full_bar[stage], phase);
            const double4 *__restrict__ tile = &tiles[stage][0];
#pragma unroll 4
            for (int j = 0; j < nb; ++j) {
                double4 b;
                if (UNI) { // x,y | z: the mass is not read
                    const double2 xy = *reinterpret_cast<const double2 *>(&tile[j]);
                    b = make_double4(xy.x, xy.y, tile[j].z, 0.0);
                } else {
                    b = tile[j];
                }
#pragma unroll
                for (int r = 0; r < IPT; ++r) {
                    const double dx = b.x - xi[r];
                    const double dy = b.y - yi[r];
                    const double dz = b.z - zi[r];
                    double r2 = fma(dx, dx, p.eps2);
                    r2 = fma(dy, dy, r2);
                    r2 = fma(dz, dz, r2);
                    const double s = UNI ? inv_r3(r2) : inv_r3_times_m(r2, b.w);
                    ax[r] = fma(dx, s, ax[r]);
                    ay[r] = fma(dy, s, ay[r]);
                    az[r] = fma(dz, s, az[r]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[stage]);
            if (++stage == STAGES64) { stage = 0; phase ^= 1u; }
        }
        w += cnt;

        bool do_epilogue = true;
        if (cnt != n_jtiles) {
            const unsigned long long x0 = (unsigned long long)ib * n_jtiles;
            const unsigned c_first = p.ws.owner(x0);
            const unsigned c_last = p.ws.owner(x0 + n_jtiles - 1);
            const unsigned nslots = c_last - c_first + 1;
            const unsigned slot = blockIdx.x - c_first;
            constexpr int SLOT = TPB * 3 * HALF; // double2 words per slot
            double2 *ws = reinterpret_cast<double2 *>(p.workspace) +
                          ((size_t)ib * p.max_slots) * (size_t)SLOT;
            double2 *mine = ws + (size_t)slot * SLOT;
#pragma unroll
            for (int h = 0; h < HALF; ++h) {
                mine[(0 * HALF + h) * TPB + tidc] = make_double2(ax[2 * h], ax[2 * h + 1]);
                mine[(1 * HALF + h) * TPB + tidc] = make_double2(ay[2 * h], ay[2 * h + 1]);
                mine[(2 * HALF + h) * TPB + tidc] = make_double2(az[2 * h], az[2 * h + 1]);
            }
            __threadfence();
            named_sync64(1, TPB);
            if (tidc == 0) {
                const unsigned old = atomicAdd(&p.counters[ib], 1u);
                const unsigned last = (old == nslots - 1);
                if (last) p.counters[ib] = 0;
                s_last = last;
            }
            named_sync64(1, TPB);
            do_epilogue = (s_last != 0);
            if (do_epilogue) {
                __threadfence();
#pragma unroll
                for (int r = 0; r < IPT; ++r) ax[r] = ay[r] = az[r] = 0.0;
                for (unsigned s = 0; s < nslots; ++s) {
                    const double2 *src = ws + (size_t)s * SLOT;
#pragma unroll
                    for (int h = 0; h < HALF; ++h) {
                        double2 v;
                        v = __ldcg(&src[(0 * HALF + h) * TPB + tidc]); ax[2 * h] += v.x; ax[2 * h + 1] += v.y;
                        v = __ldcg(&src[(1 * HALF + h) * TPB + tidc]); ay[2 * h] += v.x; ay[2 * h + 1] += v.y;
                        v = __ldcg(&src[(2 * HALF + h) * TPB + tidc]); az[2 * h] += v.x; az[2 * h + 1] += v.y;
                    }
                }
            }
            named_sync64(1, TPB);
        }

        // multi-rank: sums parked between the own-slice and the gathered j-ranges
        if (do_epilogue && (p.flags & (NB_F_PARTIAL_IN | NB_F_PARTIAL_OUT))) {
            double2 *park = reinterpret_cast<double2 *>(p.partial) + (size_t)ib * (TPB * 3 * HALF);
#pragma unroll
            for (int h = 0; h < HALF; ++h) {
                if (p.flags & NB_F_PARTIAL_IN) {
                    double2 v;
                    v = park[(0 * HALF + h) * TPB + tidc]; ax[2 * h] += v.x; ax[2 * h + 1] += v.y;
                    v = park[(1 * HALF + h) * TPB + tidc]; ay[2 * h] += v.x; ay[2 * h + 1] += v.y;
                    v = park[(2 * HALF + h) * TPB + tidc]; az[2 * h] += v.x; az[2 * h + 1] += v.y;
                } else {
                    park[(0 * HALF + h) * TPB + tidc] = make_double2(ax[2 * h], ax[2 * h + 1]);
                    park[(1 * HALF + h) * TPB + tidc] = make_double2(ay[2 * h], ay[2 * h + 1]);
                    park[(2 * HALF + h) * TPB + tidc] = make_double2(az[2 * h], az[2 * h + 1]);
                }
            }
            if (!(p.flags & NB_F_PARTIAL_IN)) do_epilogue = false;
        }

        if (do_epilogue) {
#pragma unroll
            for (int r = 0; r < IPT; ++r) {
                if (!valid[r]) continue;
                const unsigned b = bi[r];
                const double a3[3] = {p.G * ax[r], p.G * ay[r], p.G * az[r]};
                if (p.flags & NB_F_STORE_ACC)
                    static_cast<double4 *>(p.acc)[b] = make_double4(a3[0], a3[1], a3[2], 0.0);
                if (p.flags & (NB_F_KICK2 | NB_F_KICK1_DRIFT)) {
                    double4 *vel = static_cast<double4 *>(p.vel);
                    double4 V = vel[b];
                    double v3[3] = {V.x, V.y, V.z};
                    double x3[3] = {xi[r], yi[r], zi[r]};
                    const bool real = b < p.n_real;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        if (p.flags & NB_F_KICK2) v3[c] = kick64(v3[c], a3[c], p.dt);
                        if (p.flags & NB_F_KICK1_DRIFT) {
                            v3[c] = kick64(v3[c], a3[c], p.dt);
                            x3[c] = drift64(x3[c], v3[c], p.dt);
                        }
                        if (!real) { v3[c] = 0.0; x3[c] = (c == 0 ? xi[r] : c == 1 ? yi[r] : zi[r]); }
                    }
                    vel[b] = make_double4(v3[0], v3[1], v3[2], 0.0);
                    if (p.flags & NB_F_KICK1_DRIFT) {
                        // mass rides along: q_next[b].w was set at upload and never changes
                        double *qn = reinterpret_cast<double *>(static_cast<double4 *>(p.q_next) + b);
                        qn[0] = x3[0]; qn[1] = x3[1]; qn[2] = x3[2];
                    }
                }
            }
        }
    }
}

struct VariantF64 {
    int nw, tb, minb, ipt;
};
#ifndef NB_F64_LARGE_VARIANT
#define NB_F64_LARGE_VARIANT 4
#endif
constexpr int kNumVariants64 = 8;
constexpr VariantF64 kVariants64[kNumVariants64] = {
    {8, 128, 2, 2}, // 0 round 1's large class: i-block 512 bodies
    {4, 128, 4, 2}, // 1 medium: i-block 256 bodies
    {1, 32, 8, 2},  // 2 small : i-block 64 bodies
    {8, 128, 2, 4}, // 3 large, 4 i-bodies per thread: i-block 1024 bodies, <= 128 registers
    {8, 128, 1, 4}, // 4 the same with one CTA per SM
    {4, 128, 4, 4}, // 5 medium, 4 i-bodies per thread
    {16, 128, 1, 4}, // 6 one CTA of 16 warps per SM
    {4, 128, 2, 4}, // 7 medium, two CTAs of 4 warps, registers uncapped
};

// X(index, NW, TB, MINB, IPT)
#define NB_F64_VARIANTS(X) \
    X(0, 8, 128, 2, 2)     \
    X(1, 4, 128, 4, 2)     \
    X(2, 1, 32, 8, 2)      \
    X(3, 8, 128, 2, 4)     \
    X(4, 8, 128, 1, 4)     \
    X(5, 4, 128, 4, 4)     \
    X(6, 16, 128, 1, 4)    \
    X(7, 4, 128, 2, 4)

template <int NW, int TB, int MINB, int IPT>
int occupancy64()
{
    int n = 0, m = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, direct_f64_kernel<NW, TB, MINB, false, IPT>,
                                                  NW * 32, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&m, direct_f64_kernel<NW, TB, MINB, true, IPT>,
                                                  NW * 32, 0);
    return n < m ? n : m;
}
} // namespace

cudaError_t plan_direct_f64(DirectPlan *plan, const SliceGeom &sg, int sm_count, int ctas_per_sm,
                            int tune_flags)
{
    const unsigned slice_units = sg.slice_units, n_units = sg.n_units;
    // by work (i x j); measured (round 2, profiles/r2_direct_f64_variants_*.jsonl): the large
    // class (one CTA of 8 warps per SM, 4 i-bodies per thread, 174 registers) wins from 20k^2 up
    // (20k: 0.404 vs 0.412 ms medium; 30k: 0.859 vs 0.901; 100k: 9.24 vs 9.56 for round 1's
    // large class), small at 10k^2 and below
    const double work = (double)slice_units * (double)n_units;
    // (later scan, profiles/r2_mid_size_scan.txt: 10k^2 121.7 us large / 127.1 small / 129.3 medium;
    // 15k^2 229.6 / 268.0 / 242.0 -- the large class from 10k^2)
    int v = work >= 9.0e7 ? NB_F64_LARGE_VARIANT : 2;
    const int forced = (tune_flags >> 8) & 0xf;
    if (forced >= 1 && forced <= kNumVariants64) v = forced - 1;
    const VariantF64 &k = kVariants64[v];
    plan->precision = 0;
    plan->variant = v;
    plan->units_per_iblock = (unsigned)(k.nw * 32 * k.ipt);
    plan->units_per_tile = (unsigned)k.tb;
    plan->n_iblocks = (slice_units + plan->units_per_iblock - 1) / plan->units_per_iblock;
    plan->n_jtiles = (n_units + plan->units_per_tile - 1) / plan->units_per_tile;
    plan->block = (unsigned)k.nw * 32u;
    plan->slot_doubles = (size_t)k.nw * 32 * 3 * k.ipt;
    int occ = 0;
    switch (v) {
#define X(I, NW, TB, MINB, IPT) \
    case I: occ = occupancy64<NW, TB, MINB, IPT>(); break;
        NB_F64_VARIANTS(X)
#undef X
    default: break;
    }
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (occ < 1) occ = 1;
    int per_sm = ctas_per_sm > 0 ? ctas_per_sm : k.minb;
    if (per_sm > occ) per_sm = occ;
    plan->grid = (unsigned)(sm_count * per_sm);
    finish_plan(plan, sg, tune_flags);
    return cudaSuccess;
}

cudaError_t launch_direct_f64(const DirectPlan &plan, int geom, StepParams p, int tune_flags,
                              cudaStream_t stream)
{
    (void)tune_flags;
    apply_geom(plan, geom, &p);
    const bool uni = p.uniform_mass != 0;
    switch (plan.variant) {
#define X(I, NW, TB, MINB, IPT)                                                                      \
    case I:                                                                                          \
        if (uni) direct_f64_kernel<NW, TB, MINB, true, IPT><<<p.ws.n_ctas, plan.block, 0, stream>>>(p); \
        else direct_f64_kernel<NW, TB, MINB, false, IPT><<<p.ws.n_ctas, plan.block, 0, stream>>>(p);  \
        break;
        NB_F64_VARIANTS(X)
#undef X
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

} // namespace nb
