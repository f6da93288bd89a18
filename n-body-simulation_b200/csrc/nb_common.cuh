// nb_common.cuh -- shared device helpers: mbarrier / bulk-async-copy (TMA) PTX,
// the (i-block x j-tile) work partition, and launch parameter blocks.
// sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nb {

// ------------------------------------------------------------------ PTX ----
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NB_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NB_DONE_%=;\n"
        "bra NB_WAIT_%=;\n"
        "NB_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk asynchronous copy global -> shared (TMA without a tensor map: the
// body arrays are linear).  SASS: UBLKCP.  bytes % 16 == 0, both 16 B aligned.
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                            uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ float rsqrt_approx(float x)
{
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ double rsqrt_approx_f64(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}

// ------------------------------------------------------- work partition ----
// Work = n_iblocks x n_jtiles tile pairs, enumerated i-block major.  CTA c of
// G owns the contiguous range [bound(c), bound(c+1)).  An i-block whose j
// range is cut by a CTA boundary is finished by the last CTA to arrive: every
// contributor parks its partial sums in the workspace slot (c - first
// contributor) and the last one adds the slots in slot order, so the result
// does not depend on arrival order.
struct WorkSplit {
    unsigned long long total; // n_iblocks * n_jtiles
    unsigned n_iblocks, n_jtiles, n_ctas;
    __host__ __device__ unsigned long long bound(unsigned c) const
    {
        return (total * c) / n_ctas;
    }
    // CTA that owns tile-pair index w
    __host__ __device__ unsigned owner(unsigned long long w) const
    {
        unsigned c = (unsigned)((w * n_ctas) / total);
        while (c + 1 < n_ctas && bound(c + 1) <= w) ++c;
        while (c > 0 && bound(c) > w) --c;
        return c;
    }
};

// flags for the fused force + integrate launch
enum : unsigned {
    NB_F_KICK2 = 1u,      // finish the previous step: v += a*dt/2
    NB_F_KICK1_DRIFT = 2u, // start the next step:     v += a*dt/2 ; x' = x + v*dt
    NB_F_STORE_ACC = 4u,  // keep a(x) for a later cheap kick+drift / for nb_accelerations
    // multi-rank overlap: the force sum is cut into "own slice j-tiles" (needs no
    // communication) and "everybody else's j-tiles" (after the all-gather landed)
    NB_F_PARTIAL_OUT = 8u, // park the raw FP64 sums in `partial`, no epilogue
    NB_F_PARTIAL_IN = 16u, // start from the sums parked in `partial`
};

struct StepParams {
    const void *q_cur; // positions + masses, all bodies (layout depends on precision)
    void *q_next;      // next positions (own slice written)
    void *vel;         // velocities, all bodies (own slice touched)
    void *acc;         // stored accelerations (own slice), or nullptr
    double *workspace; // partial-sum slots
    double *partial;    // parked sums of a PARTIAL_OUT launch (own slice), or nullptr
    unsigned *counters; // one per i-block, zero between launches
    unsigned n_real;    // real bodies (global)
    unsigned n_units;   // j extent in layout units (quads for FP32, bodies for FP64), padded
    unsigned i_begin;   // first i unit of this rank's slice
    unsigned i_end;     // one past the last i unit of the slice
    unsigned max_slots; // workspace slots per i-block
    unsigned jt_begin;  // first j-tile of this launch (global tile index); tiles wrap modulo jt_total
    unsigned jt_total;  // j-tiles covering all n_units; ws.n_jtiles is this launch's share
    unsigned flags;
    unsigned uniform_mass; // FP32: all masses equal -> kernels skip m_j; G below carries G * m
    double dt, eps2, G;
    WorkSplit ws;
};

} // namespace nb
