// nb_direct.cuh -- host-side plan / launcher interface of the direct-sum kernels.
#pragma once
#include "nb_common.cuh"

namespace nb {

// Device body layouts
//   FP32 "quad-planar": one 64-byte record per 4 consecutive bodies,
//        float4 X, float4 Y, float4 Z, float4 M   (unit = quad)
//        velocities: float4 VX, VY, VZ per quad; stored accelerations likewise.
//   FP64: double4 {x,y,z,m} per body (unit = body); velocities / accelerations
//        double4 {x,y,z,0}.
// One launch covers the j-tiles [jt_begin, jt_begin + n_jtiles) modulo the
// total; its (i-block x j-tile) list is cut evenly over `grid` CTAs.
struct LaunchGeom {
    unsigned n_jtiles, jt_begin, grid, max_slots;
};
enum { NB_GEOM_FULL = 0, NB_GEOM_OWN = 1, NB_GEOM_REST = 2 };

struct DirectPlan {
    int precision;          // NB_PREC_*
    int variant;            // index into the kernel table (size class)
    unsigned units_per_iblock;
    unsigned units_per_tile;
    unsigned n_iblocks, n_jtiles; // n_jtiles: tiles covering all j
    unsigned grid, block;         // grid: CTAs of a full machine (sm_count x CTAs per SM)
    // geom[FULL]: all j in one launch.  Multi-rank steady state uses two
    // launches so the all-gather of the other ranks' positions overlaps compute:
    // geom[OWN] = the j-tiles of this rank's own slice, geom[REST] = the others,
    // starting right after the own slice and wrapping around.
    LaunchGeom geom[3];
    bool can_split;
    size_t slot_doubles;     // doubles per workspace slot (= per parked i-block)
    size_t workspace_bytes;
    size_t partial_bytes;
    size_t counter_bytes;
};

// Where this rank's slice sits in the padded j extent (all in layout units).
struct SliceGeom {
    unsigned slice_units; // i extent with real bodies owned by this rank
    unsigned n_units;     // padded j extent (all ranks)
    unsigned own_begin;   // first unit of the rank's padded slice
    unsigned own_units;   // padded slice length (== n_units when world == 1)
};
// slice alignment (units) that keeps slice boundaries on j-tile boundaries in every size class
constexpr unsigned kSliceAlignF32 = 64, kSliceAlignF64 = 128;

// ctas_per_sm == 0 -> default. tune_flags: NB_TUNE_* from nbody_engine.h.
cudaError_t plan_direct_f32(DirectPlan *plan, const SliceGeom &sg, int sm_count, int ctas_per_sm,
                            int tune_flags);
cudaError_t launch_direct_f32(const DirectPlan &plan, int geom, StepParams p, int tune_flags,
                              cudaStream_t stream);

cudaError_t plan_direct_f64(DirectPlan *plan, const SliceGeom &sg, int sm_count, int ctas_per_sm,
                            int tune_flags);
cudaError_t launch_direct_f64(const DirectPlan &plan, int geom, StepParams p, int tune_flags,
                              cudaStream_t stream);

// Fills geom[], workspace/partial sizes of a plan from (n_iblocks, n_jtiles, grid).
void finish_plan(DirectPlan *plan, const SliceGeom &sg, int tune_flags);
// p.ws / p.max_slots / p.jt_* of one launch
void apply_geom(const DirectPlan &plan, int geom, StepParams *p);

} // namespace nb
