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
struct DirectPlan {
    int precision;          // NB_PREC_*
    int variant;            // index into the kernel table (size class)
    unsigned units_per_iblock;
    unsigned units_per_tile;
    unsigned n_iblocks, n_jtiles;
    unsigned grid, block;
    unsigned max_slots;      // workspace slots per i-block
    size_t slot_doubles;     // doubles per slot
    size_t workspace_bytes;
    size_t counter_bytes;
};

// slice_units: i extent owned by this rank; n_units: padded j extent.
// ctas_per_sm == 0 -> default. tune_flags: NB_TUNE_* from nbody_engine.h.
cudaError_t plan_direct_f32(DirectPlan *plan, unsigned slice_units, unsigned n_units, int sm_count,
                            int ctas_per_sm, int tune_flags);
cudaError_t launch_direct_f32(const DirectPlan &plan, StepParams p, int tune_flags,
                              cudaStream_t stream);

cudaError_t plan_direct_f64(DirectPlan *plan, unsigned slice_units, unsigned n_units, int sm_count,
                            int ctas_per_sm, int tune_flags);
cudaError_t launch_direct_f64(const DirectPlan &plan, StepParams p, int tune_flags,
                              cudaStream_t stream);

// Fills ws/max_slots of a plan from (n_iblocks, n_jtiles, grid).
void finish_plan(DirectPlan *plan, int tune_flags);

} // namespace nb
