// nb_aux.cuh -- launchers of the O(N) helper kernels (nb_aux.cu).
#pragma once
#include <cuda_runtime.h>

namespace nb {

// host packing (device copies of pos_xyzm[4n], vel_xyz[3n] or nullptr) -> device layout;
// both position buffers are filled (masses and padding never change afterwards).
// cls (may be nullptr): the four words of mass_class; cls[0] != 0 parks the padding bodies at
// 1e18 (FP32) / 1e150 (FP64) instead of leaving them massless at the origin.
cudaError_t pack_bodies(int precision, const double *pos, const double *vel, unsigned n,
                        unsigned n_units, void *q0, void *q1, void *v, const unsigned *cls, cudaStream_t s);
// cls[0] = 1 iff `enabled` and all n masses of the staged host packing are equal (as floats, or
// as doubles with f64) and positive, cls[1] = bits of (float)m_0, cls[2..3] = bits of m_0.
// Device-side: the host never scans the masses.
cudaError_t mass_class(const double *pos, unsigned n, unsigned *cls, bool enabled, bool f64,
                       cudaStream_t s);
// device layout -> host packing; planes = 4 (pos+mass) or 3 (vel / acc)
// scatter (optional): body b of the device order goes to row scatter[b] of `out`
cudaError_t unpack_bodies(int precision, const void *src, unsigned n, int planes, double *out,
                          const unsigned *scatter, cudaStream_t s);
// dst row k = src row order[k] (rows of `planes` doubles)
cudaError_t permute_rows(const double *src, double *dst, const unsigned *order, unsigned n,
                         int planes, cudaStream_t s);
// out[chunk_base[m / chunk] + m % chunk] = in[m]: deals chunks of a sequence out to new places
cudaError_t remap_chunks(const unsigned *in, unsigned *out, unsigned n, unsigned chunk,
                         const unsigned *chunk_base, cudaStream_t s);
// v += (a*dt)/2 ; x' = x + v*dt from stored accelerations, units [i_begin, i_end)
cudaError_t kickdrift(int precision, const void *qc, void *qn, void *vel, const void *acc,
                      unsigned i_begin, unsigned i_end, unsigned n_real, double dt, cudaStream_t s);
unsigned energy_blocks(unsigned n);
// partial[2*b] = sum m v^2, partial[2*b+1] = sum_{i<j} m_i m_j / r_ij of block b
cudaError_t energy_partials(const double *pos_xyzm, const double *vel_xyz, unsigned n,
                            double *partial, cudaStream_t s);

// energy_kernel with its partials in a file-scope device array (no allocation); device pointers in
// the host packing, result {KE, PE} written to d_ke_pe on stream s.  cudaErrorInvalidValue when n
// exceeds NB_ENERGY_STATIC_MAX_BODIES.
cudaError_t energy_static(int device, const double *d_pos_xyzm, const double *d_vel_xyz, unsigned n,
                          double G, double *d_ke_pe, cudaStream_t s);

} // namespace nb
