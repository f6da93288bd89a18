/*
 * nbody_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's exhaustive (direct-sum) N-body path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this.  The product (include/nbody_engine.h and the
 * host programs) never links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below
 * (a) against the reference's only known-answer test
 *     (common/get_energy.c:49-59 -> KE 1629932, PE -207.735345789164597),
 * (b) bit-for-bit against the reference's own C sources compiled in place
 *     from /root/reference into oracle/_ref/ (oracle/Makefile `ref` target),
 * (c) against the tests/golden npz files, produced by tests/golden/make_golden.py
 *     from those compiled reference sources.
 *
 * Flat array conventions (same packing as the reference's CUDA host code,
 * cuda/exhaustive.cu:240-251):
 *     pos_xyzm[4*i + 0..3] = x, y, z, mass        vel_xyz[3*i + 0..2] = vx, vy, vz
 */
#ifndef NBODY_ORACLE_H
#define NBODY_ORACLE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Error codes mirror the reference's exit codes (serial/exhaustive.c:84-86,
 * 124-127, 136-142): 1 = cannot open / not well formed, 2 = out of memory. */
#define ORC_OK 0
#define ORC_EFILE 1
#define ORC_ENOMEM 2

/* serial/exhaustive.c:79-146 (count_entities_file + get_entities). */
int orc_read_bodies(const char *path, unsigned *n_out, double **pos_xyzm, double **vel_xyz);
void orc_free(void *p);

/* serial/exhaustive.c:43-53: strtof() into float, n_steps = (end-start)/dt in float. */
int orc_parse_times(const char *start_s, const char *end_s, const char *dt_s,
                    float *start, float *end, float *dt);

/* serial/exhaustive.c:155-181.  acc_xyz[3*i..] ; j == i is included. */
void orc_acceleration(unsigned n, const double *pos_xyzm, double *acc_xyz);
/* openmp/exhaustive.c:166-200 (omp for over i, per-i local accumulator). */
void orc_acceleration_mt(unsigned n, const double *pos_xyzm, double *acc_xyz, int threads);
/* Slice form used by the multi-rank host tests: rows [i0, i1) against all n. */
void orc_acceleration_slice(unsigned n, const double *pos_xyzm, unsigned i0, unsigned i1,
                            double *acc_xyz_slice);

/* serial/exhaustive.c:192-255.  Advances pos/vel in place by n_steps of
 * kick-drift-kick.  If frames != NULL the reference's RESULTS output is written
 * (frame 0 + one frame per drift, "%lu,%lf,%lf,%lf,%lf\n").  threads <= 1 is
 * the serial program, threads > 1 the OpenMP one (identical results). */
int orc_propagate(unsigned n, double *pos_xyzm, double *vel_xyz, int n_steps, float dt,
                  const char *frames_path, int threads);

/* common/get_energy.c:7-47 on the packed layout (pos stride 4). */
void orc_get_energy(unsigned n, const double *pos_xyzm, const double *vel_xyz, double G,
                    double *KE, double *PE);
/* common/get_energy.c:7-47 on its own flat layout pos[3n], vel[3n], mass[n]. */
void orc_get_energy_flat(const double *pos, const double *vel, const double *mass, double G,
                         int sz, double *KE, double *PE);

/* FP32 twin of orc_acceleration: inputs rounded to float, per-pair arithmetic
 * in float, accumulation in double.  Not a reference function (the reference
 * has no FP32 path); used to bound the FP32 engine's rounding model. */
void orc_acceleration_f32(unsigned n, const double *pos_xyzm, double *acc_xyz);

/* ---- Barnes-Hut (serial/barnes-hut.c), see bh_oracle.c ---- */
/* get_bounding_box, serial/barnes-hut.c:326-338 */
double orc_bh_bounding_box(unsigned n, const double *pos_xyzm);
/* bbox + tree + centre of mass + THETA=0.5 traversal, serial/barnes-hut.c:470-473.
 * stats (may be NULL): [0] internal cells, [1] depth, [2] root border,
 * [3] interactions per body, [4] node visits per body. */
void orc_bh_acceleration(unsigned n, const double *pos_xyzm, double *acc_xyz, double *stats);
/* serial/barnes-hut.c:445-517 */
int orc_bh_propagate(unsigned n, double *pos_xyzm, double *vel_xyz, int n_steps, float dt);

#ifdef __cplusplus
}
#endif
#endif
