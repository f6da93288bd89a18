/*
 * nbody_engine.h -- C ABI of the B200-native gravitational N-body engine.
 *
 * This is the drop-in boundary for the reference's exhaustive (direct-sum)
 * path and its Barnes-Hut variant.  The reference (Metgraph/N-Body-simulation)
 * has no plugin/FFI layer of its own: inside each program the boundary is the
 * pair of C functions propagation()/acceleration().  Each entry point below
 * names the reference code it stands in for.  Plain pointers and sizes only;
 * no torch / C++ types.  The library is libnbody_b200.so (CUDA, sm_100a);
 * there is no CPU fallback: without a usable CUDA device nb_init fails with
 * NB_E_NODEVICE.
 *
 * Array packing is the reference CUDA host's (cuda/exhaustive.cu:240-251):
 *     pos_xyzm[4*i + 0..3] = x, y, z, mass          vel_xyz[3*i + 0..2] = vx, vy, vz
 * All host arrays are double (the reference is FP64 everywhere); an FP32
 * engine rounds them to float at nb_init / nb_upload.
 *
 * Threading: one host thread per engine.  nb_step only enqueues work on the
 * engine's stream; nb_gather / nb_sync wait for it.
 *
 * Multi-rank engines (cfg.world > 1): nb_init, nb_upload(_slice), nb_step,
 * nb_gather with velocities, nb_gather_slice, nb_accelerations and nb_energy are
 * COLLECTIVE -- they enqueue an ncclAllGather, so every rank must call them, in
 * the same order (host/main_exh.c does, one thread per rank).  A rank that calls
 * one of them alone blocks for ever in the collective.  nb_gather with
 * vel_xyz == NULL, nb_frame_begin/wait, nb_get_stats, nb_tree_stats, nb_sync and
 * nb_free are local.
 */
#ifndef NBODY_ENGINE_H
#define NBODY_ENGINE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nb_engine nb_engine;

/* return codes: 0 on success, negative otherwise (never exit() inside the
 * library; the host maps failures to the reference's exit(EXIT_FAILURE),
 * cuda/exhaustive.cu:78-84) */
#define NB_OK 0
#define NB_E_INVALID (-1)  /* bad argument / configuration */
#define NB_E_CUDA (-2)     /* CUDA runtime error (see nb_last_error) */
#define NB_E_NOMEM (-3)    /* host or device allocation failed */
#define NB_E_COMM (-4)     /* NCCL missing or failed */
#define NB_E_NODEVICE (-5) /* no CUDA device: there is no CPU fallback */
#define NB_E_STATE (-6)    /* call not valid in the engine's current state */

#define NB_PREC_FP64 0 /* the reference's precision */
#define NB_PREC_FP32 1 /* production / roofline precision */

#define NB_ALGO_DIRECT 0     /* O(N^2) exhaustive sum: cuda/exhaustive.cu, serial/exhaustive.c */
#define NB_ALGO_BARNES_HUT 1 /* THETA-criterion octree: cuda/barnes-hut.cu, serial/barnes-hut.c */

#define NB_COMM_ID_BYTES 128 /* sizeof(ncclUniqueId) */

typedef struct nb_config {
    int precision; /* NB_PREC_* */
    int algorithm; /* NB_ALGO_* */
    double dt;     /* time step; the host passes (double)strtof(argv[4]) as the reference does
                      (serial/exhaustive.c:43,51; cuda/exhaustive.cu:49,56) */
    double eps2;   /* softening, reference literal 0.01 (serial/exhaustive.c:172).  The direct sum
                      needs eps2 > 0: it includes the j = i term like the reference (:163) and the
                      softening is what makes that term exactly 0 (nb_init returns NB_E_INVALID
                      otherwise).  Barnes-Hut skips the body's own leaf by index and accepts 0. */
    double G;      /* BIG_G = 1.0 (serial/exhaustive.c:24) */
    double theta;  /* THETA = 0.5 (serial/barnes-hut.c:40); Barnes-Hut only */
    int device;    /* CUDA device ordinal this engine (rank) runs on */
    int rank;      /* i-slice owner index, 0 <= rank < world */
    int world;     /* number of ranks sharing the bodies (1 = single GPU) */
    const void *comm_id; /* world > 1: NB_COMM_ID_BYTES from nb_comm_id() of rank 0 */
    int tune_ctas_per_sm; /* 0 = engine default; launch geometry is ours (the reference's
                             grid_size/block_size argv are accepted and ignored) */
    int tune_flags;       /* 0 = default; NB_TUNE_* bits, for experiments only */
} nb_config;

#define NB_TUNE_SCALAR_FP32 1 /* FP32 kernel without packed f32x2 instructions */
#define NB_TUNE_NO_SPLIT 2    /* one CTA per i-block over all j (no j-range splitting) */
#define NB_TUNE_NO_OVERLAP 4  /* world > 1: wait for the all-gather before any force work */
#define NB_TUNE_NO_UNIFORM 16  /* direct sum: keep m_j in the inner loop even when all masses are equal */
#define NB_TUNE_COUNT_VISITS 8 /* Barnes-Hut: count interactions / node visits (nb_tree_stats) */
#define NB_TUNE_NO_REORDER 32  /* Barnes-Hut: keep the caller's body order on the device */
#define NB_TUNE_NO_INTERLEAVE 64 /* Barnes-Hut, world > 1: rank slices = contiguous pieces of the Morton curve */
#define NB_TUNE_NO_GRAPH 128   /* Barnes-Hut: launch the ~75 kernels of an evaluation one by one, not as a CUDA graph */
#define NB_TUNE_REPLICATED_SORT (1 << 28) /* Barnes-Hut, world > 1: every rank sorts all keys itself (round 1)
                                             instead of sorting its slice and merging the all-gathered runs */
/* bits 8..15: force a kernel size class (value - 1); bits 16..23: Barnes-Hut, world > 1: log2 of the
 * pieces per rank the Morton curve is dealt in, + 1 (0 = the default, 32 pieces); bits 24..27:
 * Barnes-Hut, world == 1: sort the keys as that many slices + merge, the way a multi-rank engine
 * does (rehearsal of the multi-rank sort on one GPU) */

typedef struct nb_stats {
    double last_step_ms;        /* device time (CUDA events on the engine stream) of the last nb_step */
    double last_force_ms;       /* device time of the last force kernel alone */
    unsigned long long force_launches; /* force-kernel launches since nb_init */
    unsigned long long kernel_launches; /* all engine kernel launches since nb_init */
    unsigned long long interactions;   /* pairwise interactions evaluated by this rank since nb_init */
    unsigned long long steps_done;     /* completed drifts */
    int grid, block, smem_bytes;       /* geometry of the last force launch */
    size_t n, n_padded, slice_begin, slice_end;
    unsigned long long collectives;       /* all-gathers issued since nb_init (world > 1) */
    unsigned long long split_evaluations; /* force evaluations issued as own-slice + gathered launches */
    int uniform_mass; /* 1: the direct kernel runs with the mass multiply hoisted (all masses equal) */
    int kernel_variant; /* direct sum: index of the size class the plan chose (FP32: 0..2 j-packed,
                         * 3..8 i-packed, nb_direct_f32.cu; FP64: 0..2 two, 3..5 four i-bodies per thread, nb_direct_f64.cu); -1 for Barnes-Hut */
} nb_stats;

/* Fills cfg with the reference's constants: FP64, direct, eps2 = 0.01, G = 1,
 * theta = 0.5, dt = (double)0.01f, device 0, world 1. */
void nb_config_default(nb_config *cfg);

/* Replaces the allocation + H2D prologue of propagation()
 * (cuda/exhaustive.cu:116-129).  Copies the n bodies in; the caller keeps
 * ownership of the host arrays.  Every rank passes ALL n bodies. */
int nb_init(nb_engine **e, const nb_config *cfg, size_t n, const double *pos_xyzm,
            const double *vel_xyz);

/* Re-loads positions/velocities of an existing engine (same n), resetting the
 * step counter: the H2D half of the end-to-end benchmark. */
int nb_upload(nb_engine *e, const double *pos_xyzm, const double *vel_xyz);

/* Host I/O by slices, for multi-rank runs (no reference counterpart: the reference is single-GPU;
 * these stand where cuda/exhaustive.cu:125-129 and :153-165 copy everything).  nb_upload and
 * nb_gather move all n bodies over every rank's PCIe link; here rank r moves only rows
 * [begin, begin + count) of the caller's order (nb_io_slice: count = ceil(n / world), the last
 * rank's share clipped) and the ranks exchange the rest over NVLink (ncclAllGather of the staged
 * rows / of the velocities).  pos_rows = count x 4 doubles (x,y,z,m), vel_rows = count x 3 (may
 * be NULL like in nb_upload / nb_gather).  Collective when world > 1; with world == 1 they are
 * nb_upload / nb_gather.  The state after nb_upload_slice is the state after nb_upload of the
 * concatenated rows, bit for bit. */
int nb_io_slice(nb_engine *e, size_t *begin, size_t *count);
int nb_upload_slice(nb_engine *e, const double *pos_rows, const double *vel_rows);
int nb_gather_slice(nb_engine *e, double *pos_rows, double *vel_rows);

/* Advances n_steps of kick-drift-kick leapfrog: the time loop of propagation()
 * (serial/exhaustive.c:218-250, cuda/exhaustive.cu:135-151).  One fused
 * force+integrate launch per step; a run of n steps followed by nb_gather
 * with velocities costs n+1 force evaluations, like the reference.
 * A Barnes-Hut engine with world > 1 lays its state out again every 32 steps (environment
 * NB_BH_RELAYOUT_EVERY=K, the same on every rank; 0: never): bodies are dealt to the ranks along
 * the Morton curve of the moment.  No value changes, only where a body lives. */
int nb_step(nb_engine *e, int n_steps);

/* Copies the current state out: what the reference reads back at
 * cuda/exhaustive.cu:153-165 (positions+mass of the current frame) plus the
 * synchronised velocities v_k.  vel_xyz may be NULL (frame output needs
 * positions only).  Blocks until the engine's stream is idle.  With
 * velocities it is collective when world > 1 (closing force evaluation +
 * all-gather of the velocities). */
int nb_gather(nb_engine *e, double *pos_xyzm, double *vel_xyz);

/* Frame pipeline for the RESULTS loop (serial/exhaustive.c:208-238 writes one
 * frame per step; cuda/exhaustive.cu:153-165 reads every frame back with a
 * blocking copy).  nb_frame_begin snapshots the current positions+mass
 * (4n doubles) and starts copying them to pos_xyzm on a side stream, then
 * returns; steps enqueued afterwards run while the copy lands and while the
 * host formats the previous frame.  nb_frame_wait blocks until the snapshot
 * headed for pos_xyzm is complete (NULL: every outstanding snapshot).  Two
 * snapshots can be in flight.  pos_xyzm should come from nb_host_alloc
 * (page-locked), otherwise the copy is staged by the driver and does not
 * overlap. */
int nb_frame_begin(nb_engine *e, double *pos_xyzm);
int nb_frame_wait(nb_engine *e, const double *pos_xyzm);
int nb_host_alloc(void **p, size_t bytes);
int nb_host_free(void *p);

/* acceleration() on the current positions (serial/exhaustive.c:155-181,
 * cuda/exhaustive.cu:273-338, or the Barnes-Hut evaluation): acc_xyz[3n].
 * Does not advance the state.  Collective when world > 1 (all-gathers the
 * ranks' slices of the accelerations). */
int nb_accelerations(nb_engine *e, double *acc_xyz);

/* get_energy() of common/get_energy.c:7-47 on the current state, on the GPU
 * (KE, unsoftened PE over i <= j).  Collective when world > 1 (closes the step
 * and all-gathers the velocities). */
int nb_energy(nb_engine *e, double *KE, double *PE);

/* Barnes-Hut engines: facts about the octree of the last force evaluation --
 * what create_tree/add_ents/center_of_mass left behind (serial/barnes-hut.c:
 * 272-354).  nodes = internal cells (root included), depth counts the root as 1,
 * root_border = get_bounding_box's max (:326-338).  interactions / node_visits
 * are totals over this rank's bodies when NB_TUNE_COUNT_VISITS is set.  Returns
 * NB_E_NOMEM if the node pool or the traversal stack overflowed (the reference
 * exits in that case, openmp/barnes-hut.c:158-161). */
typedef struct nb_tree_info {
    unsigned nodes, depth;
    double root_border;
    unsigned long long interactions, node_visits;
    int pool_overflow, stack_overflow;
    /* NB_TUNE_COUNT_VISITS: child records tested by a warp of the walk (17 FP64 warp
     * instructions each: the FP64-pipe work of the traversal) */
    unsigned long long child_tests;
    /* NB_TUNE_NO_GRAPH: device time (ms) of the stages of the last evaluation -- cube + keys,
     * radix sort (world > 1: of the own slice), merge of the ranks' sorted runs (world > 1),
     * bodies in Morton order (+ own list), build, centre of mass, traversal; else 0.
     * [7] (world > 1, always): the last position exchange on the side stream, from the moment this
     * rank was ready to the end of the all-gather -- waiting for the slowest rank included */
    double stage_ms[8];
} nb_tree_info;
int nb_tree_stats(nb_engine *e, nb_tree_info *out);

/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).
 * The energy of nb_energy -- KE = 1/2 sum m v^2, unsoftened PE = -G sum_{i<j} m_i m_j / r_ij -- of n
 * bodies the CALLER holds on `device`: d_pos_xyzm (4n doubles x,y,z,m) and d_vel_xyz (3n doubles)
 * are device pointers in the host packing of nb_init; {KE, PE} is written to the device pointer
 * d_ke_pe (2 doubles) by work enqueued on `stream` (a cudaStream_t; NULL = the default stream).
 * The call does not synchronise and allocates nothing -- no cudaMalloc / cudaFree, no engine: the
 * kernel's block partials live in a file-scope device array of the library, 32 KB on every GPU
 * that loads it, which bounds n: n > NB_ENERGY_STATIC_MAX_BODIES is refused with NB_E_INVALID and
 * a message in nb_last_error().  All calls on a device share that array (it keeps the last call's
 * contents; each call overwrites what it reads, nothing needs clearing): calls that overlap on
 * the GPU -- different streams, no dependency -- race; order them.  Same kernel and same
 * summation order as nb_energy, so the two agree bit for bit on the same bodies in the same order.
 * It selects `device` as the calling thread's current device. */
#define NB_ENERGY_STATIC_MAX_BODIES 1048576
int nb_energy_static(int device, const double *d_pos_xyzm, const double *d_vel_xyz, size_t n,
                     double G, double *d_ke_pe, void *stream);

int nb_sync(nb_engine *e);
int nb_get_stats(nb_engine *e, nb_stats *out);

/* Releases every device allocation, stream and communicator
 * (cuda/exhaustive.cu:167-169). */
void nb_free(nb_engine *e);

/* Multi-rank plumbing: rank 0 creates an id, the host distributes it. */
int nb_comm_id(void *id_out, size_t bytes);

const char *nb_strerror(int code);
/* Detail of the last failure on this thread (CUDA / NCCL message). */
const char *nb_last_error(void);
int nb_device_count(void);
const char *nb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_ENGINE_H */
