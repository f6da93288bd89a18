/*
 * main_exh.c -- host program of the exhaustive (direct-sum) simulation.
 *
 * Drop-in for the reference's exh_cuda / exhaustive-cuda (cuda/exhaustive.cu:28-171,
 * Makefile:61) and, with five arguments, for exh_serial (serial/exhaustive.c:32-71):
 *
 *     exh_cuda input_filename start_time end_time delta_time output_filename [grid_size block_size]
 *
 * grid_size/block_size are accepted and ignored (launch geometry belongs to the
 * engine).  Plain C; everything on the GPU goes through the C ABI of
 * include/nbody_engine.h (nb_init / nb_step / nb_gather / nb_free), called where
 * the reference calls propagation() (cuda/exhaustive.cu:56).
 *
 * Like the reference, the frame file is written only by the RESULTS build
 * (exh_cuda_results; the reference's *_debug targets, Makefile:17-18) or when
 * NB_RESULTS=1 is set; otherwise the output name is accepted and ignored.
 * Compiled with -DNB_BARNES_HUT this is barnes-hut-cuda (cuda/barnes-hut.cu:814-1186):
 * the same five arguments, the THETA = 0.5 octree evaluation in place of the
 * direct sum, and -- as in that program, which force-defines RESULTS (:50) and
 * writes its first frame after the first drift (:1064) -- the frame file is
 * written by default and carries n_steps frames, no initial one.
 * Optional knobs (environment, so the reference command line is unchanged):
 *   NB_PRECISION=fp64|fp32 (default fp64, the reference's precision)
 *   NB_DEVICE=<ordinal>    NB_CHECKPOINT=<file>: final state in INPUT format
 *   NB_ENERGY=1: get_energy() of common/get_energy.c:7-47 on the initial and the final state, on
 *   the GPU, printed to stderr in the format of its test ("KE: %.15f, PE: %.15f", :57)
 *   NB_GPUS=<n>: one engine per GPU (devices NB_DEVICE .. NB_DEVICE+n-1), i-slices + NCCL
 *   all-gather per step; output identical in form, rank 0 writes it
 *   NB_CHECKPOINT_EVERY=<k>: also every k steps (no-RESULTS runs), written to <file>.tmp and
 *   renamed, so a restart (`input = <file>`, start_time = the time reached) never sees a torn file
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/nbody_engine.h"
#include "nb_host.h"

#define BILLION 1000000000.0

static void die_engine(const char *what, int rc)
{
    /* cuda_check_error behaviour, cuda/exhaustive.cu:78-84 */
    fprintf(stderr, "Error: %s (%s)\nLine: %s\n", nb_strerror(rc), nb_last_error(), what);
    exit(EXIT_FAILURE);
}

#define NB_MAX_RANKS 16

static int write_checkpoint(const char *path, size_t n, const double *pos, const double *vel);

/* what every rank of a run shares */
struct run {
    nb_config cfg;
    int world;
    size_t n;
    double *positions, *velocities; /* in: initial state (all ranks read); out: final state (rank 0 writes) */
    int n_steps, results, initial_frame;
    const char *output;
    float start, dt;
    char comm_id[NB_COMM_ID_BYTES];
    pthread_barrier_t barrier;
    unsigned long long interactions, force_launches; /* summed over the ranks / of rank 0 */
    pthread_mutex_t lock;
};
struct rank {
    struct run *run;
    int id;
};

/* every rank calls it (velocities of a multi-GPU run are gathered collectively), rank 0 prints */
static void report_energy(nb_engine *eng, int lead, const char *when)
{
    double ke = 0, pe = 0;
    int rc = nb_energy(eng, &ke, &pe);
    if (rc)
        die_engine("nb_energy", rc);
    if (lead)
        fprintf(stderr, "%s KE: %.15f, PE: %.15f\n", when, ke, pe);
}

/*
 * The time loop of one rank.  Rank 0 owns the files; the other ranks of a multi-GPU run do
 * the same engine calls (nb_step; nb_gather with velocities is collective) into scratch.
 */
static void *rank_main(void *arg)
{
    struct rank *me = arg;
    struct run *R = me->run;
    const int lead = me->id == 0, multi = R->world > 1;
    const size_t n = R->n;
    int rc;
    nb_config cfg = R->cfg;
    cfg.rank = me->id;
    cfg.world = R->world;
    cfg.device = R->cfg.device + me->id;
    cfg.comm_id = multi ? R->comm_id : NULL;
    nb_engine *eng = NULL;
    if ((rc = nb_init(&eng, &cfg, n, R->positions, R->velocities)))
        die_engine("nb_init", rc);
    double *pos = R->positions, *vel = R->velocities;
    const char *en = getenv("NB_ENERGY");
    const int energy = en && atoi(en) != 0;
    if (energy)
        report_energy(eng, lead, "initial");
    if (multi) {
        pthread_barrier_wait(&R->barrier); /* every rank has read the initial state */
        if (!lead) {
            pos = malloc((n ? n : 1) * 4 * sizeof(double));
            vel = malloc((n ? n : 1) * 3 * sizeof(double));
            if (!pos || !vel) {
                fprintf(stderr, "Error during memory allocation\n");
                exit(2);
            }
        }
    }

    if (R->results) {
        FILE *fpt = NULL;
        double *frame[2] = {NULL, NULL};
        if (lead) {
            fpt = fopen(R->output, "w");
            if (!fpt) {
                fprintf(stderr, "Error opening file '%s'\n", R->output);
                exit(1);
            }
            if (R->initial_frame)
                nbh_write_frame(fpt, n, R->positions); /* frame 0 = initial state */
            for (int k = 0; k < 2; k++)
                if ((rc = nb_host_alloc((void **)&frame[k], (n ? n : 1) * 4 * sizeof(double))))
                    die_engine("nb_host_alloc", rc);
        }
        /* The reference reads every frame back with a blocking copy and formats it
         * while the GPU idles (cuda/exhaustive.cu:135-165).  Here step t+1 is
         * enqueued before frame t is waited for, so the kernel, the copy (side
         * stream, page-locked buffers) and the formatting overlap. */
        for (int t = 0; t <= R->n_steps; t++) {
            if (t < R->n_steps) {
                if ((rc = nb_step(eng, 1)))
                    die_engine("nb_step", rc);
                if (lead && (rc = nb_frame_begin(eng, frame[t & 1])))
                    die_engine("nb_frame_begin", rc);
            }
            if (lead && t > 0) {
                if ((rc = nb_frame_wait(eng, frame[(t - 1) & 1])))
                    die_engine("nb_frame_wait", rc);
                nbh_write_frame(fpt, n, frame[(t - 1) & 1]);
            }
        }
        if (lead) {
            nb_host_free(frame[0]);
            nb_host_free(frame[1]);
            fclose(fpt);
        }
    } else {
        const char *ck = getenv("NB_CHECKPOINT"), *ev = getenv("NB_CHECKPOINT_EVERY");
        const int every = (ck && ev) ? atoi(ev) : 0;
        for (int done = 0; done < R->n_steps;) {
            const int k = (every > 0 && every < R->n_steps - done) ? every : R->n_steps - done;
            if ((rc = nb_step(eng, k)))
                die_engine("nb_step", rc);
            done += k;
            if (done < R->n_steps) { /* intermediate checkpoint; the final one is written by main */
                if ((rc = nb_gather(eng, pos, vel)))
                    die_engine("nb_gather", rc);
                if (lead && write_checkpoint(ck, n, pos, vel) == 0)
                    fprintf(stderr, "checkpoint after step %d (t = %f): %s\n", done,
                            (double)(R->start + (float)done * R->dt), ck);
            }
        }
    }
    /* closing half kick + read-back of the final state */
    if ((rc = nb_gather(eng, pos, vel)))
        die_engine("nb_gather", rc);
    if (energy)
        report_energy(eng, lead, "final");

    nb_stats st;
    if (nb_get_stats(eng, &st) == NB_OK) {
        pthread_mutex_lock(&R->lock);
        R->interactions += st.interactions;
        if (lead)
            R->force_launches = st.force_launches;
        pthread_mutex_unlock(&R->lock);
    }
    nb_free(eng);
    if (!lead) {
        free(pos);
        free(vel);
    }
    return NULL;
}

/* state in INPUT format (%.17g: restarts are bit-exact), via a temporary file and rename() */
static int write_checkpoint(const char *path, size_t n, const double *pos, const double *vel)
{
    char tmp[4096];
    if (snprintf(tmp, sizeof tmp, "%s.tmp", path) >= (int)sizeof tmp ||
        nbh_write_bodies(tmp, n, pos, vel) != NBH_OK || rename(tmp, path) != 0) {
        fprintf(stderr, "Error opening file '%s'\n", path);
        return -1;
    }
    return 0;
}

int main(int argc, char *argv[])
{
#ifdef NB_BARNES_HUT
    /* cuda/barnes-hut.cu:841-844 takes five arguments and returns -1 otherwise; the reference's
     * README.md:67 documents an optional sixth (cache_sz_MB), accepted and ignored here */
    if (argc != 6 && argc != 7) {
        fprintf(stderr, "Usage: %s input_filename start_time end_time delta_time output_filename\n",
                argv[0]);
        return -1;
    }
#else
    if (argc != 6 && argc != 8) {
        fprintf(stderr,
                "Usage: %s input_filename start_time end_time delta_time output_filename "
                "[grid_size block_size]\n",
                argv[0]);
        exit(1);
    }
#endif
    size_t n_ents = 0;
    double *positions = NULL, *velocities = NULL;
    int rc = nbh_read_bodies(argv[1], &n_ents, &positions, &velocities);
    if (rc == NBH_EOPEN) {
        fprintf(stderr, "Error opening file '%s'\n", argv[1]);
        exit(1);
    } else if (rc == NBH_EFORMAT) {
        fprintf(stderr, "Error while reading file '%s': file is not well formed\n", argv[1]);
        exit(1);
    } else if (rc == NBH_ENOMEM) {
        fprintf(stderr, "Error during memory allocation\n");
        exit(2);
    }

    float start, end, dt;
    int n_steps = nbh_parse_times(argv[2], argv[3], argv[4], &start, &end, &dt);
    printf("Start: %f, end: %f, delta time: %f, time steps: %d, ents: %d\n", start, end, dt,
           n_steps, (int)n_ents);

    int results = 0, initial_frame = 1;
#ifdef RESULTS
    results = 1;
#endif
#ifdef NB_BARNES_HUT
    results = 1;
    initial_frame = 0;
#endif
    const char *env = getenv("NB_RESULTS");
    if (env)
        results = atoi(env) != 0;

    nb_config cfg;
    nb_config_default(&cfg);
    cfg.dt = (double)dt; /* the float promoted to double, cuda/exhaustive.cu:38,56 */
#ifdef NB_BARNES_HUT
    cfg.algorithm = NB_ALGO_BARNES_HUT; /* FP64 only, like the reference */
#endif
    env = getenv("NB_PRECISION");
    if (env && strcmp(env, "fp32") == 0)
        cfg.precision = NB_PREC_FP32;
    env = getenv("NB_DEVICE");
    if (env)
        cfg.device = atoi(env);

    int world = 1;
    env = getenv("NB_GPUS");
    if (env)
        world = atoi(env);
    if (world < 1 || world > NB_MAX_RANKS) {
        fprintf(stderr, "NB_GPUS must be 1..%d\n", NB_MAX_RANKS);
        exit(1);
    }

    struct timespec s, e;
    clock_gettime(CLOCK_REALTIME, &s);

    struct run run = {0};
    pthread_mutex_init(&run.lock, NULL);
    run.cfg = cfg;
    run.world = world;
    run.n = n_ents;
    run.positions = positions;
    run.velocities = velocities;
    run.n_steps = n_steps;
    run.results = results;
    run.initial_frame = initial_frame;
    run.output = argv[5];
    run.start = start;
    run.dt = dt;
    struct rank ranks[NB_MAX_RANKS];
    if (world == 1) {
        ranks[0].run = &run;
        ranks[0].id = 0;
        rank_main(&ranks[0]);
    } else {
        /* one engine per GPU, one host thread per engine (include/nbody_engine.h); devices
         * NB_DEVICE .. NB_DEVICE + world - 1 */
        if ((rc = nb_comm_id(run.comm_id, sizeof run.comm_id)))
            die_engine("nb_comm_id", rc);
        pthread_barrier_init(&run.barrier, NULL, (unsigned)world);
        pthread_t th[NB_MAX_RANKS];
        for (int r = 0; r < world; r++) {
            ranks[r].run = &run;
            ranks[r].id = r;
            if (pthread_create(&th[r], NULL, rank_main, &ranks[r]) != 0) {
                fprintf(stderr, "Error creating thread %d\n", r);
                exit(2);
            }
        }
        for (int r = 0; r < world; r++)
            pthread_join(th[r], NULL);
        pthread_barrier_destroy(&run.barrier);
    }
    clock_gettime(CLOCK_REALTIME, &e);

    env = getenv("NB_CHECKPOINT");
    if (env)
        write_checkpoint(env, n_ents, positions, velocities);

    printf("Completed. Output file: %s\n", argv[5]);
    double time_spent = (e.tv_sec - s.tv_sec) + (e.tv_nsec - s.tv_nsec) / BILLION;
    printf("Elapsed wall time: %f s\n", time_spent);

    if (time_spent > 0 && cfg.algorithm == NB_ALGO_DIRECT)
        printf("Pairwise interactions: %llu (%.3e /s), force launches: %llu%s\n", run.interactions,
               (double)run.interactions / time_spent, run.force_launches,
               world > 1 ? " per GPU" : "");
    free(positions);
    free(velocities);
    return 0;
}
