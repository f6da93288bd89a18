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
 *   NB_CHECKPOINT_EVERY=<k>: also every k steps (no-RESULTS runs), written to <file>.tmp and
 *   renamed, so a restart (`input = <file>`, start_time = the time reached) never sees a torn file
 */
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
    if (argc != 6 && argc != 8) {
        fprintf(stderr,
                "Usage: %s input_filename start_time end_time delta_time output_filename "
                "[grid_size block_size]\n",
                argv[0]);
        exit(1);
    }
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

    struct timespec s, e;
    clock_gettime(CLOCK_REALTIME, &s);

    nb_engine *eng = NULL;
    rc = nb_init(&eng, &cfg, n_ents, positions, velocities);
    if (rc)
        die_engine("nb_init", rc);

    FILE *fpt = NULL;
    if (results) {
        fpt = fopen(argv[5], "w");
        if (!fpt) {
            fprintf(stderr, "Error opening file '%s'\n", argv[5]);
            exit(1);
        }
        if (initial_frame)
            nbh_write_frame(fpt, n_ents, positions); /* frame 0 = initial state */
        /* The reference reads every frame back with a blocking copy and formats it
         * while the GPU idles (cuda/exhaustive.cu:135-165).  Here step t+1 is
         * enqueued before frame t is waited for, so the kernel, the copy (side
         * stream, page-locked buffers) and the formatting overlap. */
        double *frame[2] = {NULL, NULL};
        for (int k = 0; k < 2; k++)
            if ((rc = nb_host_alloc((void **)&frame[k], (n_ents ? n_ents : 1) * 4 * sizeof(double))))
                die_engine("nb_host_alloc", rc);
        for (int t = 0; t <= n_steps; t++) {
            if (t < n_steps) {
                if ((rc = nb_step(eng, 1)))
                    die_engine("nb_step", rc);
                if ((rc = nb_frame_begin(eng, frame[t & 1])))
                    die_engine("nb_frame_begin", rc);
            }
            if (t > 0) {
                if ((rc = nb_frame_wait(eng, frame[(t - 1) & 1])))
                    die_engine("nb_frame_wait", rc);
                nbh_write_frame(fpt, n_ents, frame[(t - 1) & 1]);
            }
        }
        nb_host_free(frame[0]);
        nb_host_free(frame[1]);
        fclose(fpt);
    } else {
        const char *ck = getenv("NB_CHECKPOINT"), *ev = getenv("NB_CHECKPOINT_EVERY");
        const int every = (ck && ev) ? atoi(ev) : 0;
        for (int done = 0; done < n_steps;) {
            const int k = (every > 0 && every < n_steps - done) ? every : n_steps - done;
            if ((rc = nb_step(eng, k)))
                die_engine("nb_step", rc);
            done += k;
            if (done < n_steps) { /* intermediate checkpoint; the final one is written below */
                if ((rc = nb_gather(eng, positions, velocities)))
                    die_engine("nb_gather", rc);
                if (write_checkpoint(ck, n_ents, positions, velocities) == 0)
                    fprintf(stderr, "checkpoint after step %d (t = %f): %s\n", done,
                            (double)(start + (float)done * dt), ck);
            }
        }
    }
    /* closing half kick + read-back of the final state */
    if ((rc = nb_gather(eng, positions, velocities)))
        die_engine("nb_gather", rc);
    clock_gettime(CLOCK_REALTIME, &e);

    env = getenv("NB_CHECKPOINT");
    if (env)
        write_checkpoint(env, n_ents, positions, velocities);

    printf("Completed. Output file: %s\n", argv[5]);
    double time_spent = (e.tv_sec - s.tv_sec) + (e.tv_nsec - s.tv_nsec) / BILLION;
    printf("Elapsed wall time: %f s\n", time_spent);

    nb_stats st;
    if (nb_get_stats(eng, &st) == NB_OK && time_spent > 0 && cfg.algorithm == NB_ALGO_DIRECT)
        printf("Pairwise interactions: %llu (%.3e /s), force launches: %llu\n", st.interactions,
               (double)st.interactions / time_spent, st.force_launches);
    nb_free(eng);
    free(positions);
    free(velocities);
    return 0;
}
