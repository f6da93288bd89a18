/*
 * ref_harness_exh.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Wraps the reference's OWN exhaustive sources, compiled where they lie under
 * /root/reference (never copied into this repo), so tests can call its
 * functions in-process at full precision (the reference's output file keeps
 * only 6 decimals).  Built by oracle/Makefile into oracle/_ref/ as
 *   libref_exh_serial.so  (-DREF_SRC=".../serial/exhaustive.c")
 *   libref_exh_omp.so     (-DREF_SRC=".../openmp/exhaustive.c" -DREF_OMP -fopenmp)
 * The reference's main() is renamed out of the way; everything else is used
 * unmodified.
 */
#define main ref_program_main
#include REF_SRC
#undef main

#include <string.h>

#ifdef REF_OMP
void ref_set_threads(int t) { thread_count = t; }
#else
void ref_set_threads(int t) { (void)t; }
#endif

/* get_entities (serial/exhaustive.c:110, openmp/exhaustive.c:111) -> packed arrays.
 * NOTE: the reference exit()s on a bad file; tests exercise that through the
 * stand-alone binaries instead. */
int ref_read_bodies(const char *path, unsigned *n, double **pos_xyzm, double **vel_xyz)
{
    Entity *ents;
    uint cnt;
#ifdef REF_OMP
    RVec3 *vel;
    get_entities((char *)path, &ents, &vel, &cnt);
#else
    get_entities((char *)path, &ents, &cnt);
#endif
    double *p = malloc((size_t)cnt * 4 * sizeof(double));
    double *v = malloc((size_t)cnt * 3 * sizeof(double));
    for (size_t i = 0; i < cnt; i++) {
        p[4 * i] = ents[i].pos.x;
        p[4 * i + 1] = ents[i].pos.y;
        p[4 * i + 2] = ents[i].pos.z;
        p[4 * i + 3] = ents[i].mass;
#ifdef REF_OMP
        v[3 * i] = vel[i].x;
        v[3 * i + 1] = vel[i].y;
        v[3 * i + 2] = vel[i].z;
#else
        v[3 * i] = ents[i].vel.x;
        v[3 * i + 1] = ents[i].vel.y;
        v[3 * i + 2] = ents[i].vel.z;
#endif
    }
    free(ents);
#ifdef REF_OMP
    free(vel);
#endif
    *n = cnt;
    *pos_xyzm = p;
    *vel_xyz = v;
    return 0;
}

void ref_free(void *p) { free(p); }

static Entity *pack_in(unsigned n, const double *pm, const double *v)
{
    Entity *e = calloc(n, sizeof(Entity));
    for (size_t i = 0; i < n; i++) {
        e[i].pos.x = pm[4 * i];
        e[i].pos.y = pm[4 * i + 1];
        e[i].pos.z = pm[4 * i + 2];
        e[i].mass = pm[4 * i + 3];
#ifndef REF_OMP
        if (v) {
            e[i].vel.x = v[3 * i];
            e[i].vel.y = v[3 * i + 1];
            e[i].vel.z = v[3 * i + 2];
        }
#endif
    }
    (void)v;
    return e;
}

/* acceleration() -- serial/exhaustive.c:155 / openmp/exhaustive.c:166 */
void ref_acceleration(unsigned n, const double *pm, double *acc_xyz)
{
    Entity *e = pack_in(n, pm, NULL);
    RVec3 *a = malloc((size_t)n * sizeof(RVec3));
#ifdef REF_OMP
#pragma omp parallel num_threads(thread_count)
#endif
    acceleration(n, e, a);
    memcpy(acc_xyz, a, (size_t)n * sizeof(RVec3));
    free(a);
    free(e);
}

/* propagation() -- serial/exhaustive.c:192 / openmp/exhaustive.c:212.
 * The frame file is only written if the harness was built with -DRESULTS. */
void ref_propagate(unsigned n, double *pm, double *v, int n_steps, float dt, const char *output)
{
    Entity *e = pack_in(n, pm, v);
#ifdef REF_OMP
    RVec3 *vel = malloc((size_t)n * sizeof(RVec3));
    memcpy(vel, v, (size_t)n * sizeof(RVec3));
    propagation(e, vel, n, n_steps, dt, output);
    memcpy(v, vel, (size_t)n * sizeof(RVec3));
    free(vel);
#else
    propagation(e, n, n_steps, dt, output);
#endif
    for (size_t i = 0; i < n; i++) {
        pm[4 * i] = e[i].pos.x;
        pm[4 * i + 1] = e[i].pos.y;
        pm[4 * i + 2] = e[i].pos.z;
#ifndef REF_OMP
        v[3 * i] = e[i].vel.x;
        v[3 * i + 1] = e[i].vel.y;
        v[3 * i + 2] = e[i].vel.z;
#endif
    }
    free(e);
}
