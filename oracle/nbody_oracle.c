/*
 * nbody_oracle.c -- TEST INFRASTRUCTURE ONLY (see nbody_oracle.h).
 *
 * Plain-C restatement of the reference's exhaustive path on flat arrays.
 * Every routine names the reference lines it follows; the arithmetic keeps
 * the reference's operation order so that, compiled without FMA contraction,
 * it reproduces the reference bit for bit (tests/test_oracle.py checks this
 * against oracle/_ref/, the reference's own sources compiled in place).
 *
 * Build: see oracle/Makefile (-O2 -ffp-contract=off -fopenmp).
 */
#include "nbody_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* serial/exhaustive.c:24 -- BIG_G is 1.0 in every program of the reference. */
static const double ORC_G = 1.0;
/* serial/exhaustive.c:171-172 -- Plummer softening eps^2, a source literal. */
static const double ORC_EPS2 = 0.01;

void orc_free(void *p) { free(p); }

/*
 * serial/exhaustive.c:79-101 then :110-146.
 * Pass 1 counts '\n' and adds one (files without a trailing newline); that
 * number only sizes the allocation.  Pass 2 scans records of seven doubles in
 * the order x,y,z,vx,vy,vz,mass; a comma must follow each number directly,
 * blanks after the comma are optional.  Scanning stops at the first record
 * that does not yield seven numbers; if any byte is left after that point the
 * file "is not well formed".  The body count is the number of full records.
 */
int orc_read_bodies(const char *path, unsigned *n_out, double **pos_xyzm, double **vel_xyz)
{
    FILE *f = fopen(path, "r");
    if (!f)
        return ORC_EFILE;
    size_t lines = 0;
    for (int c = getc(f); c != EOF; c = getc(f))
        lines += (c == '\n');
    lines += 1;
    rewind(f);

    double *p = malloc(lines * 4 * sizeof(double));
    double *v = malloc(lines * 3 * sizeof(double));
    if (!p || !v) {
        free(p);
        free(v);
        fclose(f);
        return ORC_ENOMEM;
    }
    size_t k = 0;
    double r[7];
    while (k < lines &&
           fscanf(f, "%lf, %lf, %lf, %lf, %lf, %lf, %lf\n", &r[0], &r[1], &r[2], &r[3], &r[4],
                  &r[5], &r[6]) == 7) {
        p[4 * k + 0] = r[0];
        p[4 * k + 1] = r[1];
        p[4 * k + 2] = r[2];
        p[4 * k + 3] = r[6];
        v[3 * k + 0] = r[3];
        v[3 * k + 1] = r[4];
        v[3 * k + 2] = r[5];
        k++;
    }
    int leftover = fgetc(f) != EOF;
    fclose(f);
    if (leftover) {
        free(p);
        free(v);
        return ORC_EFILE;
    }
    *n_out = (unsigned)k;
    *pos_xyzm = p;
    *vel_xyz = v;
    return ORC_OK;
}

/* serial/exhaustive.c:43-53: the three times are floats; the step count is
 * the float quotient truncated toward zero. */
int orc_parse_times(const char *start_s, const char *end_s, const char *dt_s, float *start,
                    float *end, float *dt)
{
    *start = strtof(start_s, NULL);
    *end = strtof(end_s, NULL);
    *dt = strtof(dt_s, NULL);
    int n_steps = (*end - *start) / *dt;
    return n_steps;
}

/* One row of serial/exhaustive.c:155-181: a_i = sum_j G * d * (d.d + eps2)^-1.5 * m_j,
 * d = x_j - x_i, summed in j order, including j == i (which adds exactly 0). */
static inline void orc_row(unsigned n, const double *pm, unsigned i, double *a)
{
    const double xi = pm[4 * i], yi = pm[4 * i + 1], zi = pm[4 * i + 2];
    double ax = 0, ay = 0, az = 0;
    for (unsigned j = 0; j < n; j++) {
        double dx = pm[4 * j] - xi;
        double dy = pm[4 * j + 1] - yi;
        double dz = pm[4 * j + 2] - zi;
        double w = dx * dx + dy * dy + dz * dz + ORC_EPS2;
        w = pow(w, -1.5);
        ax += ORC_G * dx * w * pm[4 * j + 3];
        ay += ORC_G * dy * w * pm[4 * j + 3];
        az += ORC_G * dz * w * pm[4 * j + 3];
    }
    a[0] = ax;
    a[1] = ay;
    a[2] = az;
}

void orc_acceleration(unsigned n, const double *pm, double *acc)
{
    for (unsigned i = 0; i < n; i++)
        orc_row(n, pm, i, acc + 3 * (size_t)i);
}

void orc_acceleration_slice(unsigned n, const double *pm, unsigned i0, unsigned i1, double *acc)
{
    for (unsigned i = i0; i < i1; i++)
        orc_row(n, pm, i, acc + 3 * (size_t)(i - i0));
}

/* openmp/exhaustive.c:166-200: rows are independent, so any row partition
 * gives the same bits as the serial loop. */
void orc_acceleration_mt(unsigned n, const double *pm, double *acc, int threads)
{
    if (threads < 1)
        threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (long i = 0; i < (long)n; i++)
        orc_row(n, pm, (unsigned)i, acc + 3 * (size_t)i);
}

/* The reference's RESULTS frame line, serial/exhaustive.c:212-214, 235-237. */
static void orc_write_frame(FILE *fp, unsigned n, const double *pm)
{
    for (size_t i = 0; i < n; i++)
        fprintf(fp, "%lu,%lf,%lf,%lf,%lf\n", (unsigned long)i, pm[4 * i], pm[4 * i + 1],
                pm[4 * i + 2], pm[4 * i + 3]);
}

/*
 * serial/exhaustive.c:192-255 (threads <= 1) and openmp/exhaustive.c:212-287.
 * a = F(x); repeat n_steps times { v += a*dt/2 ; x += v*dt ; [frame] ;
 * a = F(x) ; v += a*dt/2 }.  dt is a float promoted to double in every
 * product; the half kick is (a*dt)/2.0, two rounded operations.
 */
int orc_propagate(unsigned n, double *pm, double *v, int n_steps, float dt,
                  const char *frames_path, int threads)
{
    double *acc = malloc((size_t)n * 3 * sizeof(double));
    if (!acc)
        return ORC_ENOMEM;
    FILE *fp = NULL;
    if (frames_path) {
        fp = fopen(frames_path, "w");
        if (!fp) {
            free(acc);
            return ORC_EFILE;
        }
    }
    orc_acceleration_mt(n, pm, acc, threads);
    if (fp)
        orc_write_frame(fp, n, pm);
    for (int t = 0; t < n_steps; t++) {
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                v[3 * i + k] += acc[3 * i + k] * dt / 2.0;
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                pm[4 * i + k] += v[3 * i + k] * dt;
        if (fp)
            orc_write_frame(fp, n, pm);
        orc_acceleration_mt(n, pm, acc, threads);
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                v[3 * i + k] += acc[3 * i + k] * dt / 2.0;
    }
    if (fp)
        fclose(fp);
    free(acc);
    return ORC_OK;
}

/*
 * common/get_energy.c:7-47.  KE = 0.5 * sum m |v|^2 ; PE = -G * sum_{i<=j}
 * m_i m_j / r_ij with NO softening; the r == 0 diagonal contributes 0 through
 * `val > 0 ? 1/val : val` (:38-40).  Loop and accumulation order as in the
 * reference.
 */
void orc_get_energy_flat(const double *pos, const double *vel, const double *mass, double G,
                         int sz, double *KE, double *PE)
{
    double ke = 0;
    for (int i = 0; i < sz; i++) {
        double s = 0;
        for (int a = 0; a < 3; a++)
            s += vel[3 * i + a] * vel[3 * i + a];
        ke += mass[i] * s;
    }
    *KE = ke * 0.5;
    double pe = 0;
    for (int i = 0; i < sz; i++) {
        for (int j = i; j < sz; j++) {
            double s = 0;
            for (int a = 0; a < 3; a++) {
                double d = pos[3 * j + a] - pos[3 * i + a];
                s += d * d;
            }
            double val = sqrt(s);
            val = val > 0 ? 1.0 / val : val;
            val *= -mass[i] * mass[j];
            pe += val;
        }
    }
    *PE = pe * G;
}

void orc_get_energy(unsigned n, const double *pm, const double *vel, double G, double *KE,
                    double *PE)
{
    double *pos = malloc((size_t)n * 3 * sizeof(double));
    double *mass = malloc((size_t)n * sizeof(double));
    for (size_t i = 0; i < n; i++) {
        pos[3 * i] = pm[4 * i];
        pos[3 * i + 1] = pm[4 * i + 1];
        pos[3 * i + 2] = pm[4 * i + 2];
        mass[i] = pm[4 * i + 3];
    }
    orc_get_energy_flat(pos, vel, mass, G, (int)n, KE, PE);
    free(pos);
    free(mass);
}

/* FP32 model of the same sum (no reference counterpart): float inputs and
 * per-pair float arithmetic with a correctly rounded float r^-3/2 obtained
 * through double, double accumulation. */
void orc_acceleration_f32(unsigned n, const double *pm, double *acc)
{
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; i++) {
        float xi = (float)pm[4 * i], yi = (float)pm[4 * i + 1], zi = (float)pm[4 * i + 2];
        double ax = 0, ay = 0, az = 0;
        for (unsigned j = 0; j < n; j++) {
            float dx = (float)pm[4 * j] - xi;
            float dy = (float)pm[4 * j + 1] - yi;
            float dz = (float)pm[4 * j + 2] - zi;
            float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, (float)ORC_EPS2)));
            float ri = (float)(1.0 / sqrt((double)r2));
            float s = (ri * ri) * ((float)pm[4 * j + 3] * ri);
            ax += (double)(dx * s);
            ay += (double)(dy * s);
            az += (double)(dz * s);
        }
        acc[3 * i] = ax;
        acc[3 * i + 1] = ay;
        acc[3 * i + 2] = az;
    }
}
