/*
 * bh_oracle.c -- TEST INFRASTRUCTURE ONLY (see nbody_oracle.h).
 *
 * CPU restatement of the reference's serial Barnes-Hut force evaluation
 * (serial/barnes-hut.c).  The reference inserts bodies one at a time into a
 * growing octree (add_ent, :169-263); the cell set that results depends only
 * on the positions (a branch exists iff its cell holds >= 2 bodies, every
 * leaf is one body), so this restatement builds the same tree top-down by
 * recursive octant partitioning and then applies the reference's
 * centre-of-mass recurrence (:285-317) and traversal (:389-434) unchanged in
 * arithmetic and visiting order -- accelerations come out bit-identical
 * (tests/test_oracle.py checks against oracle/_ref/libref_bh_serial.so).
 *
 * Parity status: PINNED by that bitwise comparison and tests/golden/bh_*.npz.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "nbody_oracle.h"

#define BH_G 1.0      /* serial/barnes-hut.c:39 */
#define BH_THETA 0.5  /* serial/barnes-hut.c:40 */
#define BH_EPS2 0.01  /* serial/barnes-hut.c:372 */

typedef struct {
    double mass, cx, cy, cz; /* total mass, centre of mass */
    int count;               /* bodies below */
    int child[8];            /* >= 0: body index; <= -2: cell index -(c+2); -1 empty */
} bh_cell;

typedef struct {
    bh_cell *cells;
    int n_cells, cap;
    const double *pm;
    int max_depth;
} bh_tree;

static int bh_new_cell(bh_tree *t)
{
    if (t->n_cells == t->cap) {
        t->cap = t->cap ? t->cap * 2 : 1024;
        t->cells = realloc(t->cells, (size_t)t->cap * sizeof(bh_cell));
    }
    bh_cell *c = &t->cells[t->n_cells];
    memset(c, 0, sizeof(*c));
    for (int k = 0; k < 8; k++)
        c->child[k] = -1;
    return t->n_cells++;
}

/* Octant rule of get_indx_loc (serial/barnes-hut.c:114-133): `>=` goes high,
 * index = z*4 + y*2 + x; the child centre is centre +- border/4. */
static int bh_build(bh_tree *t, int *idx, int cnt, double cx, double cy, double cz, double border,
                    int depth)
{
    int me = bh_new_cell(t);
    if (depth > t->max_depth)
        t->max_depth = depth;
    t->cells[me].count = cnt;
    int num[8] = {0}, start[8];
    int *oct = malloc((size_t)cnt * sizeof(int));
    int *tmp = malloc((size_t)cnt * sizeof(int));
    for (int k = 0; k < cnt; k++) {
        const double *p = t->pm + 4 * (size_t)idx[k];
        int o = (p[2] >= cz) * 4 + (p[1] >= cy) * 2 + (p[0] >= cx);
        oct[k] = o;
        num[o]++;
    }
    start[0] = 0;
    for (int o = 1; o < 8; o++)
        start[o] = start[o - 1] + num[o - 1];
    int fill[8];
    memcpy(fill, start, sizeof(fill));
    for (int k = 0; k < cnt; k++)
        tmp[fill[oct[k]]++] = idx[k];
    memcpy(idx, tmp, (size_t)cnt * sizeof(int));
    free(oct);
    free(tmp);
    double q = border / 4;
    for (int o = 0; o < 8; o++) {
        if (num[o] == 0)
            continue;
        if (num[o] == 1) {
            t->cells[me].child[o] = idx[start[o]];
        } else {
            double ncx = cx + ((o & 1) ? q : -q);
            double ncy = cy + ((o & 2) ? q : -q);
            double ncz = cz + ((o & 4) ? q : -q);
            int c = bh_build(t, idx + start[o], num[o], ncx, ncy, ncz, border / 2, depth + 1);
            t->cells[me].child[o] = -(c + 2);
        }
    }
    return me;
}

/* center_of_mass (serial/barnes-hut.c:285-317): post-order; the node starts at
 * mass 0, centre 0 and folds its children in octant order with the
 * incremental weighted mean written exactly as the reference writes it. */
static void bh_com(bh_tree *t, int ci)
{
    for (int o = 0; o < 8; o++) {
        int ch = t->cells[ci].child[o];
        if (ch <= -2)
            bh_com(t, -ch - 2);
    }
    bh_cell *c = &t->cells[ci];
    for (int o = 0; o < 8; o++) {
        int ch = c->child[o];
        if (ch == -1)
            continue;
        double m, x, y, z;
        if (ch >= 0) {
            const double *p = t->pm + 4 * (size_t)ch;
            x = p[0], y = p[1], z = p[2], m = p[3];
        } else {
            bh_cell *k = &t->cells[-ch - 2];
            x = k->cx, y = k->cy, z = k->cz, m = k->mass;
        }
        double nm = c->mass + m;
        c->cx = (x * m / nm) + (c->cx * c->mass / nm);
        c->cy = (y * m / nm) + (c->cy * c->mass / nm);
        c->cz = (z * m / nm) + (c->cz * c->mass / nm);
        c->mass = nm;
    }
}

typedef struct {
    long interactions, visits;
} bh_counts;

/* calculate_acceleration (serial/barnes-hut.c:363-378) */
static inline void bh_pair(const double *pi, double x, double y, double z, double m, double *a)
{
    double dx = x - pi[0], dy = y - pi[1], dz = z - pi[2];
    double w = dx * dx + dy * dy + dz * dz + BH_EPS2;
    w = pow(w, -1.5);
    a[0] += BH_G * dx * w * m;
    a[1] += BH_G * dy * w * m;
    a[2] += BH_G * dz * w * m;
}

/* get_distance (serial/barnes-hut.c:101-104) */
static inline double bh_dist(const double *pi, double x, double y, double z)
{
    return sqrt(pow(x - pi[0], 2) + pow(y - pi[1], 2) + pow(z - pi[2], 2));
}

/* get_acceleration_rec (serial/barnes-hut.c:389-416): accept the cell when
 * border/dist(COM) < THETA, else open it and visit children 0..7, skipping
 * the body's own leaf by index.  Leaves are always accepted. */
static void bh_walk(const bh_tree *t, int ci, int id, const double *pi, double border, double *a,
                    bh_counts *cn)
{
    const bh_cell *c = &t->cells[ci];
    cn->visits++;
    double d = bh_dist(pi, c->cx, c->cy, c->cz);
    if (border / d < BH_THETA) {
        bh_pair(pi, c->cx, c->cy, c->cz, c->mass, a);
        cn->interactions++;
        return;
    }
    for (int o = 0; o < 8; o++) {
        int ch = c->child[o];
        if (ch == -1 || ch == id)
            continue;
        if (ch >= 0) {
            const double *p = t->pm + 4 * (size_t)ch;
            cn->visits++;
            bh_pair(pi, p[0], p[1], p[2], p[3], a);
            cn->interactions++;
        } else {
            bh_walk(t, -ch - 2, id, pi, border / 2, a, cn);
        }
    }
}

/* get_bounding_box (serial/barnes-hut.c:326-338): 2 * max |coordinate|. */
double orc_bh_bounding_box(unsigned n, const double *pm)
{
    double mx = 0.0;
    for (size_t i = 0; i < n; i++)
        for (int k = 0; k < 3; k++)
            mx = fabs(pm[4 * i + k]) > mx ? fabs(pm[4 * i + k]) : mx;
    return mx * 2;
}

/*
 * One Barnes-Hut force evaluation as serial/barnes-hut.c:470-473 performs it.
 * stats (optional, 5 doubles): internal cells, depth, root border,
 * interactions per body, node visits per body.
 */
void orc_bh_acceleration(unsigned n, const double *pm, double *acc_xyz, double *stats)
{
    bh_tree t = {0};
    t.pm = pm;
    int *idx = malloc((size_t)n * sizeof(int));
    for (unsigned i = 0; i < n; i++)
        idx[i] = (int)i;
    double max = orc_bh_bounding_box(n, pm);
    bh_counts total = {0, 0};
    if (n >= 2) {
        bh_build(&t, idx, (int)n, 0.0, 0.0, 0.0, max, 1);
        bh_com(&t, 0);
#pragma omp parallel
        {
            bh_counts cn = {0, 0};
#pragma omp for schedule(dynamic, 64)
            for (long i = 0; i < (long)n; i++) {
                double a[3] = {0, 0, 0};
                bh_walk(&t, 0, (int)i, pm + 4 * i, max, a, &cn);
                acc_xyz[3 * i] = a[0];
                acc_xyz[3 * i + 1] = a[1];
                acc_xyz[3 * i + 2] = a[2];
            }
#pragma omp critical
            {
                total.interactions += cn.interactions;
                total.visits += cn.visits;
            }
        }
    } else if (n == 1) {
        acc_xyz[0] = acc_xyz[1] = acc_xyz[2] = 0.0;
    }
    if (stats) {
        stats[0] = t.n_cells;
        stats[1] = t.max_depth;
        stats[2] = max;
        stats[3] = n ? (double)total.interactions / n : 0;
        stats[4] = n ? (double)total.visits / n : 0;
    }
    free(t.cells);
    free(idx);
}

/* propagation (serial/barnes-hut.c:445-517): same KDK loop, tree rebuilt after
 * every drift. */
int orc_bh_propagate(unsigned n, double *pm, double *v, int n_steps, float dt)
{
    double *acc = malloc((size_t)n * 3 * sizeof(double));
    if (!acc)
        return ORC_ENOMEM;
    orc_bh_acceleration(n, pm, acc, NULL);
    for (int s = 0; s < n_steps; s++) {
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                v[3 * i + k] += acc[3 * i + k] * dt / 2.0;
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                pm[4 * i + k] += v[3 * i + k] * dt;
        orc_bh_acceleration(n, pm, acc, NULL);
        for (size_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                v[3 * i + k] += acc[3 * i + k] * dt / 2.0;
    }
    free(acc);
    return ORC_OK;
}
