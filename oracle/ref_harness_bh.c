/*
 * ref_harness_bh.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Wraps the reference's OWN serial Barnes-Hut source (REF_SRC =
 * /root/reference/serial/barnes-hut.c), compiled in place, so tests can ask it
 * for accelerations, tree statistics and propagated states at full precision.
 * Built by oracle/Makefile into oracle/_ref/libref_bh_serial.so.
 */
#define main ref_program_main
#include REF_SRC
#undef main

#include <string.h>

static Entity *bh_pack(int n, const double *pm, const double *v)
{
    Entity *e = calloc(n, sizeof(Entity));
    for (int i = 0; i < n; i++) {
        e[i].pos.x = pm[4 * i];
        e[i].pos.y = pm[4 * i + 1];
        e[i].pos.z = pm[4 * i + 2];
        e[i].mass = pm[4 * i + 3];
        if (v) {
            e[i].vel.x = v[3 * i];
            e[i].vel.y = v[3 * i + 1];
            e[i].vel.z = v[3 * i + 2];
        }
    }
    return e;
}

static int bh_depth(Octtree *t, int node)
{
    Octnode *nd = &t->nodes[node];
    if (nd->ents == 1)
        return 0;
    int d = 0;
    for (int c = 0; c < 8; c++)
        if (nd->children[c] != -1) {
            int k = bh_depth(t, nd->children[c]);
            d = k > d ? k : d;
        }
    return d + 1;
}

/* One force evaluation exactly as propagation() does it
 * (serial/barnes-hut.c:470-473): bbox, insert all, COM, traversal.
 * stats[0] = internal node count, stats[1] = tree depth, stats[2] = tree->max. */
void ref_bh_acceleration(int n, const double *pm, double *acc_xyz, double *stats)
{
    Entity *e = bh_pack(n, pm, NULL);
    Octtree tree;
    create_tree(n, &tree);
    RVec3 *acc = malloc((size_t)n * sizeof(RVec3));
    get_bounding_box(e, n, &tree.max);
    add_ents(&tree, e, n);
    center_of_mass(&tree, &tree.nodes[tree.root]);
    get_acceleration(&tree, acc, n);
    memcpy(acc_xyz, acc, (size_t)n * sizeof(RVec3));
    if (stats) {
        stats[0] = tree.firstfree - tree.root;
        stats[1] = bh_depth(&tree, tree.root);
        stats[2] = tree.max;
    }
    free(tree.nodes);
    free(acc);
    free(e);
}

/* Dump the internal nodes (mass, COM, body count) in the reference's storage
 * order, for cell-set comparisons.  out has room for cap rows of 5 doubles. */
int ref_bh_nodes(int n, const double *pm, double *out, int cap)
{
    Entity *e = bh_pack(n, pm, NULL);
    Octtree tree;
    create_tree(n, &tree);
    get_bounding_box(e, n, &tree.max);
    add_ents(&tree, e, n);
    center_of_mass(&tree, &tree.nodes[tree.root]);
    int cnt = tree.firstfree - tree.root;
    for (int k = 0; k < cnt && k < cap; k++) {
        Octnode *nd = &tree.nodes[tree.root + k];
        out[5 * k] = nd->mass;
        out[5 * k + 1] = nd->center.x;
        out[5 * k + 2] = nd->center.y;
        out[5 * k + 3] = nd->center.z;
        out[5 * k + 4] = nd->ents;
    }
    free(tree.nodes);
    free(e);
    return cnt;
}

/* propagation() -- serial/barnes-hut.c:445-517 */
void ref_bh_propagate(int n, double *pm, double *v, int n_steps, float dt, const char *output)
{
    Entity *e = bh_pack(n, pm, v);
    propagation(e, n, n_steps, dt, output);
    for (int i = 0; i < n; i++) {
        pm[4 * i] = e[i].pos.x;
        pm[4 * i + 1] = e[i].pos.y;
        pm[4 * i + 2] = e[i].pos.z;
        v[3 * i] = e[i].vel.x;
        v[3 * i + 1] = e[i].vel.y;
        v[3 * i + 2] = e[i].vel.z;
    }
    free(e);
}
