/*
 * ref_harness_energy.c -- TEST INFRASTRUCTURE ONLY.
 * Compiles the reference's common/get_energy.c in place (REF_SRC) and exports
 * its get_energy() plus its -DTEST_GET_ENERGY known-answer inputs' result.
 * Built by oracle/Makefile into oracle/_ref/libref_energy.so.
 */
#include REF_SRC

void ref_get_energy(double *pos, double *vel, double *mass, double G, int sz, double *KE,
                    double *PE)
{
    get_energy(pos, vel, mass, G, sz, KE, PE);
}
