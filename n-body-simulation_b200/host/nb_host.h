/*
 * nb_host.h -- host-side I/O of the N-body programs, in C like the reference's.
 *
 * Keeps the reference's process-level contract: the body CSV
 * (x, y, z, vx, vy, vz, mass per line; serial/exhaustive.c:130-132), the
 * float-typed times and step count (serial/exhaustive.c:43-53), the frame
 * file ("%lu,%lf,%lf,%lf,%lf\n" per body per frame; serial/exhaustive.c:212-214,
 * 235-237) and its exit codes.  None of it touches the GPU.
 */
#ifndef NB_HOST_H
#define NB_HOST_H

#include <stddef.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NBH_OK 0
#define NBH_EOPEN 1     /* "Error opening file"           -> exit(1), serial/exhaustive.c:84-86 */
#define NBH_EFORMAT 3   /* "file is not well formed"      -> exit(1), serial/exhaustive.c:136-142 */
#define NBH_ENOMEM 2    /* "Error during memory allocation" -> exit(2), serial/exhaustive.c:124-127 */

/* Reads a body file into the packed arrays of the reference's CUDA host
 * (cuda/exhaustive.cu:240-251): pos_xyzm[4n], vel_xyz[3n], malloc'ed here,
 * freed with nbh_free.  Acceptance rules are those of the reference's
 * fscanf("%lf, %lf, %lf, %lf, %lf, %lf, %lf\n") loop followed by its
 * fgetc() != EOF check: a comma directly after each of the first six numbers,
 * optional blanks after the comma, optional trailing newline, any unparsed
 * byte left -> NBH_EFORMAT; an incomplete last record that runs into EOF is
 * dropped silently, as the reference does. */
int nbh_read_bodies(const char *path, size_t *n, double **pos_xyzm, double **vel_xyz);
int nbh_parse_bodies(const char *text, size_t len, size_t *n, double **pos_xyzm,
                     double **vel_xyz);
void nbh_free(void *p);

/* start/end/dt are floats and the step count is their float quotient truncated
 * to int (serial/exhaustive.c:49-53).  Returns n_steps. */
int nbh_parse_times(const char *start_s, const char *end_s, const char *dt_s, float *start,
                    float *end, float *dt);

/* Appends one frame: "%lu,%lf,%lf,%lf,%lf\n" = index,x,y,z,mass per body. */
int nbh_write_frame(FILE *fp, size_t n, const double *pos_xyzm);
/* Formats one frame into buf (capacity cap); returns bytes written or 0 if it
 * does not fit.  Byte-identical to the fprintf form. */
size_t nbh_format_frame(char *buf, size_t cap, size_t n, const double *pos_xyzm);
/* "%lf" of one double, exactly as printf rounds it; returns length. */
size_t nbh_format_lf(char *out, double v);

/* Writes bodies in the INPUT format (x,y,z,vx,vy,vz,mass with %.17g), so a
 * gathered state can be fed back as an input file (restart). */
int nbh_write_bodies(const char *path, size_t n, const double *pos_xyzm, const double *vel_xyz);

#ifdef __cplusplus
}
#endif
#endif
