/*
 * nb_host.c -- body-file reader, frame writer and time parsing of the host
 * programs (see nb_host.h for the reference lines each routine follows).
 */
#include "nb_host.h"
#ifdef _OPENMP
#include <omp.h>
#endif

#include <ctype.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void nbh_free(void *p) { free(p); }

/* ------------------------------------------------------------- reader ---- */

static size_t skip_ws(const char *t, size_t p, size_t len)
{
    while (p < len && isspace((unsigned char)t[p]))
        p++;
    return p;
}

/*
 * One pass over text[p, len).  A record is seven numbers; what may sit between
 * them is exactly what the reference's scanf format allows: "<num>," then
 * optional white space before the next number, and optional white space after
 * the seventh.  `p` always points at the first byte scanf would not have
 * consumed, so "bytes left at p" is the reference's fgetc() != EOF test.
 * *partial: numbers of an unfinished record consumed at the end of the range.
 */
static int parse_range(const char *text, size_t p, size_t len, size_t *n_out, double **pos_out,
                       double **vel_out, int *partial)
{
    size_t cap = 1024, n = 0;
    double *pos = malloc(cap * 4 * sizeof(double));
    double *vel = malloc(cap * 3 * sizeof(double));
    *partial = 0;
    *pos_out = *vel_out = NULL;
    if (!pos || !vel) {
        free(pos);
        free(vel);
        return NBH_ENOMEM;
    }
    for (;;) {
        double r[7];
        int got = 0;
        for (int k = 0; k < 7; k++) {
            p = skip_ws(text, p, len); /* %lf skips leading white space */
            if (p >= len)
                break;
            char *end;
            double v = strtod(text + p, &end);
            if (end == text + p)
                break; /* matching failure: nothing consumed */
            p = (size_t)(end - text);
            r[k] = v;
            got++;
            if (k < 6) {
                if (p >= len || text[p] != ',')
                    break; /* the comma must follow the number directly */
                p++;
            }
        }
        if (got != 7) {
            *partial = got;
            break;
        }
        p = skip_ws(text, p, len); /* the "\n" directive: any white space */
        if (n == cap) {
            cap *= 2;
            double *np = realloc(pos, cap * 4 * sizeof(double));
            double *nv = np ? realloc(vel, cap * 3 * sizeof(double)) : NULL;
            if (!np || !nv) {
                free(np ? np : pos);
                free(vel);
                return NBH_ENOMEM;
            }
            pos = np;
            vel = nv;
        }
        pos[4 * n + 0] = r[0];
        pos[4 * n + 1] = r[1];
        pos[4 * n + 2] = r[2];
        pos[4 * n + 3] = r[6];
        vel[3 * n + 0] = r[3];
        vel[3 * n + 1] = r[4];
        vel[3 * n + 2] = r[5];
        n++;
    }
    if (p < len) {
        free(pos);
        free(vel);
        return NBH_EFORMAT;
    }
    *n_out = n;
    *pos_out = pos;
    *vel_out = vel;
    return NBH_OK;
}

/*
 * Large inputs are cut after a line feed into one range per thread and the
 * ranges parsed side by side.  The scanf grammar does not care about lines, so
 * a cut can fall inside a record ("1,2,3,\n4,5,6,7"): a range that is not a
 * whole number of records (or fails in any way) sends the whole text through
 * the serial parser, whose verdict is the reference's.  When every range is
 * clean, the serial parser would have reached each cut at a record boundary
 * too (a number cannot span a line feed), so the result is the same.
 */
#define NBH_PAR_MIN_BYTES (4u << 20)
#define NBH_PAR_MAX 64

int nbh_parse_bodies(const char *text, size_t len, size_t *n_out, double **pos_out,
                     double **vel_out)
{
    int partial, nt = 1;
#ifdef _OPENMP
    nt = omp_get_max_threads();
#endif
    if (nt > NBH_PAR_MAX)
        nt = NBH_PAR_MAX;
    if (len / NBH_PAR_MIN_BYTES < (size_t)nt)
        nt = (int)(len / NBH_PAR_MIN_BYTES);
    if (nt >= 2) {
        size_t cut[NBH_PAR_MAX + 1], cnt[NBH_PAR_MAX];
        double *ppos[NBH_PAR_MAX], *pvel[NBH_PAR_MAX];
        int rcs[NBH_PAR_MAX], part[NBH_PAR_MAX];
        cut[0] = 0;
        cut[nt] = len;
        for (int t = 1; t < nt; t++) {
            size_t c = len / (size_t)nt * (size_t)t;
            if (c < cut[t - 1])
                c = cut[t - 1];
            const char *nl = memchr(text + c, '\n', len - c);
            cut[t] = nl ? (size_t)(nl - text) + 1 : len;
        }
#ifdef _OPENMP
#pragma omp parallel for num_threads(nt) schedule(static, 1)
#endif
        for (int t = 0; t < nt; t++)
            rcs[t] = parse_range(text, cut[t], cut[t + 1], &cnt[t], &ppos[t], &pvel[t], &part[t]);
        int clean = 1;
        size_t total = 0;
        for (int t = 0; t < nt; t++) {
            if (rcs[t] != NBH_OK || (part[t] && t < nt - 1))
                clean = 0;
            else
                total += cnt[t];
        }
        double *pos = NULL, *vel = NULL;
        if (clean) {
            pos = malloc((total ? total : 1) * 4 * sizeof(double));
            vel = malloc((total ? total : 1) * 3 * sizeof(double));
            if (!pos || !vel) {
                free(pos);
                free(vel);
                pos = vel = NULL;
                clean = 0;
            }
        }
        if (clean) {
            size_t off[NBH_PAR_MAX];
            off[0] = 0;
            for (int t = 1; t < nt; t++)
                off[t] = off[t - 1] + cnt[t - 1];
#ifdef _OPENMP
#pragma omp parallel for num_threads(nt) schedule(static, 1)
#endif
            for (int t = 0; t < nt; t++) {
                memcpy(pos + 4 * off[t], ppos[t], cnt[t] * 4 * sizeof(double));
                memcpy(vel + 3 * off[t], pvel[t], cnt[t] * 3 * sizeof(double));
            }
        }
        for (int t = 0; t < nt; t++)
            if (rcs[t] == NBH_OK) {
                free(ppos[t]);
                free(pvel[t]);
            }
        if (clean) {
            *n_out = total;
            *pos_out = pos;
            *vel_out = vel;
            return NBH_OK;
        }
    }
    return parse_range(text, 0, len, n_out, pos_out, vel_out, &partial);
}

int nbh_read_bodies(const char *path, size_t *n, double **pos_xyzm, double **vel_xyz)
{
    FILE *f = fopen(path, "rb");
    if (!f)
        return NBH_EOPEN;
    size_t cap = 1 << 20, len = 0;
    char *buf = malloc(cap + 1);
    if (!buf) {
        fclose(f);
        return NBH_ENOMEM;
    }
    for (;;) {
        size_t got = fread(buf + len, 1, cap - len, f);
        len += got;
        if (got == 0)
            break;
        if (len == cap) {
            cap *= 2;
            char *nb = realloc(buf, cap + 1);
            if (!nb) {
                free(buf);
                fclose(f);
                return NBH_ENOMEM;
            }
            buf = nb;
        }
    }
    fclose(f);
    buf[len] = '\0'; /* strtod needs a terminator */
    int rc = nbh_parse_bodies(buf, len, n, pos_xyzm, vel_xyz);
    free(buf);
    return rc;
}

int nbh_parse_times(const char *start_s, const char *end_s, const char *dt_s, float *start,
                    float *end, float *dt)
{
    *start = strtof(start_s, NULL);
    *end = strtof(end_s, NULL);
    *dt = strtof(dt_s, NULL);
    int n_steps = (*end - *start) / *dt;
    return n_steps;
}

/* ------------------------------------------------------------- writer ---- */

static size_t put_u64(char *out, uint64_t v)
{
    char tmp[24];
    size_t k = 0;
    do {
        tmp[k++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    for (size_t i = 0; i < k; i++)
        out[i] = tmp[k - 1 - i];
    return k;
}

/*
 * printf("%lf") = six decimals, the exact binary value rounded half-to-even.
 * |v| = m * 2^e with m < 2^53; m * 10^6 < 2^73 fits unsigned __int128, so the
 * rounding of |v| * 10^6 is done exactly in integers.  Values >= 2^52 (e >= 0),
 * infinities and NaNs go through snprintf.
 */
size_t nbh_format_lf(char *out, double v)
{
    uint64_t bits;
    memcpy(&bits, &v, sizeof(bits));
    const int neg = (int)(bits >> 63);
    const int ex = (int)((bits >> 52) & 0x7ff);
    uint64_t m = bits & ((1ull << 52) - 1);
    int e;
    if (ex == 0x7ff)
        return (size_t)sprintf(out, "%lf", v);
    if (ex == 0) {
        e = -1074;
    } else {
        m |= 1ull << 52;
        e = ex - 1075;
    }
    if (e >= 0)
        return (size_t)sprintf(out, "%lf", v);
    const int sh = -e;
    unsigned __int128 q = 0;
    if (sh < 128) {
        const unsigned __int128 prod = (unsigned __int128)m * 1000000u;
        q = prod >> sh;
        const unsigned __int128 rem = prod & ((((unsigned __int128)1) << sh) - 1);
        const unsigned __int128 half = ((unsigned __int128)1) << (sh - 1);
        if (rem > half || (rem == half && (q & 1)))
            q++;
    }
    uint64_t ip, fp;
    if ((q >> 64) == 0) {
        const uint64_t q64 = (uint64_t)q;
        ip = q64 / 1000000u;
        fp = q64 % 1000000u;
    } else {
        ip = (uint64_t)(q / 1000000u);
        fp = (uint64_t)(q % 1000000u);
    }
    size_t k = 0;
    if (neg)
        out[k++] = '-';
    k += put_u64(out + k, ip);
    out[k++] = '.';
    for (int d = 5; d >= 0; d--) {
        out[k + d] = (char)('0' + fp % 10);
        fp /= 10;
    }
    k += 6;
    out[k] = '\0';
    return k;
}

/* one body line is at most 20 (index) + 4 * 32 + 5 bytes for sane values; %lf of
 * huge doubles can reach ~320 bytes per field, so check per line. */
#define NBH_LINE_MAX 1400

size_t nbh_format_frame(char *buf, size_t cap, size_t n, const double *pm)
{
    size_t k = 0;
    for (size_t i = 0; i < n; i++) {
        if (cap - k < NBH_LINE_MAX)
            return 0;
        k += put_u64(buf + k, (uint64_t)i);
        for (int c = 0; c < 4; c++) {
            buf[k++] = ',';
            k += nbh_format_lf(buf + k, pm[4 * i + c]);
        }
        buf[k++] = '\n';
    }
    return k;
}

static size_t format_lines(char *buf, size_t first, size_t last, const double *pm)
{
    size_t k = 0;
    for (size_t i = first; i < last; i++) {
        k += put_u64(buf + k, (uint64_t)i);
        for (int c = 0; c < 4; c++) {
            buf[k++] = ',';
            k += nbh_format_lf(buf + k, pm[4 * i + c]);
        }
        buf[k++] = '\n';
    }
    return k;
}

/* Lines are formatted CHUNK bodies at a time; with OpenMP the chunks of one round go to
 * different threads and are written out in body order, so the file is the same byte for
 * byte whatever the thread count (OMP_NUM_THREADS=1 gives the serial writer). */
int nbh_write_frame(FILE *fp, size_t n, const double *pm)
{
    enum { CHUNK = 1024, MAXT = 32 };
    static char *buf[MAXT];
    int nt = 1;
#ifdef _OPENMP
    nt = omp_get_max_threads();
#endif
    if (nt > MAXT)
        nt = MAXT;
    if ((size_t)nt > (n + CHUNK - 1) / CHUNK)
        nt = (int)((n + CHUNK - 1) / CHUNK);
    if (nt < 1)
        nt = 1;
    for (int t = 0; t < nt; t++)
        if (!buf[t] && !(buf[t] = malloc((size_t)CHUNK * NBH_LINE_MAX)))
            return NBH_ENOMEM;
    size_t len[MAXT];
    for (size_t base = 0; base < n; base += (size_t)nt * CHUNK) {
#ifdef _OPENMP
#pragma omp parallel for num_threads(nt) schedule(static, 1) if (nt > 1)
#endif
        for (int t = 0; t < nt; t++) {
            const size_t first = base + (size_t)t * CHUNK;
            const size_t last = first + CHUNK < n ? first + CHUNK : n;
            len[t] = first < n ? format_lines(buf[t], first, last, pm) : 0;
        }
        for (int t = 0; t < nt; t++)
            if (len[t] && fwrite(buf[t], 1, len[t], fp) != len[t])
                return NBH_EOPEN;
    }
    return NBH_OK;
}

int nbh_write_bodies(const char *path, size_t n, const double *pm, const double *v)
{
    FILE *fp = fopen(path, "w");
    if (!fp)
        return NBH_EOPEN;
    for (size_t i = 0; i < n; i++)
        fprintf(fp, "%.17g,%.17g,%.17g,%.17g,%.17g,%.17g,%.17g\n", pm[4 * i], pm[4 * i + 1],
                pm[4 * i + 2], v[3 * i], v[3 * i + 1], v[3 * i + 2], pm[4 * i + 3]);
    fclose(fp);
    return NBH_OK;
}
