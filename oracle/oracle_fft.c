/*
 * oracle_fft.c -- CPU ORACLE (test infrastructure, never shipped, never on the product path).
 *
 * Restatement of the reference's FFT arithmetic: the kissfft float build that
 * src/FFT.zig binds (lib/kissfft/kiss_fft.c, kiss_fftr.c; build flags build.zig:149-183).
 *   - decimation-in-time mixed radix, factors chosen 4 first, then 2, 3, 5 (kiss_fft.c:307-328)
 *   - twiddles exp(-+2*pi*i*k/n) evaluated in double, stored as float (kiss_fft.c:337-376)
 *   - real transform of even length n done as one complex transform of n/2 plus a
 *     split/merge pass with "super twiddles" (kiss_fftr.c:21-155)
 * The operation order inside each butterfly follows the reference so that this file agrees
 * with oracle/_ref/libkissfft_ref.so (the reference's own C, compiled unmodified) to the bit
 * when both are built with -ffp-contract=off; tests/test_oracle_fft.py pins that.
 */
#include "fvad_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef fvad_complex cpx;

struct orc_fft {
    int n;       /* complex length */
    int inverse;
    int nstage;
    int radix[32];
    int rest[32]; /* length of the sub-transform left after this stage */
    cpx *tw;      /* n twiddles */
};

static cpx cmul(cpx a, cpx b) {
    cpx m;
    m.r = a.r * b.r - a.i * b.i;
    m.i = a.r * b.i + a.i * b.r;
    return m;
}
static cpx cadd(cpx a, cpx b) { cpx m = {a.r + b.r, a.i + b.i}; return m; }
static cpx csub(cpx a, cpx b) { cpx m = {a.r - b.r, a.i - b.i}; return m; }

/* kiss_fft.c:307-328: pull out 4s, then 2s, then odd primes */
static void choose_radices(struct orc_fft *p) {
    int n = p->n, f = 4, s = 0;
    double lim = floor(sqrt((double)n));
    do {
        while (n % f) {
            if (f == 4) f = 2;
            else if (f == 2) f = 3;
            else f += 2;
            if (f > lim) f = n;
        }
        n /= f;
        p->radix[s] = f;
        p->rest[s] = n;
        ++s;
    } while (n > 1);
    p->nstage = s;
}

orc_fft *orc_fft_new(int n, int inverse) {
    struct orc_fft *p = (struct orc_fft *)calloc(1, sizeof(*p));
    if (!p) return NULL;
    p->n = n;
    p->inverse = inverse;
    p->tw = (cpx *)malloc(sizeof(cpx) * (size_t)n);
    for (int k = 0; k < n; ++k) {
        const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
        double ph = -2 * pi * k / n;
        if (inverse) ph = -ph;
        p->tw[k].r = (float)cos(ph);
        p->tw[k].i = (float)sin(ph);
    }
    choose_radices(p);
    return p;
}

void orc_fft_free(orc_fft *p) {
    if (!p) return;
    free(p->tw);
    free(p);
}

/* One recombination pass of radix r over `m` columns (kiss_fft.c:15-189). */
static void combine(const struct orc_fft *p, cpx *o, size_t stride, int r, int m) {
    const cpx *tw = p->tw;
    if (r == 2) {
        for (int k = 0; k < m; ++k) {
            cpx t = cmul(o[m + k], tw[k * stride]);
            o[m + k] = csub(o[k], t);
            o[k] = cadd(o[k], t);
        }
    } else if (r == 4) {
        for (int k = 0; k < m; ++k) {
            cpx a = cmul(o[k + m], tw[k * stride]);
            cpx b = cmul(o[k + 2 * m], tw[k * stride * 2]);
            cpx c = cmul(o[k + 3 * m], tw[k * stride * 3]);
            cpx d = csub(o[k], b);
            o[k] = cadd(o[k], b);
            cpx s = cadd(a, c);
            cpx q = csub(a, c);
            o[k + 2 * m] = csub(o[k], s);
            o[k] = cadd(o[k], s);
            if (p->inverse) {
                o[k + m].r = d.r - q.i;
                o[k + m].i = d.i + q.r;
                o[k + 3 * m].r = d.r + q.i;
                o[k + 3 * m].i = d.i - q.r;
            } else {
                o[k + m].r = d.r + q.i;
                o[k + m].i = d.i - q.r;
                o[k + 3 * m].r = d.r - q.i;
                o[k + 3 * m].i = d.i + q.r;
            }
        }
    } else if (r == 3) {
        const cpx e3 = tw[stride * m];
        for (int k = 0; k < m; ++k) {
            cpx a = cmul(o[k + m], tw[k * stride]);
            cpx b = cmul(o[k + 2 * m], tw[k * stride * 2]);
            cpx s = cadd(a, b);
            cpx d = csub(a, b);
            o[k + m].r = o[k].r - s.r * .5f;
            o[k + m].i = o[k].i - s.i * .5f;
            d.r *= e3.i;
            d.i *= e3.i;
            o[k] = cadd(o[k], s);
            o[k + 2 * m].r = o[k + m].r + d.i;
            o[k + 2 * m].i = o[k + m].i - d.r;
            o[k + m].r -= d.i;
            o[k + m].i += d.r;
        }
    } else if (r == 5) {
        const cpx ya = tw[stride * m], yb = tw[stride * 2 * m];
        for (int u = 0; u < m; ++u) {
            cpx x0 = o[u];
            cpx x1 = cmul(o[u + m], tw[u * stride]);
            cpx x2 = cmul(o[u + 2 * m], tw[2 * u * stride]);
            cpx x3 = cmul(o[u + 3 * m], tw[3 * u * stride]);
            cpx x4 = cmul(o[u + 4 * m], tw[4 * u * stride]);
            cpx s14 = cadd(x1, x4), d14 = csub(x1, x4);
            cpx s23 = cadd(x2, x3), d23 = csub(x2, x3);
            o[u].r += s14.r + s23.r;
            o[u].i += s14.i + s23.i;
            cpx e, f;
            e.r = x0.r + s14.r * ya.r + s23.r * yb.r;
            e.i = x0.i + s14.i * ya.r + s23.i * yb.r;
            f.r = d14.i * ya.i + d23.i * yb.i;
            f.i = -(d14.r * ya.i) - d23.r * yb.i;
            o[u + m] = csub(e, f);
            o[u + 4 * m] = cadd(e, f);
            cpx g, h;
            g.r = x0.r + s14.r * yb.r + s23.r * ya.r;
            g.i = x0.i + s14.i * yb.r + s23.i * ya.r;
            h.r = -(d14.i * yb.i) + d23.i * ya.i;
            h.i = d14.r * yb.i - d23.r * ya.i;
            o[u + 2 * m] = cadd(g, h);
            o[u + 3 * m] = csub(g, h);
        }
    } else {
        /* generic radix: plain O(r^2) DFT per column (kiss_fft.c:191-234) */
        cpx tmp[64];
        for (int u = 0; u < m; ++u) {
            for (int q = 0; q < r; ++q) tmp[q] = o[u + q * m];
            for (int q = 0; q < r; ++q) {
                int k = u + q * m;
                size_t twi = 0;
                o[k] = tmp[0];
                for (int j = 1; j < r; ++j) {
                    twi += stride * (size_t)k;
                    if (twi >= (size_t)p->n) twi %= (size_t)p->n;
                    o[k] = cadd(o[k], cmul(tmp[j], tw[twi]));
                }
            }
        }
    }
}

/* kiss_fft.c:236-300 */
static void descend(const struct orc_fft *p, cpx *o, const cpx *in, size_t stride, int stage) {
    const int r = p->radix[stage], m = p->rest[stage];
    if (m == 1) {
        for (int q = 0; q < r; ++q) o[q] = in[q * stride];
    } else {
        for (int q = 0; q < r; ++q) descend(p, o + (size_t)q * m, in + q * stride, stride * r, stage + 1);
    }
    combine(p, o, stride, r, m);
}

void orc_fft_run(const orc_fft *p, const fvad_complex *in, fvad_complex *out) {
    descend(p, out, in, 1, 0);
}

/* ---------------- real transforms (kiss_fftr.c) ---------------- */

struct orc_rfft {
    int n; /* real length (even) */
    int inverse;
    orc_fft *sub;
    cpx *tmp;   /* n/2 */
    cpx *super; /* n/4 */
};

orc_rfft *orc_rfft_new(int n, int inverse) {
    if (n <= 0 || (n & 1)) return NULL;
    struct orc_rfft *p = (struct orc_rfft *)calloc(1, sizeof(*p));
    const int h = n / 2;
    p->n = n;
    p->inverse = inverse;
    p->sub = orc_fft_new(h, inverse);
    p->tmp = (cpx *)malloc(sizeof(cpx) * (size_t)h);
    p->super = (cpx *)malloc(sizeof(cpx) * (size_t)(h / 2 + 1));
    for (int i = 0; i < h / 2; ++i) {
        /* kiss_fftr.c:52-58 */
        double ph = -3.14159265358979323846264338327 * ((double)(i + 1) / h + .5);
        if (inverse) ph = -ph;
        p->super[i].r = (float)cos(ph);
        p->super[i].i = (float)sin(ph);
    }
    return p;
}

void orc_rfft_free(orc_rfft *p) {
    if (!p) return;
    orc_fft_free(p->sub);
    free(p->tmp);
    free(p->super);
    free(p);
}

int orc_rfft_size(const orc_rfft *p) { return p->n; }

/* kiss_fftr.c:63-117: n reals -> n/2+1 bins, unnormalised */
void orc_rfft_forward(orc_rfft *p, const float *x, fvad_complex *X) {
    const int h = p->n / 2;
    orc_fft_run(p->sub, (const cpx *)x, p->tmp);
    const cpx dc = p->tmp[0];
    X[0].r = dc.r + dc.i;
    X[h].r = dc.r - dc.i;
    X[0].i = X[h].i = 0;
    for (int k = 1; k <= h / 2; ++k) {
        cpx a = p->tmp[k];
        cpx b = {p->tmp[h - k].r, -p->tmp[h - k].i};
        cpx e = cadd(a, b);
        cpx o = cmul(csub(a, b), p->super[k - 1]);
        X[k].r = (e.r + o.r) * .5f;
        X[k].i = (e.i + o.i) * .5f;
        X[h - k].r = (e.r - o.r) * .5f;
        X[h - k].i = (o.i - e.i) * .5f;
    }
}

/* kiss_fftr.c:119-155: n/2+1 bins -> n reals, unnormalised (output = n * true inverse) */
void orc_rfft_inverse(orc_rfft *p, const fvad_complex *X, float *x) {
    const int h = p->n / 2;
    p->tmp[0].r = X[0].r + X[h].r;
    p->tmp[0].i = X[0].r - X[h].r;
    for (int k = 1; k <= h / 2; ++k) {
        cpx a = X[k];
        cpx b = {X[h - k].r, -X[h - k].i};
        cpx e = cadd(a, b);
        cpx o = cmul(csub(a, b), p->super[k - 1]);
        p->tmp[k] = cadd(e, o);
        p->tmp[h - k] = csub(e, o);
        p->tmp[h - k].i *= -1;
    }
    orc_fft_run(p->sub, p->tmp, (cpx *)x);
}
