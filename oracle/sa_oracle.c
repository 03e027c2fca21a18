/*
 * sa_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, flat arrays, double precision) of signalAlign's banded forward/backward
 * posterior-match DP.  Written from the behaviour of the reference, citing the file:line each piece
 * follows (all paths relative to /root/reference):
 *
 *   logAdd / cubic lookup ............... impl/pairwiseAligner.c:298-318
 *   band construction ................... impl/pairwiseAligner.c:161-246
 *   split points ........................ impl/pairwiseAligner.c:1886-1937
 *   ambiguous-kmer path enumeration ..... impl/pairwiseAligner.c:723-801, legality :595-621
 *   3-state cell recurrence ............. impl/stateMachine.c:1306-1437
 *   emissions ........................... impl/stateMachine.c:285-306, 344-348, 527-701; hdp.c:2588-2612;
 *                                          hdp_math_utils.c:471-495
 *   start/end state values .............. impl/stateMachine.c:1134-1173
 *   total probability ................... impl/pairwiseAligner.c:1335-1353, 1167-1180, 828-843, 879-885
 *   posterior / expectations callbacks .. impl/pairwiseAligner.c:1355-1443, 914-981
 *   banded driver + blocked traceback ... impl/pairwiseAligner.c:1450-1590
 *   split driver + coordinate shift ..... impl/pairwiseAligner.c:1953-2080
 *
 * Parity is PINNED: tests/test_oracle_vs_reference.py compares this file bit-for-bit (same pair list, same
 * floor(p*1e7)) with the compiled reference on real fixtures and random inputs, and tests/golden/ holds the
 * reference's outputs for machines without /root/reference.
 *
 * Compile with -ffp-contract=off so the arithmetic matches the reference's x86-64 SSE2 build.
 */
#include "sa_oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define LOG_ZERO (-INFINITY)
#define ST_M 0
#define ST_X 1
#define ST_Y 2

/* ------------------------------------------------------------------------------------------- logAdd */

static inline double cubic_lookup(double x) { /* pairwiseAligner.c:301-312: float literals promoted to double */
    if (x <= 1.00f)
        return ((-0.009350833524763f * x + 0.130659527668286f) * x + 0.498799810682272f) * x + 0.693203116424741f;
    if (x <= 2.50f)
        return ((-0.014532321752540f * x + 0.139942324101744f) * x + 0.495635523139337f) * x + 0.692140569840976f;
    if (x <= 4.50f)
        return ((-0.004605031767994f * x + 0.063427417320019f) * x + 0.695956496475118f) * x + 0.514272634594009f;
    return ((-0.000458661602210f * x + 0.009695946122598f) * x + 0.930734667215156f) * x + 0.168037164329057f;
}

double sao_logAdd(double x, double y) { /* pairwiseAligner.c:314-318 */
    if (x < y) return (x == LOG_ZERO || y - x >= 7.5) ? y : cubic_lookup(y - x) + x;
    return (y == LOG_ZERO || x - y >= 7.5) ? x : cubic_lookup(x - y) + y;
}
#define logAdd sao_logAdd

#ifdef SAO_FP32_STUDY
/* ---- FP32 STUDY (never part of the pinned oracle: built as libsa_oracle_fp32.so by `make fp32study`).
 * What an FP32 band kernel with per-diagonal renormalisation would compute: every DP value is a FLOAT relative to an
 * integer offset of its diagonal (forward: round(max f(d-1)), backward: round(max b(d+1)); kept in double, exact), every
 * logAdd, dot product and posterior runs in float arithmetic on the relative values.  The cubic logAdd only sees
 * differences, so the offsets never change which of its branches is taken. */
#ifndef SAO_FP32_EVERY
#define SAO_FP32_EVERY 16
#endif
static double *g_offF = NULL, *g_offB = NULL;
static int64_t g_d = 0;      /* diagonal of the cell being computed */
static int g_nbr = 1;        /* its neighbour in the running transition: diagonal g_d - 1 or g_d - 2 */
static inline float lookupF(float x) {
    if (x <= 1.00f) return ((-0.009350833524763f * x + 0.130659527668286f) * x + 0.498799810682272f) * x + 0.693203116424741f;
    if (x <= 2.50f) return ((-0.014532321752540f * x + 0.139942324101744f) * x + 0.495635523139337f) * x + 0.692140569840976f;
    if (x <= 4.50f) return ((-0.004605031767994f * x + 0.063427417320019f) * x + 0.695956496475118f) * x + 0.514272634594009f;
    return ((-0.000458661602210f * x + 0.009695946122598f) * x + 0.930734667215156f) * x + 0.168037164329057f;
}
static inline float logAddF(float x, float y) {
    if (x < y) return (x == -INFINITY || y - x >= 7.5f) ? y : lookupF(y - x) + x;
    return (y == -INFINITY || x - y >= 7.5f) ? x : lookupF(x - y) + y;
}
static double roundedMax(const double *v, int64_t n) {
    double m = -INFINITY;
    for (int64_t i = 0; i < n; i++) if (v[i] > m) m = v[i];
    return (m == -INFINITY || m != m) ? 0.0 : rint(m);
}
#endif

/* ------------------------------------------------------------------------------------------- ambiguity */

void sao_ambig_default(sao_ambig *a) { /* pairwiseAligner.c:32-65 */
    memset(a, 0, sizeof(*a));
    static const char *tab[][2] = {
        {"R", "AG"}, {"Y", "CT"}, {"S", "CG"}, {"W", "AT"}, {"K", "GT"}, {"M", "AC"}, {"B", "CGT"}, {"D", "AGT"},
        {"H", "ACT"}, {"V", "ACG"}, {"X", "ACGT"}, {"L", "CEO"}, {"P", "CE"}, {"Q", "AI"}, {"f", "AF"},
        {"U", "ACEGOT"}, {"Z", "JT"}, {"j", "Tp"}, {"k", "Gb"}, {"l", "Gd"}, {"m", "Ce"}, {"n", "Th"}, {"o", "Ai"}};
    for (size_t i = 0; i < sizeof(tab) / sizeof(tab[0]); i++)
        strncpy(a->opts[(unsigned char) tab[i][0][0]], tab[i][1], 15);
}

int sao_ambig_from_file(sao_ambig *a, const char *path) { /* pairwiseAligner.c:68-92 */
    memset(a, 0, sizeof(*a));
    FILE *f = fopen(path, "r");
    if (!f) return -1;
    char line[100], enc[100], bases[100];
    int i = 0;
    while (i < 300 && fgets(line, sizeof(line), f) != NULL) {
        if (sscanf(line, "%s\t%s", enc, bases) == 2 && strlen(enc) == 1) { /* only 1-char codes can ever match */
            memset(a->opts[(unsigned char) enc[0]], 0, 16);
            strncpy(a->opts[(unsigned char) enc[0]], bases, 15);
        }
        i++;
    }
    fclose(f);
    return 0;
}

/* ------------------------------------------------------------------------------------------- kmers */

int64_t sao_kmer_id(const char *kmer, const char *alphabet, int64_t alphabetSize, int64_t k) {
    int64_t id = 0; /* nanopore_hdp.c:377-410: big-endian base-alphabetSize index over the sorted alphabet */
    for (int64_t i = 0; i < k; i++) {
        int64_t j = 0;
        while (j < alphabetSize && kmer[i] != alphabet[j]) j++;
        if (j == alphabetSize) return -1; /* the reference exits the process here */
        id = id * alphabetSize + j;
    }
    return id;
}

int64_t sao_paths(const char *kmer, const sao_model *m, const sao_ambig *a, int32_t *idxOut, int64_t cap) {
    /* hdCell_construct2 (pairwiseAligner.c:723-801): position-major, option-minor expansion */
    int64_t k = m->kmerLength, n = 1;
    char (*list)[16] = malloc(sizeof(char[16]));
    memset(list[0], 0, 16);
    memcpy(list[0], kmer, k);
    for (int64_t i = 0; i < k; i++) {
        const char *opts = a->opts[(unsigned char) kmer[i]];
        int64_t no = strlen(opts);
        if (no == 0) continue;
        char (*nl)[16] = malloc(sizeof(char[16]) * n * no);
        for (int64_t j = 0; j < n; j++)
            for (int64_t o = 0; o < no; o++) {
                memcpy(nl[j * no + o], list[j], 16);
                nl[j * no + o][i] = opts[o];
            }
        free(list);
        list = nl;
        n *= no;
    }
    int64_t bad = 0;
    for (int64_t j = 0; j < n && j < cap; j++) {
        int64_t id = sao_kmer_id(list[j], m->alphabet, m->alphabetSize, k);
        if (id < 0) bad = 1;
        idxOut[j] = (int32_t) id;
    }
    free(list);
    return bad ? -1 : n;
}

/* ------------------------------------------------------------------------------------------- emissions */

static inline double logGaussPdf(double x, double mu, double sigma) { /* stateMachine.c:296-306 */
    if (sigma == 0.0) return LOG_ZERO;
    double log_inv_sqrt_2pi = -0.91893853320467267;
    double l_sigma = log(sigma);
    double a = (x - mu) / sigma;
    return log_inv_sqrt_2pi - l_sigma + (-0.5 * a * a);
}

static inline double logInvGaussPdf(double eventNoise, double modelNoiseMean, double modelNoiseLambda) {
    double l_twoPi = 1.8378770664093453; /* stateMachine.c:285-294 */
    double l_eventNoise = log(eventNoise);
    double a = (eventNoise - modelNoiseMean) / modelNoiseMean;
    double l_lambda = log(modelNoiseLambda);
    return (l_lambda - l_twoPi - 3 * l_eventNoise - modelNoiseLambda * a * a / eventNoise) / 2;
}

static inline double descale(double x, double levelMean, double scale, double shift, double var) {
    return (x + var * levelMean - scale * levelMean - shift) / var; /* stateMachine.c:344-348 */
}

static double grid_spline(double q, const double *x, const double *y, const double *s, int64_t n) {
    /* hdp_math_utils.c:471-495 */
    if (q <= x[0]) return y[0] - s[0] * (x[0] - q);
    if (q >= x[n - 1]) return y[n - 1] + s[n - 1] * (q - x[n - 1]);
    double dx = x[1] - x[0];
    int64_t il = (int64_t) ((q - x[0]) / dx);
    int64_t ir = il + 1;
    double dy = y[ir] - y[il];
    double a = s[il] * dx - dy;
    double b = dy - s[ir] * dx;
    double tl = (q - x[il]) / dx;
    double tr = 1.0 - tl;
    return tr * y[il] + tl * y[ir] + tl * tr * (a * tr + b * tl);
}

double sao_emission(const sao_model *m, int64_t ki, double eventMean, double eventNoise, int match) {
    if (ki < 0) return LOG_ZERO; /* NULL kmer (x == 0 column) */
    const double *tab = match ? m->matchTable : m->gapYTable;
    /* `kmerIndex > max_index ? 0.0 : table[...]` (stateMachine.c:259-265) cannot trigger for a valid kmer */
    switch (m->kind) {
    case SAO_EMIT_GAUSS_MEANONLY_DESCALE: { /* stateMachine.c:557-605 */
        double mu = tab[ki * 5], sd = tab[ki * 5 + 1];
        double x = descale(eventMean, mu, m->scale, m->shift, m->var);
        return log((1 / m->var)) + logGaussPdf(x, mu, sd);
    }
    case SAO_EMIT_HDP: { /* stateMachine.c:527-554, nanopore_hdp.c:420-429, hdp.c:2588-2612 */
        double mu = m->matchTable[ki * 5];
        double x = descale(eventMean, mu, m->scale, m->shift, m->var);
        int32_t row = m->kmerToRow[ki];
        double interp = grid_spline(x, m->grid, m->postPred + (int64_t) row * m->gridLength,
                                    m->slopes + (int64_t) row * m->gridLength, m->gridLength);
        double dens = interp > 0.0 ? interp : 0.0;
        return log((1 / m->var) * dens);
    }
    case SAO_EMIT_GAUSS_NOISE_SCALED: { /* stateMachine.c:659-701 */
        double mu = tab[ki * 5], sd = tab[ki * 5 + 1], nm = tab[ki * 5 + 2], nl = tab[ki * 5 + 4];
        return logGaussPdf(eventMean, mu, sd) + logInvGaussPdf(eventNoise, nm, nl);
    }
    case SAO_EMIT_GAUSS_NOISE_DESCALE: { /* stateMachine.c:607-657 */
        if (eventNoise == 0) eventNoise = 0.000000001;
        double mu = tab[ki * 5], sd = tab[ki * 5 + 1], nm = tab[ki * 5 + 2], nl = tab[ki * 5 + 4];
        double x = descale(eventMean, mu, m->scale, m->shift, m->var);
        return logGaussPdf(x, mu, sd) + logInvGaussPdf(eventNoise, nm, nl);
    }
    }
    return LOG_ZERO;
}

/* ------------------------------------------------------------------------------------------- band */

typedef struct { int64_t xay, xmyL, xmyR; } Diag;

static inline int64_t diagX(int64_t xay, int64_t xmy) { return (xay + xmy) / 2; }
static inline int64_t diagY(int64_t xay, int64_t xmy) { return (xay - xmy) / 2; }
static inline int64_t diagWidth(Diag d) { return (d.xmyR - d.xmyL) / 2 + 1; }
static inline int64_t avoidOffByOne(int64_t xay, int64_t xmy) { return (xay + xmy) % 2 == 0 ? xmy : xmy + 1; }
static inline void bandP(int64_t *xmy, int64_t i, int64_t j, int64_t k) {
    if (i < j) *xmy += 2 * (j - i) * k;
}
static inline int64_t boundCoord(int64_t z, int64_t lZ) { return z < 0 ? 0 : (z > lZ ? lZ : z); }

static Diag bandDiag(int64_t xay, int64_t xL, int64_t yL, int64_t xU, int64_t yU) { /* :171-189 */
    int64_t xmyL = avoidOffByOne(xay, xL - yL);
    int64_t xmyR = avoidOffByOne(xay, xU - yU);
    bandP(&xmyL, diagX(xay, xmyL), xL, 1);
    bandP(&xmyL, yL, diagY(xay, xmyL), 1);
    bandP(&xmyR, xU, diagX(xay, xmyR), -1);
    bandP(&xmyR, diagY(xay, xmyR), yU, -1);
    Diag d = {xay, xmyL, xmyR};
    return d;
}

static Diag *band_build(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t e) { /* :195-246 */
    Diag *ds = malloc(sizeof(Diag) * (lX + lY + 1));
    int64_t ai = 0, xay = 0, pxay = 0, pxmy = 0, nxay = 0, nxmy = 0, xL = 0, yL = 0, xU = 0, yU = 0;
    while (xay <= lX + lY) {
        ds[xay] = bandDiag(xay, xL, yL, xU, yU);
        if (nxay == xay++) {
            pxay = nxay;
            pxmy = nxmy;
            int64_t x = lX, y = lY;
            if (ai < nA) {
                x = anchors[2 * ai] + 1;
                y = anchors[2 * ai + 1] + 1;
                ai++;
            }
            nxay = x + y;
            nxmy = x - y;
            /* C integer division truncates toward zero, exactly as the reference's (xay + xmy) / 2 */
            xL = boundCoord(diagX(pxay, pxmy - e), lX);
            yL = boundCoord(diagY(nxay, nxmy - e), lY);
            xU = boundCoord(diagX(nxay, nxmy + e), lX);
            yU = boundCoord(diagY(pxay, pxmy + e), lY);
        }
    }
    return ds;
}

void sao_band(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t e, int64_t *out) {
    Diag *ds = band_build(anchors, nA, lX, lY, e);
    for (int64_t i = 0; i <= lX + lY; i++) {
        out[3 * i] = ds[i].xay; out[3 * i + 1] = ds[i].xmyL; out[3 * i + 2] = ds[i].xmyR;
    }
    free(ds);
}

/* ------------------------------------------------------------------------------------------- split */

typedef struct { int64_t *v; int64_t n, cap; } I64Vec;
static void vpush(I64Vec *v, int64_t x) {
    if (v->n == v->cap) { v->cap = v->cap ? v->cap * 2 : 64; v->v = realloc(v->v, sizeof(int64_t) * v->cap); }
    v->v[v->n++] = x;
}

static int splitP(int64_t *x1, int64_t *y1, int64_t x2, int64_t y2, int64_t x3, int64_t y3, I64Vec *sp,
                  int64_t bigger, int skip) { /* :1886-1908 */
    int64_t lX2 = x3 - x2, lY2 = y3 - y2;
    if (lX2 * lY2 > bigger) {
        int64_t maxLen = sqrt(bigger);
        int64_t hX = lX2 / 2 > maxLen ? maxLen : lX2 / 2;
        int64_t hY = lY2 / 2 > maxLen ? maxLen : lY2 / 2;
        if (!skip) { vpush(sp, *x1); vpush(sp, *y1); vpush(sp, x2 + hX); vpush(sp, y2 + hY); }
        *x1 = x3 - hX;
        *y1 = y3 - hY;
        return 1;
    }
    return 0;
}

static void split_build(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t bigger, int rl, int rr,
                        I64Vec *sp) { /* :1910-1937 */
    int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    for (int64_t i = 0; i < nA; i++) {
        int64_t x3 = anchors[2 * i], y3 = anchors[2 * i + 1];
        splitP(&x1, &y1, x2, y2, x3, y3, sp, bigger, rl && i == 0);
        x2 = x3 + 1;
        y2 = y3 + 1;
    }
    if (!splitP(&x1, &y1, x2, y2, lX, lY, sp, bigger, rl && nA == 0) || !rr) {
        vpush(sp, x1); vpush(sp, y1); vpush(sp, lX); vpush(sp, lY);
    }
}

int64_t sao_split_points(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t bigger, int rl, int rr,
                         int64_t *out, int64_t cap) {
    I64Vec sp = {0};
    split_build(anchors, nA, lX, lY, bigger, rl, rr, &sp);
    int64_t n = sp.n / 4;
    for (int64_t i = 0; i < n && i < cap; i++) memcpy(out + 4 * i, sp.v + 4 * i, sizeof(int64_t) * 4);
    free(sp.v);
    return n;
}

/* ------------------------------------------------------------------------------------------- DP state */

typedef struct {
    const sao_model *m;
    int64_t lX, lY;
    const double *events;       /* lY*4, already offset to the sub-region */
    /* per matrix column x in [0, lX]: path list.  x == 0 has the single NULL-kmer path (kmerIdx -1) */
    int64_t *pathOff;           /* lX+2 */
    int32_t *pathKmer;          /* concrete kmer index per path */
    int64_t suffixMod;          /* alphabetSize^(k-1) */
} Ctx;

typedef struct {
    Diag d;
    int64_t *off;   /* width+1 prefix of path counts */
    double *v;      /* 3 doubles per path */
} DpDiag;

static DpDiag *dd_new(const Ctx *c, Diag d) { /* dpDiagonal_construct, :988-1034 */
    DpDiag *dd = malloc(sizeof(DpDiag));
    dd->d = d;
    int64_t w = diagWidth(d);
    dd->off = malloc(sizeof(int64_t) * (w + 1));
    dd->off[0] = 0;
    for (int64_t i = 0; i < w; i++) {
        int64_t x = diagX(d.xay, d.xmyL + 2 * i);
        dd->off[i + 1] = dd->off[i] + (c->pathOff[x + 1] - c->pathOff[x]);
    }
    dd->v = malloc(sizeof(double) * 3 * dd->off[w]);
    return dd;
}
static void dd_free(DpDiag *dd) { if (dd) { free(dd->off); free(dd->v); free(dd); } }
static void dd_fill(DpDiag *dd, const double s[3]) {
    int64_t n = dd->off[diagWidth(dd->d)];
    for (int64_t i = 0; i < n; i++) { dd->v[3 * i] = s[0]; dd->v[3 * i + 1] = s[1]; dd->v[3 * i + 2] = s[2]; }
}
static const double ZERO3[3] = {-INFINITY, -INFINITY, -INFINITY};
/* dpDiagonal_getCell (:1114-1120): NULL outside [xmyL, xmyR] */
static inline double *dd_cell(DpDiag *dd, int64_t xmy, int64_t *nPaths) {
    if (dd == NULL || xmy < dd->d.xmyL || xmy > dd->d.xmyR) return NULL;
    int64_t i = (xmy - dd->d.xmyL) / 2;
    *nPaths = dd->off[i + 1] - dd->off[i];
    return dd->v + 3 * dd->off[i];
}

static void stateValues(const sao_model *m, int which, double s[3]) { /* stateMachine.c:1134-1173 */
    const double *t = m->trans;
    switch (which) {
    case 0: s[0] = 0; s[1] = LOG_ZERO; s[2] = LOG_ZERO; break;                       /* start */
    case 1: s[0] = LOG_ZERO; s[1] = 0; s[2] = 0; break;                              /* ragged start */
    case 2: s[0] = t[SAO_T_MATCH_CONTINUE]; s[1] = t[SAO_T_MATCH_FROM_GAP_X]; s[2] = t[SAO_T_MATCH_FROM_GAP_Y]; break;
    case 3: s[0] = (t[SAO_T_GAP_OPEN_X] + t[SAO_T_GAP_OPEN_Y]) / 2.0; s[1] = t[SAO_T_GAP_EXTEND_X];
            s[2] = t[SAO_T_GAP_EXTEND_Y]; break;                                     /* ragged end */
    }
}

/* path_checkLegal (:595-621) on kmer indices: suffix(k-1) of q == prefix(k-1) of p; NULL kmer legal with all */
static inline int legal(const Ctx *c, int32_t kq, int32_t kp) {
    if (kq < 0 || kp < 0) return 1;
    return (kq % c->suffixMod) == (kp / c->m->alphabetSize);
}
/* stString_eq on path kmers (NULL == NULL true) */
static inline int sameKmer(int32_t a, int32_t b) { return a == b; }

typedef struct {
    int mode;            /* 0 forward, 1 backward, 2 expectations */
    double total;        /* expectations */
    double threshold;
    double *trans9;
    I64Vec *assignK, *assignE;
    int64_t xIndexForAssign, yIndexForAssign;
    int isHdp;
} Op;

static inline void doTransition(const Op *op, double *from, double *to, int f, int t, double eP, double tP) {
    switch (op->mode) {
#ifdef SAO_FP32_STUDY
    case 0: {   /* `to` on diagonal g_d, `from` on g_d - g_nbr, both relative to their forward offsets */
        const double oC = g_offF[g_d], oN = g_offF[g_d - g_nbr];
        const float c = ((float) eP + (float) tP) + (float) (oN - oC);
        to[t] = oC + (double) logAddF((float) (to[t] - oC), (float) (from[f] - oN) + c);
        break;
    }
    case 1: {   /* `to` on the source diagonal g_d, `from` is the target on g_d - g_nbr: backward offsets */
        const double oC = g_offB[g_d], oN = g_offB[g_d - g_nbr];
        const float c = ((float) eP + (float) tP) + (float) (oC - oN);
        from[f] = oN + (double) logAddF((float) (from[f] - oN), (float) (to[t] - oC) + c);
        break;
    }
#else
    case 0: to[t] = logAdd(to[t], from[f] + (eP + tP)); break;       /* :852-858 */
    case 1: from[f] = logAdd(from[f], to[t] + (eP + tP)); break;     /* :866-871 */
#endif
    default: {                                                       /* :914-968 */
        double p = exp(from[f] + to[t] + (eP + tP) - op->total);
        op->trans9[f * 3 + t] += p;
        if (op->isHdp && t == ST_M && p >= op->threshold) {
            vpush(op->assignK, op->xIndexForAssign);
            vpush(op->assignE, op->yIndexForAssign);
        }
    }
    }
}

/* stateMachine3_cellCalculate / stateMachine3HDP_cellCalculate (stateMachine.c:1306-1437) */
static void cellCalculate(const Ctx *c, const Op *op, int64_t x, int64_t y, double *cur, int64_t nCur,
                          double *lower, int64_t nLower, double *middle, int64_t nMiddle, double *upper,
                          int64_t nUpper) {
    const sao_model *m = c->m;
    const double *t = m->trans;
    const int32_t *kC = c->pathKmer + c->pathOff[x];
    int hdp = m->kind == SAO_EMIT_HDP;
    /* event pointer: index y-1, or NULLEVENT {-inf, 0} when y == 0 (pairwiseAligner.c:325, 509-512) */
    double eMean = y > 0 ? c->events[(y - 1) * 4] : LOG_ZERO;
    double eNoise = y > 0 ? c->events[(y - 1) * 4 + 1] : 0.0;
#ifdef SAO_FP32_STUDY
    g_nbr = 1;
#endif
    if (lower != NULL) {
        const int32_t *kL = c->pathKmer + c->pathOff[x - 1];
        for (int64_t p = 0; p < nCur; p++)
            for (int64_t q = 0; q < nLower; q++)
                if (legal(c, kL[q], kC[p])) {
                    double eP = hdp ? -2.3025850929940455 : (kC[p] < 0 ? LOG_ZERO : m->gapXTable[kC[p]]);
                    doTransition(op, lower + 3 * q, cur + 3 * p, ST_M, ST_X, eP, t[SAO_T_GAP_OPEN_X]);
                    doTransition(op, lower + 3 * q, cur + 3 * p, ST_X, ST_X, eP, t[SAO_T_GAP_EXTEND_X]);
                    doTransition(op, lower + 3 * q, cur + 3 * p, ST_Y, ST_X, eP, t[SAO_T_GAP_SWITCH_TO_X]);
                }
    }
#ifdef SAO_FP32_STUDY
    g_nbr = 2;
#endif
    if (middle != NULL) {
        const int32_t *kM = c->pathKmer + c->pathOff[x - 1];
        for (int64_t p = 0; p < nCur; p++)
            for (int64_t q = 0; q < nMiddle; q++)
                if (legal(c, kM[q], kC[p])) {
                    double eP = sao_emission(m, kC[p], eMean, eNoise, 1);
                    doTransition(op, middle + 3 * q, cur + 3 * p, ST_M, ST_M, eP, t[SAO_T_MATCH_CONTINUE]);
                    doTransition(op, middle + 3 * q, cur + 3 * p, ST_X, ST_M, eP, t[SAO_T_MATCH_FROM_GAP_X]);
                    doTransition(op, middle + 3 * q, cur + 3 * p, ST_Y, ST_M, eP, t[SAO_T_MATCH_FROM_GAP_Y]);
                }
    }
#ifdef SAO_FP32_STUDY
    g_nbr = 1;
#endif
    if (upper != NULL) {
        for (int64_t p = 0; p < nCur; p++)
            for (int64_t q = 0; q < nUpper; q++)
                if (sameKmer(kC[p], kC[q])) { /* upper cell has the same x, hence the same path list */
                    double eP = sao_emission(m, kC[p], eMean, eNoise, hdp ? 1 : 0);
                    doTransition(op, upper + 3 * q, cur + 3 * p, ST_M, ST_Y, eP, t[SAO_T_GAP_OPEN_Y]);
                    doTransition(op, upper + 3 * q, cur + 3 * p, ST_Y, ST_Y, eP, t[SAO_T_GAP_EXTEND_Y]);
                }
    }
}

/* diagonalCalculation (:1280-1311): cells in ascending xmy */
static void diagonalCalculation(const Ctx *c, Op *op, DpDiag *cur, DpDiag *m1, DpDiag *m2) {
    Diag d = cur->d;
#ifdef SAO_FP32_STUDY
    g_d = d.xay;
#endif
    for (int64_t xmy = d.xmyL; xmy <= d.xmyR; xmy += 2) {
        int64_t x = diagX(d.xay, xmy), y = diagY(d.xay, xmy);
        int64_t nC = 0, nL = 0, nM = 0, nU = 0;
        double *cc = dd_cell(cur, xmy, &nC);
        double *lo = dd_cell(m1, xmy - 1, &nL);
        double *mi = dd_cell(m2, xmy, &nM);
        double *up = dd_cell(m1, xmy + 1, &nU);
        /* cX = sX->get(elements, x-1) clamps index -1 to 0 (:497-499); cY = event y-1 (may be NULLEVENT) */
        op->xIndexForAssign = x - 1 >= 0 ? x - 1 : 0;
        op->yIndexForAssign = y - 1;
        cellCalculate(c, op, x, y, cc, nC, lo, nL, mi, nM, up, nU);
    }
}

/* dpDiagonal_dotProduct (:1167-1180) with hdCell_totalProbability (:828-843), cell_dotProduct (:879-885) */
static double dotProduct(const Ctx *c, DpDiag *a, DpDiag *b) {
    double total = LOG_ZERO;
    Diag d = a->d;
    for (int64_t xmy = d.xmyL; xmy <= d.xmyR; xmy += 2) {
        int64_t n1 = 0, n2 = 0;
        double *c1 = dd_cell(a, xmy, &n1), *c2 = dd_cell(b, xmy, &n2);
        const int32_t *km = c->pathKmer + c->pathOff[diagX(d.xay, xmy)];
        double cellTotal = LOG_ZERO;
        for (int64_t p = 0; p < n1; p++)
            for (int64_t q = 0; q < n2; q++)
                if (sameKmer(km[p], km[q])) {
#ifdef SAO_FP32_STUDY
                    const double oF = g_offF[d.xay], oB = g_offB[d.xay];
                    float t = (float) (c1[3 * p] - oF) + (float) (c2[3 * q] - oB);
                    t = logAddF(t, (float) (c1[3 * p + 1] - oF) + (float) (c2[3 * q + 1] - oB));
                    t = logAddF(t, (float) (c1[3 * p + 2] - oF) + (float) (c2[3 * q + 2] - oB));
                    cellTotal = (double) logAddF((float) cellTotal, t);
#else
                    double t = c1[3 * p] + c2[3 * q];
                    t = logAdd(t, c1[3 * p + 1] + c2[3 * q + 1]);
                    t = logAdd(t, c1[3 * p + 2] + c2[3 * q + 2]);
                    cellTotal = logAdd(cellTotal, t);
#endif
                }
#ifdef SAO_FP32_STUDY
        total = (double) logAddF((float) total, (float) cellTotal);
    }
    return total + (g_offF[d.xay] + g_offB[d.xay]);      /* (relative inside the loop, absolute outside) */
#else
        total = logAdd(total, cellTotal);
    }
    return total;
#endif
}

/* diagonalCalculationTotalProbability (:1335-1353) */
static double totalProbability(const Ctx *c, int64_t xay, DpDiag **f, DpDiag **b, int64_t nDiag) {
    double total = dotProduct(c, f[xay], b[xay]);
    DpDiag *fm1 = xay - 1 >= 0 ? f[xay - 1] : NULL;
    DpDiag *bp1 = xay + 1 <= nDiag ? b[xay + 1] : NULL;
    if (bp1 != NULL && fm1 != NULL) {
        DpDiag *md = dd_new(c, bp1->d);
        dd_fill(md, ZERO3);
        Op op = {0};
        diagonalCalculation(c, &op, md, NULL, fm1);
#ifdef SAO_FP32_STUDY
        {   /* both totals relative to the first one's offset, combined in float */
            const double O1 = g_offF[xay] + g_offB[xay], t2 = dotProduct(c, md, bp1);
            total = O1 + (double) logAddF((float) (total - O1), (float) (t2 - O1));
        }
#else
        total = logAdd(total, dotProduct(c, md, bp1));
#endif
        dd_free(md);
    }
    return total;
}

typedef struct {
    int expectations;
    double threshold;
    I64Vec *pairs;        /* 4 per pair: p, x, y, kmerIdx (sub-region coordinates) */
    double *trans9, *likelihood;
    I64Vec *assignK, *assignE;
    double stats[4];
} Sink;

/* diagonalCalculationPosteriorMatchProbs (:1355-1421) */
static void posteriorMatchProbs(const Ctx *c, int64_t xay, DpDiag **f, DpDiag **b, double total, Sink *s) {
    DpDiag *fd = f[xay], *bd = b[xay];
    Diag d = fd->d;
    for (int64_t xmy = d.xmyL; xmy <= d.xmyR; xmy += 2) {
        int64_t x = diagX(d.xay, xmy), y = diagY(d.xay, xmy);
        if (x > 0 && y > 0) {
            int64_t nf = 0, nb = 0;
            double *cf = dd_cell(fd, xmy, &nf), *cb = dd_cell(bd, xmy, &nb);
            const int32_t *km = c->pathKmer + c->pathOff[x];
            for (int64_t fp = 0; fp < nf; fp++)
                for (int64_t bp = 0; bp < nb; bp++)
                    if (sameKmer(km[fp], km[bp])) {
#ifdef SAO_FP32_STUDY
                        const double oF = g_offF[d.xay], oB = g_offB[d.xay];
                        double pp = exp((double) (((float) (cf[3 * fp] - oF) + (float) (cb[3 * bp] - oB)) - (float) (total - (oF + oB))));
#else
                        double pp = exp((cf[3 * fp] + cb[3 * bp]) - total);
#endif
                        if (pp >= s->threshold) {
                            if (pp > 1.0) pp = 1.0;
                            pp = floor(pp * 10000000.0); /* PAIR_ALIGNMENT_PROB_1 */
                            vpush(s->pairs, (int64_t) pp); vpush(s->pairs, x - 1); vpush(s->pairs, y - 1);
                            vpush(s->pairs, km[fp]);
                        }
                    }
        }
    }
}

/* diagonalCalculation_Expectations (:1423-1443) */
static void expectationsCallback(const Ctx *c, int64_t xay, DpDiag **f, DpDiag **b, double total, Sink *s) {
    *s->likelihood += total;
    Op op = {0};
    op.mode = 2; op.total = total; op.threshold = s->threshold; op.trans9 = s->trans9;
    op.assignK = s->assignK; op.assignE = s->assignE; op.isHdp = c->m->kind == SAO_EMIT_HDP;
    diagonalCalculation(c, &op, b[xay], xay - 1 >= 0 ? f[xay - 1] : NULL, xay - 2 >= 0 ? f[xay - 2] : NULL);
}

/* getPosteriorProbsWithBanding (:1450-1590) */
static void bandedDriver(const Ctx *c, const int64_t *anchors, int64_t nA, const sao_params *p, int raggedL,
                         int raggedR, Sink *s) {
    int64_t nDiag = c->lX + c->lY;
    if (nDiag == 0) return;
    Diag *band = band_build(anchors, nA, c->lX, c->lY, p->diagonalExpansion);
    DpDiag **f = calloc(nDiag + 1, sizeof(DpDiag *)), **b = calloc(nDiag + 1, sizeof(DpDiag *));
#ifdef SAO_FP32_STUDY
    free(g_offF); free(g_offB);
    g_offF = calloc(nDiag + 2, sizeof(double)); g_offB = calloc(nDiag + 2, sizeof(double));
#endif
    double sv[3];
    f[0] = dd_new(c, band[0]);
    stateValues(c->m, raggedL ? 1 : 0, sv);
    dd_fill(f[0], sv);
    s->stats[0] += diagWidth(band[0]);
    s->stats[1] += f[0]->off[diagWidth(band[0])];
    int64_t tracedBackTo = 0, xay = 0;
    Op fwd = {0}, bwd = {0};
    bwd.mode = 1;
    while (1) {
        xay++;
        Diag d = band[xay];
        f[xay] = dd_new(c, d);
        dd_fill(f[xay], ZERO3);
#ifdef SAO_FP32_STUDY
        /* SAO_FP32_EVERY: the offset is refreshed only on every n-th diagonal (what a kernel would do to avoid a block-wide
         * maximum per step); default 16 (what sa_wide32.cuh does) */
        g_offF[xay] = (xay % SAO_FP32_EVERY == 0) ? roundedMax(f[xay - 1]->v, 3 * f[xay - 1]->off[diagWidth(f[xay - 1]->d)])
                                                  : g_offF[xay - 1];
#endif
        s->stats[0] += diagWidth(d);
        s->stats[1] += f[xay]->off[diagWidth(d)];
        diagonalCalculation(c, &fwd, f[xay], f[xay - 1], xay - 2 >= 0 ? f[xay - 2] : NULL);
        int atEnd = xay == nDiag;
        int tracebackPoint = xay >= tracedBackTo + p->minDiagsBetweenTraceBack &&
                             diagWidth(d) <= p->diagonalExpansion * 2 + 1;
        if (atEnd || tracebackPoint) {
            b[xay] = dd_new(c, d);
            stateValues(c->m, (atEnd && raggedR) ? 3 : 2, sv);
            dd_fill(b[xay], sv);
#ifdef SAO_FP32_STUDY
            g_offB[xay] = 0.0;
            if (xay >= 1) g_offB[xay - 1] = 0.0;
#endif
            if (xay > tracedBackTo + 1) {
                b[xay - 1] = dd_new(c, band[xay - 1]);
                dd_fill(b[xay - 1], ZERO3);
            }
            int64_t tracedBackFrom = xay - (atEnd ? 0 : p->traceBackDiagonals + 1);
            double total = LOG_ZERO;
            int64_t nPost = 0;
            for (int64_t d2 = xay; d2 > tracedBackTo; d2--) {
                if (d2 > tracedBackTo + 2) {
                    b[d2 - 2] = dd_new(c, band[d2 - 2]);
                    dd_fill(b[d2 - 2], ZERO3);
                }
#ifdef SAO_FP32_STUDY
                /* offset of the diagonal about to be completed two steps from now: the level of the newest complete one */
                if (d2 >= 2) g_offB[d2 - 2] = ((d2 - 2) % SAO_FP32_EVERY == 0) ? roundedMax(b[d2]->v, 3 * b[d2]->off[diagWidth(b[d2]->d)])
                                                                                : g_offB[d2 - 1];
#endif
                if (d2 > tracedBackTo + 1)
                    diagonalCalculation(c, &bwd, b[d2], b[d2 - 1], d2 - 2 >= 0 ? b[d2 - 2] : NULL);
                if (d2 <= tracedBackFrom) {
                    if (nPost++ % 10 == 0) {
                        total = totalProbability(c, d2, f, b, nDiag);
                        s->stats[3] = total;
                    }
                    if (s->expectations) expectationsCallback(c, d2, f, b, total, s);
                    else posteriorMatchProbs(c, d2, f, b, total, s);
                    if (d2 < tracedBackFrom || atEnd) { dd_free(f[d2]); f[d2] = NULL; }
                }
                if (d2 + 1 <= nDiag) { dd_free(b[d2 + 1]); b[d2 + 1] = NULL; }
            }
            dd_free(b[tracedBackTo + 1]); b[tracedBackTo + 1] = NULL;
            dd_free(f[tracedBackTo]); f[tracedBackTo] = NULL;
            tracedBackTo = tracedBackFrom;
            s->stats[2] += 1;
        }
        if (atEnd) break;
    }
    for (int64_t i = 0; i <= nDiag; i++) { dd_free(f[i]); dd_free(b[i]); }
    free(f); free(b); free(band);
}

static int ctx_init(Ctx *c, const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                    const sao_ambig *ambig) {
    c->m = m; c->lX = lX; c->lY = lY; c->events = events;
    c->suffixMod = 1;
    for (int i = 0; i < m->kmerLength - 1; i++) c->suffixMod *= m->alphabetSize;
    c->pathOff = malloc(sizeof(int64_t) * (lX + 2));
    int64_t capPaths = 2 * lX + 16, n = 0;
    c->pathKmer = malloc(sizeof(int32_t) * capPaths);
    c->pathOff[0] = 0;
    c->pathKmer[n++] = -1; /* x == 0: the NULL kmer */
    c->pathOff[1] = n;
    int64_t maxP = 1;
    for (int i = 0; i < m->kmerLength; i++) maxP *= 16;
    int32_t *tmp = malloc(sizeof(int32_t) * 4096);
    int64_t tmpCap = 4096;
    for (int64_t x = 1; x <= lX; x++) {
        int64_t np = sao_paths(target + (x - 1), m, ambig, tmp, tmpCap);
        if (np > tmpCap) {
            tmpCap = np;
            tmp = realloc(tmp, sizeof(int32_t) * tmpCap);
            np = sao_paths(target + (x - 1), m, ambig, tmp, tmpCap);
        }
        if (np < 0) { free(tmp); return -1; }
        if (n + np > capPaths) { capPaths = 2 * (n + np); c->pathKmer = realloc(c->pathKmer, sizeof(int32_t) * capPaths); }
        memcpy(c->pathKmer + n, tmp, sizeof(int32_t) * np);
        n += np;
        c->pathOff[x + 1] = n;
    }
    free(tmp);
    return 0;
}
static void ctx_free(Ctx *c) { free(c->pathOff); free(c->pathKmer); }

/* getPosteriorProbsWithBandingSplittingAlignmentsByLargeGaps (:1953-2016) + list handling of :2043-2080 */
static int64_t splitDriver(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                           const int64_t *anchors, int64_t nA, const sao_params *p, int raggedL, int raggedR,
                           const sao_ambig *ambig, Sink *s, I64Vec *allPairs) {
    sao_ambig defAmbig;
    if (ambig == NULL) { sao_ambig_default(&defAmbig); ambig = &defAmbig; }
    I64Vec sp = {0};
    split_build(anchors, nA, lX, lY, p->splitMatrixBiggerThanThis, raggedL, raggedR, &sp);
    int64_t nSplit = sp.n / 4, j = 0;
    for (int64_t i = 0; i < nSplit; i++) {
        int64_t x1 = sp.v[4 * i], y1 = sp.v[4 * i + 1], x2 = sp.v[4 * i + 2], y2 = sp.v[4 * i + 3];
        I64Vec sub = {0};
        while (j < nA) {
            int64_t x = anchors[2 * j], y = anchors[2 * j + 1];
            if (x + y >= x2 + y2) break;
            vpush(&sub, x - x1); vpush(&sub, y - y1);
            j++;
        }
        Ctx c;
        if (ctx_init(&c, m, target + x1, x2 - x1, events + 4 * y1, y2 - y1, ambig) != 0) {
            free(sub.v); free(sp.v);
            return -1;
        }
        I64Vec regionPairs = {0};
        s->pairs = &regionPairs;
        int64_t assignStart = s->assignK ? s->assignK->n : 0;
        bandedDriver(&c, sub.v, sub.n / 2, p, raggedL || i > 0, raggedR || i < nSplit - 1, s);
        /* alignedPairCoordinateCorrectionFn (:2043-2050): shift, then pop from the END of the sub-list */
        if (allPairs)
            for (int64_t q = regionPairs.n / 4 - 1; q >= 0; q--) {
                vpush(allPairs, regionPairs.v[4 * q]); vpush(allPairs, regionPairs.v[4 * q + 1] + x1);
                vpush(allPairs, regionPairs.v[4 * q + 2] + y1); vpush(allPairs, regionPairs.v[4 * q + 3]);
            }
        if (s->assignK) /* assignments are pointers into the full sequences: convert to global offsets */
            for (int64_t q = assignStart; q < s->assignK->n; q++) { s->assignK->v[q] += x1; s->assignE->v[q] += y1; }
        free(regionPairs.v);
        free(sub.v);
        ctx_free(&c);
    }
    free(sp.v);
    return 0;
}

int64_t sao_align(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                  const int64_t *anchors, int64_t nA, const sao_params *p, int raggedL, int raggedR,
                  const sao_ambig *ambig, int64_t *out, int64_t cap, double *stats) {
    Sink s = {0};
    s.threshold = p->threshold;
    I64Vec all = {0};
    if (splitDriver(m, target, lX, events, lY, anchors, nA, p, raggedL, raggedR, ambig, &s, &all) != 0) {
        free(all.v);
        return -1;
    }
    int64_t n = all.n / 4;
    for (int64_t i = 0; i < n && i < cap; i++) memcpy(out + 4 * i, all.v + 4 * i, sizeof(int64_t) * 4);
    if (stats) memcpy(stats, s.stats, sizeof(s.stats));
    free(all.v);
    return n;
}

int64_t sao_expectations(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                         const int64_t *anchors, int64_t nA, const sao_params *p, int raggedL, int raggedR,
                         const sao_ambig *ambig, double *transitions9, double *likelihood,
                         int64_t *assignKmerOff, int64_t *assignEvent, int64_t cap) {
    Sink s = {0};
    s.expectations = 1;
    s.threshold = p->threshold;
    s.trans9 = transitions9;
    s.likelihood = likelihood;
    I64Vec ak = {0}, ae = {0};
    s.assignK = &ak; s.assignE = &ae;
    if (splitDriver(m, target, lX, events, lY, anchors, nA, p, raggedL, raggedR, ambig, &s, NULL) != 0) {
        free(ak.v); free(ae.v);
        return -1;
    }
    int64_t n = ak.n;
    for (int64_t i = 0; i < n && i < cap; i++) { assignKmerOff[i] = ak.v[i]; assignEvent[i] = ae.v[i]; }
    free(ak.v); free(ae.v);
    return n;
}

int64_t sao_align_noband(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                         const sao_params *p, int raggedL, int raggedR, const sao_ambig *ambig,
                         int64_t *out, int64_t cap) { /* :2104-2162 */
    sao_ambig defAmbig;
    if (ambig == NULL) { sao_ambig_default(&defAmbig); ambig = &defAmbig; }
    Ctx c;
    if (ctx_init(&c, m, target, lX, events, lY, ambig) != 0) return -1;
    int64_t nDiag = lX + lY;
    Diag *band = band_build(NULL, 0, lX, lY, 2);
    DpDiag **f = calloc(nDiag + 1, sizeof(DpDiag *)), **b = calloc(nDiag + 1, sizeof(DpDiag *));
    for (int64_t i = 0; i <= nDiag; i++) {
        f[i] = dd_new(&c, band[i]); dd_fill(f[i], ZERO3);
        b[i] = dd_new(&c, band[i]); dd_fill(b[i], ZERO3);
    }
    double sv[3];
    stateValues(m, raggedL ? 1 : 0, sv); dd_fill(f[0], sv);
    stateValues(m, raggedR ? 3 : 2, sv); dd_fill(b[nDiag], sv);
    Op fwd = {0}, bwd = {0};
    bwd.mode = 1;
    for (int64_t i = 0; i <= nDiag; i++)
        diagonalCalculation(&c, &fwd, f[i], i >= 1 ? f[i - 1] : NULL, i >= 2 ? f[i - 2] : NULL);
    for (int64_t i = nDiag; i > 0; i--)
        diagonalCalculation(&c, &bwd, b[i], b[i - 1], i >= 2 ? b[i - 2] : NULL);
    double total = totalProbability(&c, nDiag, f, b, nDiag);
    I64Vec pairs = {0};
    Sink s = {0};
    s.threshold = p->threshold; s.pairs = &pairs;
    for (int64_t i = 0; i <= nDiag; i++) posteriorMatchProbs(&c, i, f, b, total, &s);
    int64_t n = pairs.n / 4;
    for (int64_t i = 0; i < n && i < cap; i++) memcpy(out + 4 * i, pairs.v + 4 * i, sizeof(int64_t) * 4);
    for (int64_t i = 0; i <= nDiag; i++) { dd_free(f[i]); dd_free(b[i]); }
    free(f); free(b); free(band); free(pairs.v);
    ctx_free(&c);
    return n;
}

/* ------------------------------------------------------------------------------------------- per-read scaling
 * signalUtils_estimateNanoporeParams (impl/signalMachineUtils.c:186-225) as signalMachine runs it before every alignment:
 *   assignments ......... nanopore_getOneDAssignmentsFromRead, impl/nanopore.c:601-631: kmer i of the base-called read is
 *                         paired with event map[i] whenever that event index exceeds every one taken so far
 *   mean parameters ..... nanopore_compute_mean_scale_params(drift_out, var_out), :756-826: weighted least squares of the
 *                         event mean on (1, level_mean, time), weights 1 / level_sd^2 -> shift, scale, drift; var = root of
 *                         the mean squared weighted residual
 *   noise parameters .... nanopore_compute_noise_scale_params, :878-946: the same in two unknowns on the event sd
 *   solver .............. nanopore_lineq_solve, :691-748 (Gaussian elimination; the pivot swap is kept as written,
 *                         including `x_out[swap] = x_out[i]`)
 *   drift adjustment .... nanopore_adjustEventsForDriftP, :633-638: mean -= time * drift for EVERY event of the strand
 * model: K x 5 match table (level_mean, level_sd, noise_mean, noise_sd, noise_lambda); events: n x 4 (mean, sd, duration,
 * start time); out7 = scale, shift, var, shift_sd, scale_sd, var_sd, drift (NanoporeReadAdjustmentParameters order).
 * Returns the number of assignments, or -1 when there are none / a kmer is outside the alphabet (the reference aborts). */
#define SAO_MACHEP 1.11022302462515654042E-16 /* impl/hdp_math_utils.c / nanopore.c MACHEP */
static int lineq_solve(const double *A, const double *b, double *x, int64_t n) {
    double aux[9];
    for (int64_t i = 0; i < n; i++) {
        x[i] = b[i];
        for (int64_t j = 0; j < n; j++) aux[i * n + j] = A[i * n + j];
    }
    for (int64_t i = 0; i < n; i++) {
        if (fabs(aux[i * n + i]) < SAO_MACHEP) {
            int64_t swap = i + 1;
            if (swap >= n) return -1;
            while (aux[swap * n + i] < SAO_MACHEP) {
                swap++;
                if (swap >= n) return -1;
            }
            for (int64_t j = 0; j < n; j++) {
                double t = aux[i * n + j];
                aux[i * n + j] = aux[swap * n + j];
                aux[swap * n + j] = t;
            }
            x[i] = x[swap];
            x[swap] = x[i];
        }
        double factor = 1.0 / aux[i * n + i];
        x[i] *= factor;
        for (int64_t j = 0; j < n; j++) aux[i * n + j] *= factor;
        for (int64_t sub = i + 1; sub < n; sub++) {
            factor = aux[sub * n + i];
            x[sub] -= factor * x[i];
            for (int64_t j = 0; j < n; j++) aux[sub * n + j] -= factor * aux[i * n + j];
        }
    }
    for (int64_t i = n - 1; i >= 0; i--)
        for (int64_t sub = i - 1; sub >= 0; sub--) {
            double factor = aux[sub * n + i];
            x[sub] -= factor * x[i];
            for (int64_t j = 0; j < n; j++) aux[sub * n + j] -= factor * aux[i * n + j];
        }
    return 0;
}

int64_t sao_estimate_params(const double *model, const char *alphabet, int64_t alphabetSize, int64_t k, const char *read,
                            int64_t readLength, const int64_t *eventMap, double *events, int64_t nEvents, int adjustEvents,
                            double *out7) {
    int64_t rows = readLength - (k - 1), n = 0, prev = -1;
    if (rows <= 0) return -1;
    double *ev = malloc(sizeof(double) * rows), *sd = malloc(sizeof(double) * rows), *tm = malloc(sizeof(double) * rows);
    int64_t *id = malloc(sizeof(int64_t) * rows);
    for (int64_t i = 0; i < rows; i++) {
        int64_t e = eventMap[i];
        int64_t ki = sao_kmer_id(read + i, alphabet, alphabetSize, k);
        if (ki < 0) { free(ev); free(sd); free(tm); free(id); return -1; }
        if (e > prev) {
            ev[n] = events[4 * e]; sd[n] = events[4 * e + 1]; tm[n] = events[4 * e + 3]; id[n] = ki;
            n++;
            prev = e;
        }
    }
    if (n == 0) { free(ev); free(sd); free(tm); free(id); return -1; }
    double XWX[9] = {0}, XWy[3] = {0}, beta[3];
    for (int64_t i = 0; i < n; i++) {
        double level_mean = model[id[i] * 5], level_sd = model[id[i] * 5 + 1];
        double inv_var = 1.0 / (level_sd * level_sd);
        double scaled_mean = level_mean * inv_var, scaled_time = tm[i] * inv_var;
        XWX[0] += inv_var; XWX[1] += scaled_mean; XWX[2] += scaled_time;
        XWX[4] += scaled_mean * level_mean; XWX[5] += scaled_mean * tm[i]; XWX[8] += scaled_time * tm[i];
        XWy[0] += inv_var * ev[i]; XWy[1] += scaled_mean * ev[i]; XWy[2] += scaled_time * ev[i];
    }
    XWX[3] = XWX[1]; XWX[6] = XWX[2]; XWX[7] = XWX[5];
    int bad = lineq_solve(XWX, XWy, beta, 3);
    out7[1] = beta[0]; out7[0] = beta[1]; out7[6] = beta[2];
    double disp = 0.0;
    for (int64_t i = 0; i < n; i++) {
        double level_mean = model[id[i] * 5], level_sd = model[id[i] * 5 + 1];
        double predicted = beta[0] + beta[1] * level_mean + beta[2] * tm[i];
        double residual = ev[i] - predicted;
        disp += (residual * residual) / (level_sd * level_sd);
    }
    out7[2] = sqrt(disp / n);
    double NWX[4] = {0}, NWy[2] = {0}, nb[2];
    for (int64_t i = 0; i < n; i++) {
        double noise_mean = model[id[i] * 5 + 2], noise_sd = model[id[i] * 5 + 3];
        double inv_var = 1.0 / (noise_sd * noise_sd);
        double scaled_mean = noise_mean * inv_var;
        NWX[0] += inv_var; NWX[1] += scaled_mean; NWX[3] += scaled_mean * noise_mean;
        NWy[0] += inv_var * sd[i]; NWy[1] += scaled_mean * sd[i];
    }
    NWX[2] = NWX[1];
    bad |= lineq_solve(NWX, NWy, nb, 2);
    out7[3] = nb[0]; out7[4] = nb[1];
    disp = 0.0;
    for (int64_t i = 0; i < n; i++) {
        double noise_mean = model[id[i] * 5 + 2], noise_sd = model[id[i] * 5 + 3];
        double residual = sd[i] - (nb[0] + nb[1] * noise_mean);
        disp += (residual * residual) / (noise_sd * noise_sd);
    }
    out7[5] = sqrt(disp / n);
    if (adjustEvents)
        for (int64_t i = 0; i < nEvents; i++) events[4 * i] = events[4 * i] - events[4 * i + 3] * out7[6];
    free(ev); free(sd); free(tm); free(id);
    return bad ? -1 : n;
}
