/* CPU restatement ("port") of the reference consensus round, in C.
 *
 * TEST INFRASTRUCTURE ONLY: the product library never links or calls this file.
 * It is (1) a second, independent oracle for the fixed-point form (must agree
 * bit-for-bit with oracle/fixed_oracle.py) and (2) the timed CPU baseline of
 * bench.py (`cpu_baseline.kind == "port"`), because the reference's own
 * implementation (Cairo on a Starknet VM) cannot run on this box.
 *
 * Reference lines restated (paths relative to /root/reference):
 *   signed_decimal.cairo:52-63 (truncating division), :110-116 (wsad_mul/div)
 *   math.cairo:113-126 smooth_median, :170-189 quadratic deviation,
 *   :208-222 variance, :225-238 quadratic risk, :240-269 average,
 *   :271-292 sqrt, :294-310 interval checks, :320-363 skewness / kurtosis
 *   sort.cairo:13-103 (value asc, index desc selection order)
 *   contract.cairo:345-363, :365-368, :436-439, :370-434, :442-503
 * Float form: the same evaluation order in double precision, plus the
 * notebook's true-median / lower-index-kept variant (ipynb beta :342-371).
 *
 * Build: see oracle/Makefile  (gcc -O2 -march=native -shared -fPIC -pthread); threads are plain pthreads.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <pthread.h>
#include <unistd.h>

typedef __int128 i128;

#define WSAD ((i128)1000000)
#define HALF_WSAD ((i128)500000)

enum {
    ST_OK = 0, ST_INTERVAL_INPUT = 1, ST_SQRT_REL1 = 2, ST_MAX_SPREAD_ZERO = 3,
    ST_REL1_RANGE = 4, ST_SECOND_PASS_SHAPE = 5, ST_SQRT_REL2 = 6, ST_REL2_RANGE = 7,
    ST_SKEW_PANIC = 8, ST_KURT_PANIC = 9, ST_OVERFLOW = 10, ST_FIRST_PASS_SHAPE = 11
};

/* A panic is carried in a per-round context instead of longjmp. */
typedef struct { int code; } ctx_t;
#define PANIC(c, st) do { if (!(c)->code) (c)->code = (st); } while (0)
#define DIV0 100 /* internal marker: division by zero, mapped by the caller */

static i128 add_c(ctx_t *c, i128 a, i128 b) { i128 r; if (__builtin_add_overflow(a, b, &r)) { PANIC(c, ST_OVERFLOW); return 0; } return r; }
static i128 sub_c(ctx_t *c, i128 a, i128 b) { i128 r; if (__builtin_sub_overflow(a, b, &r)) { PANIC(c, ST_OVERFLOW); return 0; } return r; }
static i128 mul_c(ctx_t *c, i128 a, i128 b) { i128 r; if (__builtin_mul_overflow(a, b, &r)) { PANIC(c, ST_OVERFLOW); return 0; } return r; }

/* signed_decimal.cairo:52-63 — C's `/` truncates toward zero, as I128Div does. */
static i128 tdiv(ctx_t *c, i128 a, i128 b) { if (b == 0) { PANIC(c, DIV0); return 0; } return a / b; }
static i128 wmul(ctx_t *c, i128 a, i128 b) { return tdiv(c, add_c(c, mul_c(c, a, b), HALF_WSAD), WSAD); }
static i128 wdiv(ctx_t *c, i128 a, i128 b) { return tdiv(c, add_c(c, mul_c(c, a, WSAD), tdiv(c, b, 2)), b); }

/* math.cairo:271-292 */
static i128 wsqrt(ctx_t *c, i128 v) {
    if (v == 0) return 0;
    i128 g = v / 2, g2 = add_c(c, g, WSAD);
    for (int i = 0; ; ++i) {
        if (g == g2 || i == 50 || c->code) return g;
        i128 n = wdiv(c, v, g);
        g2 = g;
        g = add_c(c, g, n) / 2;
    }
}

/* ---- selection helpers (int64 values) */
/* Only the two middle order statistics of a column are used (math.cairo:113-126 reads sorted[mid-1] and
 * sorted[mid] of alexandria's MergeSort output), and they depend on the multiset alone, so a selection gives
 * the same numbers as the full sort: quickselect (median-of-three pivot, Hoare partition) puts the mid-th
 * smallest at a[mid] with everything left of it <= a[mid]; sorted[mid-1] is then the maximum of the left part. */
#define SELECT_BODY(T)                                                                   \
    int lo = 0, hi = n - 1;                                                              \
    while (hi - lo > 16) {                                                               \
        int m = lo + (hi - lo) / 2;                                                      \
        T t;                                                                             \
        if (a[m] < a[lo]) { t = a[m]; a[m] = a[lo]; a[lo] = t; }                         \
        if (a[hi] < a[lo]) { t = a[hi]; a[hi] = a[lo]; a[lo] = t; }                      \
        if (a[hi] < a[m]) { t = a[hi]; a[hi] = a[m]; a[m] = t; }                         \
        T pv = a[m];                                                                     \
        int i = lo, j = hi;                                                              \
        while (i <= j) {                                                                 \
            while (a[i] < pv) ++i;                                                       \
            while (pv < a[j]) --j;                                                       \
            if (i <= j) { t = a[i]; a[i] = a[j]; a[j] = t; ++i; --j; }                   \
        }                                                                                \
        if (k <= j) hi = j; else if (k >= i) lo = i; else return;                        \
    }                                                                                    \
    for (int i = lo + 1; i <= hi; ++i) {                                                 \
        T v = a[i]; int j = i;                                                           \
        while (j > lo && a[j - 1] > v) { a[j] = a[j - 1]; --j; }                         \
        a[j] = v;                                                                        \
    }
static void select_i64(int64_t *a, int n, int k) { SELECT_BODY(int64_t) }

/* math.cairo:113-126: (s[mid-1] + s[mid]) / 2, mid = n/2; the 3-value branch is dead. */
static i128 smooth_median_i64(ctx_t *c, int64_t *tmp, int n) {
    int mid = n / 2;
    select_i64(tmp, n, mid);
    int64_t below = tmp[0];
    for (int i = 1; i < mid; ++i) if (tmp[i] > below) below = tmp[i];
    return tdiv(c, add_c(c, below, tmp[mid]), 2);
}

static i128 reliability_fixed(ctx_t *c, i128 avg_qr, int dim, int constrained, i128 max_spread, int st_sqrt) {
    ctx_t l = {0};
    i128 r;
    if (constrained) {                               /* contract.cairo:436-439 */
        i128 s = wsqrt(&l, avg_qr / dim);
        if (l.code == DIV0) { PANIC(c, st_sqrt); return 0; }
        if (l.code) { PANIC(c, l.code); return 0; }
        r = sub_c(c, WSAD, mul_c(c, s, 2));
    } else {                                         /* contract.cairo:365-368 */
        i128 s = wsqrt(&l, avg_qr);
        if (l.code == DIV0) { PANIC(c, st_sqrt); return 0; }
        if (l.code) { PANIC(c, l.code); return 0; }
        i128 m = max_spread > s ? s : max_spread;    /* math.cairo:312-318 */
        i128 q = wdiv(&l, m, max_spread);
        if (l.code == DIV0) { PANIC(c, ST_MAX_SPREAD_ZERO); return 0; }
        if (l.code) { PANIC(c, l.code); return 0; }
        r = sub_c(c, WSAD, q);
    }
    return r;
}

static int64_t clamp64(i128 v) { return (int64_t)v; }

/* One round, all N rows present. Outputs written so far stay valid after a panic. */
static int round_fixed(const int64_t *x, int N, int D, int F, int constrained, int64_t max_spread,
                       int want_moments, int64_t *essence1, int64_t *qr, int64_t *rel1,
                       uint8_t *reliable, int64_t *consensus, int64_t *rel2, int64_t *skew,
                       int64_t *kurt, int64_t *tmp /* N */, int *ord /* N */) {
    ctx_t c = {0};
    int R = N - F;
    if (constrained)                                  /* math.cairo:298-310 */
        for (int i = 0; i < N * D; ++i) if (x[i] < 0 || x[i] > 1000000) return ST_INTERVAL_INPUT;
    if (N < 2) return ST_FIRST_PASS_SHAPE;
    /* first pass essence — math.cairo:152-165 */
    for (int d = 0; d < D; ++d) {
        for (int i = 0; i < N; ++i) tmp[i] = x[i * D + d];
        essence1[d] = clamp64(smooth_median_i64(&c, tmp, N));
    }
    /* quadratic risk — math.cairo:225-238 */
    i128 sum = 0;
    for (int i = 0; i < N; ++i) {
        i128 q = 0;
        for (int d = 0; d < D; ++d) { i128 df = sub_c(&c, x[i * D + d], essence1[d]); q = add_c(&c, q, wmul(&c, df, df)); }
        qr[i] = clamp64(q);
        sum = add_c(&c, sum, q);
    }
    if (c.code) return c.code;
    i128 r1 = reliability_fixed(&c, sum / N, D, constrained, max_spread, ST_SQRT_REL1);
    if (c.code) return c.code;
    *rel1 = clamp64(r1);
    if (r1 < 0 || r1 > WSAD) return ST_REL1_RANGE;   /* math.cairo:294-296 */
    /* sort.cairo: (value asc, index desc); contract.cairo:345-363 */
    if (R < 0) return ST_SECOND_PASS_SHAPE;
    for (int i = 0; i < N; ++i) {                    /* rank by counting: exact, order-free */
        int rank = 0;
        for (int j = 0; j < N; ++j) rank += (qr[j] < qr[i]) || (qr[j] == qr[i] && j > i);
        reliable[i] = rank < R;
        ord[i] = rank;
    }
    /* second pass — contract.cairo:472-489 / :402-420 */
    if (constrained ? (R < 2) : (R < 1)) return ST_SECOND_PASS_SHAPE;
    i128 sum2 = 0;
    for (int i = 0; i < N; ++i) if (reliable[i]) sum2 = add_c(&c, sum2, qr[i]);
    for (int d = 0; d < D; ++d) {
        int k = 0;
        if (constrained) {
            for (int i = 0; i < N; ++i) if (reliable[i]) tmp[k++] = x[i * D + d];
            consensus[d] = clamp64(smooth_median_i64(&c, tmp, R));
        } else {
            i128 s = 0;
            for (int i = 0; i < N; ++i) if (reliable[i]) s = add_c(&c, s, x[i * D + d]);
            consensus[d] = clamp64(s / R);
        }
    }
    if (c.code) return c.code;
    i128 r2 = reliability_fixed(&c, sum2 / R, D, constrained, max_spread, ST_SQRT_REL2);
    if (c.code) return c.code;
    *rel2 = clamp64(r2);
    if (r2 < 0 || r2 > WSAD) return ST_REL2_RANGE;
    if (!want_moments) return ST_OK;
    /* skewness for every dimension, then kurtosis — contract.cairo:491-500 */
    for (int pass = 0; pass < 2; ++pass) {
        for (int d = 0; d < D; ++d) {
            ctx_t l = {0};
            i128 s = 0;
            for (int i = 0; i < N; ++i) if (reliable[i]) s = add_c(&l, s, x[i * D + d]);
            i128 mean = s / R;                        /* math.cairo:256-269 */
            i128 v = 0;
            for (int i = 0; i < N; ++i) if (reliable[i]) { i128 df = sub_c(&l, x[i * D + d], mean); v = add_c(&l, v, wmul(&l, df, df)); }
            i128 var = v / R;                         /* math.cairo:208-222 */
            i128 sd = wsqrt(&l, var);
            i128 acc = 0;
            for (int i = 0; i < N && !l.code; ++i) if (reliable[i]) {
                i128 z = wdiv(&l, sub_c(&l, x[i * D + d], mean), sd);
                i128 z2 = wmul(&l, z, z);
                acc = add_c(&l, acc, pass == 0 ? wmul(&l, z2, z) : wmul(&l, z2, z2));
            }
            i128 n = R, out = 0;
            if (!l.code) {
                if (pass == 0) {                      /* math.cairo:337 */
                    out = tdiv(&l, mul_c(&l, acc, n), mul_c(&l, n - 1, n - 2));
                } else {                              /* math.cairo:359-362 */
                    i128 t1 = tdiv(&l, mul_c(&l, mul_c(&l, acc, n), n + 1), n - 1);
                    i128 t2 = mul_c(&l, mul_c(&l, 3 * WSAD, n - 1), n - 1);
                    out = tdiv(&l, sub_c(&l, t1, t2), mul_c(&l, n - 2, n - 3));
                }
            }
            if (l.code == DIV0) return pass == 0 ? ST_SKEW_PANIC : ST_KURT_PANIC;
            if (l.code) return l.code;
            (pass == 0 ? skew : kurt)[d] = clamp64(out);
        }
    }
    return ST_OK;
}

/* ---- plain pthread fan-out over contiguous trial ranges */
typedef void (*range_fn)(void *arg, long lo, long hi);
typedef struct { range_fn fn; void *arg; long lo, hi; } job_t;
static void *job_main(void *p) { job_t *j = (job_t *)p; j->fn(j->arg, j->lo, j->hi); return NULL; }

int svoc_ref_max_threads(void) { long n = sysconf(_SC_NPROCESSORS_ONLN); return n < 1 ? 1 : (int)n; }

static void parallel_for(long T, int nthreads, range_fn fn, void *arg) {
    if (nthreads <= 0) nthreads = svoc_ref_max_threads();
    if (nthreads > 256) nthreads = 256;
    if ((long)nthreads > T) nthreads = T > 0 ? (int)T : 1;
    if (nthreads == 1) { fn(arg, 0, T); return; }
    pthread_t th[256]; job_t jobs[256];
    for (int k = 0; k < nthreads; ++k) {
        jobs[k].fn = fn; jobs[k].arg = arg;
        jobs[k].lo = T * k / nthreads; jobs[k].hi = T * (k + 1) / nthreads;
        pthread_create(&th[k], NULL, job_main, &jobs[k]);
    }
    for (int k = 0; k < nthreads; ++k) pthread_join(th[k], NULL);
}

typedef struct {
    const int64_t *pred; int N, D, F, constrained; int64_t max_spread; int want_moments;
    int64_t *essence1, *qr, *rel1; uint8_t *reliable; int64_t *consensus, *rel2, *skew, *kurt; int32_t *status;
} fixed_args;

static void fixed_range(void *p, long lo, long hi) {
    fixed_args *a = (fixed_args *)p;
    int N = a->N, D = a->D;
    int64_t *tmp = (int64_t *)malloc(sizeof(int64_t) * (size_t)N);
    int *ord = (int *)malloc(sizeof(int) * (size_t)N);
    for (long t = lo; t < hi; ++t)
        a->status[t] = round_fixed(a->pred + (size_t)t * N * D, N, D, a->F, a->constrained, a->max_spread,
                                   a->want_moments, a->essence1 + (size_t)t * D, a->qr + (size_t)t * N,
                                   a->rel1 + t, a->reliable + (size_t)t * N, a->consensus + (size_t)t * D,
                                   a->rel2 + t, a->skew + (size_t)t * D, a->kurt + (size_t)t * D, tmp, ord);
    free(tmp); free(ord);
}

int svoc_ref_fixed(const int64_t *pred, long T, int N, int D, int F, int constrained,
                   int64_t max_spread, int want_moments, int64_t *essence1, int64_t *qr,
                   int64_t *rel1, uint8_t *reliable, int64_t *consensus, int64_t *rel2,
                   int64_t *skew, int64_t *kurt, int32_t *status, int nthreads) {
    if (N < 1 || D < 1 || T < 0) return -1;
    fixed_args a = { pred, N, D, F, constrained, max_spread, want_moments,
                     essence1, qr, rel1, reliable, consensus, rel2, skew, kurt, status };
    parallel_for(T, nthreads, fixed_range, &a);
    return 0;
}

/* ------------------------------------------------------------------ float form (double) */
static void select_d(double *a, int n, int k) { SELECT_BODY(double) }

static double centre_d(double *tmp, int n, int kind) {
    if (kind == 2) { double s = 0; for (int i = 0; i < n; ++i) s += tmp[i]; return s / n; }
    int mid = n / 2;
    select_d(tmp, n, mid);                                                         /* tmp[mid] = sorted[mid] */
    if (kind == 1 && (n & 1)) return tmp[mid];                                     /* np.median, odd length */
    double below = tmp[0];                                                         /* sorted[mid-1] */
    for (int i = 1; i < mid; ++i) if (tmp[i] > below) below = tmp[i];
    return 0.5 * (below + tmp[mid]);                               /* np.median (even) / math.cairo:113-126 */
}

static double rel_d(double avg, int D, int constrained, double ms) {
    if (constrained) return 1.0 - 2.0 * sqrt(avg / D);
    double s = sqrt(avg); return 1.0 - (ms > s ? s : ms) / ms;
}

typedef struct {
    const double *pred; int N, D, F, constrained; double max_spread; int centre, centre2, tie, want_moments;
    double *essence1, *qr, *rel1; uint8_t *reliable; double *consensus, *rel2, *skew, *kurt;
} float_args;

static void float_range(void *p, long lo, long hi) {
    float_args *a = (float_args *)p;
    int N = a->N, D = a->D, R = a->N - a->F, tie = a->tie;
    double *tmp = (double *)malloc(sizeof(double) * (size_t)N);
    for (long t = lo; t < hi; ++t) {
        const double *x = a->pred + (size_t)t * N * D;
        double *e1 = a->essence1 + (size_t)t * D, *q = a->qr + (size_t)t * N, *co = a->consensus + (size_t)t * D;
        uint8_t *rl = a->reliable + (size_t)t * N;
        for (int d = 0; d < D; ++d) { for (int i = 0; i < N; ++i) tmp[i] = x[i * D + d]; e1[d] = centre_d(tmp, N, a->centre); }
        double s = 0;
        for (int i = 0; i < N; ++i) { double acc = 0; for (int d = 0; d < D; ++d) { double df = x[i * D + d] - e1[d]; acc += df * df; } q[i] = acc; s += acc; }
        a->rel1[t] = rel_d(s / N, D, a->constrained, a->max_spread);
        for (int i = 0; i < N; ++i) {
            int rank = 0;
            for (int j = 0; j < N; ++j) rank += (q[j] < q[i]) || (q[j] == q[i] && (tie == 0 ? j > i : j < i));
            rl[i] = rank < R;
        }
        double s2 = 0;
        for (int i = 0; i < N; ++i) if (rl[i]) s2 += q[i];
        for (int d = 0; d < D; ++d) { int k = 0; for (int i = 0; i < N; ++i) if (rl[i]) tmp[k++] = x[i * D + d]; co[d] = centre_d(tmp, R, a->centre2); }
        a->rel2[t] = rel_d(s2 / R, D, a->constrained, a->max_spread);
        if (a->want_moments) {
            for (int d = 0; d < D; ++d) {
                double m = 0, v = 0, s3 = 0, s4 = 0, r = R;
                for (int i = 0; i < N; ++i) if (rl[i]) m += x[i * D + d];
                m /= r;
                for (int i = 0; i < N; ++i) if (rl[i]) { double df = x[i * D + d] - m; v += df * df; }
                double sd = sqrt(v / r);
                for (int i = 0; i < N; ++i) if (rl[i]) { double z = (x[i * D + d] - m) / sd; s3 += z * z * z; s4 += z * z * z * z; }
                a->skew[(size_t)t * D + d] = s3 * r / ((r - 1) * (r - 2));
                a->kurt[(size_t)t * D + d] = (s4 * r * (r + 1) / (r - 1) - 3.0 * (r - 1) * (r - 1)) / ((r - 2) * (r - 3));
            }
        }
    }
    free(tmp);
}

int svoc_ref_float(const double *pred, long T, int N, int D, int F, int constrained, double max_spread,
                   int centre, int centre2, int tie, int want_moments, double *essence1, double *qr,
                   double *rel1, uint8_t *reliable, double *consensus, double *rel2, double *skew,
                   double *kurt, int nthreads) {
    if (N < 2 || D < 1 || T < 0 || F < 0 || F >= N) return -1;
    float_args a = { pred, N, D, F, constrained, max_spread, centre, centre2, tie, want_moments,
                     essence1, qr, rel1, reliable, consensus, rel2, skew, kurt };
    parallel_for(T, nthreads, float_range, &a);
    return 0;
}
