/*
 * glm_oracle.c -- CPU restatement of Photon ML's GLM objective aggregation.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing shipped in photon_ml_b200/ may link, import or execute
 * this file; it exists so that tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs have an independent answer to compare the CUDA path with.
 *
 * Parity status: PINNED against the reference's own known-answer tests and golden vectors
 * (tests/test_oracle_golden.py): MathUtilsTest log1pExp KATs, LogisticLossFunctionTest
 * identities, the two scikit-learn goldens of GameEstimatorIntegTest, the heart.txt
 * normalization metamorphic property, the closed-form Hessian diagonals of
 * DistributedOptimizationProblemTest and the finite-difference properties of
 * SingleNodeObjectiveFunctionTest.  The reference itself (Scala/Spark) cannot be compiled or
 * run here (no JVM), so there is no oracle/_ref build.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * photon-lib/src/main/scala/com/linkedin/photon/ml/  = lib/
 * photon-api/src/main/scala/com/linkedin/photon/ml/  = api/).
 *
 * The row loop is literally the reference's: one row at a time, fp64, sparse dot in storage
 * order, then +offset, then +marginShift, pointwise loss, weight*loss accumulation, axpy into a
 * dense d-vector.  Breeze 0.11.2 (org.scalanlp:breeze_2.11:0.11.2, not in the tree) supplies
 * dot/axpy on SparseVector x DenseVector: sequential fp64 accumulation over the stored entries.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

enum { ORACLE_LOGISTIC = 0, ORACLE_POISSON = 1, ORACLE_SQUARED = 2, ORACLE_SMOOTHED_HINGE = 3 };

/* lib/constants/MathConst.scala:23 */
#define POSITIVE_RESPONSE_THRESHOLD 0.5

/* lib/util/MathUtils.scala:34-40 */
double oracle_log1p_exp(double x)
{
    if (x > 0) return x + log1p(exp(-x));
    return log1p(exp(x));
}

/* api/function/glm/LogisticLossFunction.scala:54 */
static double sigmoid(double z) { return 1.0 / (1.0 + exp(-z)); }

/*
 * (l, l') for each loss:
 *   logistic       api/function/glm/LogisticLossFunction.scala:70-77
 *   poisson        api/function/glm/PoissonLossFunction.scala:40-43
 *   squared        api/function/glm/SquaredLossFunction.scala:42-45
 *   smoothed hinge api/function/svm/SmoothedHingeLossFunction.scala:40-63 (derivative already
 *                  multiplied by the +-1 label)
 */
void oracle_loss_and_dz(int loss, double margin, double label, double *l, double *dl)
{
    switch (loss) {
    case ORACLE_LOGISTIC:
        if (label > POSITIVE_RESPONSE_THRESHOLD) {
            *l = oracle_log1p_exp(-margin);
            *dl = -sigmoid(-margin);
        } else {
            *l = oracle_log1p_exp(margin);
            *dl = sigmoid(margin);
        }
        return;
    case ORACLE_POISSON: {
        double prediction = exp(margin);
        *l = prediction - margin * label;
        *dl = prediction - label;
        return;
    }
    case ORACLE_SQUARED: {
        double delta = margin - label;
        *l = delta * delta / 2.0;
        *dl = delta;
        return;
    }
    default: {
        double modified = (label < POSITIVE_RESPONSE_THRESHOLD) ? -1.0 : 1.0;
        double z = modified * margin;
        double loss_v, deriv;
        if (z <= 0) loss_v = 0.5 - z;
        else if (z < 1) loss_v = 0.5 * (1.0 - z) * (1.0 - z);
        else loss_v = 0.0;
        if (z < 0) deriv = -1.0;
        else if (z < 1) deriv = z - 1.0;
        else deriv = 0.0;
        *l = loss_v;
        *dl = deriv * modified;
        return;
    }
    }
}

/*
 * l'': LogisticLossFunction.scala:86-89, PoissonLossFunction.scala:52,
 * SquaredLossFunction.scala:54.  The smoothed hinge is not twice differentiable in the
 * reference (it only extends DiffFunction); NaN marks the misuse.
 */
double oracle_dzz(int loss, double margin, double label)
{
    (void)label;
    switch (loss) {
    case ORACLE_LOGISTIC: {
        double s = sigmoid(margin);
        return s * (1 - s);
    }
    case ORACLE_POISSON:
        return exp(margin);
    case ORACLE_SQUARED:
        return 1.0;
    default:
        return NAN;
    }
}

/* One data set in the layout LabeledPoint rows flatten to (lib/data/LabeledPoint.scala:32-36). */
typedef struct {
    int64_t n;
    int32_t d;
    const int64_t *row_ptr; /* n+1 */
    const int32_t *col;     /* nnz */
    const double *val;      /* nnz, fp64 as in the reference */
    const double *label;    /* n */
    const double *offset;   /* n or NULL (0) */
    const double *weight;   /* n or NULL (1) */
} oracle_data;

/* Breeze SparseVector dot DenseVector: sequential over stored entries. */
static double sparse_dot(const oracle_data *D, int64_t i, const double *dense)
{
    double s = 0.0;
    for (int64_t k = D->row_ptr[i]; k < D->row_ptr[i + 1]; ++k) s += D->val[k] * dense[D->col[k]];
    return s;
}

static void sparse_axpy(const oracle_data *D, int64_t i, double a, double *dense)
{
    for (int64_t k = D->row_ptr[i]; k < D->row_ptr[i + 1]; ++k) dense[D->col[k]] += a * D->val[k];
}

/*
 * ValueAndGradientAggregator.init, lib/function/glm/ValueAndGradientAggregator.scala:90-122:
 * effective coefficients ec = coef .* factors (or coef), marginShift = -ec . shifts (or 0).
 * Returns 0 on success, 1 when a `require` of the reference fails.
 */
static int effective(int32_t d, const double *coef, const double *factors, const double *shifts,
                     int32_t intercept, double *ec, double *shift_dot)
{
    if (factors) {
        if (intercept >= 0 && factors[intercept] != 1.0) return 1;
        for (int32_t j = 0; j < d; ++j) ec[j] = coef[j] * factors[j];
    } else {
        memcpy(ec, coef, sizeof(double) * (size_t)d);
    }
    double s = 0.0;
    if (shifts) {
        if (intercept < 0) return 1; /* lib/normalization/NormalizationContext.scala:44 */
        if (shifts[intercept] != 0.0) return 1;
        for (int32_t j = 0; j < d; ++j) s += ec[j] * shifts[j];
    }
    *shift_dot = s;
    return 0;
}

/* ValueAndGradientAggregator.getVector, :199-221 */
static void finalise(int32_t d, const double *vector_sum, double prefactor, const double *factors,
                     const double *shifts, double *out)
{
    for (int32_t j = 0; j < d; ++j) {
        double v = vector_sum[j];
        if (shifts) v = v - shifts[j] * prefactor;
        if (factors) v = v * factors[j];
        out[j] = v;
    }
}

/*
 * Raw sums of ValueAndGradientAggregator.add (:132-153) over rows [lo, hi):
 * valueSum, vectorShiftPrefactorSum, vectorSum[d].
 */
static void vg_rows(const oracle_data *D, int loss, const double *ec, double margin_shift,
                    int64_t lo, int64_t hi, double *value_sum, double *prefactor_sum,
                    double *vector_sum)
{
    double vs = 0.0, ps = 0.0;
    for (int64_t i = lo; i < hi; ++i) {
        double off = D->offset ? D->offset[i] : 0.0;
        double wt = D->weight ? D->weight[i] : 1.0;
        /* LabeledPoint.computeMargin (lib/data/LabeledPoint.scala:62) then + marginShift */
        double margin = (sparse_dot(D, i, ec) + off) + margin_shift;
        double l, dl;
        oracle_loss_and_dz(loss, margin, D->label[i], &l, &dl);
        vs += wt * l;
        ps += wt * dl;
        sparse_axpy(D, i, wt * dl, vector_sum);
    }
    *value_sum = vs;
    *prefactor_sum = ps;
}

/*
 * ValueAndGradientAggregator.calculateValueAndGradient, Iterable form (:261-274): a single
 * sequential aggregate.  The smoothed hinge follows SmoothedHingeLossFunction.calculateAt
 * (api/function/svm/SmoothedHingeLossFunction.scala:73-84): raw coefficients, normalization
 * ignored.
 */
int oracle_value_and_gradient(int64_t n, int32_t d, const int64_t *row_ptr, const int32_t *col,
                              const double *val, const double *label, const double *offset,
                              const double *weight, int loss, const double *coef,
                              const double *factors, const double *shifts, int32_t intercept,
                              double *out_value, double *out_grad)
{
    oracle_data D = { n, d, row_ptr, col, val, label, offset, weight };
    if (loss == ORACLE_SMOOTHED_HINGE) { factors = NULL; shifts = NULL; }
    double *ec = (double *)malloc(sizeof(double) * (size_t)(d > 0 ? d : 1));
    double *vsum = (double *)calloc((size_t)(d > 0 ? d : 1), sizeof(double));
    double sdot;
    if (effective(d, coef, factors, shifts, intercept, ec, &sdot)) { free(ec); free(vsum); return 1; }
    double value, pref;
    vg_rows(&D, loss, ec, -sdot, 0, n, &value, &pref, vsum);
    *out_value = value;
    finalise(d, vsum, pref, factors, shifts, out_grad);
    free(ec);
    free(vsum);
    return 0;
}

/* One partition's seqOp loop / one combOp, as thread bodies. */
typedef struct {
    const oracle_data *D; int loss; const double *ec; double margin_shift; int64_t lo, hi; size_t dd;
    double *vsum; double *value; double *pref; const double *merge_from; int32_t d;
} part_job;

static void *part_add_rows(void *arg)
{
    part_job *j = (part_job *)arg;
    memset(j->vsum, 0, sizeof(double) * j->dd);
    vg_rows(j->D, j->loss, j->ec, j->margin_shift, j->lo, j->hi, j->value, j->pref, j->vsum);
    return NULL;
}

static void *part_merge(void *arg)
{
    part_job *j = (part_job *)arg;
    for (int32_t k = 0; k < j->d; ++k) j->vsum[k] += j->merge_from[k];
    return NULL;
}

/*
 * The RDD form (:235-250): `parts` contiguous row partitions, one aggregator each (seqOp = add),
 * merged pairwise (combOp = merge, :161-176) the way treeAggregate does, then finalised once.
 * The partitions run on `parts` pthreads, each with its own dense d-vector -- this is
 * the shape of Spark local[N] and is what bench.py times as the CPU baseline.
 */
int oracle_value_and_gradient_partitioned(int64_t n, int32_t d, const int64_t *row_ptr,
                                          const int32_t *col, const double *val,
                                          const double *label, const double *offset,
                                          const double *weight, int loss, const double *coef,
                                          const double *factors, const double *shifts,
                                          int32_t intercept, int parts, double *out_value,
                                          double *out_grad)
{
    oracle_data D = { n, d, row_ptr, col, val, label, offset, weight };
    if (loss == ORACLE_SMOOTHED_HINGE) { factors = NULL; shifts = NULL; }
    if (parts < 1) parts = 1;
    size_t dd = (size_t)(d > 0 ? d : 1);
    double *ec = (double *)malloc(sizeof(double) * dd);
    double sdot;
    if (effective(d, coef, factors, shifts, intercept, ec, &sdot)) { free(ec); return 1; }
    double *vsums = (double *)malloc(sizeof(double) * dd * (size_t)parts);
    double *values = (double *)calloc((size_t)parts, sizeof(double));
    double *prefs = (double *)calloc((size_t)parts, sizeof(double));
    part_job *jobs = (part_job *)calloc((size_t)parts, sizeof(part_job));
    pthread_t *tids = (pthread_t *)calloc((size_t)parts, sizeof(pthread_t));
    for (int p = 0; p < parts; ++p) {
        part_job j = { &D, loss, ec, -sdot, n * p / parts, n * (p + 1) / parts, dd,
                       vsums + dd * (size_t)p, &values[p], &prefs[p], NULL, 0 };
        jobs[p] = j;
        if (p > 0) pthread_create(&tids[p], NULL, part_add_rows, &jobs[p]);
    }
    part_add_rows(&jobs[0]);
    for (int p = 1; p < parts; ++p) pthread_join(tids[p], NULL);
    /* pairwise tree merge: stride doubling, each level's merges run concurrently */
    for (int stride = 1; stride < parts; stride *= 2) {
        int first = 1;
        for (int p = 0; p < parts - stride; p += 2 * stride) {
            jobs[p].merge_from = vsums + dd * (size_t)(p + stride);
            jobs[p].d = d;
            values[p] += values[p + stride];
            prefs[p] += prefs[p + stride];
            if (!first) pthread_create(&tids[p], NULL, part_merge, &jobs[p]);
            first = 0;
        }
        part_merge(&jobs[0]);
        for (int p = 2 * stride; p < parts - stride; p += 2 * stride) pthread_join(tids[p], NULL);
    }
    free(jobs);
    free(tids);
    *out_value = values[0];
    finalise(d, vsums, prefs[0], factors, shifts, out_grad);
    free(ec);
    free(vsums);
    free(values);
    free(prefs);
    return 0;
}

/*
 * HessianVectorAggregator.init/add (lib/function/glm/HessianVectorAggregator.scala:60-85,
 * 96-121) and calcHessianVector (:164-178): ev = v .* factors, fs = ev . shifts,
 * e_i = wt * l''(z_i) * (x_i . ev - fs); prefactor += e_i; vectorSum += e_i x_i.
 */
int oracle_hessian_vector(int64_t n, int32_t d, const int64_t *row_ptr, const int32_t *col,
                          const double *val, const double *label, const double *offset,
                          const double *weight, int loss, const double *coef, const double *v,
                          const double *factors, const double *shifts, int32_t intercept,
                          double *out_hv)
{
    oracle_data D = { n, d, row_ptr, col, val, label, offset, weight };
    if (loss == ORACLE_SMOOTHED_HINGE) return 2;
    size_t dd = (size_t)(d > 0 ? d : 1);
    double *ec = (double *)malloc(sizeof(double) * dd);
    double *ev = (double *)malloc(sizeof(double) * dd);
    double *vsum = (double *)calloc(dd, sizeof(double));
    double sdot, fs;
    if (effective(d, coef, factors, shifts, intercept, ec, &sdot) ||
        effective(d, v, factors, shifts, intercept, ev, &fs)) {
        free(ec); free(ev); free(vsum);
        return 1;
    }
    double margin_shift = -sdot, pref = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double off = offset ? offset[i] : 0.0;
        double wt = weight ? weight[i] : 1.0;
        double margin = (sparse_dot(&D, i, ec) + off) + margin_shift;
        double dzz = oracle_dzz(loss, margin, label[i]);
        double e = wt * dzz * (sparse_dot(&D, i, ev) - fs);
        pref += e;
        sparse_axpy(&D, i, e, vsum);
    }
    finalise(d, vsum, pref, factors, shifts, out_hv);
    free(ec); free(ev); free(vsum);
    return 0;
}

/*
 * HessianDiagonalAggregator.add (lib/function/glm/HessianDiagonalAggregator.scala:51-61):
 * margin from the raw coefficients (no normalization), vectorSum += wt*l''(z) * (x .* x).
 */
int oracle_hessian_diagonal(int64_t n, int32_t d, const int64_t *row_ptr, const int32_t *col,
                            const double *val, const double *label, const double *offset,
                            const double *weight, int loss, const double *coef, double *out_diag)
{
    oracle_data D = { n, d, row_ptr, col, val, label, offset, weight };
    if (loss == ORACLE_SMOOTHED_HINGE) return 2;
    memset(out_diag, 0, sizeof(double) * (size_t)d);
    for (int64_t i = 0; i < n; ++i) {
        double off = offset ? offset[i] : 0.0;
        double wt = weight ? weight[i] : 1.0;
        double margin = sparse_dot(&D, i, coef) + off;
        double a = wt * oracle_dzz(loss, margin, label[i]);
        for (int64_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k) out_diag[col[k]] += a * (val[k] * val[k]);
    }
    return 0;
}

/*
 * Arbiter: the same value+gradient sums in long double with Neumaier compensation.  Used only
 * when the CUDA path and the plain restatement disagree near 1e-12, to tell which is closer to
 * the exact sum.  (Not a restatement of reference code.)
 */
typedef struct { long double s, c; } kahan;
static void kadd(kahan *k, long double x)
{
    long double t = k->s + x;
    if (fabsl(k->s) >= fabsl(x)) k->c += (k->s - t) + x;
    else k->c += (x - t) + k->s;
    k->s = t;
}

int oracle_value_and_gradient_arbiter(int64_t n, int32_t d, const int64_t *row_ptr,
                                      const int32_t *col, const double *val, const double *label,
                                      const double *offset, const double *weight, int loss,
                                      const double *coef, const double *factors,
                                      const double *shifts, int32_t intercept, double *out_value,
                                      double *out_grad)
{
    if (loss == ORACLE_SMOOTHED_HINGE) { factors = NULL; shifts = NULL; }
    size_t dd = (size_t)(d > 0 ? d : 1);
    double *ec = (double *)malloc(sizeof(double) * dd);
    double sdot_unused;
    if (effective(d, coef, factors, shifts, intercept, ec, &sdot_unused)) { free(ec); return 1; }
    long double sdot = 0.0L;
    if (shifts) for (int32_t j = 0; j < d; ++j) sdot += (long double)ec[j] * shifts[j];
    kahan *vs = (kahan *)calloc(dd, sizeof(kahan));
    kahan value = { 0, 0 }, pref = { 0, 0 };
    for (int64_t i = 0; i < n; ++i) {
        long double m = 0.0L;
        for (int64_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k) m += (long double)val[k] * ec[col[k]];
        m += (offset ? offset[i] : 0.0);
        m -= sdot;
        double wt = weight ? weight[i] : 1.0;
        double l, dl;
        oracle_loss_and_dz(loss, (double)m, label[i], &l, &dl);
        kadd(&value, (long double)wt * l);
        kadd(&pref, (long double)wt * dl);
        for (int64_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k)
            kadd(&vs[col[k]], (long double)(wt * dl) * val[k]);
    }
    *out_value = (double)(value.s + value.c);
    long double P = pref.s + pref.c;
    for (int32_t j = 0; j < d; ++j) {
        long double v = vs[j].s + vs[j].c;
        if (shifts) v -= shifts[j] * P;
        if (factors) v *= factors[j];
        out_grad[j] = (double)v;
    }
    free(ec);
    free(vs);
    return 0;
}

int oracle_max_threads(void)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
