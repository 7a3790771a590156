// photon_ml_b200.hpp -- C++ host-side mirror of the reference's objective-function plug-in API over the
// C ABI of photon_ml_b200.h (header only).
//
// The reference is Scala; where no JVM is available this is the host side a compiled caller uses, with
// the reference's names, argument meaning and error behaviour:
//   lib/function/ObjectiveFunction.scala:25-73, DiffFunction.scala:25-52, TwiceDiffFunction.scala:25-52
//   api/function/glm/DistributedGLMLossFunction.scala:49-171 / SingleNodeGLMLossFunction.scala:42-153
//   lib/function/L2Regularization.scala:25-184 (stackable, applied on the host exactly as there)
//   lib/normalization/NormalizationContext.scala:39-111
//   lib/optimization/Optimizer.scala:37-245, TRON.scala:80-339 (the in-tree optimizer, restated exactly)
// (lib/ = photon-lib/src/main/scala/com/linkedin/photon/ml/, api/ = photon-api/...).
// `Data` is a packed DataSet (device-resident CSR + CSC) instead of RDD[LabeledPoint]; `Coefficients` is
// std::vector<double>.  require(...) failures throw IllegalArgumentException, CUDA failures RuntimeError.
#ifndef PHOTON_ML_B200_HPP
#define PHOTON_ML_B200_HPP

#include <cmath>
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "photon_ml_b200.h"

namespace photon_ml_b200 {

using Vector = std::vector<double>;

struct IllegalArgumentException : std::invalid_argument {
    using std::invalid_argument::invalid_argument;
};
struct RuntimeError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void check(int rc)
{
    if (rc == PML_OK) return;
    if (rc == PML_EINVAL) throw IllegalArgumentException(pml_last_error());
    throw RuntimeError(std::string("photon_ml_b200 error ") + std::to_string(rc) + ": " + pml_last_error());
}

class Context {
  public:
    explicit Context(int device = 0) { check(pml_ctx_create(device, &h_)); }
    ~Context() { pml_ctx_destroy(h_); }
    Context(const Context &) = delete;
    Context &operator=(const Context &) = delete;
    pml_ctx *handle() const { return h_; }

  private:
    pml_ctx *h_ = nullptr;
};

// NormalizationContext.scala:39-47 (+ the space conversions :64-111)
struct NormalizationContext {
    std::shared_ptr<const Vector> factors, shifts;  // null = None
    int interceptId = -1;                           // -1 = None

    NormalizationContext() = default;
    NormalizationContext(std::shared_ptr<const Vector> f, std::shared_ptr<const Vector> s, int intercept)
        : factors(std::move(f)), shifts(std::move(s)), interceptId(intercept)
    {
        if (shifts && interceptId < 0) throw IllegalArgumentException("Shift without intercept is illegal.");
        if (factors && shifts && factors->size() != shifts->size())
            throw IllegalArgumentException("Factors and shifts vectors should have the same size");
    }
    Vector modelToOriginalSpace(Vector c) const
    {
        if (factors) for (size_t j = 0; j < c.size(); ++j) c[j] *= (*factors)[j];
        if (shifts) {
            double dot = 0.0;
            for (size_t j = 0; j < c.size(); ++j) dot += c[j] * (*shifts)[j];
            c[(size_t)interceptId] -= dot;
        }
        return c;
    }
    Vector modelToTransformedSpace(Vector c) const
    {
        if (shifts) {
            double dot = 0.0;
            for (size_t j = 0; j < c.size(); ++j) dot += c[j] * (*shifts)[j];
            c[(size_t)interceptId] += dot;
        }
        if (factors) for (size_t j = 0; j < c.size(); ++j) c[j] /= (*factors)[j];
        return c;
    }
};

// LabeledPoint rows packed once (replaces RDD[LabeledPoint] / Iterable[LabeledPoint] as Data)
class DataSet {
  public:
    DataSet(const Context &ctx, int64_t n, int32_t d, const int64_t *row_ptr, const int32_t *col, const double *val,
            const double *labels, const double *offsets = nullptr, const double *weights = nullptr)
        : n_(n), d_(d)
    {
        check(pml_dataset_create_csr(ctx.handle(), n, d, row_ptr, col, val, PML_F64, PML_STORE_AUTO, labels, offsets,
                                     weights, &h_));
    }
    ~DataSet() { pml_dataset_destroy(h_); }
    DataSet(const DataSet &) = delete;
    DataSet &operator=(const DataSet &) = delete;
    int64_t rows() const { return n_; }
    int32_t dimension() const { return d_; }
    pml_dataset *handle() const { return h_; }
    void setOffsets(const double *offsets) { check(pml_dataset_set_offsets(h_, offsets)); }  // addScoresToOffsets
    void setWeights(const double *weights) { check(pml_dataset_set_weights(h_, weights)); }
    void apply(const NormalizationContext &ctx)
    {
        if (applied_ == &ctx) return;
        if (ctx.factors && (int64_t)ctx.factors->size() != d_)
            throw IllegalArgumentException("Size mismatch. Factors vector size: " + std::to_string(ctx.factors->size()) +
                                           " != " + std::to_string(d_) + ".");
        check(pml_dataset_set_normalization(h_, ctx.factors ? ctx.factors->data() : nullptr,
                                            ctx.shifts ? ctx.shifts->data() : nullptr, ctx.interceptId));
        applied_ = &ctx;
    }

  private:
    pml_dataset *h_ = nullptr;
    int64_t n_;
    int32_t d_;
    const NormalizationContext *applied_ = nullptr;
};

// ---- the plug-in API ---------------------------------------------------------------------------
class ObjectiveFunction {
  public:
    using Data = DataSet;
    using Coefficients = Vector;
    virtual ~ObjectiveFunction() = default;
    virtual double value(Data &input, const Coefficients &coefficients, const NormalizationContext &ctx) = 0;
    virtual int domainDimension(const Data &input) const { return input.dimension(); }
};

class DiffFunction : public ObjectiveFunction {
  public:
    virtual Vector gradient(Data &input, const Coefficients &c, const NormalizationContext &ctx) { return calculate(input, c, ctx).second; }
    virtual std::pair<double, Vector> calculate(Data &input, const Coefficients &c, const NormalizationContext &ctx) = 0;
    double value(Data &input, const Coefficients &c, const NormalizationContext &ctx) override { return calculate(input, c, ctx).first; }
};

class TwiceDiffFunction : public DiffFunction {
  public:
    virtual Vector hessianVector(Data &input, const Coefficients &c, const Coefficients &multiplyVector,
                                 const NormalizationContext &ctx) = 0;
    virtual Vector hessianDiagonal(Data &input, const Coefficients &c) = 0;
};

inline void requireSize(size_t got, int64_t want, const char *what)
{
    if ((int64_t)got != want)
        throw IllegalArgumentException(std::string("Size mismatch. ") + what + " size: " + std::to_string(got) + " != " +
                                       std::to_string(want) + ".");
}

// DistributedGLMLossFunction / SingleNodeGLMLossFunction: one-line delegations to the aggregators there,
// one C-ABI call each here.
class GLMLossFunction : public TwiceDiffFunction {
  public:
    explicit GLMLossFunction(int loss) : loss_(loss) {}
    double value(Data &in, const Coefficients &c, const NormalizationContext &ctx) override
    {
        requireSize(c.size(), in.dimension(), "Coefficient");
        in.apply(ctx);
        double f = 0.0;
        check(pml_value_and_gradient(in.handle(), loss_, c.data(), &f, nullptr));
        return f;
    }
    std::pair<double, Vector> calculate(Data &in, const Coefficients &c, const NormalizationContext &ctx) override
    {
        requireSize(c.size(), in.dimension(), "Coefficient");
        in.apply(ctx);
        double f = 0.0;
        Vector g(c.size());
        check(pml_value_and_gradient(in.handle(), loss_, c.data(), &f, g.data()));
        return { f, std::move(g) };
    }
    Vector hessianVector(Data &in, const Coefficients &c, const Coefficients &v, const NormalizationContext &ctx) override
    {
        requireSize(c.size(), in.dimension(), "Coefficient");
        requireSize(v.size(), in.dimension(), "Multiply vector");
        in.apply(ctx);
        // the coefficient vector's address + contents identify it: TRON's CG loop passes the same vector up to 20 times
        if (c.data() != tagged_ptr_ || c != tagged_copy_) {
            tagged_ptr_ = c.data();
            tagged_copy_ = c;
            ++tag_;
        }
        Vector hv(c.size());
        check(pml_hessian_vector(in.handle(), loss_, c.data(), tag_, v.data(), hv.data()));
        return hv;
    }
    Vector hessianDiagonal(Data &in, const Coefficients &c) override
    {
        requireSize(c.size(), in.dimension(), "Coefficient");
        Vector diag(c.size());
        check(pml_hessian_diagonal(in.handle(), loss_, c.data(), diag.data()));
        return diag;
    }

  private:
    int loss_;
    const double *tagged_ptr_ = nullptr;
    Vector tagged_copy_;
    uint64_t tag_ = 0;
};
using DistributedGLMLossFunction = GLMLossFunction;
using SingleNodeGLMLossFunction = GLMLossFunction;

// L2RegularizationTwiceDiff as a stackable mix-in (L2Regularization.scala:139-184)
template <class Base>
class L2RegularizationTwiceDiff : public Base {
  public:
    template <class... A>
    explicit L2RegularizationTwiceDiff(double l2RegWeight, A &&...a) : Base(std::forward<A>(a)...), l2RegWeight(l2RegWeight) {}
    double l2RegWeight;
    double value(typename Base::Data &in, const Vector &c, const NormalizationContext &ctx) override
    {
        return Base::value(in, c, ctx) + l2RegValue(c);
    }
    std::pair<double, Vector> calculate(typename Base::Data &in, const Vector &c, const NormalizationContext &ctx) override
    {
        auto r = Base::calculate(in, c, ctx);
        r.first += l2RegValue(c);
        for (size_t j = 0; j < c.size(); ++j) r.second[j] += c[j] * l2RegWeight;
        return r;
    }
    Vector hessianVector(typename Base::Data &in, const Vector &c, const Vector &v, const NormalizationContext &ctx) override
    {
        Vector hv = Base::hessianVector(in, c, v, ctx);
        for (size_t j = 0; j < v.size(); ++j) hv[j] += v[j] * l2RegWeight;
        return hv;
    }
    Vector hessianDiagonal(typename Base::Data &in, const Vector &c) override
    {
        Vector d = Base::hessianDiagonal(in, c);
        for (double &x : d) x += l2RegWeight;
        return d;
    }

  private:
    double l2RegValue(const Vector &c) const
    {
        double dot = 0.0;
        for (double x : c) dot += x * x;
        return l2RegWeight * dot / 2;
    }
};

// ---- Optimizer + TRON (Optimizer.scala:136-196, TRON.scala:152-339) ------------------------------
struct OptimizerState {
    Vector coefficients;
    double loss = 0.0;
    Vector gradient;
    int iter = 0;
};

inline double dot(const Vector &a, const Vector &b)
{
    double s = 0.0;
    for (size_t j = 0; j < a.size(); ++j) s += a[j] * b[j];
    return s;
}
inline double norm2(const Vector &a) { return std::sqrt(dot(a, a)); }

class TRON {
  public:
    enum Reason { NotConverged, MaxIterations, ObjectiveNotImproving, FunctionValuesConverged, GradientConverged };
    static constexpr int MAX_CG_ITERATIONS = 20;

    explicit TRON(NormalizationContext ctx, int maxNumImprovementFailures = 5, double tolerance = 1.0e-5, int maxNumIterations = 15)
        : ctx_(std::move(ctx)), maxFailures_(maxNumImprovementFailures), relTolerance_(tolerance), maxIter_(maxNumIterations)
    {
    }

    std::pair<Vector, double> optimize(TwiceDiffFunction &f, const Vector &initialCoefficients, DataSet &data)
    {
        const Vector x0 = ctx_.modelToTransformedSpace(initialCoefficients);
        // absolute tolerances from the state at ZERO coefficients (Optimizer.scala:182)
        OptimizerState zero = calculateState(f, Vector(x0.size(), 0.0), data, 0);
        lossAbsTol_ = zero.loss * relTolerance_;
        gradAbsTol_ = norm2(zero.gradient) * relTolerance_;
        OptimizerState cur = calculateState(f, x0, data, 0);
        delta_ = norm2(cur.gradient);  // TRON.init
        OptimizerState prev = cur;
        bool havePrev = false;
        do {
            prev = cur;
            havePrev = true;
            cur = runOneIteration(f, prev, data);
        } while (reason(cur, prev, havePrev) == NotConverged);
        convergenceReason = reason(cur, prev, true);
        iterations = cur.iter;
        return { cur.coefficients, cur.loss };
    }

    Reason convergenceReason = NotConverged;
    int iterations = 0;
    int evaluations = 0;
    int hessianVectorEvaluations = 0;
    std::vector<int> cgIterations;

  private:
    OptimizerState calculateState(TwiceDiffFunction &f, const Vector &c, DataSet &data, int iter)
    {
        auto r = f.calculate(data, c, ctx_);
        ++evaluations;
        return OptimizerState{ c, r.first, std::move(r.second), iter };
    }
    Reason reason(const OptimizerState &cur, const OptimizerState &prev, bool havePrev) const
    {
        if (cur.iter >= maxIter_) return MaxIterations;
        if (havePrev && cur.iter == prev.iter) return ObjectiveNotImproving;
        if (havePrev && std::fabs(cur.loss - prev.loss) <= lossAbsTol_) return FunctionValuesConverged;
        if (norm2(cur.gradient) <= gradAbsTol_) return GradientConverged;
        return NotConverged;
    }
    OptimizerState runOneIteration(TwiceDiffFunction &f, const OptimizerState &st, DataSet &data)
    {
        const double eta0 = 1e-4, eta1 = 0.25, eta2 = 0.75, sigma1 = 0.25, sigma2 = 0.5, sigma3 = 4.0;
        const bool first = st.iter == 0;
        bool improved = false;
        int failures = 0;
        OptimizerState fin = st;
        do {
            Vector step, residual;
            const int cg = truncatedCG(f, st.gradient, delta_, data, st.coefficients, step, residual);
            cgIterations.push_back(cg);
            Vector updated(st.coefficients);
            for (size_t j = 0; j < updated.size(); ++j) updated[j] += step[j];
            const double gs = dot(st.gradient, step);
            const double predicted = -0.5 * (gs - dot(step, residual));
            auto r = f.calculate(data, updated, ctx_);
            ++evaluations;
            const double value = r.first;
            const double actual = st.loss - value;
            const double stepNorm = norm2(step);
            if (first) delta_ = std::min(delta_, stepNorm);
            const double alpha = (value - st.loss - gs <= 0) ? sigma3 : std::max(sigma1, -0.5 * (gs / (value - st.loss - gs)));
            if (actual < eta0 * predicted) delta_ = std::min(std::max(alpha, sigma1) * stepNorm, sigma2 * delta_);
            else if (actual < eta1 * predicted) delta_ = std::max(sigma1 * delta_, std::min(alpha * stepNorm, sigma2 * delta_));
            else if (actual < eta2 * predicted) delta_ = std::max(sigma1 * delta_, std::min(alpha * stepNorm, sigma3 * delta_));
            else delta_ = std::max(delta_, std::min(alpha * stepNorm, sigma3 * delta_));
            if (actual > eta0 * predicted) {
                improved = true;
                fin = OptimizerState{ std::move(updated), value, std::move(r.second), st.iter + 1 };
            } else {
                ++failures;
            }
        } while (!improved && failures < maxFailures_);
        return fin;
    }
    int truncatedCG(TwiceDiffFunction &f, const Vector &gradient, double boundary, DataSet &data, const Vector &coefficients,
                    Vector &step, Vector &residual)
    {
        const size_t d = gradient.size();
        step.assign(d, 0.0);
        residual.resize(d);
        for (size_t j = 0; j < d; ++j) residual[j] = gradient[j] * -1.0;
        Vector direction(residual);
        const double tol = 0.1 * norm2(gradient);
        int iteration = 0;
        bool done = false;
        double rTr = dot(residual, residual);
        while (iteration < MAX_CG_ITERATIONS && !done) {
            if (norm2(residual) <= tol) {
                done = true;
            } else {
                ++iteration;
                const Vector Hd = f.hessianVector(data, coefficients, direction, ctx_);
                ++hessianVectorEvaluations;
                double alpha = rTr / dot(direction, Hd);
                for (size_t j = 0; j < d; ++j) step[j] += direction[j] * alpha;
                if (norm2(step) > boundary) {
                    alpha = -alpha;
                    for (size_t j = 0; j < d; ++j) step[j] += direction[j] * alpha;
                    const double std_ = dot(step, direction), sts = dot(step, step), dtd = dot(direction, direction);
                    const double dsq = boundary * boundary;
                    const double rad = std::sqrt(std_ * std_ + dtd * (dsq - sts));
                    alpha = std_ >= 0 ? (dsq - sts) / (std_ + rad) : (rad - std_) / dtd;
                    for (size_t j = 0; j < d; ++j) step[j] += direction[j] * alpha;
                    alpha = -alpha;
                    for (size_t j = 0; j < d; ++j) residual[j] += Hd[j] * alpha;
                    done = true;
                } else {
                    alpha = -alpha;
                    for (size_t j = 0; j < d; ++j) residual[j] += Hd[j] * alpha;
                    const double rnew = dot(residual, residual);
                    const double beta = rnew / rTr;
                    for (size_t j = 0; j < d; ++j) direction[j] = direction[j] * beta + residual[j];
                    rTr = rnew;
                }
            }
        }
        return iteration;
    }

    NormalizationContext ctx_;
    int maxFailures_;
    double relTolerance_;
    int maxIter_;
    double delta_ = std::numeric_limits<double>::max();
    double lossAbsTol_ = 0.0, gradAbsTol_ = 0.0;
};

// ---- random-effect coordinates: all per-entity problems of a partition as one batch ---------------------------
// What RandomEffectCoordinate.updateModel hands to SingleNodeGLMLossFunction entity by entity
// (api/algorithm/RandomEffectCoordinate.scala:108-119); coefficients / gradients / multiply vectors of all entities
// are concatenated in entity order, coefPtr() gives the offsets.
class Batch {
  public:
    Batch(const Context &ctx, const std::vector<int64_t> &entRowPtr, const std::vector<int32_t> &entDim, const int64_t *row_ptr,
          const int32_t *col_local, const double *val, const double *labels, const double *offsets = nullptr,
          const double *weights = nullptr)
        : entities_((int64_t)entDim.size())
    {
        if (entRowPtr.size() != entDim.size() + 1) throw IllegalArgumentException("entRowPtr must hold one more entry than entDim");
        check(pml_batch_create_csr(ctx.handle(), entities_, entRowPtr.data(), entDim.data(), row_ptr, col_local, val, PML_F64,
                                   PML_STORE_AUTO, labels, offsets, weights, &h_));
        coefPtr_.resize((size_t)entities_ + 1);
        check(pml_batch_coef_ptr(h_, coefPtr_.data()));
    }
    ~Batch() { pml_batch_destroy(h_); }
    Batch(const Batch &) = delete;
    Batch &operator=(const Batch &) = delete;
    int64_t entities() const { return entities_; }
    int64_t totalDimension() const { return coefPtr_.back(); }
    const std::vector<int64_t> &coefPtr() const { return coefPtr_; }
    pml_batch *handle() const { return h_; }

    // per-entity objective values; gradients (concatenated) when wanted
    Vector calculate(int loss, const Vector &coefs, Vector *gradients) const
    {
        requireTotal(coefs.size(), "Coefficient");
        Vector values((size_t)entities_);
        if (gradients) gradients->resize(coefs.size());
        check(pml_batch_value_and_gradient(h_, loss, coefs.data(), values.data(), gradients ? gradients->data() : nullptr));
        return values;
    }
    Vector hessianVector(int loss, const Vector &coefs, const Vector &multiply, uint64_t coefTag = 0) const
    {
        requireTotal(coefs.size(), "Coefficient");
        requireTotal(multiply.size(), "Multiply vector");
        Vector out(coefs.size());
        check(pml_batch_hessian_vector(h_, loss, coefs.data(), coefTag, multiply.data(), out.data()));
        return out;
    }
    void setOffsets(const double *offsets) { check(pml_batch_set_offsets(h_, offsets)); }   // LocalDataSet.addScoresToOffsets
    void setActive(const std::vector<uint8_t> *active)
    {
        if (active && (int64_t)active->size() != entities_) throw IllegalArgumentException("Size mismatch. active size");
        check(pml_batch_set_active(h_, active ? active->data() : nullptr));
    }
    // one NormalizationContext per entity, concatenated like the coefficients; interceptLocal[e] = -1 for none
    void setNormalization(const Vector *factors, const Vector *shifts, const std::vector<int32_t> *interceptLocal)
    {
        if (factors) requireTotal(factors->size(), "Factors vector");
        if (shifts) requireTotal(shifts->size(), "Shifts vector");
        check(pml_batch_set_normalization(h_, factors ? factors->data() : nullptr, shifts ? shifts->data() : nullptr,
                                          interceptLocal ? interceptLocal->data() : nullptr));
    }

  private:
    void requireTotal(size_t got, const char *what) const
    {
        if ((int64_t)got != totalDimension())
            throw IllegalArgumentException(std::string("Size mismatch. ") + what + " size: " + std::to_string(got) + " != " +
                                           std::to_string(totalDimension()) + ".");
    }
    pml_batch *h_ = nullptr;
    int64_t entities_;
    std::vector<int64_t> coefPtr_;
};

// ---- several processes, one GPU each: the merge of treeAggregate over NVLink peer memory -----------------------
// `exchange` is the caller's only host-side communication: it takes this rank's 64-byte handle and returns the handles
// of all ranks concatenated in rank order (an MPI_Allgather, a Spark collect + broadcast, ...).
class Peer {
  public:
    template <class Exchange>
    Peer(const Context &ctx, int rank, int world, int32_t d, Exchange &&exchange)
    {
        check(pml_peer_create(ctx.handle(), rank, world, d, &h_));
        std::vector<char> mine(PML_PEER_HANDLE_BYTES);
        int rc = pml_peer_handle(h_, mine.data());
        if (rc == PML_OK) {
            const std::vector<char> all = exchange(mine);
            if (all.size() != (size_t)world * PML_PEER_HANDLE_BYTES) rc = -1;
            else rc = pml_peer_connect(h_, all.data());
        }
        if (rc != PML_OK) {
            pml_peer_destroy(h_);
            h_ = nullptr;
            if (rc == -1) throw IllegalArgumentException("exchange must return world * PML_PEER_HANDLE_BYTES bytes");
            check(rc);
        }
    }
    ~Peer() { pml_peer_destroy(h_); }
    Peer(const Peer &) = delete;
    Peer &operator=(const Peer &) = delete;
    pml_peer *handle() const { return h_; }
    // DistributedGLMLossFunction.calculate / hessianVector over all ranks' rows; device pointers, asynchronous
    void calculate(const DataSet &data, int loss, const double *coef_dev, double *out_value_dev, double *out_grad_dev) const
    {
        check(pml_value_and_gradient_allreduce_dev(data.handle(), h_, loss, coef_dev, out_value_dev, out_grad_dev));
    }
    void hessianVector(const DataSet &data, int loss, const double *coef_dev, uint64_t coefTag, const double *v_dev,
                       double *out_hv_dev) const
    {
        check(pml_hessian_vector_allreduce_dev(data.handle(), h_, loss, coef_dev, coefTag, v_dev, out_hv_dev));
    }
    // The same with HOST vectors, the shape DistributedObjectiveFunction hands over (DistributedObjectiveFunction.scala:
    // 57-58, 73): every rank passes the same coefficients and uploads only its slice; gatherOutput = false writes only
    // out[slice()] (ranks sharing one host buffer assemble the vector in it), true downloads all of it.  Synchronous.
    std::pair<int64_t, int64_t> slice() const
    {
        int64_t lo = 0, hi = 0;
        check(pml_peer_slice(h_, &lo, &hi));
        return { lo, hi };
    }
    double calculate(const DataSet &data, int loss, const std::vector<double> &coef, std::vector<double> &outGrad,
                     bool gatherOutput = true) const
    {
        double value = 0.0;
        outGrad.resize(coef.size());
        check(pml_value_and_gradient_allreduce(data.handle(), h_, loss, coef.data(), &value, outGrad.data(), gatherOutput ? 1 : 0));
        return value;
    }
    void hessianVector(const DataSet &data, int loss, const std::vector<double> &coef, uint64_t coefTag,
                       const std::vector<double> &v, std::vector<double> &outHv, bool gatherOutput = true) const
    {
        outHv.resize(v.size());
        check(pml_hessian_vector_allreduce(data.handle(), h_, loss, coef.empty() ? nullptr : coef.data(), coefTag, v.data(),
                                           outHv.data(), gatherOutput ? 1 : 0));
    }
    bool timedOut() const
    {
        int st = 0;
        check(pml_peer_status(h_, &st));
        return st != 0;
    }

  private:
    pml_peer *h_ = nullptr;
};

}  // namespace photon_ml_b200

#endif  // PHOTON_ML_B200_HPP
