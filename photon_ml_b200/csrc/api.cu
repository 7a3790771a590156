// api.cu -- the C ABI of include/photon_ml_b200.h: contexts, packed data sets, and the
// evaluation entry points that string the kernels of kernels.cu together.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "internal.h"

namespace pml {

static thread_local std::string g_error;

void set_error(const std::string &msg) { g_error = msg; }

int fail(int code, const std::string &msg)
{
    g_error = msg;
    return code;
}

namespace {

struct TmpDev {
    void *p = nullptr;
    ~TmpDev() { if (p) cudaFree(p); }
};

template <typename T>
int dev_alloc(T **out, size_t count, int64_t *bytes)
{
    void *p = nullptr;
    const size_t b = (count ? count : 1) * sizeof(T);
    cudaError_t e = cudaMalloc(&p, b);
    if (e != cudaSuccess) return fail(PML_ENOMEM, std::string("cudaMalloc of ") + std::to_string(b) + " bytes: " + cudaGetErrorString(e));
    *out = static_cast<T *>(p);
    if (bytes) *bytes += (int64_t)b;
    return PML_OK;
}

int upload_optional(cudaStream_t s, double **slot, const double *src, int64_t n, bool from_device, int64_t *bytes)
{
    if (src == nullptr) {
        if (*slot) {
            cudaFree(*slot);
            *slot = nullptr;
            if (bytes) *bytes -= n * 8;
        }
        return PML_OK;
    }
    if (*slot == nullptr) PML_TRY(dev_alloc(slot, (size_t)n, bytes));
    if (n > 0) {
        PML_CUDA(cudaMemcpyAsync(*slot, src, sizeof(double) * (size_t)n,
                                 from_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaStreamSynchronize(s));
    }
    return PML_OK;
}

// nnz/row -> CSR margin-kernel variant (DESIGN.md "row pass variants").  Measured on B200
// (profiles/): with ~50 entries per row 8 lanes per row is fastest (1.28 ms vs 1.6 / 1.3 for 4 /
// 16 lanes and 2.1-2.4 ms for the segment-per-warp and segment-per-CTA variants), i.e. about six
// entries per lane; very short rows go to the entry-chunked variant, which does not care about
// row boundaries.
int pick_variant(int64_t n, int64_t nnz)
{
    const double mean = n > 0 ? (double)nnz / (double)n : 0.0;
    int lanes = 2;
    while (lanes < 32 && lanes * 2 * 6.0 <= mean) lanes *= 2;
    return lanes;
}

bool valid_loss(int loss) { return loss >= PML_LOSS_LOGISTIC && loss <= PML_LOSS_SMOOTHED_HINGE; }

struct Guard {
    pml_ctx *ctx;
    explicit Guard(pml_ctx *c) : ctx(c) { ctx->mu.lock(); }
    ~Guard() { ctx->mu.unlock(); }
};

int count(pml_ctx *ctx, int launched)
{
    if (launched < 0) {
        cudaError_t e = cudaGetLastError();
        return fail(PML_ECUDA, std::string("kernel launch failed: ") + cudaGetErrorString(e));
    }
    ctx->launches += launched;
    return PML_OK;
}

// Launch through this wrapper so that, with profiling on, each kernel class is bracketed by
// CUDA events on the context's stream (bench.py's live per-kernel durations).
template <typename F>
int run(pml_ctx *ctx, int kind, F &&launch)
{
    if (!ctx->profiling) return count(ctx, launch());
    cudaEvent_t a, b;
    PML_CUDA(cudaEventCreate(&a));
    PML_CUDA(cudaEventCreate(&b));
    PML_CUDA(cudaEventRecord(a, ctx->stream));
    const int launched = launch();
    PML_CUDA(cudaEventRecord(b, ctx->stream));
    if (launched > 0) ctx->spans.push_back({ kind, a, b });
    else { cudaEventDestroy(a); cudaEventDestroy(b); }
    return count(ctx, launched);
}

// -------------------------------------------------------------------------------------------
// evaluation pipelines (device pointers; the caller holds the context lock)
// -------------------------------------------------------------------------------------------
int run_col_and_fixup(pml_dataset *ds, bool square, double *out, bool with_scalars, bool write_value)
{
    pml_ctx *ctx = ds->ctx;
    cudaStream_t s = ctx->stream;
    const TiledSide &hot = ds->t.hot;
    if (hot.n_chunks > 0) {
        // hot columns: r slices staged in shared memory, per-(tile, column) sums into `partial` (zeroed when packed:
        // every run is assigned afresh by each pass, slots without a run are never written)
        PML_TRY(run(ctx, PML_K_COL_TILED, [&] {
            return launch_seg_tiled(s, ctx->sm_count, ds->value_bytes, square, hot, ds->r, ds->n);
        }));
        PML_TRY(run(ctx, PML_K_FIXUP, [&] {
            return launch_fixup(s, hot.fix, hot.n_fix, hot.n_fix_long, hot.carry, hot.partial, nullptr, nullptr, 0, nullptr,
                                0, false, false);
        }));
        PML_TRY(run(ctx, PML_K_FINALIZE, [&] { 
            return launch_reduce_partial(s, hot.n_tiles, hot.n_hot_cols, hot.partial, ds->t.nzcol_id, out);
        }));
    }
    SegParams cp;
    std::memset(&cp, 0, sizeof(cp));
    cp.nnz = ds->t.cold_nnz;
    cp.idx = ds->t.csc_row;
    cp.val = ds->t.csc_val;
    cp.run_id = ds->t.cold_run_id;
    cp.chunk_lo = ds->t.chunk_col_lo;
    cp.table = ds->r;
    cp.carry = ds->carry_head;
    cp.n_chunks = ds->t.n_chunks;
    cp.active = ds->col_chunk_active;
    // cold columns cut into super blocks of rows: per-(super block, column) sums (slots without a run stay zero from
    // pack time), then one reduction over the blocks
    const bool super = ds->t.n_super > 1;
    double *col_out = super ? ds->t.cold_partial : out;
    cp.out = col_out;
    PML_TRY(run(ctx, PML_K_COL, [&] {
        return launch_seg_spmv(s, ctx->sm_count, ds->value_bytes, false, square, cp, 0);
    }));
    if (super) {
        PML_TRY(run(ctx, PML_K_FIXUP, [&] {
            return launch_fixup(s, ds->t.fix, ds->t.n_fix, ds->t.n_fix_long, ds->carry_head, col_out, nullptr, nullptr, 0,
                                nullptr, 0, false, false);
        }));
        PML_TRY(run(ctx, PML_K_FINALIZE, [&] {
            return launch_reduce_partial(s, ds->t.n_super, ds->t.cold_cols, ds->t.cold_partial,
                                         ds->t.nzcol_id + hot.n_hot_cols, out);
        }));
    }
    PML_TRY(run(ctx, PML_K_FIXUP, [&] {
        return launch_fixup(s, super ? nullptr : ds->t.fix, super ? 0 : ds->t.n_fix, super ? 0 : ds->t.n_fix_long,
                            ds->carry_head, out, nullptr, nullptr, ds->d, with_scalars ? ds->partials : nullptr,
                            ds->row_grid, write_value, true);
    }));
    return PML_OK;
}

// The per-row half of an evaluation: either one of the warp-oriented row_pass_kernel variants, or
// (variant 1) the entry-chunked seg_spmv_kernel followed by its fix-up and the row epilogue.
int run_rows(pml_dataset *ds, RowPassParams &rp, bool two)
{
    pml_ctx *ctx = ds->ctx;
    cudaStream_t s = ctx->stream;
    // With a sliced-ELL copy of the rows every pass runs one lane per row on it (sell_rows_kernel).  Without (batches,
    // rows too uneven; measured, profiles/): single-vector passes (value+gradient, cached Hessian-vector, diagonal,
    // scores) are fastest on the entry-chunked kernel (1.06 ms vs 1.29 ms at 50 entries/row); the paired
    // (ec, ev) pass of a cold Hessian-vector product is fastest with lanes-per-row (1.59 vs 2.35 ms).
    const int variant = ds->variant_forced ? ds->lanes_per_row : (two ? ds->lanes_per_row : 1);
    rp.block_active = rp.mode == MODE_SCORE ? nullptr : ds->row_block_active;
    const bool sell_pairs = std::getenv("PML_NO_SELL_PAIRS") == nullptr;
    if (ds->sell.idx != nullptr && !ds->variant_forced && (!two || sell_pairs)) {
        // one lane per row over the sliced-ELL copy: every row's dot is written, nothing to fix up
        if (two && ds->zb == nullptr) PML_TRY(dev_alloc(&ds->zb, (size_t)ds->n, &ds->bytes));
        PML_TRY(run(ctx, PML_K_ROW, [&] {
            return launch_sell_rows(s, ctx->sm_count, ds->value_bytes, ds->sell, ds->n, rp.vec, ds->t.m, ds->za,
                                    two ? ds->zb : nullptr);
        }));
        return run(ctx, PML_K_EPILOGUE, [&] { return launch_row_epilogue(s, ctx->sm_count, rp, ds->za, ds->zb, &ds->row_grid); });
    }
    if (variant != 1) {
        return run(ctx, PML_K_ROW, [&] {
            return launch_row_pass(s, ctx->sm_count, ds->value_bytes, variant, two, rp, &ds->row_grid);
        });
    }
    if (two && ds->zb == nullptr) PML_TRY(dev_alloc(&ds->zb, (size_t)ds->n, &ds->bytes));
    if (two && ds->carry_row2 == nullptr) PML_TRY(dev_alloc(&ds->carry_row2, (size_t)ds->rs.n_chunks + 1, &ds->bytes));
    if (ds->rs.n_runs < ds->n) {  // empty rows have no run: their dot product is 0
        PML_CUDA(cudaMemsetAsync(ds->za, 0, sizeof(double) * (size_t)ds->n, s));
        if (two) PML_CUDA(cudaMemsetAsync(ds->zb, 0, sizeof(double) * (size_t)ds->n, s));
    }
    SegParams sp;
    std::memset(&sp, 0, sizeof(sp));
    sp.nnz = ds->nnz;
    sp.idx = ds->col;
    sp.val = ds->val;
    sp.run_id = ds->rs.nzrow_id;
    sp.chunk_lo = ds->rs.chunk_lo;
    sp.table = rp.vec;
    sp.out = ds->za;
    sp.out2 = two ? ds->zb : nullptr;
    sp.carry = ds->carry_row;
    sp.carry2 = two ? ds->carry_row2 : nullptr;
    sp.n_chunks = ds->rs.n_chunks;
    // (scores are wanted for every row, masked entities included)
    sp.active = rp.mode == MODE_SCORE ? nullptr : ds->row_chunk_active;
    PML_TRY(run(ctx, PML_K_ROW, [&] {
        return launch_seg_spmv(s, ctx->sm_count, ds->value_bytes, two, false, sp, ds->t.m);
    }));
    PML_TRY(run(ctx, PML_K_FIXUP, [&] {
        return launch_fixup(s, ds->rs.fix, ds->rs.n_fix, ds->rs.n_fix_long, ds->carry_row, ds->za, two ? ds->carry_row2 : nullptr,
                            two ? ds->zb : nullptr, 0, nullptr, 0, false, false);
    }));
    return run(ctx, PML_K_EPILOGUE, [&] { return launch_row_epilogue(s, ctx->sm_count, rp, ds->za, ds->zb, &ds->row_grid); });
}

// want_vector = false: DistributedGLMLossFunction.value -- only valueSum (and prefactorSum) are wanted, so the
// column pass is skipped and out_raw_dev[0..d) is left zero.
int vg_raw_launches(pml_dataset *ds, int loss, const double *coef_dev, double *out_raw_dev, bool want_vector)
{
    pml_ctx *ctx = ds->ctx;
    cudaStream_t s = ctx->stream;
    const bool hinge = loss == PML_LOSS_SMOOTHED_HINGE;  // ignores normalization (SmoothedHingeLossFunction.scala:73-84)
    const bool norm = !hinge && (ds->factors || ds->shifts);
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    if (!norm && ds->keep_column_order && ds->t.m == ds->d) {
        // a batch of per-entity problems without empty columns: internal ids ARE the caller's ids and there is
        // nothing to scale, so the kernels read the caller's vector directly (one pass over d-sized arrays less)
        rp.vec = coef_dev;
    } else {
        // besides ec = c .* f this moves the coefficients to the columns' internal ids
        int np = 0;
        PML_TRY(run(ctx, PML_K_PREPARE, [&] {
            return launch_prepare_vectors(s, ds->d, coef_dev, nullptr, norm ? ds->factors : nullptr,
                                          norm ? ds->shifts : nullptr, ds->t.rank, ds->vec, ds->shift_partials, &np);
        }));
        rp.vec = ds->vec;
        if (norm && ds->shifts) {
            rp.shift_partials = ds->shift_partials;
            rp.n_shift_partials = np;
        }
    }
    PML_CUDA(cudaMemsetAsync(out_raw_dev, 0, sizeof(double) * (size_t)(ds->d + 2), s));
    rp.n = ds->n;
    rp.row_ptr = ds->row_ptr;
    rp.col = ds->col;
    rp.val = ds->val;
    rp.label = ds->label;
    rp.offset = ds->offset;
    rp.weight = ds->weight;
    rp.r = ds->r;
    rp.partials = ds->partials;
    rp.row_loss = ds->want_row_loss ? ds->row_loss : nullptr;
    rp.mode = MODE_VG;
    rp.loss = loss;
    PML_TRY(run_rows(ds, rp, false));
    if (!want_vector) {
        return run(ctx, PML_K_FIXUP, [&] {
            return launch_fixup(s, nullptr, 0, 0, nullptr, out_raw_dev, nullptr, nullptr, ds->d, ds->partials, ds->row_grid,
                                true, true);
        });
    }
    return run_col_and_fixup(ds, false, out_raw_dev, true, true);
}

// The value+gradient sequence is 8-11 launches and 2-3 memsets with nothing data-dependent on the host side, so
// it is captured once per (data-set state, loss, buffers) into a CUDA graph and replayed: on small problems and on
// small per-GPU shards the gaps between launches are a visible share of the evaluation (profiles/: config 1 and the
// 8-GPU steps).  Any change to the data set (offsets, weights, labels, normalization, variant) bumps `version`
// and the next call re-captures.  Profiling (per-kernel events) and PML_NO_GRAPH=1 bypass the graph; so does a
// caller that hands over different buffers on every call (re-capturing costs more than the launches it saves).
template <typename F>
int replay(pml_dataset *ds, pml_graph &g, const pml_graph_key &key, F &&launches)
{
    pml_ctx *ctx = ds->ctx;
    static const bool disabled = std::getenv("PML_NO_GRAPH") != nullptr;
    if (disabled || ctx->profiling || ds->graph_failed) return launches();
    if (g.exec == nullptr || !(g.key == key)) {
        if (g.misses >= 4 && !(g.last == key)) {  // buffers keep changing: plain launches until a key repeats
            g.last = key;
            return launches();
        }
        const bool same_but_buffers = g.exec != nullptr && g.key.version == key.version && g.key.loss == key.loss &&
                                      g.key.flags == key.flags;
        g.misses = (same_but_buffers && !(g.last == key)) ? g.misses + 1 : 0;
        g.last = key;
        if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
        cudaGraph_t graph = nullptr;
        const int64_t before = ctx->launches;
        if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
            cudaGetLastError();
            ds->graph_failed = true;
            return launches();
        }
        const int rc = launches();
        const cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
        g.kernels = (int)(ctx->launches - before);
        ctx->launches = before;
        if (rc != PML_OK || e != cudaSuccess || graph == nullptr ||
            cudaGraphInstantiate(&g.exec, graph, 0) != cudaSuccess) {
            cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            g.exec = nullptr;
            ds->graph_failed = true;  // e.g. a launch that cannot be captured: stay on plain launches
            if (rc != PML_OK) return rc;
            return launches();
        }
        cudaGraphDestroy(graph);
        g.key = key;
    }
    else g.misses = 0;
    PML_CUDA(cudaGraphLaunch(g.exec, ctx->stream));
    ctx->launches += g.kernels;
    return PML_OK;
}

int vg_raw(pml_dataset *ds, int loss, const double *coef_dev, double *out_raw_dev, bool want_vector = true)
{
    pml_graph_key key;
    key.version = ds->version;
    key.loss = loss;
    key.coef = coef_dev;
    key.out = out_raw_dev;
    key.flags = (want_vector ? 1 : 0) | (ds->want_row_loss ? 2 : 0);
    return replay(ds, ds->vg_graph, key, [&] { return vg_raw_launches(ds, loss, coef_dev, out_raw_dev, want_vector); });
}

// The launches of one Hessian-vector product.  cached = the curvature wt*l''(z) of every row is taken from dcache
// (written by an earlier product at the same coefficients), so only ev is gathered; keep = write dcache.
int hv_raw_launches(pml_dataset *ds, int loss, const double *coef_dev, const double *v_dev, double *out_raw_dev, bool cached,
                    bool keep)
{
    pml_ctx *ctx = ds->ctx;
    cudaStream_t s = ctx->stream;
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.n = ds->n;
    rp.row_ptr = ds->row_ptr;
    rp.col = ds->col;
    rp.val = ds->val;
    rp.label = ds->label;
    rp.offset = ds->offset;
    rp.weight = ds->weight;
    rp.r = ds->r;
    rp.partials = ds->partials;
    rp.loss = loss;
    if (cached && !ds->factors && !ds->shifts && ds->keep_column_order && ds->t.m == ds->d) {
        rp.vec = v_dev;   // as in vg_raw_launches: identity renumbering, nothing to scale
    } else {
        int np = 0;
        PML_TRY(run(ctx, PML_K_PREPARE, [&] {
            return cached ? launch_prepare_vectors(s, ds->d, v_dev, nullptr, ds->factors, ds->shifts, ds->t.rank, ds->vec,
                                                   ds->shift_partials, &np)
                          : launch_prepare_vectors(s, ds->d, coef_dev, v_dev, ds->factors, ds->shifts, ds->t.rank, ds->vec,
                                                   ds->shift_partials, &np);
        }));
        rp.vec = ds->vec;
        if (ds->shifts) {
            rp.shift_partials = ds->shift_partials;
            rp.n_shift_partials = np;
        }
    }
    rp.dcache = (cached || keep) ? ds->dcache : nullptr;
    rp.mode = cached ? MODE_HV_CACHED : MODE_HV_COLD;
    PML_TRY(run_rows(ds, rp, !cached));
    PML_CUDA(cudaMemsetAsync(out_raw_dev, 0, sizeof(double) * (size_t)(ds->d + 2), s));
    return run_col_and_fixup(ds, false, out_raw_dev, true, false);
}

// TRON asks for tens of Hessian-vector products per outer iteration, on small problems each a handful of
// launch-bound kernels: both sequences (with and without the margin pass) are replayed as CUDA graphs like vg_raw.
int hv_raw(pml_dataset *ds, int loss, const double *coef_dev, uint64_t tag, const double *v_dev, double *out_raw_dev)
{
    if (loss == PML_LOSS_SMOOTHED_HINGE)
        return fail(PML_EINVAL, "the smoothed hinge loss is not twice differentiable (it only implements DiffFunction)");
    const bool cached = tag != 0 && ds->cache_valid && ds->cached_tag == tag && ds->cached_loss == loss;
    const bool keep = tag != 0;
    if (!cached) {
        if (coef_dev == nullptr) return fail(PML_EINVAL, "coefficients are required when no cached pass matches the tag");
        // allocate outside the captured sequence
        if (keep && ds->dcache == nullptr) PML_TRY(dev_alloc(&ds->dcache, (size_t)ds->n, &ds->bytes));
        // (the paired pass over the sliced-ELL copy or on the entry-chunked kernel, run_rows)
        if ((ds->sell.idx != nullptr || ds->lanes_per_row == 1) && ds->zb == nullptr)
            PML_TRY(dev_alloc(&ds->zb, (size_t)ds->n, &ds->bytes));
        if (ds->lanes_per_row == 1 && ds->carry_row2 == nullptr)
            PML_TRY(dev_alloc(&ds->carry_row2, (size_t)ds->rs.n_chunks + 1, &ds->bytes));
    }
    pml_graph_key key;
    key.version = ds->version;
    key.loss = loss;
    key.coef = cached ? nullptr : coef_dev;
    key.v = v_dev;
    key.out = out_raw_dev;
    key.flags = keep ? 1 : 0;
    PML_TRY(replay(ds, cached ? ds->hv_cached_graph : ds->hv_cold_graph, key,
                   [&] { return hv_raw_launches(ds, loss, coef_dev, v_dev, out_raw_dev, cached, keep); }));
    if (!cached) {
        ds->cache_valid = keep;
        ds->cached_tag = tag;
        ds->cached_loss = loss;
    }
    return PML_OK;
}

// Device-pointer callers (torch tensors) tend to hand over a different buffer on every call, which would force a
// re-capture each time.  Up to STAGE_DEV_DIM coefficients the vector is copied (device to device, a few us) into the
// data set's own buffer once the pointer differs from the one the graph holds, so the replay keeps hitting.
constexpr int STAGE_DEV_DIM = 1 << 20;

const double *stable_input(pml_dataset *ds, const pml_graph &g, const double *held, const double *p, double *stage)
{
    static const bool disabled = std::getenv("PML_NO_GRAPH") != nullptr;
    if (p == nullptr || p == held || p == stage || g.exec == nullptr || ds->d > STAGE_DEV_DIM || disabled ||
        ds->ctx->profiling || ds->graph_failed)
        return p;
    if (cudaMemcpyAsync(stage, p, sizeof(double) * (size_t)ds->d, cudaMemcpyDeviceToDevice, ds->ctx->stream) != cudaSuccess) {
        cudaGetLastError();
        return p;
    }
    return stage;
}

int finalize(pml_dataset *ds, int loss, const double *raw_dev, double *out_value_dev, double *out_vec_dev)
{
    pml_ctx *ctx = ds->ctx;
    const bool hinge = loss == PML_LOSS_SMOOTHED_HINGE;
    return run(ctx, PML_K_FINALIZE, [&] { return launch_finalize(ctx->stream, ds->d, raw_dev, hinge ? nullptr : ds->factors,
                                      hinge ? nullptr : ds->shifts, out_value_dev, out_vec_dev); });
}

int diag_dev(pml_dataset *ds, int loss, const double *coef_dev, double *out_dev /* d+2 */)
{
    pml_ctx *ctx = ds->ctx;
    cudaStream_t s = ctx->stream;
    if (loss == PML_LOSS_SMOOTHED_HINGE)
        return fail(PML_EINVAL, "the smoothed hinge loss is not twice differentiable (it only implements DiffFunction)");
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.n = ds->n;
    rp.row_ptr = ds->row_ptr;
    rp.col = ds->col;
    rp.val = ds->val;
    rp.label = ds->label;
    rp.offset = ds->offset;
    rp.weight = ds->weight;
    rp.r = ds->r;
    {
        int np = 0;  // raw coefficients (HessianDiagonalAggregator ignores normalization), moved to internal ids
        PML_TRY(run(ctx, PML_K_PREPARE, [&] {
            return launch_prepare_vectors(s, ds->d, coef_dev, nullptr, nullptr, nullptr, ds->t.rank, ds->vec,
                                          ds->shift_partials, &np);
        }));
    }
    rp.vec = ds->vec;
    rp.mode = MODE_DIAG;
    rp.loss = loss;
    PML_TRY(run_rows(ds, rp, false));
    PML_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(double) * (size_t)(ds->d + 2), s));
    return run_col_and_fixup(ds, true, out_dev, false, false);
}

int check_ds(const pml_dataset *ds)
{
    if (ds == nullptr || ds->ctx == nullptr) return fail(PML_EINVAL, "null data set handle");
    return PML_OK;
}

}  // namespace
}  // namespace pml

using namespace pml;

// ===========================================================================================
extern "C" {

const char *pml_last_error(void) { return g_error.c_str(); }

int pml_ctx_create(int device, pml_ctx **out)
{
    if (out == nullptr) return fail(PML_EINVAL, "out is null");
    *out = nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(PML_ECUDA, std::string("no CUDA device available (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return fail(PML_EINVAL, "device index out of range");
    PML_CUDA(cudaSetDevice(device));
    pml_ctx *ctx = new (std::nothrow) pml_ctx();
    if (!ctx) return fail(PML_ENOMEM, "host allocation failed");
    ctx->device = device;
    cudaDeviceProp prop;
    PML_CUDA(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    if (prop.major < 10) {
        delete ctx;
        return fail(PML_ECUDA, "this library is built for sm_100a (B200) only");
    }
    PML_CUDA(cudaStreamCreate(&ctx->stream));
    PML_CUDA(cudaMallocHost(&ctx->h_scalar, 64));
    PML_CUDA(cudaMallocHost(&ctx->h_stage, sizeof(double) * 3 * (SMALL_IO_DIM + 2)));
    *out = ctx;
    return PML_OK;
}

int pml_ctx_destroy(pml_ctx *ctx)
{
    if (!ctx) return PML_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamDestroy(ctx->stream);
    cudaFreeHost(ctx->h_scalar);
    cudaFreeHost(ctx->h_stage);
    delete ctx;
    return PML_OK;
}

int pml_ctx_synchronize(pml_ctx *ctx)
{
    if (!ctx) return fail(PML_EINVAL, "null context");
    PML_CUDA(cudaSetDevice(ctx->device));
    PML_CUDA(cudaStreamSynchronize(ctx->stream));
    {
        // An exchange over peer memory that gave up waiting for another rank has poisoned its result with NaN and
        // raised its status word: the caller that reads results after this synchronize must hear about it.
        std::lock_guard<std::mutex> lock(ctx->mu);
        for (const int *st : ctx->peer_status)
            if (*st != 0) return fail(PML_ECUDA, "a merge over peer memory timed out waiting for another rank");
    }
    return PML_OK;
}

int pml_ctx_stream(pml_ctx *ctx, void **out_stream)
{
    if (!ctx || !out_stream) return fail(PML_EINVAL, "null argument");
    *out_stream = (void *)ctx->stream;
    return PML_OK;
}

int pml_ctx_launch_count(pml_ctx *ctx, int64_t *out)
{
    if (!ctx || !out) return fail(PML_EINVAL, "null argument");
    *out = ctx->launches;
    return PML_OK;
}

int pml_ctx_set_profiling(pml_ctx *ctx, int on)
{
    if (!ctx) return fail(PML_EINVAL, "null context");
    Guard g(ctx);
    ctx->profiling = on != 0;
    return PML_OK;
}

int pml_ctx_read_profile(pml_ctx *ctx, double *out_ms, int64_t *out_launches)
{
    if (!ctx || !out_ms || !out_launches) return fail(PML_EINVAL, "null argument");
    Guard g(ctx);
    PML_CUDA(cudaSetDevice(ctx->device));
    PML_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < PML_K_COUNT; ++k) { out_ms[k] = 0.0; out_launches[k] = 0; }
    for (auto &sp : ctx->spans) {
        float ms = 0.f;
        PML_CUDA(cudaEventElapsedTime(&ms, sp.a, sp.b));
        out_ms[sp.kind] += ms;
        out_launches[sp.kind] += 1;
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    ctx->spans.clear();
    return PML_OK;
}

int pml_host_alloc(void **out, int64_t bytes)
{
    if (!out || bytes < 0) return fail(PML_EINVAL, "bad argument");
    PML_CUDA(cudaMallocHost(out, (size_t)(bytes ? bytes : 1)));
    return PML_OK;
}

int pml_host_free(void *p)
{
    if (p) PML_CUDA(cudaFreeHost(p));
    return PML_OK;
}

int pml_dataset_destroy(pml_dataset *ds)
{
    if (!ds) return PML_OK;
    if (ds->ctx) {
        cudaSetDevice(ds->ctx->device);
        cudaStreamSynchronize(ds->ctx->stream);
    }
    for (pml_graph *g : { &ds->vg_graph, &ds->hv_cold_graph, &ds->hv_cached_graph })
        if (g->exec) cudaGraphExecDestroy(g->exec);
    void *ptrs[] = { ds->row_ptr, ds->col, ds->val, ds->label, ds->offset, ds->weight, ds->t.idx_alloc, ds->t.val_alloc, ds->t.hot.idx16, ds->t.hot.uniform, ds->t.hot.run_id, ds->t.hot.chunk_lo, ds->t.hot.tile_first_chunk, ds->t.hot.fix,
                     ds->t.hot.partial, ds->t.hot.carry, ds->t.cold_slot, ds->t.cold_partial,
                     ds->t.nzcol_ptr, ds->t.nzcol_id, ds->t.rank, ds->t.chunk_col_lo, ds->t.fix, ds->rs.nzrow_id, ds->rs.chunk_lo,
                     ds->rs.fix, ds->za, ds->zb, ds->carry_row, ds->carry_row2, ds->row_loss, ds->row_chunk_active, ds->col_chunk_active, ds->row_block_active, ds->factors, ds->shifts, ds->r,
                     ds->dcache, ds->partials, ds->carry_head, ds->raw, ds->vec, ds->shift_partials, ds->io_a, ds->io_b,
                     ds->io_n, ds->sell.idx, ds->sell.val, ds->sell.base, ds->sell.perm };
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete ds;
    return PML_OK;
}

int pml_dataset_create_csr_dev(pml_ctx *ctx, int64_t n, int32_t d, const int64_t *row_ptr_dev,
                               const int32_t *col_idx_dev, const void *values_dev, int values_dtype,
                               int storage, const double *labels_dev, const double *offsets_dev,
                               const double *weights_dev, pml_dataset **out)
{
    if (!ctx || !out) return fail(PML_EINVAL, "null argument");
    *out = nullptr;
    if (n < 0 || d < 0) return fail(PML_EINVAL, "negative size");
    if (n > 0 && (!row_ptr_dev || !labels_dev)) return fail(PML_EINVAL, "row_ptr and labels are required");
    if (values_dtype != PML_F32 && values_dtype != PML_F64) return fail(PML_EINVAL, "values_dtype must be PML_F32 or PML_F64");
    const int storage_arg = storage;
    storage &= ~PML_KEEP_COLUMN_ORDER_FLAG;
    if (storage < PML_STORE_F32 || storage > PML_STORE_ONES) return fail(PML_EINVAL, "unknown storage mode");
    Guard g(ctx);
    PML_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = ctx->stream;

    pml_dataset *ds = new (std::nothrow) pml_dataset();
    if (!ds) return fail(PML_ENOMEM, "host allocation failed");
    ds->ctx = ctx;
    ds->n = n;
    ds->d = d;
    struct Cleanup {
        pml_dataset *ds;
        pml_ctx *ctx;
        ~Cleanup() {
            if (ds) {
                // pml_dataset_destroy takes no lock, so it is safe under the Guard
                pml_dataset_destroy(ds);
            }
        }
    } cleanup{ ds, ctx };

    PML_TRY(dev_alloc(&ds->row_ptr, (size_t)n + 1, &ds->bytes));
    if (n > 0) {
        int64_t bad = 0;
        PML_TRY(convert_row_ptr(s, row_ptr_dev, ds->row_ptr, n + 1, &bad));
        if (bad) return fail(PML_EINVAL, "row_ptr must start at 0, be non-decreasing and stay below 4e9 entries per device");
        uint32_t last = 0;
        PML_CUDA(cudaMemcpyAsync(&last, ds->row_ptr + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        PML_CUDA(cudaStreamSynchronize(s));
        ds->nnz = last;
    } else {
        PML_CUDA(cudaMemsetAsync(ds->row_ptr, 0, sizeof(uint32_t), s));
        ds->nnz = 0;
    }
    const int64_t nnz = ds->nnz;
    if (nnz > 0 && (!col_idx_dev || !values_dev)) return fail(PML_EINVAL, "col_idx and values are required");
    if (nnz > 0 && d == 0) return fail(PML_EINVAL, "entries present but d == 0");

    int64_t bad_cols = 0;
    PML_TRY(check_columns(s, col_idx_dev, nnz, d, &bad_cols));
    if (bad_cols)  // features.size == dim (ValueAndGradientAggregator.scala:141-143)
        return fail(PML_EINVAL, "column index outside [0, d): " + std::to_string(bad_cols) + " entries");

    const int src_bytes = values_dtype == PML_F32 ? 4 : 8;
    int dst_bytes = src_bytes;
    bool ones = false;
    if (storage == PML_STORE_ONES || storage == PML_STORE_AUTO) {
        PML_TRY(ones_check(s, values_dev, src_bytes, nnz, &ones));
        if (!ones && storage == PML_STORE_ONES)
            return fail(PML_EINVAL, "PML_STORE_ONES requested but some stored value is not exactly 1.0");
    }
    if (ones) dst_bytes = 0;  // indicator features: the indices say it all
    else if (storage == PML_STORE_F64) dst_bytes = 8;
    else if (src_bytes == 8) {
        bool exact = true;
        PML_TRY(narrow_check(s, static_cast<const double *>(values_dev), nnz, &exact));
        if (exact) dst_bytes = 4;
        else if (storage == PML_STORE_F32)
            return fail(PML_EINVAL, "PML_STORE_F32 requested but some values are not exactly representable in fp32; "
                                    "use PML_STORE_F64 or PML_STORE_AUTO");
    }
    ds->value_bytes = dst_bytes;

    // 16 entries of zeroed padding: seg_spmv_kernel's 128-bit loads may read up to 7 entries past nnz
    PML_TRY(dev_alloc(&ds->col, (size_t)nnz + 16, &ds->bytes));
    PML_CUDA(cudaMemsetAsync(ds->col + nnz, 0, sizeof(int32_t) * 16, s));
    if (dst_bytes > 0) {
        void *v = nullptr;
        cudaError_t e = cudaMalloc(&v, (size_t)(nnz + 16) * dst_bytes);
        if (e != cudaSuccess) return fail(PML_ENOMEM, std::string("cudaMalloc values: ") + cudaGetErrorString(e));
        ds->val = v;
        ds->bytes += (nnz + 16) * dst_bytes;
        PML_CUDA(cudaMemsetAsync((char *)v + (size_t)nnz * dst_bytes, 0, (size_t)16 * dst_bytes, s));
    }
    if (nnz > 0) {
        PML_CUDA(cudaMemcpyAsync(ds->col, col_idx_dev, sizeof(int32_t) * (size_t)nnz, cudaMemcpyDeviceToDevice, s));
        if (dst_bytes > 0) PML_TRY(convert_values(s, values_dev, src_bytes, ds->val, dst_bytes, nnz));
    }
    PML_TRY(upload_optional(s, &ds->label, n > 0 ? labels_dev : nullptr, n, true, &ds->bytes));
    if (ds->label == nullptr) PML_TRY(dev_alloc(&ds->label, 1, &ds->bytes));
    PML_TRY(upload_optional(s, &ds->offset, offsets_dev, n, true, &ds->bytes));
    PML_TRY(upload_optional(s, &ds->weight, weights_dev, n, true, &ds->bytes));

    const int chunk_entries = SEG_CHUNK;
    PML_TRY(build_transpose(s, ctx->sm_count, n, d, nnz, ds->row_ptr, ds->col, ds->val, dst_bytes, chunk_entries,
                            (storage_arg & PML_KEEP_COLUMN_ORDER_FLAG) == 0, &ds->t));
    ds->keep_column_order = (storage_arg & PML_KEEP_COLUMN_ORDER_FLAG) != 0;
    ds->bytes += ds->t.bytes;
    // after the transpose (which sorts by the column ids): from here on the CSR column array carries the row-start
    // flags in bit 31
    PML_TRY(build_row_side(s, n, nnz, ds->row_ptr, ds->col, chunk_entries, &ds->rs));
    ds->bytes += ds->rs.bytes;
    // (The rows of the CSR stream keep their ascending column order: re-dealing a row's entries over the lanes for
    // the shared-memory banks, as pack.cu does inside the runs of the tiled CSC region, was measured to slow the pass
    // from 1.03 to 1.12 ms -- in sorted order the lanes that hold row heads read the same few hottest coefficients at
    // the same item, which is a broadcast, not a conflict.)
    // The rows once more as sliced ELL for the one-lane-per-row margin pass (not for batches of per-entity problems:
    // their masks work on chunks of the entry stream; not when rows are so uneven that padding would cost more than
    // 15 % of the entries, PML_SELL_MAX_OVERHEAD; PML_NO_SELL keeps every pass on the entry-chunked stream).
    if (!ds->keep_column_order && std::getenv("PML_NO_SELL") == nullptr) {
        const char *lim = std::getenv("PML_SELL_MAX_OVERHEAD"), *sig = std::getenv("PML_SELL_SIGMA_SHIFT");
        PML_TRY(build_sell(s, n, nnz, ds->row_ptr, ds->col, ds->val, dst_bytes, lim ? std::atof(lim) : 0.15,
                           sig ? std::atoi(sig) : 16,
                           std::getenv("PML_NO_SELL_BANK_ORDER") ? 0 : (int32_t)std::min<int64_t>(ds->t.m, SEG_TABLE_BYTES / 8 - 1),
                           &ds->sell));
        ds->bytes += ds->sell.bytes;
    }
    PML_TRY(dev_alloc(&ds->za, (size_t)n, &ds->bytes));
    PML_TRY(dev_alloc(&ds->carry_row, (size_t)ds->rs.n_chunks + 1, &ds->bytes));

    PML_TRY(dev_alloc(&ds->r, (size_t)n, &ds->bytes));
    PML_TRY(dev_alloc(&ds->partials, 2 * (size_t)MAX_PARTIALS, &ds->bytes));
    PML_TRY(dev_alloc(&ds->shift_partials, 2 * (size_t)MAX_PARTIALS, &ds->bytes));
    PML_TRY(dev_alloc(&ds->carry_head, (size_t)ds->t.n_chunks + 1, &ds->bytes));
    PML_TRY(dev_alloc(&ds->raw, (size_t)d + 2, &ds->bytes));
    PML_TRY(dev_alloc(&ds->vec, 2 * (size_t)d + 2, &ds->bytes));
    PML_TRY(dev_alloc(&ds->io_a, (size_t)d + 2, &ds->bytes));
    PML_TRY(dev_alloc(&ds->io_b, (size_t)d + 2, &ds->bytes));
    ds->lanes_per_row = pick_variant(n, nnz);
    PML_CUDA(cudaStreamSynchronize(s));
    cleanup.ds = nullptr;
    *out = ds;
    return PML_OK;
}

int pml_dataset_create_csr(pml_ctx *ctx, int64_t n, int32_t d, const int64_t *row_ptr, const int32_t *col_idx,
                           const void *values, int values_dtype, int storage, const double *labels,
                           const double *offsets, const double *weights, pml_dataset **out)
{
    if (!ctx || !out) return fail(PML_EINVAL, "null argument");
    *out = nullptr;
    if (n < 0 || d < 0) return fail(PML_EINVAL, "negative size");
    if (n > 0 && (!row_ptr || !labels)) return fail(PML_EINVAL, "row_ptr and labels are required");
    if (values_dtype != PML_F32 && values_dtype != PML_F64) return fail(PML_EINVAL, "values_dtype must be PML_F32 or PML_F64");
    const int64_t nnz = n > 0 ? row_ptr[n] : 0;
    if (n > 0 && (row_ptr[0] != 0 || nnz < 0)) return fail(PML_EINVAL, "row_ptr must start at 0 and be non-decreasing");
    if (nnz > 0 && (!col_idx || !values)) return fail(PML_EINVAL, "col_idx and values are required");
    const int vb = values_dtype == PML_F32 ? 4 : 8;
    TmpDev t_rp, t_col, t_val, t_lab, t_off, t_wt;
    {
        Guard g(ctx);
        PML_CUDA(cudaSetDevice(ctx->device));
        PML_CUDA(cudaMalloc(&t_rp.p, sizeof(int64_t) * (size_t)(n + 1)));
        PML_CUDA(cudaMalloc(&t_col.p, sizeof(int32_t) * (size_t)(nnz + 1)));
        PML_CUDA(cudaMalloc(&t_val.p, (size_t)vb * (size_t)(nnz + 1)));
        PML_CUDA(cudaMalloc(&t_lab.p, sizeof(double) * (size_t)(n + 1)));
        if (n > 0) {
            PML_CUDA(cudaMemcpy(t_rp.p, row_ptr, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice));
            PML_CUDA(cudaMemcpy(t_lab.p, labels, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
        }
        if (nnz > 0) {
            PML_CUDA(cudaMemcpy(t_col.p, col_idx, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice));
            PML_CUDA(cudaMemcpy(t_val.p, values, (size_t)vb * (size_t)nnz, cudaMemcpyHostToDevice));
        }
        if (offsets && n > 0) {
            PML_CUDA(cudaMalloc(&t_off.p, sizeof(double) * (size_t)n));
            PML_CUDA(cudaMemcpy(t_off.p, offsets, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
        }
        if (weights && n > 0) {
            PML_CUDA(cudaMalloc(&t_wt.p, sizeof(double) * (size_t)n));
            PML_CUDA(cudaMemcpy(t_wt.p, weights, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
        }
    }
    return pml_dataset_create_csr_dev(ctx, n, d, (const int64_t *)t_rp.p, (const int32_t *)t_col.p, t_val.p, values_dtype,
                                      storage, (const double *)t_lab.p, (const double *)t_off.p, (const double *)t_wt.p,
                                      out);
}

int pml_dataset_get_info(const pml_dataset *ds, pml_dataset_info *out)
{
    PML_TRY(check_ds(ds));
    if (!out) return fail(PML_EINVAL, "out is null");
    out->n = ds->n;
    out->nnz = ds->nnz;
    out->d = ds->d;
    out->value_bytes = ds->value_bytes;
    out->margin_variant = ds->variant_forced ? ds->lanes_per_row : 1;  // variant of single-vector passes
    out->n_chunks = ds->t.n_chunks;
    out->nonempty_cols = ds->t.m;
    out->device_bytes = ds->bytes;
    out->tiled_hot_cols = ds->t.hot.n_hot_cols;
    out->tiled_tiles = ds->t.hot.n_tiles;
    out->tiled_tile_rows = ds->t.hot.tile_rows;
    out->sliced_ell_padding_permille =
        ds->sell.idx && ds->nnz > 0 ? (int32_t)((ds->sell.len - ds->nnz) * 1000 / ds->nnz) : -1;
    const int64_t V = ds->value_bytes, n = ds->n, d = ds->d, nnz = ds->nnz;
    // CSR stream + CSC stream + (y, o, wt) + r written and read + pointers + (read w, write g)
    out->vg_algorithmic_bytes = 2 * nnz * (V + 4) + n * 24 + n * 16 + (n + 1) * 4 + ((int64_t)d + 1) * 4 + (int64_t)d * 16;
    out->hv_algorithmic_bytes = 2 * nnz * (V + 4) + n * 16 + n * 16 + (n + 1) * 4 + ((int64_t)d + 1) * 4 + (int64_t)d * 24;
    return PML_OK;
}

static int set_rows(pml_dataset *ds, double **slot, const double *src, bool required)
{
    PML_TRY(check_ds(ds));
    if (required && !src && ds->n > 0) return fail(PML_EINVAL, "null array");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    ds->cache_valid = false;
    ++ds->version;
    return upload_optional(ds->ctx->stream, slot, src, ds->n, false, &ds->bytes);
}

int pml_dataset_set_offsets(pml_dataset *ds, const double *offsets) { return ds ? set_rows(ds, &ds->offset, offsets, false) : check_ds(ds); }
int pml_dataset_set_weights(pml_dataset *ds, const double *weights) { return ds ? set_rows(ds, &ds->weight, weights, false) : check_ds(ds); }
int pml_dataset_set_labels(pml_dataset *ds, const double *labels) { return ds ? set_rows(ds, &ds->label, labels, true) : check_ds(ds); }

int pml_dataset_set_normalization(pml_dataset *ds, const double *factors, const double *shifts, int32_t intercept_id)
{
    PML_TRY(check_ds(ds));
    if (intercept_id < -1 || intercept_id >= ds->d) return fail(PML_EINVAL, "intercept id outside [0, d)");
    if (shifts && intercept_id < 0) return fail(PML_EINVAL, "Shift without intercept is illegal.");
    if (factors && intercept_id >= 0 && factors[intercept_id] != 1.0)
        return fail(PML_EINVAL, "The intercept should not be transformed. Intercept scaling factor: " +
                                    std::to_string(factors[intercept_id]));
    if (shifts && intercept_id >= 0 && shifts[intercept_id] != 0.0)
        return fail(PML_EINVAL, "The intercept should not be transformed. Intercept shift: " + std::to_string(shifts[intercept_id]));
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    ds->cache_valid = false;
    ++ds->version;
    ds->intercept = intercept_id;
    cudaStream_t s = ds->ctx->stream;
    auto put = [&](double **slot, const double *src) -> int {
        if (!src) {
            if (*slot) { cudaFree(*slot); *slot = nullptr; }
            return PML_OK;
        }
        if (!*slot) PML_TRY(dev_alloc(slot, (size_t)ds->d, &ds->bytes));
        PML_CUDA(cudaMemcpyAsync(*slot, src, sizeof(double) * (size_t)ds->d, cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaStreamSynchronize(s));
        return PML_OK;
    };
    PML_TRY(put(&ds->factors, factors));
    PML_TRY(put(&ds->shifts, shifts));
    return PML_OK;
}

int pml_dataset_set_margin_variant(pml_dataset *ds, int lanes_per_row)
{
    PML_TRY(check_ds(ds));
    Guard g(ds->ctx);
    if (lanes_per_row == -1) {
        ds->lanes_per_row = pick_variant(ds->n, ds->nnz);
        ds->variant_forced = false;
        ++ds->version;
        return PML_OK;
    }
    ++ds->version;
    ds->variant_forced = true;
    switch (lanes_per_row) {
    case 0: case 1: case 2: case 4: case 8: case 16: case 32:
        ds->lanes_per_row = lanes_per_row;
        return PML_OK;
    default:
        return fail(PML_EINVAL, "variant must be one of -1, 0, 1, 2, 4, 8, 16, 32");
    }
}

// ---- device-buffer entry points -------------------------------------------------------------
int pml_value_and_gradient_raw_dev(pml_dataset *ds, int loss, const double *coef_dev, double *out_raw_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef_dev || !out_raw_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    return vg_raw(ds, loss, coef_dev, out_raw_dev);
}

int pml_hessian_vector_raw_dev(pml_dataset *ds, int loss, const double *coef_dev, uint64_t coef_tag,
                               const double *v_dev, double *out_raw_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!v_dev || !out_raw_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    return hv_raw(ds, loss, coef_dev, coef_tag, v_dev, out_raw_dev);
}

int pml_finalize_dev(pml_dataset *ds, int loss, const double *raw_dev, double *out_value_dev, double *out_vec_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!raw_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    return finalize(ds, loss, raw_dev, out_value_dev, out_vec_dev);
}

int pml_value_and_gradient_dev(pml_dataset *ds, int loss, const double *coef_dev, double *out_value_dev,
                               double *out_grad_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    coef_dev = stable_input(ds, ds->vg_graph, ds->vg_graph.key.coef, coef_dev, ds->io_a);
    PML_TRY(vg_raw(ds, loss, coef_dev, ds->raw, out_grad_dev != nullptr));
    return finalize(ds, loss, ds->raw, out_value_dev, out_grad_dev);
}

int pml_hessian_vector_dev(pml_dataset *ds, int loss, const double *coef_dev, uint64_t coef_tag, const double *v_dev,
                           double *out_hv_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!v_dev || !out_hv_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    const bool cached = coef_tag != 0 && ds->cache_valid && ds->cached_tag == coef_tag && ds->cached_loss == loss;
    const pml_graph &hg = cached ? ds->hv_cached_graph : ds->hv_cold_graph;
    if (!cached) coef_dev = stable_input(ds, hg, hg.key.coef, coef_dev, ds->io_a);
    v_dev = stable_input(ds, hg, hg.key.v, v_dev, ds->io_b);
    PML_TRY(hv_raw(ds, loss, coef_dev, coef_tag, v_dev, ds->raw));
    return finalize(ds, loss, ds->raw, nullptr, out_hv_dev);
}

// ---- host-buffer entry points (the drop-in calls) ------------------------------------------
int pml_value_and_gradient(pml_dataset *ds, int loss, const double *coef, double *out_value, double *out_grad)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef && ds->d > 0) return fail(PML_EINVAL, "null coefficients");
    if (!out_value) return fail(PML_EINVAL, "null out_value");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const size_t db = sizeof(double) * (size_t)ds->d;
    const bool norm = loss != PML_LOSS_SMOOTHED_HINGE && (ds->factors || ds->shifts);
    // Small coefficient vectors go through the context's pinned staging buffers: a copy from/to pageable memory
    // makes the driver stage and synchronise each transfer, which on a latency-bound problem (config 1) costs more
    // than the kernels.  Large vectors are handed to the driver directly (an extra host copy of MBs would cost more).
    const bool staged = ds->d <= SMALL_IO_DIM;
    double *h_in = ds->ctx->h_stage, *h_out = ds->ctx->h_stage + SMALL_IO_DIM + 2;
    if (ds->d > 0) {
        if (staged) std::memcpy(h_in, coef, db);
        PML_CUDA(cudaMemcpyAsync(ds->io_a, staged ? h_in : coef, db, cudaMemcpyHostToDevice, s));
    }
    PML_TRY(vg_raw(ds, loss, ds->io_a, ds->raw, out_grad != nullptr));
    const double *src = ds->raw;
    if (norm && out_grad) {
        PML_TRY(finalize(ds, loss, ds->raw, nullptr, ds->io_b));
        src = ds->io_b;
    }
    if (out_grad && ds->d > 0) PML_CUDA(cudaMemcpyAsync(staged ? h_out : out_grad, src, db, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaMemcpyAsync(ds->ctx->h_scalar, ds->raw + ds->d, sizeof(double), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    if (staged && out_grad && ds->d > 0) std::memcpy(out_grad, h_out, db);
    *out_value = ds->ctx->h_scalar[0];
    return PML_OK;
}

int pml_hessian_vector(pml_dataset *ds, int loss, const double *coef, uint64_t coef_tag, const double *v, double *out_hv)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if ((!v || !out_hv) && ds->d > 0) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const size_t db = sizeof(double) * (size_t)ds->d;
    const bool cached = coef_tag != 0 && ds->cache_valid && ds->cached_tag == coef_tag && ds->cached_loss == loss;
    // small vectors through pinned staging, as in pml_value_and_gradient
    const bool staged = ds->d <= SMALL_IO_DIM;
    double *h_coef = ds->ctx->h_stage, *h_out = h_coef + SMALL_IO_DIM + 2, *h_v = h_out + SMALL_IO_DIM + 2;
    if (!cached) {
        if (!coef && ds->d > 0) return fail(PML_EINVAL, "null coefficients");
        if (ds->d > 0) {
            if (staged) std::memcpy(h_coef, coef, db);
            PML_CUDA(cudaMemcpyAsync(ds->io_a, staged ? h_coef : coef, db, cudaMemcpyHostToDevice, s));
        }
    }
    if (ds->d > 0) {
        if (staged) std::memcpy(h_v, v, db);
        PML_CUDA(cudaMemcpyAsync(ds->io_b, staged ? h_v : v, db, cudaMemcpyHostToDevice, s));
    }
    PML_TRY(hv_raw(ds, loss, ds->io_a, coef_tag, ds->io_b, ds->raw));
    const double *src = ds->raw;
    if (ds->factors || ds->shifts) {
        PML_TRY(finalize(ds, loss, ds->raw, nullptr, ds->io_b));
        src = ds->io_b;
    }
    if (ds->d > 0) PML_CUDA(cudaMemcpyAsync(staged ? h_out : out_hv, src, db, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    if (staged && ds->d > 0) std::memcpy(out_hv, h_out, db);
    return PML_OK;
}

int pml_hessian_diagonal(pml_dataset *ds, int loss, const double *coef, double *out_diag)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if ((!coef || !out_diag) && ds->d > 0) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const size_t db = sizeof(double) * (size_t)ds->d;
    if (ds->d > 0) PML_CUDA(cudaMemcpyAsync(ds->io_a, coef, db, cudaMemcpyHostToDevice, s));
    PML_TRY(diag_dev(ds, loss, ds->io_a, ds->raw));
    if (ds->d > 0) PML_CUDA(cudaMemcpyAsync(out_diag, ds->raw, db, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    return PML_OK;
}

int pml_scores(pml_dataset *ds, const double *coef, double *out_scores)
{
    PML_TRY(check_ds(ds));
    if ((!coef && ds->d > 0) || (!out_scores && ds->n > 0)) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    if (ds->d > 0) PML_CUDA(cudaMemcpyAsync(ds->io_a, coef, sizeof(double) * (size_t)ds->d, cudaMemcpyHostToDevice, s));
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.n = ds->n;
    rp.row_ptr = ds->row_ptr;
    rp.col = ds->col;
    rp.val = ds->val;
    rp.label = ds->label;
    rp.r = ds->r;
    {
        int np = 0;
        PML_TRY(run(ds->ctx, PML_K_PREPARE, [&] {
            return launch_prepare_vectors(s, ds->d, ds->io_a, nullptr, nullptr, nullptr, ds->t.rank, ds->vec,
                                          ds->shift_partials, &np);
        }));
    }
    rp.vec = ds->vec;
    rp.mode = MODE_SCORE;
    PML_TRY(run_rows(ds, rp, false));
    if (ds->n > 0) PML_CUDA(cudaMemcpyAsync(out_scores, ds->r, sizeof(double) * (size_t)ds->n, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    return PML_OK;
}

// ---- coordinate-descent glue on the device ---------------------------------
int pml_scores_dev(pml_dataset *ds, const double *coef_dev, double *out_scores_dev)
{
    PML_TRY(check_ds(ds));
    if ((!coef_dev && ds->d > 0) || (!out_scores_dev && ds->n > 0)) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.n = ds->n;
    rp.row_ptr = ds->row_ptr;
    rp.col = ds->col;
    rp.val = ds->val;
    rp.label = ds->label;
    rp.r = ds->n > 0 ? out_scores_dev : ds->r;
    {
        int np = 0;
        PML_TRY(run(ds->ctx, PML_K_PREPARE, [&] {
            return launch_prepare_vectors(s, ds->d, coef_dev, nullptr, nullptr, nullptr, ds->t.rank, ds->vec,
                                          ds->shift_partials, &np);
        }));
    }
    rp.vec = ds->vec;
    rp.mode = MODE_SCORE;
    return run_rows(ds, rp, false);
}

int pml_loss_from_scores_dev(pml_dataset *ds, int loss, const double *scores_dev, double *out_value_dev)
{
    PML_TRY(check_ds(ds));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if ((!scores_dev && ds->n > 0) || !out_value_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    // the scalar half of `add` on given margins: z_i = score_i + offset_i, sum_i wt_i * l(z_i, y_i)
    RowPassParams rp;
    std::memset(&rp, 0, sizeof(rp));
    rp.n = ds->n;
    rp.label = ds->label;
    rp.offset = ds->offset;
    rp.weight = ds->weight;
    rp.r = ds->r;
    rp.partials = ds->partials;
    rp.mode = MODE_VG;
    rp.loss = loss;
    PML_TRY(run(ds->ctx, PML_K_EPILOGUE, [&] { return launch_row_epilogue(s, ds->ctx->sm_count, rp, scores_dev, nullptr, &ds->row_grid); }));
    PML_TRY(run(ds->ctx, PML_K_FIXUP, [&] {
        return launch_fixup(s, nullptr, 0, 0, nullptr, ds->raw, nullptr, nullptr, 0, ds->partials, ds->row_grid, true, true);
    }));
    PML_CUDA(cudaMemcpyAsync(out_value_dev, ds->raw, sizeof(double), cudaMemcpyDeviceToDevice, s));
    return PML_OK;
}

int pml_dataset_set_offsets_dev(pml_dataset *ds, const double *offsets_dev)
{
    PML_TRY(check_ds(ds));
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    ds->cache_valid = false;
    ++ds->version;
    if (offsets_dev == nullptr) {
        if (ds->offset) { cudaStreamSynchronize(ds->ctx->stream); cudaFree(ds->offset); ds->offset = nullptr; }
        return PML_OK;
    }
    if (ds->offset == nullptr) PML_TRY(dev_alloc(&ds->offset, (size_t)ds->n, &ds->bytes));
    if (ds->n > 0)
        PML_CUDA(cudaMemcpyAsync(ds->offset, offsets_dev, sizeof(double) * (size_t)ds->n, cudaMemcpyDeviceToDevice,
                                 ds->ctx->stream));
    return PML_OK;
}

int pml_offsets_from_scores_dev(pml_ctx *ctx, int64_t n, const double *base_offsets_dev, const double *scores_dev,
                                const int64_t *row_map_dev, double *out_offsets_dev)
{
    if (!ctx) return fail(PML_EINVAL, "null context");
    if (n < 0 || (n > 0 && (!scores_dev || !out_offsets_dev))) return fail(PML_EINVAL, "bad argument");
    Guard g(ctx);
    PML_CUDA(cudaSetDevice(ctx->device));
    return count(ctx, launch_offsets_from_scores(ctx->stream, n, base_offsets_dev, scores_dev, row_map_dev, out_offsets_dev));
}

// ---- batched per-entity problems (random effects) -------------------------------------------
int pml_batch_create_csr(pml_ctx *ctx, int64_t n_entities, const int64_t *ent_row_ptr, const int32_t *ent_dim,
                         const int64_t *row_ptr, const int32_t *col_local, const void *values, int values_dtype,
                         int storage, const double *labels, const double *offsets, const double *weights,
                         pml_batch **out)
{
    if (!ctx || !out) return fail(PML_EINVAL, "null argument");
    *out = nullptr;
    if (n_entities < 0) return fail(PML_EINVAL, "negative size");
    if (n_entities > 0 && (!ent_row_ptr || !ent_dim)) return fail(PML_EINVAL, "ent_row_ptr and ent_dim are required");
    if (n_entities > 0 && ent_row_ptr[0] != 0) return fail(PML_EINVAL, "ent_row_ptr must start at 0");
    const int64_t n = n_entities > 0 ? ent_row_ptr[n_entities] : 0;
    if (n > 0 && (!row_ptr || !labels)) return fail(PML_EINVAL, "row_ptr and labels are required");
    const int64_t nnz = n > 0 ? row_ptr[n] : 0;
    if (values_dtype != PML_F32 && values_dtype != PML_F64) return fail(PML_EINVAL, "values_dtype must be PML_F32 or PML_F64");
    std::vector<int64_t> coef_ptr((size_t)n_entities + 1, 0);
    for (int64_t e = 0; e < n_entities; ++e) {
        if (ent_dim[e] < 0) return fail(PML_EINVAL, "negative entity dimension");
        if (ent_row_ptr[e + 1] < ent_row_ptr[e]) return fail(PML_EINVAL, "ent_row_ptr must be non-decreasing");
        coef_ptr[(size_t)e + 1] = coef_ptr[(size_t)e] + ent_dim[e];
    }
    const int64_t d_total = coef_ptr[(size_t)n_entities];
    if (d_total > 2000000000LL) return fail(PML_EINVAL, "sum of entity dimensions exceeds 2e9; shard entities over more devices");

    const int vb = values_dtype == PML_F32 ? 4 : 8;
    TmpDev t_rp, t_col, t_val, t_lab, t_off, t_wt, t_erp, t_cp, t_dim, t_bad;
    pml_batch *b = new (std::nothrow) pml_batch();
    if (!b) return fail(PML_ENOMEM, "host allocation failed");
    struct Cleanup {
        pml_batch *b;
        ~Cleanup() { if (b) pml_batch_destroy(b); }
    } cleanup{ b };
    b->n_entities = n_entities;
    b->coef_ptr = coef_ptr;
    {
        Guard g(ctx);
        PML_CUDA(cudaSetDevice(ctx->device));
        cudaStream_t s = ctx->stream;
        PML_CUDA(cudaMalloc(&t_rp.p, sizeof(int64_t) * (size_t)(n + 1)));
        PML_CUDA(cudaMalloc(&t_col.p, sizeof(int32_t) * (size_t)(nnz + 1)));
        PML_CUDA(cudaMalloc(&t_val.p, (size_t)vb * (size_t)(nnz + 1)));
        PML_CUDA(cudaMalloc(&t_lab.p, sizeof(double) * (size_t)(n + 1)));
        PML_CUDA(cudaMalloc(&t_cp.p, sizeof(int64_t) * (size_t)(n_entities + 1)));
        PML_CUDA(cudaMalloc(&t_dim.p, sizeof(int32_t) * (size_t)(n_entities + 1)));
        PML_CUDA(cudaMalloc(&t_bad.p, sizeof(unsigned long long)));
        PML_CUDA(cudaMalloc((void **)&b->ent_row_ptr, sizeof(int64_t) * (size_t)(n_entities + 1)));
        PML_CUDA(cudaMalloc((void **)&b->ent_values, sizeof(double) * (size_t)(n_entities + 1)));
        PML_CUDA(cudaMemsetAsync(t_bad.p, 0, sizeof(unsigned long long), s));
        if (n_entities > 0) {
            PML_CUDA(cudaMemcpyAsync(b->ent_row_ptr, ent_row_ptr, sizeof(int64_t) * (size_t)(n_entities + 1), cudaMemcpyHostToDevice, s));
            PML_CUDA(cudaMemcpyAsync(t_cp.p, coef_ptr.data(), sizeof(int64_t) * (size_t)(n_entities + 1), cudaMemcpyHostToDevice, s));
            PML_CUDA(cudaMemcpyAsync(t_dim.p, ent_dim, sizeof(int32_t) * (size_t)n_entities, cudaMemcpyHostToDevice, s));
        } else {
            PML_CUDA(cudaMemsetAsync(b->ent_row_ptr, 0, sizeof(int64_t), s));
        }
        if (n > 0) {
            // row_ptr must be valid before it is dereferenced on the device
            for (int64_t i = 0; i < n; ++i)
                if (row_ptr[i + 1] < row_ptr[i] || row_ptr[i] < 0) return fail(PML_EINVAL, "row_ptr must start at 0 and be non-decreasing");
            if (row_ptr[0] != 0) return fail(PML_EINVAL, "row_ptr must start at 0 and be non-decreasing");
            PML_CUDA(cudaMemcpyAsync(t_rp.p, row_ptr, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice, s));
            PML_CUDA(cudaMemcpyAsync(t_lab.p, labels, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, s));
        }
        if (nnz > 0) {
            if (!col_local || !values) return fail(PML_EINVAL, "col_idx and values are required");
            PML_CUDA(cudaMemcpyAsync(t_col.p, col_local, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, s));
            PML_CUDA(cudaMemcpyAsync(t_val.p, values, (size_t)vb * (size_t)nnz, cudaMemcpyHostToDevice, s));
        }
        if (offsets && n > 0) {
            PML_CUDA(cudaMalloc(&t_off.p, sizeof(double) * (size_t)n));
            PML_CUDA(cudaMemcpyAsync(t_off.p, offsets, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, s));
        }
        if (weights && n > 0) {
            PML_CUDA(cudaMalloc(&t_wt.p, sizeof(double) * (size_t)n));
            PML_CUDA(cudaMemcpyAsync(t_wt.p, weights, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, s));
        }
        // local column j of entity e -> global column coef_ptr[e] + j (checks j < ent_dim[e])
        if (nnz > 0) {
            PML_TRY(count(ctx, launch_globalize_cols(s, n_entities, b->ent_row_ptr, (const int64_t *)t_cp.p,
                                                     (const int32_t *)t_dim.p, (const int64_t *)t_rp.p, (int32_t *)t_col.p,
                                                     (unsigned long long *)t_bad.p)));
        }
        unsigned long long bad = 0;
        PML_CUDA(cudaMemcpyAsync(&bad, t_bad.p, sizeof(bad), cudaMemcpyDeviceToHost, s));
        PML_CUDA(cudaStreamSynchronize(s));
        if (bad) return fail(PML_EINVAL, "local column index outside [0, ent_dim[e]): " + std::to_string(bad) + " entries");
    }
    // block-diagonal: keep the column order (entity by entity), so an entity's coefficients stay contiguous
    PML_TRY(pml_dataset_create_csr_dev(ctx, n, (int32_t)d_total, (const int64_t *)t_rp.p, (const int32_t *)t_col.p, t_val.p,
                                       values_dtype, storage | PML_KEEP_COLUMN_ORDER_FLAG, (const double *)t_lab.p, (const double *)t_off.p,
                                       (const double *)t_wt.p, &b->ds));
    {
        Guard g(ctx);
        PML_TRY(dev_alloc(&b->ds->row_loss, (size_t)n, &b->ds->bytes));
    }
    cleanup.b = nullptr;
    *out = b;
    return PML_OK;
}

static void batch_clear_normalization(pml_batch *b);

int pml_batch_destroy(pml_batch *b)
{
    if (!b) return PML_OK;
    if (b->ds) {
        cudaSetDevice(b->ds->ctx->device);
        cudaStreamSynchronize(b->ds->ctx->stream);
    }
    if (b->ent_row_ptr) cudaFree(b->ent_row_ptr);
    if (b->ent_values) cudaFree(b->ent_values);
    batch_clear_normalization(b);
    if (b->coef_ptr_dev) cudaFree(b->coef_ptr_dev);
    if (b->ent_of_coef) cudaFree(b->ent_of_coef);
    if (b->active_dev) cudaFree(b->active_dev);
    if (b->ds) pml_dataset_destroy(b->ds);
    delete b;
    return PML_OK;
}

int pml_batch_dataset(pml_batch *b, pml_dataset **out)
{
    if (!b || !out) return fail(PML_EINVAL, "null argument");
    *out = b->ds;
    return PML_OK;
}

int pml_batch_coef_ptr(pml_batch *b, int64_t *out)
{
    if (!b || !out) return fail(PML_EINVAL, "null argument");
    std::memcpy(out, b->coef_ptr.data(), sizeof(int64_t) * b->coef_ptr.size());
    return PML_OK;
}

// ---- per-entity NormalizationContext of a batch ---------------------------------------------------------------
static bool batch_normalized(const pml_batch *b, int loss)
{
    return loss != PML_LOSS_SMOOTHED_HINGE && (b->factors != nullptr || b->shifts != nullptr);
}

// device copy of the coefficient offsets (and, on request, the entity of every coefficient), made on first use
static int batch_index(pml_batch *b, bool need_ent_of)
{
    cudaStream_t s = b->ds->ctx->stream;
    if (b->coef_ptr_dev == nullptr) {
        PML_CUDA(cudaMalloc(&b->coef_ptr_dev, sizeof(int64_t) * (size_t)(b->n_entities + 1)));
        PML_CUDA(cudaMemcpyAsync(b->coef_ptr_dev, b->coef_ptr.data(), sizeof(int64_t) * (size_t)(b->n_entities + 1),
                                 cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaStreamSynchronize(s));
    }
    if (need_ent_of && b->ent_of_coef == nullptr && b->ds->d > 0) {
        PML_CUDA(cudaMalloc(&b->ent_of_coef, sizeof(int32_t) * (size_t)b->ds->d));
        PML_TRY(count(b->ds->ctx, launch_entity_of_coef(s, b->n_entities, b->coef_ptr_dev, b->ent_of_coef)));
    }
    return PML_OK;
}

// What the block-diagonal data set is evaluated at: x itself, or x .* factors with the per-entity shift term folded
// into the intercept entry (see the kernels' comment), left in `scratch`.
static int batch_effective(pml_batch *b, const double *x_dev, double *scratch, const double **out)
{
    pml_ctx *ctx = b->ds->ctx;
    cudaStream_t s = ctx->stream;
    PML_TRY(run(ctx, PML_K_PREPARE, [&] { return launch_scale(s, b->ds->d, x_dev, b->factors, scratch); }));
    if (b->shifts) {
        PML_TRY(run(ctx, PML_K_PREPARE, [&] {
            return launch_segment_dot(s, b->n_entities, b->coef_ptr_dev, scratch, b->shifts, b->dots);
        }));
        PML_TRY(run(ctx, PML_K_PREPARE, [&] { return launch_absorb_shift(s, b->n_entities, b->icpt_pos, b->dots, scratch); }));
    }
    *out = scratch;
    return PML_OK;
}

static int batch_get_vector(pml_batch *b, const double *V, double *out_dev)
{
    pml_ctx *ctx = b->ds->ctx;
    return run(ctx, PML_K_FINALIZE, [&] {
        return launch_batch_finalize(ctx->stream, b->ds->d, b->ent_of_coef, V, b->factors, b->shifts, b->icpt_pos, out_dev);
    });
}

static void batch_clear_normalization(pml_batch *b)
{
    for (double **p : { &b->factors, &b->shifts, &b->ec, &b->ev, &b->dots })
        if (*p) { cudaFree(*p); *p = nullptr; }
    if (b->icpt_pos) { cudaFree(b->icpt_pos); b->icpt_pos = nullptr; }
}

int pml_batch_set_normalization(pml_batch *b, const double *factors, const double *shifts, const int32_t *intercept_local)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    const int64_t E = b->n_entities, D = ds->d;
    // the `require`s of NormalizationContext / ValueAndGradientAggregator.init, entity by entity
    std::vector<int64_t> pos((size_t)E, -1);
    for (int64_t e = 0; e < E; ++e) {
        const int64_t lo = b->coef_ptr[(size_t)e], dim = b->coef_ptr[(size_t)e + 1] - lo;
        const int32_t ic = intercept_local ? intercept_local[e] : -1;
        if (ic < -1 || ic >= dim) return fail(PML_EINVAL, "intercept id outside [0, ent_dim) for entity " + std::to_string(e));
        if (shifts && dim > 0 && ic < 0) return fail(PML_EINVAL, "Shift without intercept is illegal.");
        if (ic < 0) continue;
        pos[(size_t)e] = lo + ic;
        if (factors && factors[lo + ic] != 1.0)
            return fail(PML_EINVAL, "The intercept should not be transformed. Intercept scaling factor: " + std::to_string(factors[lo + ic]));
        if (shifts && shifts[lo + ic] != 0.0)
            return fail(PML_EINVAL, "The intercept should not be transformed. Intercept shift: " + std::to_string(shifts[lo + ic]));
    }
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    PML_CUDA(cudaStreamSynchronize(s));
    ds->cache_valid = false;
    batch_clear_normalization(b);
    if ((!factors && !shifts) || D == 0 || E == 0) return PML_OK;
    PML_TRY(batch_index(b, true));
    const size_t db = sizeof(double) * (size_t)D;
    if (factors) {
        PML_CUDA(cudaMalloc(&b->factors, db));
        PML_CUDA(cudaMemcpyAsync(b->factors, factors, db, cudaMemcpyHostToDevice, s));
    }
    if (shifts) {
        PML_CUDA(cudaMalloc(&b->shifts, db));
        PML_CUDA(cudaMemcpyAsync(b->shifts, shifts, db, cudaMemcpyHostToDevice, s));
    }
    PML_CUDA(cudaMalloc(&b->icpt_pos, sizeof(int64_t) * (size_t)E));
    PML_CUDA(cudaMemcpyAsync(b->icpt_pos, pos.data(), sizeof(int64_t) * (size_t)E, cudaMemcpyHostToDevice, s));
    PML_CUDA(cudaMalloc(&b->ec, db));
    PML_CUDA(cudaMalloc(&b->ev, db));
    PML_CUDA(cudaMalloc(&b->dots, sizeof(double) * (size_t)E));
    if (shifts) {
        // the shift term is folded into the intercept coefficient: every row must hold (intercept, 1.0)
        unsigned long long *bad_dev = nullptr, bad = 0;
        PML_CUDA(cudaMalloc(&bad_dev, sizeof(bad)));
        PML_CUDA(cudaMemsetAsync(bad_dev, 0, sizeof(bad), s));
        const int rc = count(ds->ctx, launch_check_intercepts(s, ds->value_bytes, E, b->ent_row_ptr, ds->row_ptr, ds->col, ds->val,
                                                              ds->t.rank, b->icpt_pos, bad_dev));
        cudaMemcpyAsync(&bad, bad_dev, sizeof(bad), cudaMemcpyDeviceToHost, s);
        cudaStreamSynchronize(s);
        cudaFree(bad_dev);
        if (rc != PML_OK || bad != 0) {
            batch_clear_normalization(b);
            if (rc != PML_OK) return rc;
            return fail(PML_EINVAL, "shifts need the intercept feature with value 1 in every row of its entity: " +
                                        std::to_string(bad) + " rows lack it");
        }
    }
    PML_CUDA(cudaStreamSynchronize(s));
    return PML_OK;
}

int pml_batch_value_and_gradient_dev(pml_batch *b, int loss, const double *coefs_dev, double *out_values_dev,
                                     double *out_grads_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if ((!coefs_dev && ds->d > 0) || (!out_values_dev && b->n_entities > 0)) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    double *raw = ds->raw;
    const bool normalized = batch_normalized(b, loss);
    if (normalized) PML_TRY(batch_effective(b, coefs_dev, b->ec, &coefs_dev));
    ds->want_row_loss = true;
    const int rc = vg_raw(ds, loss, coefs_dev, raw, out_grads_dev != nullptr);
    ds->want_row_loss = false;
    PML_TRY(rc);
    PML_TRY(run(ds->ctx, PML_K_ENTITY, [&] {
        return launch_entity_values(s, b->n_entities, b->ent_row_ptr, ds->row_loss, out_values_dev);
    }));
    if (out_grads_dev && ds->d > 0) {
        if (normalized) PML_TRY(batch_get_vector(b, ds->raw, out_grads_dev));
        else PML_CUDA(cudaMemcpyAsync(out_grads_dev, ds->raw, sizeof(double) * (size_t)ds->d, cudaMemcpyDeviceToDevice, s));
    }
    if (b->active_dev)
        PML_TRY(run(ds->ctx, PML_K_ENTITY, [&] {
            return launch_mask_entities(s, b->n_entities, b->coef_ptr_dev, b->active_dev, out_values_dev, out_grads_dev);
        }));
    return PML_OK;
}

int pml_batch_value_and_gradient(pml_batch *b, int loss, const double *coefs, double *out_values, double *out_grads)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if ((!coefs && ds->d > 0) || (!out_values && b->n_entities > 0)) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const size_t db = sizeof(double) * (size_t)ds->d;
    if (ds->d > 0) PML_CUDA(cudaMemcpyAsync(ds->io_a, coefs, db, cudaMemcpyHostToDevice, s));
    const bool normalized = batch_normalized(b, loss);
    const double *coefs_dev = ds->io_a;
    if (normalized) PML_TRY(batch_effective(b, coefs_dev, b->ec, &coefs_dev));
    ds->want_row_loss = true;
    const int rc = vg_raw(ds, loss, coefs_dev, ds->raw);
    ds->want_row_loss = false;
    PML_TRY(rc);
    PML_TRY(run(ds->ctx, PML_K_ENTITY, [&] {
        return launch_entity_values(s, b->n_entities, b->ent_row_ptr, ds->row_loss, b->ent_values);
    }));
    double *grads_dev = ds->raw;
    if (normalized && out_grads && ds->d > 0) {
        PML_TRY(batch_get_vector(b, ds->raw, ds->io_b));
        grads_dev = ds->io_b;
    }
    if (b->active_dev)
        PML_TRY(run(ds->ctx, PML_K_ENTITY, [&] {
            return launch_mask_entities(s, b->n_entities, b->coef_ptr_dev, b->active_dev, b->ent_values,
                                        out_grads ? grads_dev : nullptr);
        }));
    if (out_grads && ds->d > 0) PML_CUDA(cudaMemcpyAsync(out_grads, grads_dev, db, cudaMemcpyDeviceToHost, s));
    if (b->n_entities > 0)
        PML_CUDA(cudaMemcpyAsync(out_values, b->ent_values, sizeof(double) * (size_t)b->n_entities, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    return PML_OK;
}

int pml_batch_hessian_vector_dev(pml_batch *b, int loss, const double *coefs_dev, uint64_t coef_tag, const double *vs_dev,
                                 double *out_hvs_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    // entities are the diagonal blocks of one data set: the product of the block-diagonal Hessian with the
    // concatenated vector is the concatenation of the per-entity products
    if (batch_normalized(b, loss)) {
        if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
        if (!vs_dev || !out_hvs_dev) return fail(PML_EINVAL, "null buffer");
        Guard g(ds->ctx);
        PML_CUDA(cudaSetDevice(ds->ctx->device));
        const bool cached = coef_tag != 0 && ds->cache_valid && ds->cached_tag == coef_tag && ds->cached_loss == loss;
        if (!cached) {
            if (!coefs_dev) return fail(PML_EINVAL, "coefficients are required when no cached pass matches the tag");
            PML_TRY(batch_effective(b, coefs_dev, b->ec, &coefs_dev));
        }
        PML_TRY(batch_effective(b, vs_dev, b->ev, &vs_dev));
        PML_TRY(hv_raw(ds, loss, coefs_dev, coef_tag, vs_dev, ds->raw));
        PML_TRY(batch_get_vector(b, ds->raw, out_hvs_dev));
    } else {
        PML_TRY(pml_hessian_vector_dev(ds, loss, coefs_dev, coef_tag, vs_dev, out_hvs_dev));
    }
    if (b->active_dev) {
        Guard g(ds->ctx);
        PML_TRY(run(ds->ctx, PML_K_ENTITY, [&] {
            return launch_mask_entities(ds->ctx->stream, b->n_entities, b->coef_ptr_dev, b->active_dev, nullptr, out_hvs_dev);
        }));
    }
    return PML_OK;
}

int pml_batch_hessian_vector(pml_batch *b, int loss, const double *coefs, uint64_t coef_tag, const double *vs, double *out_hvs)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    if (batch_normalized(b, loss)) {
        if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
        if ((!coefs || !vs || !out_hvs) && ds->d > 0) return fail(PML_EINVAL, "null buffer");
        const size_t db = sizeof(double) * (size_t)ds->d;
        {
            Guard g(ds->ctx);
            PML_CUDA(cudaSetDevice(ds->ctx->device));
            PML_CUDA(cudaMemcpyAsync(ds->io_a, coefs, db, cudaMemcpyHostToDevice, ds->ctx->stream));
            PML_CUDA(cudaMemcpyAsync(ds->io_b, vs, db, cudaMemcpyHostToDevice, ds->ctx->stream));
        }
        // the device form does the rest (it takes the lock itself, and applies the active mask); the product lands
        // in io_a, which the coefficients no longer need by then
        PML_TRY(pml_batch_hessian_vector_dev(b, loss, ds->io_a, coef_tag, ds->io_b, ds->io_a));
        Guard g(ds->ctx);
        PML_CUDA(cudaMemcpyAsync(out_hvs, ds->io_a, db, cudaMemcpyDeviceToHost, ds->ctx->stream));
        PML_CUDA(cudaStreamSynchronize(ds->ctx->stream));
        return PML_OK;
    } else {
        PML_TRY(pml_hessian_vector(ds, loss, coefs, coef_tag, vs, out_hvs));
    }
    if (b->active_dev && ds->d > 0) {  // rare on the host path: mask on the host copy
        std::vector<uint8_t> act((size_t)b->n_entities);
        PML_CUDA(cudaMemcpy(act.data(), b->active_dev, act.size(), cudaMemcpyDeviceToHost));
        for (int64_t e = 0; e < b->n_entities; ++e)
            if (!act[(size_t)e])
                for (int64_t j = b->coef_ptr[(size_t)e]; j < b->coef_ptr[(size_t)e + 1]; ++j) out_hvs[j] = 0.0;
    }
    return PML_OK;
}

int pml_batch_set_offsets(pml_batch *b, const double *offsets)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    return pml_dataset_set_offsets(b->ds, offsets);
}

int pml_batch_segment_dot_dev(pml_batch *b, const double *x_dev, const double *y_dev, double *out_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    if (b->n_entities > 0 && (!x_dev || !y_dev || !out_dev)) return fail(PML_EINVAL, "null buffer");
    Guard g(b->ds->ctx);
    PML_CUDA(cudaSetDevice(b->ds->ctx->device));
    PML_TRY(batch_index(b, false));
    return run(b->ds->ctx, PML_K_ENTITY, [&] {
        return launch_segment_dot(b->ds->ctx->stream, b->n_entities, b->coef_ptr_dev, x_dev, y_dev, out_dev);
    });
}

int pml_batch_cg_step_dev(pml_batch *b, const uint8_t *run_dev, double l2, double *direction_dev, const double *hd_dev,
                          double *step_dev, double *residual_dev, double *rtr_dev, const double *delta_dev,
                          uint8_t *out_over_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    if (b->n_entities > 0 && (!run_dev || !direction_dev || !hd_dev || !step_dev || !residual_dev || !rtr_dev || !delta_dev ||
                              !out_over_dev))
        return fail(PML_EINVAL, "null buffer");
    Guard g(b->ds->ctx);
    PML_CUDA(cudaSetDevice(b->ds->ctx->device));
    PML_TRY(batch_index(b, false));
    return run(b->ds->ctx, PML_K_ENTITY, [&] {
        return launch_batch_cg_step(b->ds->ctx->stream, b->n_entities, b->coef_ptr_dev, run_dev, l2, direction_dev, hd_dev,
                                    step_dev, residual_dev, rtr_dev, delta_dev, out_over_dev);
    });
}

int pml_batch_lbfgs_direction_dev(pml_batch *b, const uint8_t *go_dev, int m, const double *S_dev, const double *Y_dev,
                                  const double *rho_dev, const int32_t *head_dev, const int32_t *len_dev,
                                  const double *grad_dev, double *out_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    if (m < 1 || m > 16) return fail(PML_EINVAL, "1 to 16 correction pairs");
    if (b->n_entities > 0 && (!S_dev || !Y_dev || !rho_dev || !head_dev || !len_dev || !grad_dev || !out_dev))
        return fail(PML_EINVAL, "null buffer");
    Guard g(b->ds->ctx);
    PML_CUDA(cudaSetDevice(b->ds->ctx->device));
    PML_TRY(batch_index(b, false));
    return run(b->ds->ctx, PML_K_ENTITY, [&] {
        return launch_batch_lbfgs_direction(b->ds->ctx->stream, b->n_entities, b->ds->d, b->coef_ptr_dev, go_dev, m, S_dev, Y_dev,
                                            rho_dev, head_dev, len_dev, grad_dev, out_dev);
    });
}

int pml_batch_lbfgs_push_dev(pml_batch *b, const uint8_t *mask_dev, int m, const double *rho_new_dev, const double *s_dev,
                             const double *y_dev, double *S_dev, double *Y_dev, double *rho_dev, int32_t *head_dev,
                             int32_t *len_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    if (m < 1 || m > 16) return fail(PML_EINVAL, "1 to 16 correction pairs");
    if (b->n_entities > 0 && (!mask_dev || !rho_new_dev || !s_dev || !y_dev || !S_dev || !Y_dev || !rho_dev || !head_dev || !len_dev))
        return fail(PML_EINVAL, "null buffer");
    Guard g(b->ds->ctx);
    PML_CUDA(cudaSetDevice(b->ds->ctx->device));
    PML_TRY(batch_index(b, true));
    return run(b->ds->ctx, PML_K_ENTITY, [&] {
        return launch_batch_lbfgs_push(b->ds->ctx->stream, b->n_entities, b->ds->d, b->ent_of_coef, mask_dev, m, rho_new_dev,
                                       s_dev, y_dev, S_dev, Y_dev, rho_dev, head_dev, len_dev);
    });
}

int pml_batch_segment_axpby_dev(pml_batch *b, const double *a_dev, const double *x_dev, const double *c_dev,
                                const double *y_dev, const uint8_t *active_dev, double *out_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    if (b->ds->d > 0 && (!x_dev || !out_dev)) return fail(PML_EINVAL, "null buffer");
    Guard g(b->ds->ctx);
    PML_CUDA(cudaSetDevice(b->ds->ctx->device));
    PML_TRY(batch_index(b, true));
    return run(b->ds->ctx, PML_K_ENTITY, [&] {
        return launch_segment_axpby(b->ds->ctx->stream, b->ds->d, b->ent_of_coef, a_dev, x_dev, c_dev, y_dev, active_dev,
                                    out_dev);
    });
}

// After the mask changed: which chunks of the CSR / CSC streams and which 32-row blocks still hold an active entity
// (the evaluation kernels skip the others).  PML_NO_SKIP_INACTIVE=1 keeps streaming everything (A/B measurements).
static int batch_refresh_flags(pml_batch *b)
{
    pml_dataset *ds = b->ds;
    cudaStream_t s = ds->ctx->stream;
    static const bool off = std::getenv("PML_NO_SKIP_INACTIVE") != nullptr;
    const bool had = ds->row_chunk_active != nullptr;
    if (b->active_dev == nullptr || off || ds->n == 0 || ds->nnz == 0 || ds->t.hot.n_chunks > 0 || ds->t.n_super > 1) {
        if (had) {
            PML_CUDA(cudaStreamSynchronize(s));
            cudaFree(ds->row_chunk_active);
            cudaFree(ds->col_chunk_active);
            cudaFree(ds->row_block_active);
            ds->row_chunk_active = ds->col_chunk_active = ds->row_block_active = nullptr;
            ++ds->version;  // the captured sequences hold the old pointers
        }
        return PML_OK;
    }
    PML_TRY(batch_index(b, true));
    if (!had) {
        PML_CUDA(cudaMalloc(&ds->row_chunk_active, (size_t)ds->rs.n_chunks + 1));
        PML_CUDA(cudaMalloc(&ds->col_chunk_active, (size_t)ds->t.n_chunks + 1));
        PML_CUDA(cudaMalloc(&ds->row_block_active, (size_t)((ds->n + 31) / 32) + 1));
        ++ds->version;
    }
    return count(ds->ctx, launch_active_flags(s, b->n_entities, b->active_dev, b->ent_row_ptr, ds->n, ds->nnz, ds->row_ptr,
                                              ds->rs.n_chunks, ds->row_chunk_active, ds->row_block_active, ds->t.m,
                                              ds->t.nzcol_ptr, ds->t.nzcol_id, b->ent_of_coef, ds->t.cold_nnz, ds->t.n_chunks,
                                              ds->col_chunk_active));
}

int pml_batch_set_active(pml_batch *b, const uint8_t *active)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    if (active == nullptr) {
        if (b->active_dev) {
            PML_CUDA(cudaStreamSynchronize(s));
            cudaFree(b->active_dev);
            b->active_dev = nullptr;
        }
        return batch_refresh_flags(b);
    }
    if (b->n_entities == 0) return PML_OK;
    PML_TRY(batch_index(b, false));
    if (b->active_dev == nullptr) PML_CUDA(cudaMalloc(&b->active_dev, (size_t)b->n_entities));
    PML_CUDA(cudaMemcpyAsync(b->active_dev, active, (size_t)b->n_entities, cudaMemcpyHostToDevice, s));
    PML_TRY(batch_refresh_flags(b));
    PML_CUDA(cudaStreamSynchronize(s));
    return PML_OK;
}

int pml_batch_set_active_dev(pml_batch *b, const uint8_t *active_dev)
{
    if (!b || !b->ds) return fail(PML_EINVAL, "null batch handle");
    pml_dataset *ds = b->ds;
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    if (active_dev == nullptr) {
        if (b->active_dev) {
            PML_CUDA(cudaStreamSynchronize(s));
            cudaFree(b->active_dev);
            b->active_dev = nullptr;
        }
        return batch_refresh_flags(b);
    }
    if (b->n_entities == 0) return PML_OK;
    PML_TRY(batch_index(b, false));
    if (b->active_dev == nullptr) PML_CUDA(cudaMalloc(&b->active_dev, (size_t)b->n_entities));
    PML_CUDA(cudaMemcpyAsync(b->active_dev, active_dev, (size_t)b->n_entities, cudaMemcpyDeviceToDevice, s));
    return batch_refresh_flags(b);
}

// ---- the merge step across GPUs over peer memory (peer.cu) ------------------------------------
static int check_peer(const pml_dataset *ds, const pml_peer *peer)
{
    if (peer == nullptr) return fail(PML_EINVAL, "null peer handle");
    if (peer->ctx != ds->ctx) return fail(PML_EINVAL, "the peer handle belongs to another context");
    if (peer->d != ds->d) return fail(PML_EINVAL, "the peer handle was created for another dimension");
    if (!peer->connected) return fail(PML_ESTATE, "pml_peer_connect has not been called");
    if (*peer->h_status != 0)
        return fail(PML_ECUDA, "an earlier merge over peer memory timed out waiting for another rank");
    return PML_OK;
}

int pml_value_and_gradient_allreduce_dev(pml_dataset *ds, pml_peer *peer, int loss, const double *coef_dev,
                                         double *out_value_dev, double *out_grad_dev)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const bool hinge = loss == PML_LOSS_SMOOTHED_HINGE;
    coef_dev = stable_input(ds, ds->vg_graph, ds->vg_graph.key.coef, coef_dev, ds->io_a);
    PML_TRY(vg_raw(ds, loss, coef_dev, peer_raw(peer), true));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] {
        return launch_peer_merge(s, ds->ctx->sm_count, peer, hinge ? nullptr : ds->factors, hinge ? nullptr : ds->shifts);
    }));
    const double *res = peer_result(peer);
    if (out_grad_dev && ds->d > 0)
        PML_CUDA(cudaMemcpyAsync(out_grad_dev, res, sizeof(double) * (size_t)ds->d, cudaMemcpyDeviceToDevice, s));
    if (out_value_dev) PML_CUDA(cudaMemcpyAsync(out_value_dev, res + ds->d, sizeof(double), cudaMemcpyDeviceToDevice, s));
    return PML_OK;
}

int pml_peer_merge_dev(pml_dataset *ds, pml_peer *peer, int loss)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    const bool hinge = loss == PML_LOSS_SMOOTHED_HINGE;
    return run(ds->ctx, PML_K_PEER, [&] {
        return launch_peer_merge(ds->ctx->stream, ds->ctx->sm_count, peer, hinge ? nullptr : ds->factors,
                                 hinge ? nullptr : ds->shifts);
    });
}

int pml_hessian_vector_allreduce_dev(pml_dataset *ds, pml_peer *peer, int loss, const double *coef_dev, uint64_t coef_tag,
                                     const double *v_dev, double *out_hv_dev)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!v_dev) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const bool cached = coef_tag != 0 && ds->cache_valid && ds->cached_tag == coef_tag && ds->cached_loss == loss;
    const pml_graph &hg = cached ? ds->hv_cached_graph : ds->hv_cold_graph;
    if (!cached) coef_dev = stable_input(ds, hg, hg.key.coef, coef_dev, ds->io_a);
    v_dev = stable_input(ds, hg, hg.key.v, v_dev, ds->io_b);
    PML_TRY(hv_raw(ds, loss, coef_dev, coef_tag, v_dev, peer_raw(peer)));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_merge(s, ds->ctx->sm_count, peer, ds->factors, ds->shifts); }));
    if (out_hv_dev && ds->d > 0)
        PML_CUDA(cudaMemcpyAsync(out_hv_dev, peer_result(peer), sizeof(double) * (size_t)ds->d, cudaMemcpyDeviceToDevice, s));
    return PML_OK;
}


// ---- host-buffer forms of the two exchanges (the shape the JNI natives bind: DistributedObjectiveFunction.scala:57-58,73)
// Every rank is handed the same coefficient vector in host memory and uploads only ITS 1/N slice over its own PCIe
// link; the slices are exchanged over NVLink (peer_gather_kernel), the evaluation and the merge run as in the
// *_allreduce_dev forms, and every rank downloads its slice of the result (gather_output = 0: out_grad[lo, hi) is
// written, [lo, hi) from pml_peer_slice -- ranks that share one host buffer assemble the full vector in it) or all of
// it (gather_output = 1).  The objective value is returned on every rank.
static int upload_slices(pml_dataset *ds, pml_peer *peer, const double *coef, const double *v, int *n_vec)
{
    cudaStream_t s = ds->ctx->stream;
    const bool whole = peer->world == 1 || peer->mc_va != nullptr;  // no pairwise mappings: every rank uploads everything
    int64_t lo = 0, hi = ds->d;
    if (!whole) peer_slice(peer, &lo, &hi);
    if (coef && hi > lo)
        PML_CUDA(cudaMemcpyAsync(peer_vec(peer, 0) + lo, coef + lo, sizeof(double) * (size_t)(hi - lo), cudaMemcpyHostToDevice, s));
    if (v && hi > lo)
        PML_CUDA(cudaMemcpyAsync(peer_vec(peer, 1) + lo, v + lo, sizeof(double) * (size_t)(hi - lo), cudaMemcpyHostToDevice, s));
    *n_vec = coef && v ? 2 : (v ? -2 : 1);
    return PML_OK;
}

static int download_result(pml_dataset *ds, pml_peer *peer, double *out_value, double *out_vec, int gather_output)
{
    cudaStream_t s = ds->ctx->stream;
    const double *res = peer_result(peer);
    int64_t lo = 0, hi = ds->d;
    if (!gather_output) peer_slice(peer, &lo, &hi);
    if (out_vec && hi > lo)
        PML_CUDA(cudaMemcpyAsync(out_vec + lo, res + lo, sizeof(double) * (size_t)(hi - lo), cudaMemcpyDeviceToHost, s));
    if (out_value) PML_CUDA(cudaMemcpyAsync(ds->ctx->h_scalar, res + ds->d, sizeof(double), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    if (out_value) *out_value = ds->ctx->h_scalar[0];
    if (*peer->h_status != 0)
        return fail(PML_ECUDA, "the exchange over peer memory timed out waiting for another rank");
    return PML_OK;
}

int pml_value_and_gradient_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, double *out_value,
                                     double *out_grad, int gather_output)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const bool hinge = loss == PML_LOSS_SMOOTHED_HINGE;
    int n_vec = 1;
    PML_TRY(upload_slices(ds, peer, coef, nullptr, &n_vec));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_gather(s, ds->ctx->sm_count, peer, n_vec); }));
    PML_TRY(vg_raw(ds, loss, peer_vec(peer, 0), peer_raw(peer), true));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] {
        return launch_peer_merge(s, ds->ctx->sm_count, peer, hinge ? nullptr : ds->factors, hinge ? nullptr : ds->shifts);
    }));
    return download_result(ds, peer, out_value, out_grad, gather_output);
}

// HessianDiagonalAggregator over all ranks' rows (DistributedGLMLossFunction.hessianDiagonal, :134-137): every rank's
// diagonal sums, merged like the other two (no normalization: the reference's diagonal ignores it).  Host vectors.
int pml_hessian_diagonal_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, double *out_diag,
                                   int gather_output)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!coef) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    int n_vec = 1;
    PML_TRY(upload_slices(ds, peer, coef, nullptr, &n_vec));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_gather(s, ds->ctx->sm_count, peer, n_vec); }));
    PML_TRY(diag_dev(ds, loss, peer_vec(peer, 0), peer_raw(peer)));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_merge(s, ds->ctx->sm_count, peer, nullptr, nullptr); }));
    return download_result(ds, peer, nullptr, out_diag, gather_output);
}

int pml_hessian_vector_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, uint64_t coef_tag,
                                 const double *v, double *out_hv, int gather_output)
{
    PML_TRY(check_ds(ds));
    PML_TRY(check_peer(ds, peer));
    if (!valid_loss(loss)) return fail(PML_EINVAL, "unknown loss");
    if (!v) return fail(PML_EINVAL, "null buffer");
    Guard g(ds->ctx);
    PML_CUDA(cudaSetDevice(ds->ctx->device));
    cudaStream_t s = ds->ctx->stream;
    const bool cached = coef_tag != 0 && ds->cache_valid && ds->cached_tag == coef_tag && ds->cached_loss == loss;
    if (!cached && !coef) return fail(PML_EINVAL, "coefficients are required when no cached pass matches the tag");
    int n_vec = 2;
    PML_TRY(upload_slices(ds, peer, cached ? nullptr : coef, v, &n_vec));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_gather(s, ds->ctx->sm_count, peer, n_vec); }));
    PML_TRY(hv_raw(ds, loss, peer_vec(peer, 0), coef_tag, peer_vec(peer, 1), peer_raw(peer)));
    PML_TRY(run(ds->ctx, PML_K_PEER, [&] { return launch_peer_merge(s, ds->ctx->sm_count, peer, ds->factors, ds->shifts); }));
    return download_result(ds, peer, nullptr, out_hv, gather_output);
}

}  // extern "C"
