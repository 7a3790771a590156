// kernels.cu -- the per-evaluation kernels of the GLM objective path (sm_100a).
//
//   prepare_vectors_kernel  a1/a6  ValueAndGradientAggregator.init / HessianVectorAggregator.init
//   row_pass_kernel         a2/a7/a9  the per-row half of `add`: sparse dot, pointwise loss, scalars
//   col_pass_kernel         a2/a7/a9  the axpy half of `add`, done column-wise over the CSC copy
//   fixup_kernel            a3  the part of `merge` that joins columns split across chunks
//   finalize_kernel         a4  getVector
//
// (aN = row of SURVEY.md 8a; reference file:line in DESIGN.md and at each kernel.)
// All accumulation is fp64.  No atomics anywhere: every output word has exactly one writer, and
// every reduction has a fixed shape, so results are bit-reproducible for a given packing.
//
// Columns carry "internal ids" assigned at pack time by descending entry count (pack.cu), so the
// coefficients of the hottest columns sit at the front of the effective-coefficient vector and
// can be staged in shared memory: a scattered 8-byte gather costs one L1 tag lookup per lane
// (one per cycle per SM -- the measured limiter of the first version of these kernels), while
// shared memory serves 32 lanes in a few cycles.
#include <cuda_runtime.h>
#include <stdint.h>

#include "internal.h"
#include "pointwise.cuh"

namespace pml {

// ------------------------------------------------------------------------------------------
// prepare_vectors_kernel -- ValueAndGradientAggregator.init (lib/function/glm/
// ValueAndGradientAggregator.scala:90-122) and HessianVectorAggregator.init
// (HessianVectorAggregator.scala:60-85): ec = coef .* factors, ev = v .* factors, and the two
// dots with `shifts` as per-CTA partials (the row kernel adds them up in its prologue, so the
// scalar needs no launch of its own).  The results are written at the columns' internal ids
// (`rank`), and when v != null interleaved as (ec_j, ev_j) so that the row kernel fetches both
// with one 16-byte access.
// ------------------------------------------------------------------------------------------
constexpr int PREP_THREADS = 256;

__global__ void __launch_bounds__(PREP_THREADS)
prepare_vectors_kernel(int d, const double *__restrict__ coef, const double *__restrict__ v,
                       const double *__restrict__ factors, const double *__restrict__ shifts,
                       const int32_t *__restrict__ rank, double *__restrict__ out_vec,
                       double *__restrict__ shift_partials)
{
    __shared__ double red[2 * (PREP_THREADS / 32)];
    double s0 = 0.0, s1 = 0.0;
    for (int j = blockIdx.x * PREP_THREADS + threadIdx.x; j < d; j += gridDim.x * PREP_THREADS) {
        const double f = factors ? factors[j] : 1.0;
        const double ec = factors ? __dmul_rn(coef[j], f) : coef[j];
        const double sh = shifts ? shifts[j] : 0.0;
        const int32_t r = rank[j];
        s0 += __dmul_rn(ec, sh);
        if (v) {
            const double ev = factors ? __dmul_rn(v[j], f) : v[j];
            s1 += __dmul_rn(ev, sh);
            reinterpret_cast<double2 *>(out_vec)[r] = make_double2(ec, ev);
        } else {
            out_vec[r] = ec;
        }
    }
    block_sum2<PREP_THREADS / 32>(s0, s1, red);
    if (threadIdx.x == 0) {
        shift_partials[2 * blockIdx.x] = s0;
        shift_partials[2 * blockIdx.x + 1] = s1;
    }
}

int launch_prepare_vectors(cudaStream_t s, int d, const double *coef, const double *v, const double *factors,
                           const double *shifts, const int32_t *rank, double *out_vec, double *shift_partials,
                           int *n_partials)
{
    int grid = (d + PREP_THREADS * 4 - 1) / (PREP_THREADS * 4);
    if (grid < 1) grid = 1;
    if (grid > 1024) grid = 1024;
    prepare_vectors_kernel<<<grid, PREP_THREADS, 0, s>>>(d, coef, v, factors, shifts, rank, out_vec, shift_partials);
    *n_partials = grid;
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// row_pass_kernel -- the per-row half of ValueAndGradientAggregator.add
// (ValueAndGradientAggregator.scala:132-153), HessianVectorAggregator.add
// (HessianVectorAggregator.scala:96-121) and HessianDiagonalAggregator.add
// (HessianDiagonalAggregator.scala:51-61): margin z_i = x_i.ec + o_i + marginShift
// (LabeledPoint.computeMargin, lib/data/LabeledPoint.scala:62), the pointwise loss, the scalar
// sums, and the per-row multiplier r_i that the column pass turns into the axpy.
//
// Persistent CTAs (one per SM when the hot table is large).  The first `n_hot` entries of the
// effective-coefficient vector are copied to shared memory once per CTA; a lane whose column id
// is below n_hot reads shared memory, the rest gather through L1.
//
// A warp owns 32 consecutive rows at a time.
//   LPR in {2,4,8,16,32}: "lanes per row" -- 32/LPR rows are reduced concurrently by sub-warps
//     of LPR lanes, LPR passes cover the 32 rows (LPR = 32 is warp-per-row).
//   LPR == 0: segment-per-warp -- the warp streams the contiguous entry range of its 32 rows
//     with fully coalesced loads, parks the products in shared memory, and each lane then adds
//     up its own row in storage order (the reference's order, so margins match bit for bit).
// Either way lane L ends up holding row L's dot product, so labels/offsets/weights are read and
// r is written coalesced and the fp64 exp/log1p work runs with all 32 lanes active.
// ------------------------------------------------------------------------------------------
template <bool TWO>
__device__ __forceinline__ void lookup(const double *__restrict__ hot, int n_hot, const double *__restrict__ vec,
                                       int32_t c, double &w0, double &w1)
{
    if (TWO) {
        double2 g;
        if (c < n_hot) g = reinterpret_cast<const double2 *>(hot)[c];
        else g = __ldg(reinterpret_cast<const double2 *>(vec) + c);
        w0 = g.x;
        w1 = g.y;
    } else {
        w0 = (c < n_hot) ? hot[c] : __ldg(vec + c);
        w1 = 0.0;
    }
}

template <typename VT, int LPR, bool TWO>
__global__ void __launch_bounds__(R_THREADS, 1) row_pass_kernel(const RowPassParams p)
{
    constexpr int CAPE = TWO ? R_STREAM_CAP / 2 : R_STREAM_CAP;
    extern __shared__ __align__(16) double dyn_smem[];
    __shared__ double red[2 * R_WARPS];
    // dynamic shared memory: [stage (LPR == 0 only)] [hot table]
    double *stage = dyn_smem;
    double *hot = dyn_smem + ((LPR == 0) ? R_WARPS * R_STREAM_CAP : 0);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const VT *__restrict__ val = static_cast<const VT *>(p.val);
    const int32_t *__restrict__ col = p.col;
    const double *__restrict__ vec = p.vec;
    const int n_hot = p.n_hot;

    {
        const int words = TWO ? 2 * n_hot : n_hot;
        for (int i = threadIdx.x; i < words; i += R_THREADS) hot[i] = vec[i];
    }

    // marginShift = -(ec . shifts); featureVectorProductShift = ev . shifts
    double shift0 = 0.0, shift1 = 0.0;
    if (p.shift_partials != nullptr) {
        for (int i = threadIdx.x; i < p.n_shift_partials; i += R_THREADS) {
            shift0 += p.shift_partials[2 * i];
            shift1 += p.shift_partials[2 * i + 1];
        }
        block_sum2<R_WARPS>(shift0, shift1, red);
    }
    __syncthreads();  // hot table visible
    const double margin_shift = (p.mode == MODE_HV_CACHED) ? 0.0 : -shift0;
    const double fs = (p.mode == MODE_HV_CACHED) ? shift0 : shift1;

    double accF = 0.0, accP = 0.0;
    const int64_t n_blk = (p.n + 31) >> 5;
    for (int64_t blk = (int64_t)blockIdx.x * R_WARPS + warp; blk < n_blk; blk += (int64_t)gridDim.x * R_WARPS) {
        const int64_t row0 = blk << 5;
        const int64_t my_row = row0 + lane;
        const int64_t clamp_lo = my_row < p.n ? my_row : p.n;
        const int64_t clamp_hi = my_row + 1 < p.n ? my_row + 1 : p.n;
        const uint32_t rp_lo = __ldg(p.row_ptr + clamp_lo);
        const uint32_t rp_hi = __ldg(p.row_ptr + clamp_hi);
        double za = 0.0, zb = 0.0;

        if (LPR == 0) {
            double *buf = stage + warp * R_STREAM_CAP;
            const uint32_t base = __shfl_sync(0xffffffffu, rp_lo, 0);
            const uint32_t lim = __shfl_sync(0xffffffffu, rp_hi, 31);
            for (uint32_t pb = base; pb < lim; pb += CAPE) {
                const uint32_t pe = (lim - pb > (uint32_t)CAPE) ? pb + CAPE : lim;
#pragma unroll 4
                for (uint32_t k = pb + lane; k < pe; k += 32) {
                    const int32_t c = __ldcs(col + k);
                    const double x = widen(__ldcs(val + k));
                    double w0, w1;
                    lookup<TWO>(hot, n_hot, vec, c, w0, w1);
                    buf[k - pb] = __dmul_rn(x, w0);
                    if (TWO) buf[CAPE + (k - pb)] = __dmul_rn(x, w1);
                }
                __syncwarp();
                const uint32_t lo = rp_lo > pb ? rp_lo : pb;
                const uint32_t hi = rp_hi < pe ? rp_hi : pe;
                for (uint32_t k = lo; k < hi; ++k) {
                    za += buf[k - pb];
                    if (TWO) zb += buf[CAPE + (k - pb)];
                }
                __syncwarp();
            }
        } else {
            constexpr int L = (LPR == 0) ? 1 : LPR;  // keeps the dead branch compilable
            constexpr int RPP = 32 / L;              // rows per pass
            const int sub = lane / L, sl = lane % L;
#pragma unroll 1
            for (int pass = 0; pass < L; ++pass) {
                const int rloc = pass * RPP + sub;
                const uint32_t start = __shfl_sync(0xffffffffu, rp_lo, rloc);
                const uint32_t end = __shfl_sync(0xffffffffu, rp_hi, rloc);
                double a = 0.0, b = 0.0;
                uint32_t k = start + sl;
                for (; k + 3 * L < end; k += 4 * L) {
                    const int32_t c0 = __ldcs(col + k), c1 = __ldcs(col + k + L);
                    const int32_t c2 = __ldcs(col + k + 2 * L), c3 = __ldcs(col + k + 3 * L);
                    const double x0 = widen(__ldcs(val + k)), x1 = widen(__ldcs(val + k + L));
                    const double x2 = widen(__ldcs(val + k + 2 * L)), x3 = widen(__ldcs(val + k + 3 * L));
                    double w00, w01, w10, w11, w20, w21, w30, w31;
                    lookup<TWO>(hot, n_hot, vec, c0, w00, w01);
                    lookup<TWO>(hot, n_hot, vec, c1, w10, w11);
                    lookup<TWO>(hot, n_hot, vec, c2, w20, w21);
                    lookup<TWO>(hot, n_hot, vec, c3, w30, w31);
                    a += (x0 * w00 + x1 * w10) + (x2 * w20 + x3 * w30);
                    if (TWO) b += (x0 * w01 + x1 * w11) + (x2 * w21 + x3 * w31);
                }
                for (; k < end; k += L) {
                    const int32_t c = __ldcs(col + k);
                    const double x = widen(__ldcs(val + k));
                    double w0, w1;
                    lookup<TWO>(hot, n_hot, vec, c, w0, w1);
                    a = fma(x, w0, a);
                    if (TWO) b = fma(x, w1, b);
                }
#pragma unroll
                for (int o = L / 2; o > 0; o >>= 1) {
                    a += __shfl_xor_sync(0xffffffffu, a, o);
                    if (TWO) b += __shfl_xor_sync(0xffffffffu, b, o);
                }
                // lane `lane` wants row `lane`: produced in pass lane/RPP by sub-warp lane%RPP
                const double ga = __shfl_sync(0xffffffffu, a, (lane % RPP) * L);
                const double gb = TWO ? __shfl_sync(0xffffffffu, b, (lane % RPP) * L) : 0.0;
                if (lane / RPP == pass) {
                    za = ga;
                    zb = gb;
                }
            }
        }

        if (my_row < p.n) {
            const double o = p.offset ? __ldcs(p.offset + my_row) : 0.0;
            const double wt = p.weight ? __ldcs(p.weight + my_row) : 1.0;
            switch (p.mode) {
            case MODE_VG: {
                const double z = (za + o) + margin_shift;
                double l, dl;
                loss_and_dz(p.loss, z, __ldcs(p.label + my_row), l, dl);
                const double wdl = __dmul_rn(wt, dl);
                p.r[my_row] = wdl;
                accF += __dmul_rn(wt, l);
                accP += wdl;
                break;
            }
            case MODE_HV_COLD: {
                const double z = (za + o) + margin_shift;
                const double dq = __dmul_rn(wt, dzz_loss(p.loss, z));
                if (p.dcache) p.dcache[my_row] = dq;
                const double e = __dmul_rn(dq, zb - fs);
                p.r[my_row] = e;
                accP += e;
                break;
            }
            case MODE_HV_CACHED: {
                const double e = __dmul_rn(__ldcs(p.dcache + my_row), za - fs);
                p.r[my_row] = e;
                accP += e;
                break;
            }
            case MODE_DIAG: {
                const double z = za + o;
                p.r[my_row] = __dmul_rn(wt, dzz_loss(p.loss, z));
                break;
            }
            default:
                p.r[my_row] = za;
                break;
            }
        }
    }

    if (p.partials != nullptr) {
        block_sum2<R_WARPS>(accF, accP, red);
        if (threadIdx.x == 0) {
            p.partials[2 * blockIdx.x] = accF;
            p.partials[2 * blockIdx.x + 1] = accP;
        }
    }
}

template <typename VT, int LPR, bool TWO>
static int launch_row_one(cudaStream_t s, int grid, size_t smem, const RowPassParams &p)
{
    // per device and per instantiation; cheap enough to repeat on every launch
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(row_pass_kernel<VT, LPR, TWO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             R_MAX_DYN_SMEM) != cudaSuccess)
        return -1;
    row_pass_kernel<VT, LPR, TWO><<<grid, R_THREADS, smem, s>>>(p);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

template <typename VT, bool TWO>
static int launch_row_variant(cudaStream_t s, int grid, int lpr, size_t smem, const RowPassParams &p)
{
    switch (lpr) {
    case 0: return launch_row_one<VT, 0, TWO>(s, grid, smem, p);
    case 2: return launch_row_one<VT, 2, TWO>(s, grid, smem, p);
    case 4: return launch_row_one<VT, 4, TWO>(s, grid, smem, p);
    case 8: return launch_row_one<VT, 8, TWO>(s, grid, smem, p);
    case 16: return launch_row_one<VT, 16, TWO>(s, grid, smem, p);
    case 32: return launch_row_one<VT, 32, TWO>(s, grid, smem, p);
    default: return -1;
    }
}

int row_pass_hot_capacity(int lanes_per_row, bool two)
{
    const size_t stage = lanes_per_row == 0 ? sizeof(double) * R_WARPS * R_STREAM_CAP : 0;
    return (int)((R_MAX_DYN_SMEM - stage) / (two ? 16 : 8));
}

int launch_row_pass(cudaStream_t s, int sm_count, int value_bytes, int lanes_per_row, bool two,
                    RowPassParams p, int64_t n_nonempty_cols, int *grid_out)
{
    const int64_t n_blk = (p.n + 31) / 32;
    int64_t want = (n_blk + R_WARPS - 1) / R_WARPS;
    int grid = (int)(want < sm_count ? want : sm_count);
    if (grid < 1) grid = 1;
    if (grid > MAX_PARTIALS) grid = MAX_PARTIALS;
    *grid_out = grid;
    const int cap = row_pass_hot_capacity(lanes_per_row, two);
    p.n_hot = (int)(n_nonempty_cols < cap ? n_nonempty_cols : cap);
    const size_t stage = lanes_per_row == 0 ? sizeof(double) * R_WARPS * R_STREAM_CAP : 0;
    const size_t smem = stage + (size_t)p.n_hot * (two ? 16 : 8);
    if (value_bytes == 4)
        return two ? launch_row_variant<float, true>(s, grid, lanes_per_row, smem, p)
                   : launch_row_variant<float, false>(s, grid, lanes_per_row, smem, p);
    return two ? launch_row_variant<double, true>(s, grid, lanes_per_row, smem, p)
               : launch_row_variant<double, false>(s, grid, lanes_per_row, smem, p);
}

// ------------------------------------------------------------------------------------------
// col_pass_kernel -- the axpy half of `add`, transposed: out_j = sum_i val_ij * r_i over the CSC
// copy (vectorSum += (wt*l') x_i summed column by column instead of scattered row by row;
// ValueAndGradientAggregator.scala:150, HessianVectorAggregator.scala:118; with SQUARE,
// HessianDiagonalAggregator.scala:59).
//
// Work is split by ENTRIES, not by columns (column lengths are power-law): CTA c owns entries
// [c*CHUNK, (c+1)*CHUNK), CHUNK = 256*ITEMS.  Each thread loads ITEMS consecutive entries
// (row index with the column-start flag in bit 31, value) with 128-bit loads, gathers r for all
// of them at once, adds up its own run, and a CTA-wide segmented scan carries partial sums
// across threads.  Columns that start in the chunk are written by this CTA (single writer); the
// part of the chunk that precedes its first column start belongs to a column begun in an earlier
// chunk and goes to carry_head[c] for fixup_kernel.
// ------------------------------------------------------------------------------------------
struct SegScan {
    int cnt;     // column starts seen
    int flag;    // any column start seen
    double val;  // running sum since the last column start
};

__device__ __forceinline__ SegScan seg_combine(const SegScan &l, const SegScan &r)
{
    SegScan o;
    o.cnt = l.cnt + r.cnt;
    o.flag = l.flag | r.flag;
    o.val = r.flag ? r.val : l.val + r.val;
    return o;
}

template <typename VT, bool SQUARE, int ITEMS>
__global__ void __launch_bounds__(T_THREADS) col_pass_kernel(const ColPassParams p)
{
    constexpr int CHUNK = T_THREADS * ITEMS;
    __shared__ SegScan wagg[T_THREADS / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = blockIdx.x;
    const int64_t base = (int64_t)c * CHUNK;
    const int cnt = (p.nnz - base) < CHUNK ? (int)(p.nnz - base) : CHUNK;
    const int col_lo = p.chunk_col_lo[c], col_hi = p.chunk_col_lo[c + 1];
    const VT *__restrict__ val = static_cast<const VT *>(p.val) + base + tid * ITEMS;
    const int32_t *__restrict__ row = p.row + base + tid * ITEMS;
    const double *__restrict__ r = p.r;

    // ---- load ITEMS consecutive entries (the arrays are padded by 16 entries past nnz) ----
    int32_t rw[ITEMS];
    double pr[ITEMS];
    const int mine_cnt = cnt - tid * ITEMS;  // entries of this thread that exist (may be <= 0 in the last chunk)
#pragma unroll
    for (int i = 0; i < ITEMS / 4; ++i) {
        if (mine_cnt <= 0) {
            rw[4 * i] = rw[4 * i + 1] = rw[4 * i + 2] = rw[4 * i + 3] = 0;
            pr[4 * i] = pr[4 * i + 1] = pr[4 * i + 2] = pr[4 * i + 3] = 0.0;
            continue;
        }
        const int4 rr = __ldg(reinterpret_cast<const int4 *>(row) + i);
        rw[4 * i] = rr.x; rw[4 * i + 1] = rr.y; rw[4 * i + 2] = rr.z; rw[4 * i + 3] = rr.w;
        if (sizeof(VT) == 4) {
            const float4 vv = __ldg(reinterpret_cast<const float4 *>(val) + i);
            pr[4 * i] = vv.x; pr[4 * i + 1] = vv.y; pr[4 * i + 2] = vv.z; pr[4 * i + 3] = vv.w;
        } else {
            const double2 va = __ldg(reinterpret_cast<const double2 *>(val) + 2 * i);
            const double2 vb = __ldg(reinterpret_cast<const double2 *>(val) + 2 * i + 1);
            pr[4 * i] = va.x; pr[4 * i + 1] = va.y; pr[4 * i + 2] = vb.x; pr[4 * i + 3] = vb.y;
        }
    }
    unsigned mask = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const bool live = i < mine_cnt;
        if (live && rw[i] < 0) mask |= 1u << i;
        const int32_t ri = live ? (rw[i] & 0x7fffffff) : 0;
        double x = live ? pr[i] : 0.0;
        if (SQUARE) x *= x;
        pr[i] = __dmul_rn(__ldg(r + ri), x);
    }

    // ---- blocked segmented reduction ----
    SegScan mine;
    mine.cnt = __popc(mask);
    mine.flag = mask != 0;
    {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            if ((mask >> i) & 1u) s = 0.0;
            s += pr[i];
        }
        mine.val = s;  // sum after the last column start (the whole run if there is none)
    }
    SegScan inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        SegScan up;
        up.cnt = __shfl_up_sync(0xffffffffu, inc.cnt, o);
        up.flag = __shfl_up_sync(0xffffffffu, inc.flag, o);
        up.val = __shfl_up_sync(0xffffffffu, inc.val, o);
        if (lane >= o) inc = seg_combine(up, inc);
    }
    if (lane == 31) wagg[warp] = inc;
    SegScan exc;
    exc.cnt = __shfl_up_sync(0xffffffffu, inc.cnt, 1);
    exc.flag = __shfl_up_sync(0xffffffffu, inc.flag, 1);
    exc.val = __shfl_up_sync(0xffffffffu, inc.val, 1);
    if (lane == 0) { exc.cnt = 0; exc.flag = 0; exc.val = 0.0; }
    __syncthreads();
    SegScan pre;
    pre.cnt = 0; pre.flag = 0; pre.val = 0.0;
    for (int w = 0; w < warp; ++w) pre = seg_combine(pre, wagg[w]);
    exc = seg_combine(pre, exc);

    // walk again, closing columns
    if (mine.flag) {
        double s = 0.0;
        int ord = col_lo + exc.cnt;  // internal id of the next column start this thread meets
        bool first = true;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            if ((mask >> i) & 1u) {
                const double total = first ? exc.val + s : s;
                if (first && exc.cnt == 0) p.carry_head[c] = total;  // column began in an earlier chunk
                else p.out[p.nzcol_id[ord - 1]] = total;
                first = false;
                ++ord;
                s = 0.0;
            }
            s += pr[i];
        }
    }
    if (tid == T_THREADS - 1) {
        const SegScan tot = seg_combine(exc, mine);
        if (tot.flag) p.out[p.nzcol_id[col_hi - 1]] = tot.val;  // last column started here: owner
        else p.carry_head[c] = tot.val;                         // the whole chunk is one column's middle
    }
}

template <typename VT, bool SQUARE>
static int launch_col_items(cudaStream_t s, int items, const ColPassParams &p)
{
    switch (items) {
    case 8: col_pass_kernel<VT, SQUARE, 8><<<p.n_chunks, T_THREADS, 0, s>>>(p); break;
    case 16: col_pass_kernel<VT, SQUARE, 16><<<p.n_chunks, T_THREADS, 0, s>>>(p); break;
    default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

int launch_col_pass(cudaStream_t s, int value_bytes, bool square, int chunk_entries, const ColPassParams &p)
{
    if (p.n_chunks <= 0) return 0;
    const int items = chunk_entries / T_THREADS;
    if (value_bytes == 4)
        return square ? launch_col_items<float, true>(s, items, p) : launch_col_items<float, false>(s, items, p);
    return square ? launch_col_items<double, true>(s, items, p) : launch_col_items<double, false>(s, items, p);
}

// ------------------------------------------------------------------------------------------
// fixup_kernel -- the `merge` of partial aggregates (ValueAndGradientAggregator.scala:161-176)
// as far as one GPU is concerned: columns cut by chunk boundaries get their heads added (one
// warp per column, fixed order), and the last CTA adds up the row kernel's per-CTA (F, P).
// ------------------------------------------------------------------------------------------
constexpr int F_THREADS = 256;

__global__ void __launch_bounds__(F_THREADS)
fixup_kernel(const FixEntry *__restrict__ fix, int n_fix, const double *__restrict__ carry_head,
             double *__restrict__ out, int d, const double *__restrict__ partials, int n_partials,
             int write_value, int write_prefactor)
{
    __shared__ double red[2 * (F_THREADS / 32)];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (blockIdx.x == gridDim.x - 1 && partials != nullptr) {
        double f = 0.0, pp = 0.0;
        for (int i = threadIdx.x; i < n_partials; i += F_THREADS) {
            f += partials[2 * i];
            pp += partials[2 * i + 1];
        }
        block_sum2<F_THREADS / 32>(f, pp, red);
        if (threadIdx.x == 0) {
            out[d] = write_value ? f : 0.0;
            out[d + 1] = write_prefactor ? pp : 0.0;
        }
        return;
    }
    const int q = blockIdx.x * (F_THREADS / 32) + warp;
    if (q >= n_fix) return;
    const FixEntry e = fix[q];
    double s = 0.0;
    for (int cc = e.chunk_begin + lane; cc < e.chunk_end; cc += 32) s += carry_head[cc];
    s = warp_sum(s);
    if (lane == 0) out[e.col] += s;
}

int launch_fixup(cudaStream_t s, const FixEntry *fix, int n_fix, const double *carry_head, double *out, int d,
                 const double *partials, int n_partials, bool write_value, bool write_prefactor)
{
    int grid = (n_fix + (F_THREADS / 32) - 1) / (F_THREADS / 32) + (partials ? 1 : 0);
    if (grid == 0) return 0;
    fixup_kernel<<<grid, F_THREADS, 0, s>>>(fix, n_fix, carry_head, out, d, partials, n_partials,
                                            write_value ? 1 : 0, write_prefactor ? 1 : 0);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// finalize_kernel -- ValueAndGradientAggregator.getVector (:199-221):
//   result_j = (vectorSum_j - shifts_j * vectorShiftPrefactorSum) * factors_j   (4 cases)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
finalize_kernel(int d, const double *__restrict__ raw, const double *__restrict__ factors,
                const double *__restrict__ shifts, double *__restrict__ out_value, double *__restrict__ out_vec)
{
    const double P = raw[d + 1];
    if (blockIdx.x == 0 && threadIdx.x == 0 && out_value != nullptr) *out_value = raw[d];
    if (out_vec == nullptr) return;
    for (int j = blockIdx.x * 256 + threadIdx.x; j < d; j += gridDim.x * 256) {
        double v = raw[j];
        if (shifts) v = __dsub_rn(v, __dmul_rn(shifts[j], P));
        if (factors) v = __dmul_rn(v, factors[j]);
        out_vec[j] = v;
    }
}

int launch_finalize(cudaStream_t s, int d, const double *raw, const double *factors, const double *shifts,
                    double *out_value, double *out_vec)
{
    int grid = (d + 1023) / 1024;
    if (grid < 1) grid = 1;
    if (grid > 2048) grid = 2048;
    finalize_kernel<<<grid, 256, 0, s>>>(d, raw, factors, shifts, out_value, out_vec);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace pml
