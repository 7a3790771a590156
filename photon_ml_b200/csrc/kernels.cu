// kernels.cu -- the per-evaluation kernels of the GLM objective path (sm_100a).
//
//   prepare_vectors_kernel  a1/a6  ValueAndGradientAggregator.init / HessianVectorAggregator.init
//   row_pass_kernel         a2/a7/a9  the per-row half of `add` (lanes-per-row and segment-per-warp variants)
//   seg_spmv_warp_kernel    a2/a7/a9  entry-chunked segmented product: the CSR stream (default for single-vector
//                                     passes) and the cold columns of the CSC copy
//   seg_spmv_tiled_kernel   a2/a7/a9  the axpy half of `add` over the hot columns of the CSC copy, r staged in
//                                     shared memory tile by tile
//   row_epilogue_kernel     a2/a7/a9  pointwise loss and scalar sums after seg_spmv_warp_kernel on the CSR stream
//   fixup_kernel, reduce_partial_kernel  a3  the part of `merge` that joins runs split across chunks / tiles
//   finalize_kernel         a4  getVector
//
// (aN = row of SURVEY.md 8a; reference file:line in DESIGN.md and at each kernel.)
// All accumulation is fp64.  No atomics anywhere: every output word has exactly one writer, and
// every reduction has a fixed shape, so results are bit-reproducible for a given packing.
//
// Columns carry "internal ids" assigned at pack time by descending entry count (pack.cu), so the
// coefficients of the hottest columns sit at the front of the effective-coefficient vector and
// can be staged in shared memory: a scattered 8-byte gather from global memory costs an L1 tag
// lookup per lane and a 32-byte L2 sector (the measured limiter of the first version of these
// kernels), while shared memory serves 32 lanes in a few cycles.
#include <cuda_runtime.h>

#include <cstdlib>
#include <type_traits>
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
// This is the warp-oriented form of the pass (persistent 256-thread CTAs, gathers through L1); by default it
// serves the paired (ec, ev) pass of a cold Hessian-vector product, and any pass when a variant is forced.
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
__device__ __forceinline__ void gather(const double *__restrict__ vec, int32_t c, double &w0, double &w1)
{
    if (TWO) {
        const double2 g = __ldg(reinterpret_cast<const double2 *>(vec) + c);
        w0 = g.x;
        w1 = g.y;
    } else {
        w0 = __ldg(vec + c);
        w1 = 0.0;
    }
}

template <typename VT, int LPR, bool TWO>
__global__ void __launch_bounds__(R_THREADS, 5) row_pass_kernel(const RowPassParams p)
{
    constexpr int CAPE = TWO ? R_STREAM_CAP / 2 : R_STREAM_CAP;
    __shared__ double red[2 * R_WARPS];
    __shared__ double stage[(LPR == 0) ? R_WARPS * R_STREAM_CAP : 1];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const VT *__restrict__ val = static_cast<const VT *>(p.val);
    const int32_t *__restrict__ col = p.col;
    const double *__restrict__ vec = p.vec;

    // marginShift = -(ec . shifts); featureVectorProductShift = ev . shifts
    double shift0 = 0.0, shift1 = 0.0;
    if (p.shift_partials != nullptr) {
        for (int i = threadIdx.x; i < p.n_shift_partials; i += R_THREADS) {
            shift0 += p.shift_partials[2 * i];
            shift1 += p.shift_partials[2 * i + 1];
        }
        block_sum2<R_WARPS>(shift0, shift1, red);
    }
    const double margin_shift = (p.mode == MODE_HV_CACHED) ? 0.0 : -shift0;
    const double fs = (p.mode == MODE_HV_CACHED) ? shift0 : shift1;

    double accF = 0.0, accP = 0.0;
    const int64_t n_blk = (p.n + 31) >> 5;
    for (int64_t blk = (int64_t)blockIdx.x * R_WARPS + warp; blk < n_blk; blk += (int64_t)gridDim.x * R_WARPS) {
        if (p.block_active != nullptr && !p.block_active[blk]) continue;  // all 32 rows belong to masked entities
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
                    const int32_t c = __ldcs(col + k) & 0x7fffffff;
                    const double x = load_value(val, k);
                    double w0, w1;
                    gather<TWO>(vec, c, w0, w1);
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
                    const int32_t c0 = __ldcs(col + k) & 0x7fffffff, c1 = __ldcs(col + k + L) & 0x7fffffff;
                    const int32_t c2 = __ldcs(col + k + 2 * L) & 0x7fffffff, c3 = __ldcs(col + k + 3 * L) & 0x7fffffff;
                    const double x0 = load_value(val, k), x1 = load_value(val, k + L);
                    const double x2 = load_value(val, k + 2 * L), x3 = load_value(val, k + 3 * L);
                    double w00, w01, w10, w11, w20, w21, w30, w31;
                    gather<TWO>(vec, c0, w00, w01);
                    gather<TWO>(vec, c1, w10, w11);
                    gather<TWO>(vec, c2, w20, w21);
                    gather<TWO>(vec, c3, w30, w31);
                    a += (x0 * w00 + x1 * w10) + (x2 * w20 + x3 * w30);
                    if (TWO) b += (x0 * w01 + x1 * w11) + (x2 * w21 + x3 * w31);
                }
                for (; k < end; k += L) {
                    const int32_t c = __ldcs(col + k) & 0x7fffffff;
                    const double x = load_value(val, k);
                    double w0, w1;
                    gather<TWO>(vec, c, w0, w1);
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
                const double wl = __dmul_rn(wt, l);
                if (p.row_loss) p.row_loss[my_row] = wl;
                accF += wl;
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

template <typename VT, bool TWO>
static int launch_row_variant(cudaStream_t s, int grid, int lpr, const RowPassParams &p)
{
    switch (lpr) {
    case 0: row_pass_kernel<VT, 0, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    case 2: row_pass_kernel<VT, 2, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    case 4: row_pass_kernel<VT, 4, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    case 8: row_pass_kernel<VT, 8, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    case 16: row_pass_kernel<VT, 16, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    case 32: row_pass_kernel<VT, 32, TWO><<<grid, R_THREADS, 0, s>>>(p); break;
    default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// 256-thread CTAs, five per SM: the gathers live on L1 (a shared-memory table of hot coefficients was
// measured for this kernel and lost more L1 than it saved).
int launch_row_pass(cudaStream_t s, int sm_count, int value_bytes, int lanes_per_row, bool two, const RowPassParams &p,
                    int *grid_out)
{
    const int64_t n_blk = (p.n + 31) / 32;
    int64_t want = (n_blk + R_WARPS - 1) / R_WARPS;
    int64_t cap = (int64_t)sm_count * 5;
    if (cap > MAX_PARTIALS) cap = MAX_PARTIALS;
    int grid = (int)(want < cap ? want : cap);
    if (grid < 1) grid = 1;
    *grid_out = grid;
    if (value_bytes == 0)
        return two ? launch_row_variant<One, true>(s, grid, lanes_per_row, p)
                   : launch_row_variant<One, false>(s, grid, lanes_per_row, p);
    if (value_bytes == 4)
        return two ? launch_row_variant<float, true>(s, grid, lanes_per_row, p)
                   : launch_row_variant<float, false>(s, grid, lanes_per_row, p);
    return two ? launch_row_variant<double, true>(s, grid, lanes_per_row, p)
               : launch_row_variant<double, false>(s, grid, lanes_per_row, p);
}

// ------------------------------------------------------------------------------------------
// seg_spmv_warp_kernel -- one kernel for both sparse products of an evaluation, chunked by ENTRIES.
//
//   CSC side ("col pass"): out_j = sum_i val_ij * r_i -- the axpy half of `add`, transposed
//     (vectorSum += (wt*l') x_i summed column by column instead of scattered row by row;
//     ValueAndGradientAggregator.scala:150, HessianVectorAggregator.scala:118; with SQUARE,
//     HessianDiagonalAggregator.scala:59).  idx = row index, table = r.
//   CSR side (row-pass variant 1, "segment-per-CTA"): z_i = sum_j val_ij * ec_j
//     (LabeledPoint.computeMargin, lib/data/LabeledPoint.scala:62).  idx = internal column id,
//     table = effective coefficients, the first n_hot of them staged in shared memory; with TWO
//     the table holds (ec_j, ev_j) pairs and two sums are produced (Hessian-vector, cold).
//
// Work is split by entries, not by rows/columns (their lengths are power-law): a chunk is 256
// consecutive entries owned by one warp, persistent CTAs hand chunks to their warps round-robin.
// Each lane loads 8 consecutive entries (index with the run-start flag in bit 31, value) with 128-bit
// loads -- the next chunk's loads are issued before the current chunk is reduced, so the stream
// stays in flight -- looks all of them up at once, adds up its own run, and a warp-wide segmented
// scan (shuffles only) carries partial sums across lanes.  Runs that start in the chunk are written
// by this warp (single writer); the part of the chunk that precedes its first run start belongs to a
// run begun in an earlier chunk and goes to carry[c] for fixup_kernel.
// ------------------------------------------------------------------------------------------
template <bool TWO>
struct SegScan {
    int cnt;      // run starts seen (0 = none)
    double val;   // running sum since the last run start
    double val2;  // second sum (TWO only)
};

template <bool TWO>
__device__ __forceinline__ SegScan<TWO> seg_identity()
{
    SegScan<TWO> o;
    o.cnt = 0;
    o.val = 0.0;
    o.val2 = 0.0;
    return o;
}

template <bool TWO>
__device__ __forceinline__ SegScan<TWO> seg_combine(const SegScan<TWO> &l, const SegScan<TWO> &r)
{
    SegScan<TWO> o;
    o.cnt = l.cnt + r.cnt;
    o.val = r.cnt ? r.val : l.val + r.val;
    o.val2 = TWO ? (r.cnt ? r.val2 : l.val2 + r.val2) : 0.0;
    return o;
}

template <bool TWO>
__device__ __forceinline__ SegScan<TWO> seg_shfl_up(const SegScan<TWO> &v, int o)
{
    SegScan<TWO> up;
    up.cnt = __shfl_up_sync(0xffffffffu, v.cnt, o);
    up.val = __shfl_up_sync(0xffffffffu, v.val, o);
    up.val2 = TWO ? __shfl_up_sync(0xffffffffu, v.val2, o) : 0.0;
    return up;
}

template <bool TWO>
__device__ __forceinline__ SegScan<TWO> seg_shfl(const SegScan<TWO> &v, int src)
{
    SegScan<TWO> o;
    o.cnt = __shfl_sync(0xffffffffu, v.cnt, src);
    o.val = __shfl_sync(0xffffffffu, v.val, src);
    o.val2 = TWO ? __shfl_sync(0xffffffffu, v.val2, src) : 0.0;
    return o;
}

// One table look-up.  LOOKUP_GLOBAL: the read-only global path; LOOKUP_MIXED: shared memory for the hot prefix
// [0, n_hot), global beyond (two predicated loads, so a warp mixing both does not diverge); LOOKUP_SHARED: every
// index is a shared-memory index (the tiled CSC pass).  The shared table is addressed through its 32-bit
// shared-window address, computed once per kernel: handed over as a generic pointer, the compiler rebuilt the
// window base (S2R SR_CgaCtaId, MOV, LEA) and kept a bounds test and a dead global path for every single entry --
// a dozen instructions per look-up where two do.
enum { LOOKUP_GLOBAL = 0, LOOKUP_MIXED = 1, LOOKUP_SHARED = 2 };

__device__ __forceinline__ uint32_t shared_window_address(const void *p)
{
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("" : "+r"(a));  // opaque from here on: keep it in a register instead of rematerialising it
    return a;
}

__device__ __forceinline__ double lds_f64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ double2 lds_f64x2(uint32_t addr)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}

template <int LOOKUP>
__device__ __forceinline__ double seg_lookup(uint32_t hot, int n_hot, const double *__restrict__ table, int k)
{
    if (LOOKUP == LOOKUP_GLOBAL) return __ldg(table + k);
    if (LOOKUP == LOOKUP_SHARED) return lds_f64(hot + 8u * (uint32_t)k);
    return (k < n_hot) ? lds_f64(hot + 8u * (uint32_t)k) : __ldg(table + k);
}

template <int LOOKUP>
__device__ __forceinline__ double2 seg_lookup2(uint32_t hot, int n_hot, const double *__restrict__ table, int k)
{
    if (LOOKUP == LOOKUP_GLOBAL || (LOOKUP == LOOKUP_MIXED && k >= n_hot))
        return __ldg(reinterpret_cast<const double2 *>(table) + k);
    return lds_f64x2(hot + 16u * (uint32_t)k);
}

template <typename VT, int ITEMS>
struct SegRegs {
    int32_t idx[ITEMS];
    VT val[ITEMS];
    __device__ __forceinline__ double value(int i) const { return widen(val[i]); }
};
template <int ITEMS>
struct SegRegs<One, ITEMS> {  // implicit 1.0: only the indices travel
    int32_t idx[ITEMS];
    __device__ __forceinline__ double value(int) const { return 1.0; }
};

template <typename VT, int ITEMS>
__device__ __forceinline__ void seg_load(SegRegs<VT, ITEMS> &g, const int32_t *__restrict__ idx, const VT *__restrict__ val,
                                         int64_t first, int mine_cnt)
{
    if (mine_cnt <= 0) {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            g.idx[i] = 0;
            if constexpr (!std::is_same<VT, One>::value) g.val[i] = (VT)0;
        }
        return;
    }
    // the arrays are padded by 16 entries past nnz, so a partially filled thread may read ahead
#pragma unroll
    for (int i = 0; i < ITEMS / 4; ++i) {
        const int4 rr = __ldcs(reinterpret_cast<const int4 *>(idx + first) + i);
        g.idx[4 * i] = rr.x; g.idx[4 * i + 1] = rr.y; g.idx[4 * i + 2] = rr.z; g.idx[4 * i + 3] = rr.w;
        if constexpr (std::is_same<VT, float>::value) {
            const float4 vv = __ldcs(reinterpret_cast<const float4 *>(val + first) + i);
            g.val[4 * i] = vv.x; g.val[4 * i + 1] = vv.y; g.val[4 * i + 2] = vv.z; g.val[4 * i + 3] = vv.w;
        } else if constexpr (std::is_same<VT, double>::value) {
            const double2 va = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i);
            const double2 vb = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i + 1);
            g.val[4 * i] = va.x; g.val[4 * i + 1] = va.y; g.val[4 * i + 2] = vb.x; g.val[4 * i + 3] = vb.y;
        }
    }
}

// Reduces one chunk held in `cur`.  FULL = every thread owns ITEMS live entries (all chunks but
// possibly the last), which keeps the bounds tests out of the persistent loop.
// A chunk belongs to one warp (32*ITEMS entries): the scan is warp-wide, no barrier, no shared memory.
template <typename VT, int ITEMS, bool TWO, bool SQUARE, int LOOKUP, bool FULL>
__device__ __forceinline__ void seg_chunk(const SegParams &p, const SegRegs<VT, ITEMS> &cur, int c, int mine_cnt,
                                          uint32_t hot)
{
    const int lane = threadIdx.x & 31;
    const double *__restrict__ table = p.table;
    const int n_hot = p.n_hot;

    // ---- look up and multiply; collect the run-start flags ----
    double pr[ITEMS], pr2[TWO ? ITEMS : 1];
    unsigned mask = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const bool live = FULL || i < mine_cnt;
        const int32_t raw = live ? cur.idx[i] : 0;
        mask |= ((unsigned)raw >> 31) << i;
        const int32_t k = raw & 0x7fffffff;
        double x = live ? cur.value(i) : 0.0;
        if (SQUARE) x *= x;
        if (TWO) {
            const double2 g = seg_lookup2<LOOKUP>(hot, n_hot, table, k);
            if constexpr (std::is_same<VT, One>::value) {
                pr[i] = live ? g.x : 0.0;
                pr2[i] = live ? g.y : 0.0;
            } else {
                pr[i] = __dmul_rn(g.x, x);
                pr2[i] = __dmul_rn(g.y, x);
            }
        } else {
            const double g = seg_lookup<LOOKUP>(hot, n_hot, table, k);
            if constexpr (std::is_same<VT, One>::value) pr[i] = live ? g : 0.0;
            else pr[i] = __dmul_rn(g, x);
        }
    }

    // ---- thread-local pass: head = sum before the first run start, tail = sum after the last ----
    double head = 0.0, head2 = 0.0, s = 0.0, s2 = 0.0;
    {
        bool seen = false;
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            if ((mask >> i) & 1u) {
                if (!seen) { head = s; head2 = s2; seen = true; }
                s = 0.0;
                s2 = 0.0;
            }
            s += pr[i];
            if (TWO) s2 += pr2[i];
        }
    }
    SegScan<TWO> mine;
    mine.cnt = __popc(mask);
    mine.val = s;  // the whole thread when it holds no run start
    mine.val2 = s2;

    // ---- warp-wide segmented scan of (cnt, tail) ----
    SegScan<TWO> inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const SegScan<TWO> up = seg_shfl_up<TWO>(inc, o);
        if (lane >= o) inc = seg_combine<TWO>(up, inc);
    }
    SegScan<TWO> exc = seg_shfl_up<TWO>(inc, 1);
    if (lane == 0) exc = seg_identity<TWO>();

    const int run_lo = p.chunk_lo[c];
    const int32_t *__restrict__ run_id = p.run_id;
    double *__restrict__ out = p.out;
    if (mask != 0) {
        // the first run start in this thread closes the run that was open when the thread began
        const double total = exc.val + head;
        const double total2 = TWO ? exc.val2 + head2 : 0.0;
        const int ord = run_lo + exc.cnt;  // ordinal of the run that STARTS at this thread's first flag
        if (exc.cnt == 0) {                // ... so the run it closes began in an earlier chunk
            p.carry[c] = total;
            if (TWO) p.carry2[c] = total2;
        } else {
            const int slot = run_id ? run_id[ord - 1] : ord - 1;
            out[slot] = total;
            if (TWO) p.out2[slot] = total2;
        }
        if (mask & (mask - 1)) {           // two or more starts: runs lying entirely inside this thread
            double t = 0.0, t2 = 0.0;
            int o2 = ord;
            bool first = true;
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                if ((mask >> i) & 1u) {
                    if (!first) {
                        const int slot = run_id ? run_id[o2 - 1] : o2 - 1;
                        out[slot] = t;
                        if (TWO) p.out2[slot] = t2;
                    }
                    first = false;
                    ++o2;
                    t = 0.0;
                    t2 = 0.0;
                }
                t += pr[i];
                if (TWO) t2 += pr2[i];
            }
        }
    }
    if (lane == 31) {
        const SegScan<TWO> tot = seg_combine<TWO>(exc, mine);
        if (tot.cnt) {  // the last run started here: this warp owns its slot
            const int last = run_lo + tot.cnt - 1;
            const int slot = run_id ? run_id[last] : last;
            out[slot] = tot.val;
            if (TWO) p.out2[slot] = tot.val2;
        } else {        // the whole chunk is the middle of one run
            p.carry[c] = tot.val;
            if (TWO) p.carry2[c] = tot.val2;
        }
    }
}

template <typename VT, int ITEMS>
__device__ __forceinline__ void seg_load_full(SegRegs<VT, ITEMS> &g, const int32_t *__restrict__ idx,
                                              const VT *__restrict__ val, int64_t first)
{
#pragma unroll
    for (int i = 0; i < ITEMS / 4; ++i) {
        const int4 rr = __ldcs(reinterpret_cast<const int4 *>(idx + first) + i);
        g.idx[4 * i] = rr.x; g.idx[4 * i + 1] = rr.y; g.idx[4 * i + 2] = rr.z; g.idx[4 * i + 3] = rr.w;
        if constexpr (std::is_same<VT, float>::value) {
            const float4 vv = __ldcs(reinterpret_cast<const float4 *>(val + first) + i);
            g.val[4 * i] = vv.x; g.val[4 * i + 1] = vv.y; g.val[4 * i + 2] = vv.z; g.val[4 * i + 3] = vv.w;
        } else if constexpr (std::is_same<VT, double>::value) {
            const double2 va = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i);
            const double2 vb = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i + 1);
            g.val[4 * i] = va.x; g.val[4 * i + 1] = va.y; g.val[4 * i + 2] = vb.x; g.val[4 * i + 3] = vb.y;
        }
    }
}

// Pulls the lines of a chunk two steps ahead into L2 (no registers held): the register prefetch of the next chunk
// then finds its data a few hundred cycles away instead of a DRAM round trip under load -- the first use of the
// prefetched registers was the single largest stall of these kernels (ncu source view, long scoreboard).
#ifndef PML_PF_AHEAD
#define PML_PF_AHEAD 4
#endif
constexpr int PF_AHEAD = PML_PF_AHEAD;  // chunks (per warp) between the L2 prefetch and the chunk being reduced

template <typename VT, int ITEMS>
__device__ __forceinline__ void seg_prefetch_l2(const int32_t *__restrict__ idx, const VT *__restrict__ val, int64_t chunk)
{
    constexpr int CHUNK = 32 * ITEMS;
    constexpr int IDX_LINES = CHUNK * 4 / 128;
    constexpr int VAL_LINES = std::is_same<VT, One>::value ? 0 : CHUNK * (int)sizeof(VT) / 128;
    const int lane = threadIdx.x & 31;
    const char *ptr = nullptr;
    if (lane < IDX_LINES) ptr = reinterpret_cast<const char *>(idx + chunk * CHUNK) + 128 * lane;
    else if (lane < IDX_LINES + VAL_LINES)
        ptr = reinterpret_cast<const char *>(val) + (size_t)chunk * CHUNK * sizeof(VT) + 128 * (lane - IDX_LINES);
    if (ptr != nullptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

// Persistent CTAs; warp w of CTA b takes chunks (b*WARPS + w) + k*(grid*WARPS).  With a table (the CSR stream) one
// 1024-thread CTA per SM shares the hot prefix among 32 warps; without (cold CSC columns) 256-thread CTAs.
template <typename VT, int ITEMS, bool TWO, bool SQUARE, bool HOT, int SW_THREADS, bool MASKED = false>
__global__ void __launch_bounds__(SW_THREADS, 1024 / SW_THREADS) seg_spmv_warp_kernel(const SegParams p)
{
    constexpr int CHUNK = 32 * ITEMS;
    constexpr int WARPS = SW_THREADS / 32;
    extern __shared__ __align__(16) double hot[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const VT *__restrict__ val = static_cast<const VT *>(p.val);
    const int32_t *__restrict__ idx = p.idx;
    if (HOT) {
        const int words = TWO ? 2 * p.n_hot : p.n_hot;
        for (int i = tid; i < words; i += SW_THREADS) hot[i] = p.table[i];
        __syncthreads();
    }
    constexpr int LOOKUP = HOT ? LOOKUP_MIXED : LOOKUP_GLOBAL;
    const uint32_t hot_at = shared_window_address(hot);
    const int n_full = (int)(p.nnz / CHUNK);
    const int stride = gridDim.x * WARPS;
    // Batches with masked entities (pml_batch_set_active): a chunk whose entries all belong to masked entities is
    // neither loaded nor reduced -- converged entities stop costing bandwidth.  Whatever such a chunk would have
    // written (run sums, its carry) feeds masked entities only, whose results the caller overwrites.  (A variant of
    // its own, MASKED: with the test compiled into the common loop the CSR pass of the bench ran 10 % slower.)
    const unsigned char *__restrict__ active = MASKED ? p.active : nullptr;
    SegRegs<VT, ITEMS> cur;
    int c = blockIdx.x * WARPS + warp;
    bool live = c < n_full && (active == nullptr || active[c] != 0);
    if (live) seg_load_full<VT, ITEMS>(cur, idx, val, (int64_t)c * CHUNK + lane * ITEMS);
    for (; c < n_full; c += stride) {
        SegRegs<VT, ITEMS> nxt;
        const int cn = c + stride;
        const bool live_nxt = cn < n_full && (active == nullptr || active[cn] != 0);
        if (live_nxt) seg_load_full<VT, ITEMS>(nxt, idx, val, (int64_t)cn * CHUNK + lane * ITEMS);
        // (not for the cold CSC columns: their gathers of r live in L2 and the extra lines cost them more than
        // the stream gains -- measured 0.42 -> 0.52 ms)
        if (HOT && active == nullptr && cn + (PF_AHEAD - 1) * stride < n_full)
            seg_prefetch_l2<VT, ITEMS>(idx, val, cn + (PF_AHEAD - 1) * stride);
        if (live) seg_chunk<VT, ITEMS, TWO, SQUARE, LOOKUP, true>(p, cur, c, ITEMS, hot_at);
        cur = nxt;
        live = live_nxt;
    }
    if (n_full < p.n_chunks && (int)(blockIdx.x * WARPS + warp) == n_full % stride &&
        (active == nullptr || active[n_full] != 0)) {
        const int64_t base = (int64_t)n_full * CHUNK;
        const int cnt = (int)(p.nnz - base);
        seg_load<VT, ITEMS>(cur, idx, val, base + lane * ITEMS, cnt - lane * ITEMS);
        seg_chunk<VT, ITEMS, TWO, SQUARE, LOOKUP, false>(p, cur, n_full, cnt - lane * ITEMS, hot_at);
    }
}

template <typename VT, bool TWO, bool SQUARE, int SW_THREADS>
static int launch_seg_warp(cudaStream_t s, int sm_count, size_t smem, const SegParams &p)
{
    const int warps = SW_THREADS / 32;
    int grid = sm_count * (1024 / SW_THREADS);
    const int want = (p.n_chunks + warps - 1) / warps;
    if (grid > want) grid = want;
    if (p.n_hot > 0) {
        if (smem > 48 * 1024 &&
            cudaFuncSetAttribute(seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, true, SW_THREADS>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, SEG_TABLE_BYTES) != cudaSuccess)
            return -1;
        if (p.active != nullptr) {
            if (smem > 48 * 1024 &&
                cudaFuncSetAttribute(seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, true, SW_THREADS, true>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, SEG_TABLE_BYTES) != cudaSuccess)
                return -1;
            seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, true, SW_THREADS, true><<<grid, SW_THREADS, smem, s>>>(p);
        } else {
            seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, true, SW_THREADS><<<grid, SW_THREADS, smem, s>>>(p);
        }
    } else if (p.active != nullptr) {
        seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, false, SW_THREADS, true><<<grid, SW_THREADS, 0, s>>>(p);
    } else {
        seg_spmv_warp_kernel<VT, 8, TWO, SQUARE, false, SW_THREADS><<<grid, SW_THREADS, 0, s>>>(p);
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// seg_spmv_tiled_kernel -- the CSC pass over the hot columns, "staged through shared memory": the rows
// are cut into tiles of <= 16000, a CTA (1024 threads, one per SM) copies the tile's slice of r into
// shared memory once and its 32 warps then reduce the tile's warp-chunks with every gather served from
// shared memory (a scattered 8-byte gather from global memory costs an L1 tag lookup per lane and a
// 32-byte L2 sector; r is 8*n bytes and does not stay in L2 next to the stream).  One run = one
// (tile, column) pair; its sum goes to partial[tile][column], reduced over tiles by reduce_partial_kernel.
// Padding entries at the end of a tile point at the zero slot hot[tile_rows] with value 0.
// ------------------------------------------------------------------------------------------
constexpr int ST_THREADS = 1024;

// The tiled region stores its row indices -- local to the tile, below 2^15 -- as 16-bit words with the run-start
// flag in bit 15: 6 bytes per entry instead of 8 with fp32 values (the pass is close to the HBM bound once its
// look-ups are free of bank conflicts).  Unpacked into the 32-bit form (flag in bit 31) the reduction expects.
template <typename VT, int ITEMS>
__device__ __forceinline__ void seg_load_full16(SegRegs<VT, ITEMS> &g, const uint16_t *__restrict__ idx,
                                                const VT *__restrict__ val, int64_t first)
{
    static_assert(ITEMS == 8, "one 16-byte load carries a lane's 8 indices");
    const uint4 rr = __ldcs(reinterpret_cast<const uint4 *>(idx + first));
    const uint32_t w[4] = { rr.x, rr.y, rr.z, rr.w };
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        g.idx[2 * i] = (int32_t)((w[i] & 0x7fffu) | ((w[i] << 16) & 0x80000000u));
        g.idx[2 * i + 1] = (int32_t)(((w[i] >> 16) & 0x7fffu) | (w[i] & 0x80000000u));
    }
#pragma unroll
    for (int i = 0; i < ITEMS / 4; ++i) {
        if constexpr (std::is_same<VT, float>::value) {
            const float4 vv = __ldcs(reinterpret_cast<const float4 *>(val + first) + i);
            g.val[4 * i] = vv.x; g.val[4 * i + 1] = vv.y; g.val[4 * i + 2] = vv.z; g.val[4 * i + 3] = vv.w;
        } else if constexpr (std::is_same<VT, double>::value) {
            const double2 va = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i);
            const double2 vb = __ldcs(reinterpret_cast<const double2 *>(val + first) + 2 * i + 1);
            g.val[4 * i] = va.x; g.val[4 * i + 1] = va.y; g.val[4 * i + 2] = vb.x; g.val[4 * i + 3] = vb.y;
        }
    }
}

template <typename VT, int ITEMS>
__device__ __forceinline__ void seg_prefetch_l2_16(const uint16_t *__restrict__ idx, const VT *__restrict__ val, int64_t chunk)
{
    constexpr int CHUNK = 32 * ITEMS;
    constexpr int IDX_LINES = CHUNK * 2 / 128;
    constexpr int VAL_LINES = std::is_same<VT, One>::value ? 0 : CHUNK * (int)sizeof(VT) / 128;
    const int lane = threadIdx.x & 31;
    const char *ptr = nullptr;
    if (lane < IDX_LINES) ptr = reinterpret_cast<const char *>(idx + chunk * CHUNK) + 128 * lane;
    else if (lane < IDX_LINES + VAL_LINES)
        ptr = reinterpret_cast<const char *>(val) + (size_t)chunk * CHUNK * sizeof(VT) + 128 * (lane - IDX_LINES);
    if (ptr != nullptr) asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr));
}

template <typename VT, bool SQUARE>
__global__ void __launch_bounds__(ST_THREADS, 1)
seg_spmv_tiled_kernel(const SegParams p, const uint16_t *__restrict__ idx, const unsigned char *__restrict__ uniform,
                      const int32_t *__restrict__ tile_first_chunk, int n_tiles, int tile_rows,
                      const double *__restrict__ r, int64_t n)
{
    constexpr int ITEMS = 8;
    constexpr int CHUNK = 32 * ITEMS;
    constexpr int WARPS = ST_THREADS / 32;
    extern __shared__ __align__(16) double hot[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const VT *__restrict__ val = static_cast<const VT *>(p.val);
    const uint32_t hot_at = shared_window_address(hot);
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        __syncthreads();  // every warp is done with the previous tile's slice
        const int64_t row0 = (int64_t)t * tile_rows;
        for (int i = tid; i <= tile_rows; i += ST_THREADS) hot[i] = (i < tile_rows && row0 + i < n) ? r[row0 + i] : 0.0;
        __syncthreads();
        const int c_end = tile_first_chunk[t + 1];
        int c = tile_first_chunk[t] + warp;
        SegRegs<VT, ITEMS> cur;
        unsigned char uni = 0;
        if (c < c_end) {
            seg_load_full16<VT, ITEMS>(cur, idx, val, (int64_t)c * CHUNK + lane * ITEMS);
            uni = uniform[c];
        }
        for (; c < c_end; c += WARPS) {
            SegRegs<VT, ITEMS> nxt;
            unsigned char uni_nxt = 0;
            const int cn = c + WARPS;
            if (cn < c_end) {
                seg_load_full16<VT, ITEMS>(nxt, idx, val, (int64_t)cn * CHUNK + lane * ITEMS);
                uni_nxt = uniform[cn];
            }
            if (cn + (PF_AHEAD - 1) * WARPS < c_end) seg_prefetch_l2_16<VT, ITEMS>(idx, val, cn + (PF_AHEAD - 1) * WARPS);
            if (uni) {
                // Marked at pack time (pack.cu, narrow_tiled_kernel): no run starts in this chunk, all 256 entries
                // continue one (tile, column) run -- the long runs of the hottest columns, about half of the region on
                // Zipf data.  No flags, no scan, no run look-up: the products are added up per lane, one warp
                // reduction, and the sum joins the run through carry[c] like any other chunk head (fixup_kernel).
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    double x = cur.value(i);
                    if (SQUARE) x *= x;
                    const double g = lds_f64(hot_at + 8u * (uint32_t)cur.idx[i]);
                    if constexpr (std::is_same<VT, One>::value) s += g;
                    else s += __dmul_rn(g, x);
                }
                s = warp_sum(s);
                if (lane == 0) p.carry[c] = s;
            } else {
                seg_chunk<VT, ITEMS, false, SQUARE, LOOKUP_SHARED, true>(p, cur, c, ITEMS, hot_at);
            }
            cur = nxt;
            uni = uni_nxt;
        }
    }
}

int launch_seg_tiled(cudaStream_t s, int sm_count, int value_bytes, bool square, const TiledSide &t, const double *r,
                     int64_t n)
{
    if (t.n_chunks <= 0) return 0;
    SegParams p;
    p.nnz = t.len;
    p.idx = nullptr;  // the tiled region's indices are the 16-bit array handed over separately
    p.val = t.val;
    p.run_id = t.run_id;
    p.chunk_lo = t.chunk_lo;
    p.table = nullptr;
    p.out = t.partial;
    p.out2 = nullptr;
    p.carry = t.carry;
    p.carry2 = nullptr;
    p.n_chunks = t.n_chunks;
    p.n_hot = 0x7fffffff;  // every index is a shared-memory index
    p.active = nullptr;
    const size_t smem = sizeof(double) * (size_t)(t.tile_rows + 1);
    const int grid = t.n_tiles < sm_count ? t.n_tiles : sm_count;
#define PML_TILED(VT, SQ)                                                                                           \
    do {                                                                                                            \
        if (smem > 48 * 1024 && cudaFuncSetAttribute(seg_spmv_tiled_kernel<VT, SQ>,                                 \
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize,                   \
                                                     (int)(sizeof(double) * 28002)) != cudaSuccess)   \
            return -1;                                                                                              \
        seg_spmv_tiled_kernel<VT, SQ><<<grid, ST_THREADS, smem, s>>>(p, t.idx16, t.uniform, t.tile_first_chunk, t.n_tiles, \
                                                                     t.tile_rows, r, n);                            \
    } while (0)
    if (value_bytes == 0) {
        if (square) PML_TILED(One, true); else PML_TILED(One, false);
    } else if (value_bytes == 4) {
        if (square) PML_TILED(float, true); else PML_TILED(float, false);
    } else {
        if (square) PML_TILED(double, true); else PML_TILED(double, false);
    }
#undef PML_TILED
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// out[original id of hot column c] = sum over tiles of partial[tile][c], tiles in a fixed order:
// 8 threads per column stride over the tiles, then a fixed tree over the 8.
__global__ void __launch_bounds__(256)
reduce_partial_kernel(int n_tiles, int K, const double *__restrict__ partial, const int32_t *__restrict__ nzcol_id,
                      double *__restrict__ out)
{
    __shared__ double acc[8][33];
    const int cx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cx;
    double s = 0.0;
    if (c < K)
        for (int t = ty; t < n_tiles; t += 8) s += partial[(size_t)t * K + c];
    acc[ty][cx] = s;
    __syncthreads();
    if (ty == 0 && c < K) {
        const double total = ((acc[0][cx] + acc[1][cx]) + (acc[2][cx] + acc[3][cx])) +
                             ((acc[4][cx] + acc[5][cx]) + (acc[6][cx] + acc[7][cx]));
        out[nzcol_id[c]] = total;
    }
}

// The same for a handful of row blocks (the super blocks of the cold columns): one thread per column.
__global__ void __launch_bounds__(256)
reduce_few_kernel(int n_blocks, int n_cols, const double *__restrict__ partial, const int32_t *__restrict__ nzcol_id,
                  double *__restrict__ out)
{
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= n_cols) return;
    double s = 0.0;
    for (int t = 0; t < n_blocks; ++t) s += __ldcs(partial + (size_t)t * n_cols + c);
    out[nzcol_id[c]] = s;
}

int launch_reduce_partial(cudaStream_t s, int n_tiles, int n_cols, const double *partial, const int32_t *nzcol_id,
                          double *out)
{
    if (n_cols <= 0) return 0;
    if (n_tiles <= 16) {
        reduce_few_kernel<<<(n_cols + 255) / 256, 256, 0, s>>>(n_tiles, n_cols, partial, nzcol_id, out);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    reduce_partial_kernel<<<(n_cols + 31) / 32, 256, 0, s>>>(n_tiles, n_cols, partial, nzcol_id, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// table_len = number of leading table entries worth staging in shared memory (0 = none)
int launch_seg_spmv(cudaStream_t s, int sm_count, int value_bytes, bool two, bool square, SegParams p, int64_t table_len)
{
    if (p.n_chunks <= 0) return 0;
    const int64_t cap_entries = SEG_TABLE_BYTES / (two ? 16 : 8);
    p.n_hot = (int)(table_len < cap_entries ? table_len : cap_entries);
    const size_t smem = (size_t)p.n_hot * (two ? 16 : 8);
#define PML_SEG_WARP_DISPATCH(T)                                                                       \
    do {                                                                                               \
        if (value_bytes == 0) {                                                                        \
            if (two) return launch_seg_warp<One, true, false, T>(s, sm_count, smem, p);                \
            return square ? launch_seg_warp<One, false, true, T>(s, sm_count, smem, p)                 \
                          : launch_seg_warp<One, false, false, T>(s, sm_count, smem, p);               \
        }                                                                                              \
        if (value_bytes == 4) {                                                                        \
            if (two) return launch_seg_warp<float, true, false, T>(s, sm_count, smem, p);              \
            return square ? launch_seg_warp<float, false, true, T>(s, sm_count, smem, p)               \
                          : launch_seg_warp<float, false, false, T>(s, sm_count, smem, p);             \
        }                                                                                              \
        if (two) return launch_seg_warp<double, true, false, T>(s, sm_count, smem, p);                 \
        return square ? launch_seg_warp<double, false, true, T>(s, sm_count, smem, p)                  \
                      : launch_seg_warp<double, false, false, T>(s, sm_count, smem, p);                \
    } while (0)
    if (p.n_hot > 0) PML_SEG_WARP_DISPATCH(1024);
    PML_SEG_WARP_DISPATCH(256);
#undef PML_SEG_WARP_DISPATCH
}

// ------------------------------------------------------------------------------------------
// sell_rows_kernel -- the sparse dots x_i . c of `add` (ValueAndGradientAggregator.scala:132-153, the loop behind
// LabeledPoint.computeMargin, lib/data/LabeledPoint.scala:62) with ONE LANE PER ROW over the sliced-ELL copy of the
// rows (pack.cu build_sell).  Against the entry-chunked kernel: no run flags, no segmented scan, no carries and no
// fix-up -- a lane keeps its row's sum in a register -- and the 32 look-ups of one instruction are the same position
// of 32 rows, which the packer has dealt over the shared-memory banks.  Coefficients beyond the table come from L2, one
// line per lane, as before.  hot[0] is the zero slot padding points at; hot[1 + k] = c_k.  TWO: the table and the
// vector hold (c_k, v_k) pairs and both dots are formed (the paired pass of a cold Hessian-vector product,
// HessianVectorAggregator.scala:96-121).
// (Tried: looking the next quad's coefficients up one iteration ahead -- slower, 0.91 against 0.84 ms.)
// ------------------------------------------------------------------------------------------
constexpr int SELL_THREADS = 1024;

template <typename VT, bool TWO>
__global__ void __launch_bounds__(SELL_THREADS, 1)
sell_rows_kernel(int64_t n, int64_t n_slices, const int32_t *__restrict__ idx, const VT *__restrict__ val,
                 const int64_t *__restrict__ base, const int32_t *__restrict__ perm, const double *__restrict__ vec, int n_hot,
                 double *__restrict__ out, double *__restrict__ out2)
{
    extern __shared__ __align__(16) double hot[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int W = TWO ? 2 : 1;
    if (tid < W) hot[tid] = 0.0;
    for (int i = tid; i < W * n_hot; i += SELL_THREADS) hot[i + W] = vec[i];
    __syncthreads();
    const uint32_t hot_at = shared_window_address(hot);
    const double *__restrict__ vm1 = vec - W;   // stored ids are column + 1
    const int64_t stride = (int64_t)gridDim.x * (SELL_THREADS / 32);
    for (int64_t sl = (int64_t)blockIdx.x * (SELL_THREADS / 32) + warp; sl < n_slices; sl += stride) {
        const int64_t b = __ldg(base + sl);
        const int quads = (int)((__ldg(base + sl + 1) - b) >> 7);
        double acc = 0.0, acc2 = 0.0;
        SegRegs<VT, 4> cur;
        if (quads > 0) seg_load_full<VT, 4>(cur, idx, val, b + lane * 4);
        for (int q = 0; q < quads; ++q) {
            SegRegs<VT, 4> nxt;
            if (q + 1 < quads) seg_load_full<VT, 4>(nxt, idx, val, b + ((int64_t)(q + 1) * 32 + lane) * 4);
            if constexpr (TWO) {
                double2 g[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int32_t k = cur.idx[j];
                    g[j] = (uint32_t)k <= (uint32_t)n_hot ? lds_f64x2(hot_at + 16u * (uint32_t)k)
                                                           : __ldg(reinterpret_cast<const double2 *>(vm1) + k);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if constexpr (std::is_same<VT, One>::value) {
                        acc += g[j].x;
                        acc2 += g[j].y;
                    } else {
                        acc = fma(g[j].x, cur.value(j), acc);
                        acc2 = fma(g[j].y, cur.value(j), acc2);
                    }
                }
            } else {
                double g[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int32_t k = cur.idx[j];
                    g[j] = (uint32_t)k <= (uint32_t)n_hot ? lds_f64(hot_at + 8u * (uint32_t)k) : __ldg(vm1 + k);
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if constexpr (std::is_same<VT, One>::value) acc += g[j];
                    else acc = fma(g[j], cur.value(j), acc);
                }
            }
            cur = nxt;
        }
        const int64_t at = (sl << 5) + lane;
        if (at < n) {
            const int32_t row = __ldg(perm + at);
            out[row] = acc;
            if (TWO) out2[row] = acc2;
        }
    }
}

template <typename VT, bool TWO>
static int launch_sell(cudaStream_t s, int grid, size_t smem, const SellSide &sell, int64_t n, const double *vec, int n_hot,
                       double *out, double *out2)
{
    if (smem > 48 * 1024 && cudaFuncSetAttribute(sell_rows_kernel<VT, TWO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 SEG_TABLE_BYTES) != cudaSuccess)
        return -1;
    sell_rows_kernel<VT, TWO><<<grid, SELL_THREADS, smem, s>>>(n, sell.n_slices, sell.idx, static_cast<const VT *>(sell.val),
                                                               sell.base, sell.perm, vec, n_hot, out, out2);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// out2 != nullptr: vec holds (c, v) pairs and both dots are wanted
int launch_sell_rows(cudaStream_t s, int sm_count, int value_bytes, const SellSide &sell, int64_t n, const double *vec,
                     int64_t table_len, double *out, double *out2)
{
    if (sell.n_slices <= 0) return 0;
    const bool two = out2 != nullptr;
    const int64_t cap = SEG_TABLE_BYTES / (two ? 16 : 8) - 1;
    const int n_hot = (int)(table_len < cap ? table_len : cap);
    const size_t smem = (two ? 16 : 8) * (size_t)(n_hot + 1);
    const int64_t want = (sell.n_slices + SELL_THREADS / 32 - 1) / (SELL_THREADS / 32);
    const int grid = (int)(want < sm_count ? want : sm_count);
    if (value_bytes == 0)
        return two ? launch_sell<One, true>(s, grid, smem, sell, n, vec, n_hot, out, out2)
                   : launch_sell<One, false>(s, grid, smem, sell, n, vec, n_hot, out, out2);
    if (value_bytes == 4)
        return two ? launch_sell<float, true>(s, grid, smem, sell, n, vec, n_hot, out, out2)
                   : launch_sell<float, false>(s, grid, smem, sell, n, vec, n_hot, out, out2);
    return two ? launch_sell<double, true>(s, grid, smem, sell, n, vec, n_hot, out, out2)
               : launch_sell<double, false>(s, grid, smem, sell, n, vec, n_hot, out, out2);
}

// ------------------------------------------------------------------------------------------
// row_epilogue_kernel -- the scalar half of `add` for row-pass variant 1, where the sparse dots
// were produced by seg_spmv_kernel: z_i = dot_i + o_i + marginShift, pointwise loss, r_i and the
// per-CTA (F, P) partials.  Same arithmetic, in the same order, as row_pass_kernel's epilogue.
// ------------------------------------------------------------------------------------------
constexpr int E_THREADS = 256;

__global__ void __launch_bounds__(E_THREADS) row_epilogue_kernel(const RowPassParams p, const double *__restrict__ za,
                                                                 const double *__restrict__ zb)
{
    __shared__ double red[2 * (E_THREADS / 32)];
    double shift0 = 0.0, shift1 = 0.0;
    if (p.shift_partials != nullptr) {
        for (int i = threadIdx.x; i < p.n_shift_partials; i += E_THREADS) {
            shift0 += p.shift_partials[2 * i];
            shift1 += p.shift_partials[2 * i + 1];
        }
        block_sum2<E_THREADS / 32>(shift0, shift1, red);
    }
    const double margin_shift = (p.mode == MODE_HV_CACHED) ? 0.0 : -shift0;
    const double fs = (p.mode == MODE_HV_CACHED) ? shift0 : shift1;
    double accF = 0.0, accP = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * E_THREADS + threadIdx.x; i < p.n; i += (int64_t)gridDim.x * E_THREADS) {
        if (p.block_active != nullptr && !p.block_active[i >> 5]) continue;
        const double o = p.offset ? __ldcs(p.offset + i) : 0.0;
        const double wt = p.weight ? __ldcs(p.weight + i) : 1.0;
        const double dot = __ldcs(za + i);
        switch (p.mode) {
        case MODE_VG: {
            const double z = (dot + o) + margin_shift;
            double l, dl;
            loss_and_dz(p.loss, z, __ldcs(p.label + i), l, dl);
            const double wdl = __dmul_rn(wt, dl);
            p.r[i] = wdl;
            const double wl = __dmul_rn(wt, l);
            if (p.row_loss) p.row_loss[i] = wl;
            accF += wl;
            accP += wdl;
            break;
        }
        case MODE_HV_COLD: {
            const double z = (dot + o) + margin_shift;
            const double dq = __dmul_rn(wt, dzz_loss(p.loss, z));
            if (p.dcache) p.dcache[i] = dq;
            const double e = __dmul_rn(dq, __ldcs(zb + i) - fs);
            p.r[i] = e;
            accP += e;
            break;
        }
        case MODE_HV_CACHED: {
            const double e = __dmul_rn(__ldcs(p.dcache + i), dot - fs);
            p.r[i] = e;
            accP += e;
            break;
        }
        case MODE_DIAG:
            p.r[i] = __dmul_rn(wt, dzz_loss(p.loss, dot + o));
            break;
        default:
            p.r[i] = dot;
            break;
        }
    }
    if (p.partials != nullptr) {
        block_sum2<E_THREADS / 32>(accF, accP, red);
        if (threadIdx.x == 0) {
            p.partials[2 * blockIdx.x] = accF;
            p.partials[2 * blockIdx.x + 1] = accP;
        }
    }
}

int launch_row_epilogue(cudaStream_t s, int sm_count, const RowPassParams &p, const double *za, const double *zb,
                        int *grid_out)
{
    int64_t want = (p.n + E_THREADS - 1) / E_THREADS;
    int64_t cap = (int64_t)sm_count * 8;
    if (cap > MAX_PARTIALS) cap = MAX_PARTIALS;
    int grid = (int)(want < cap ? want : cap);
    if (grid < 1) grid = 1;
    *grid_out = grid;
    row_epilogue_kernel<<<grid, E_THREADS, 0, s>>>(p, za, zb);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// fixup_kernel -- the `merge` of partial aggregates (ValueAndGradientAggregator.scala:161-176)
// as far as one GPU is concerned: columns cut by chunk boundaries get their heads added (one
// warp per column, fixed order), and the last CTA adds up the row kernel's per-CTA (F, P).
// ------------------------------------------------------------------------------------------
constexpr int F_THREADS = 256;

__global__ void __launch_bounds__(F_THREADS)
fixup_kernel(const FixEntry *__restrict__ fix, int n_fix, int n_long, const double *__restrict__ carry_head,
             double *__restrict__ out, const double *__restrict__ carry_head2, double *__restrict__ out2, int d,
             const double *__restrict__ partials, int n_partials, int write_value, int write_prefactor)
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
    // The list is ordered long runs first (pack time).  A run spanning >= 32 chunks gets a warp (lanes
    // stride over its heads, fixed shuffle tree); short ones -- the bulk, usually one head -- get a thread.
    const int long_blocks = (n_long + (F_THREADS / 32) - 1) / (F_THREADS / 32);
    if ((int)blockIdx.x < long_blocks) {
        const int q = blockIdx.x * (F_THREADS / 32) + warp;
        if (q >= n_long) return;
        const FixEntry e = fix[q];
        double s = 0.0, s2 = 0.0;
        for (int cc = e.chunk_begin + lane; cc < e.chunk_end; cc += 32) {
            s += carry_head[cc];
            if (carry_head2) s2 += carry_head2[cc];
        }
        s = warp_sum(s);
        s2 = warp_sum(s2);
        if (lane == 0) {
            out[e.col] += s;
            if (carry_head2) out2[e.col] += s2;
        }
        return;
    }
    const int q = n_long + ((int)blockIdx.x - long_blocks) * F_THREADS + threadIdx.x;
    if (q >= n_fix) return;
    const FixEntry e = fix[q];
    double s = 0.0, s2 = 0.0;
    for (int cc = e.chunk_begin; cc < e.chunk_end; ++cc) {
        s += carry_head[cc];
        if (carry_head2) s2 += carry_head2[cc];
    }
    out[e.col] += s;
    if (carry_head2) out2[e.col] += s2;
}

int launch_fixup(cudaStream_t s, const FixEntry *fix, int n_fix, int n_long, const double *carry_head, double *out,
                 const double *carry_head2, double *out2, int d, const double *partials, int n_partials,
                 bool write_value, bool write_prefactor)
{
    const int long_blocks = (n_long + (F_THREADS / 32) - 1) / (F_THREADS / 32);
    const int short_blocks = (n_fix - n_long + F_THREADS - 1) / F_THREADS;
    const int grid = long_blocks + short_blocks + (partials ? 1 : 0);
    if (grid == 0) return 0;
    fixup_kernel<<<grid, F_THREADS, 0, s>>>(fix, n_fix, n_long, carry_head, out, carry_head2, out2, d, partials, n_partials,
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

// ------------------------------------------------------------------------------------------
// Which chunks of the two entry streams, and which 32-row blocks, still hold rows of an ACTIVE entity of a batch
// (pml_batch_set_active).  An entity's rows are contiguous, and so are its columns in the block-diagonal data set, so
// a chunk touches a contiguous range of entities: found by binary search, one thread per chunk / block.  In the
// reference a converged entity's solve simply ends (RandomEffectCoordinate.scala:112-119 ->
// SingleNodeOptimizationProblem.scala:85-103); here its rows stop being streamed.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t last_not_above_u32(const uint32_t *__restrict__ a, int64_t n, uint32_t key)
{
    int64_t lo = 0, hi = n;  // largest i in [0, n] with a[i] <= key (a[0] = 0)
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) >> 1;
        if (a[mid] <= key) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ int64_t last_not_above_i64(const int64_t *__restrict__ a, int64_t n, int64_t key)
{
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) >> 1;
        if (a[mid] <= key) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ unsigned char any_active(const unsigned char *__restrict__ active, int64_t e0, int64_t e1)
{
    for (int64_t e = e0; e <= e1; ++e)
        if (active[e]) return 1;
    return 0;
}

__global__ void __launch_bounds__(256)
row_flags_kernel(int64_t n_entities, const unsigned char *__restrict__ active, const int64_t *__restrict__ ent_row_ptr,
                 int64_t n, int64_t nnz, const uint32_t *__restrict__ row_ptr, int n_chunks,
                 unsigned char *__restrict__ chunk_active, unsigned char *__restrict__ block_active)
{
    const int64_t n_blocks = (n + 31) >> 5;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_chunks + n_blocks;
         q += (int64_t)gridDim.x * blockDim.x) {
        int64_t r0, r1;
        if (q < n_chunks) {
            const int64_t k0 = q * SEG_CHUNK, k1 = (k0 + SEG_CHUNK < nnz ? k0 + SEG_CHUNK : nnz) - 1;
            r0 = last_not_above_u32(row_ptr, n, (uint32_t)k0);
            r1 = last_not_above_u32(row_ptr, n, (uint32_t)k1);
        } else {
            r0 = (q - n_chunks) << 5;
            r1 = r0 + 31 < n ? r0 + 31 : n - 1;
        }
        if (r0 >= n) r0 = n - 1;
        if (r1 >= n) r1 = n - 1;
        const int64_t e0 = last_not_above_i64(ent_row_ptr, n_entities, r0);
        const int64_t e1 = last_not_above_i64(ent_row_ptr, n_entities, r1);
        const unsigned char a = any_active(active, e0 < n_entities ? e0 : n_entities - 1, e1 < n_entities ? e1 : n_entities - 1);
        if (q < n_chunks) chunk_active[q] = a;
        else block_active[q - n_chunks] = a;
    }
}

__global__ void __launch_bounds__(256)
col_flags_kernel(const unsigned char *__restrict__ active, int64_t m, const uint32_t *__restrict__ nzcol_ptr,
                 const int32_t *__restrict__ nzcol_id, const int32_t *__restrict__ ent_of_coef, int64_t nnz, int n_chunks,
                 unsigned char *__restrict__ chunk_active)
{
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_chunks; q += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k0 = q * SEG_CHUNK, k1 = (k0 + SEG_CHUNK < nnz ? k0 + SEG_CHUNK : nnz) - 1;
        int64_t j0 = last_not_above_u32(nzcol_ptr, m, (uint32_t)k0);
        int64_t j1 = last_not_above_u32(nzcol_ptr, m, (uint32_t)k1);
        if (j0 >= m) j0 = m - 1;
        if (j1 >= m) j1 = m - 1;
        // the batch keeps the caller's column order, so internal ids are monotone in the entity
        chunk_active[q] = any_active(active, ent_of_coef[nzcol_id[j0]], ent_of_coef[nzcol_id[j1]]);
    }
}

int launch_active_flags(cudaStream_t s, int64_t n_entities, const unsigned char *active, const int64_t *ent_row_ptr,
                        int64_t n, int64_t nnz, const uint32_t *row_ptr, int n_row_chunks, unsigned char *row_chunk_active,
                        unsigned char *row_block_active, int64_t m, const uint32_t *nzcol_ptr, const int32_t *nzcol_id,
                        const int32_t *ent_of_coef, int64_t col_nnz, int n_col_chunks, unsigned char *col_chunk_active)
{
    if (n_entities <= 0 || n <= 0) return 0;
    int launched = 0;
    {
        int64_t grid = (n_row_chunks + ((n + 31) >> 5) + 255) / 256;
        if (grid > 148 * 16) grid = 148 * 16;
        row_flags_kernel<<<(int)grid, 256, 0, s>>>(n_entities, active, ent_row_ptr, n, nnz, row_ptr, n_row_chunks,
                                                   row_chunk_active, row_block_active);
        ++launched;
    }
    if (n_col_chunks > 0 && m > 0) {
        int64_t grid = (n_col_chunks + 255) / 256;
        if (grid > 148 * 16) grid = 148 * 16;
        col_flags_kernel<<<(int)grid, 256, 0, s>>>(active, m, nzcol_ptr, nzcol_id, ent_of_coef, col_nnz, n_col_chunks,
                                                   col_chunk_active);
        ++launched;
    }
    return cudaGetLastError() == cudaSuccess ? launched : -1;
}

// Entities taken out of a batched evaluation (pml_batch_set_active): their value and their slice of the
// concatenated vector come back as 0.  One warp per entity.
__global__ void __launch_bounds__(256)
mask_entities_kernel(int64_t n_entities, const int64_t *__restrict__ coef_ptr, const uint8_t *__restrict__ active,
                     double *__restrict__ values, double *__restrict__ vec)
{
    const int lane = threadIdx.x & 31;
    int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; e < n_entities; e += nw) {
        if (active[e]) continue;
        if (values && lane == 0) values[e] = 0.0;
        if (vec)
            for (int64_t j = coef_ptr[e] + lane; j < coef_ptr[e + 1]; j += 32) vec[j] = 0.0;
    }
}

int launch_mask_entities(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const uint8_t *active, double *values,
                         double *vec)
{
    if (n_entities <= 0 || active == nullptr || (values == nullptr && vec == nullptr)) return 0;
    int64_t grid = (n_entities * 32 + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    mask_entities_kernel<<<(int)grid, 256, 0, s>>>(n_entities, coef_ptr, active, values, vec);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// Batched per-entity problems (SingleNodeGLMLossFunction over RandomEffectDataSet.activeData,
// api/algorithm/RandomEffectCoordinate.scala:108-119).  The E local problems are evaluated as ONE
// block-diagonal problem: entity e's local column j becomes global column coef_ptr[e] + j, so
// the row/column passes above produce every entity's gradient in one go; what is left is the
// per-entity objective value, a segmented sum of the per-row wt*l over each entity's rows.
// ------------------------------------------------------------------------------------------
__global__ void globalize_cols_kernel(int64_t n_entities, const int64_t *__restrict__ ent_row_ptr,
                                      const int64_t *__restrict__ coef_ptr, const int32_t *__restrict__ ent_dim,
                                      const int64_t *__restrict__ row_ptr, int32_t *__restrict__ col,
                                      unsigned long long *__restrict__ bad)
{
    // one warp per entity; lanes stride over the entity's entries
    const int lane = threadIdx.x & 31;
    int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; e < n_entities; e += nw) {
        const int64_t a = row_ptr[ent_row_ptr[e]], b = row_ptr[ent_row_ptr[e + 1]];
        const int64_t shift = coef_ptr[e];
        const int32_t dim = ent_dim[e];
        int wrong = 0;
        for (int64_t k = a + lane; k < b; k += 32) {
            const int32_t c = col[k];
            if (c < 0 || c >= dim) ++wrong;
            col[k] = (int32_t)(shift + c);
        }
        if (wrong) atomicAdd(bad, (unsigned long long)wrong);
    }
}

int launch_globalize_cols(cudaStream_t s, int64_t n_entities, const int64_t *ent_row_ptr, const int64_t *coef_ptr,
                          const int32_t *ent_dim, const int64_t *row_ptr, int32_t *col, unsigned long long *bad)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities * 32 + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    globalize_cols_kernel<<<(int)grid, 256, 0, s>>>(n_entities, ent_row_ptr, coef_ptr, ent_dim, row_ptr, col, bad);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// value_e = sum of the per-row wt*l over the entity's rows.  A warp takes 32 entities: a lane adds up its
// own entity when it is small (the common case: a handful of rows), and entities with >= 64 rows are then
// summed by the whole warp, lanes striding rows, with a fixed shuffle tree.  Fixed order either way.
__global__ void entity_values_kernel(int64_t n_entities, const int64_t *__restrict__ ent_row_ptr,
                                     const double *__restrict__ row_loss, double *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    int64_t e0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) << 5;
    const int64_t stride = (((int64_t)gridDim.x * blockDim.x) >> 5) << 5;
    for (; e0 < n_entities; e0 += stride) {
        const int64_t e = e0 + lane;
        int64_t lo = 0, hi = 0;
        if (e < n_entities) {
            lo = ent_row_ptr[e];
            hi = ent_row_ptr[e + 1];
        }
        const bool big = hi - lo >= 64;
        if (e < n_entities && !big) {
            double s = 0.0;
            for (int64_t i = lo; i < hi; ++i) s += row_loss[i];
            out[e] = s;
        }
        unsigned todo = __ballot_sync(0xffffffffu, big);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t blo = __shfl_sync(0xffffffffu, lo, src), bhi = __shfl_sync(0xffffffffu, hi, src);
            double s = 0.0;
            for (int64_t i = blo + lane; i < bhi; i += 32) s += row_loss[i];
            s = warp_sum(s);
            if (lane == 0) out[e0 + src] = s;
        }
    }
}

int launch_entity_values(cudaStream_t s, int64_t n_entities, const int64_t *ent_row_ptr, const double *row_loss,
                         double *out)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    entity_values_kernel<<<(int)grid, 256, 0, s>>>(n_entities, ent_row_ptr, row_loss, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// Vector algebra of a lock-step batched optimizer over the concatenated per-entity vectors: per-entity dot
// products and per-entity scaled sums (what Breeze's dot / axpy do inside TRON.truncatedCG and LBFGS for ONE
// entity, lib/optimization/TRON.scala:279-339, for all entities at once).  Fixed summation order per entity.
// ------------------------------------------------------------------------------------------
__global__ void entity_of_coef_kernel(int64_t n_entities, const int64_t *__restrict__ coef_ptr, int32_t *__restrict__ ent_of)
{
    const int lane = threadIdx.x & 31;
    int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; e < n_entities; e += nw)
        for (int64_t j = coef_ptr[e] + lane; j < coef_ptr[e + 1]; j += 32) ent_of[j] = (int32_t)e;
}

int launch_entity_of_coef(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, int32_t *ent_of)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities * 32 + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    entity_of_coef_kernel<<<(int)grid, 256, 0, s>>>(n_entities, coef_ptr, ent_of);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// out[e] = sum over entity e's coefficients of x_j * y_j.  Same shape as entity_values_kernel: a lane per small
// entity, the whole warp (fixed shuffle tree) for entities of 64 coefficients and more.
__global__ void segment_dot_kernel(int64_t n_entities, const int64_t *__restrict__ coef_ptr, const double *__restrict__ x,
                                   const double *__restrict__ y, double *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    int64_t e0 = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) << 5;
    const int64_t stride = (((int64_t)gridDim.x * blockDim.x) >> 5) << 5;
    for (; e0 < n_entities; e0 += stride) {
        const int64_t e = e0 + lane;
        int64_t lo = 0, hi = 0;
        if (e < n_entities) {
            lo = coef_ptr[e];
            hi = coef_ptr[e + 1];
        }
        const bool big = hi - lo >= 64;
        if (e < n_entities && !big) {
            double s = 0.0;
            for (int64_t j = lo; j < hi; ++j) s = fma(x[j], y[j], s);
            out[e] = s;
        }
        unsigned todo = __ballot_sync(0xffffffffu, big);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t blo = __shfl_sync(0xffffffffu, lo, src), bhi = __shfl_sync(0xffffffffu, hi, src);
            double s = 0.0;
            for (int64_t j = blo + lane; j < bhi; j += 32) s = fma(x[j], y[j], s);
            s = warp_sum(s);
            if (lane == 0) out[e0 + src] = s;
        }
    }
}

int launch_segment_dot(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const double *x, const double *y,
                       double *out)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    segment_dot_kernel<<<(int)grid, 256, 0, s>>>(n_entities, coef_ptr, x, y, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// out_j = a[e(j)] * x_j + c[e(j)] * y_j   (a or c null = 1; y null = no second term); out may alias x or y.
// active (may be null): entities with active[e] == 0 keep out_j as it is.
__global__ void segment_axpby_kernel(int64_t d, const int32_t *__restrict__ ent_of, const double *__restrict__ a,
                                     const double *x, const double *__restrict__ c, const double *y,
                                     const uint8_t *__restrict__ active, double *out)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x) {
        const int32_t e = ent_of[j];
        if (active != nullptr && !active[e]) continue;
        double v = a ? __dmul_rn(a[e], x[j]) : x[j];
        if (y != nullptr) v = __dadd_rn(v, c ? __dmul_rn(c[e], y[j]) : y[j]);
        out[j] = v;
    }
}

int launch_segment_axpby(cudaStream_t s, int64_t d, const int32_t *ent_of, const double *a, const double *x,
                         const double *c, const double *y, const uint8_t *active, double *out)
{
    if (d <= 0) return 0;
    int64_t grid = (d + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    segment_axpby_kernel<<<(int)grid, 256, 0, s>>>(d, ent_of, a, x, c, y, active, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// One iteration of TRON.truncatedCG (lib/optimization/TRON.scala:300-336) for every entity that is still in its CG
// loop, after the batched Hessian-vector product Hd = H direction has been computed: the whole per-entity vector
// algebra of the iteration -- alpha, the step, the trust-region boundary case, the residual, beta, the new direction
// -- in ONE pass over the concatenated vectors instead of three dots and three to five axpy launches, each of which
// streams every entity's coefficients.  A lane owns an entity (per-entity vectors are a few dozen coefficients; the
// slices of a warp's 32 entities stay in L1 between the passes).  The arithmetic is the lock-step driver's, rounding
// for rounding: dots are sequential fused multiply-adds in index order (segment_dot_kernel), every update is one
// rounded product plus one rounded sum (segment_axpby_kernel), scalar expressions are evaluated without contraction.
// The L2 term of L2RegularizationTwiceDiff (Hd += l2 * direction, lib/function/L2Regularization.scala:151-157,176)
// is applied on the fly.
// ------------------------------------------------------------------------------------------
// The per-entity work of one CG iteration on the slices [lo, hi) of the four vectors (pointers may be global or
// shared memory).  Returns 1 when the step left the trust region.
__device__ __forceinline__ int cg_step_entity(int64_t lo, int64_t hi, double l2, double *direction, const double *hd,
                                              double *step, double *residual, double *rtr_e, double dl)
{
    const bool reg = l2 != 0.0;
    // alpha = rTr / (direction . Hd)
    double dhd = 0.0;
    for (int64_t j = lo; j < hi; ++j) {
        const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
        dhd = fma(direction[j], h, dhd);
    }
    const double r0 = *rtr_e;
    const double alpha = r0 / dhd;
    // step += alpha * direction; is it still inside the trust region?
    double sts = 0.0;
    for (int64_t j = lo; j < hi; ++j) {
        const double sj = __dadd_rn(__dmul_rn(alpha, direction[j]), step[j]);
        step[j] = sj;
        sts = fma(sj, sj, sts);
    }
    if (sqrt(sts) > dl) {
        // back to the boundary along `direction`, then stop (TRON.scala:311-327)
        double std_ = 0.0, sts2 = 0.0, dtd = 0.0;
        const double na = -alpha;
        for (int64_t j = lo; j < hi; ++j) {
            const double sj = __dadd_rn(__dmul_rn(na, direction[j]), step[j]);
            step[j] = sj;
            std_ = fma(sj, direction[j], std_);
            sts2 = fma(sj, sj, sts2);
            dtd = fma(direction[j], direction[j], dtd);
        }
        const double dsq = __dmul_rn(dl, dl);
        const double rad = sqrt(__dadd_rn(__dmul_rn(std_, std_), __dmul_rn(dtd, __dsub_rn(dsq, sts2))));
        const double alpha_b = std_ >= 0.0 ? __dsub_rn(dsq, sts2) / __dadd_rn(std_, rad) : __dsub_rn(rad, std_) / dtd;
        const double nab = -alpha_b;
        for (int64_t j = lo; j < hi; ++j) {
            const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
            step[j] = __dadd_rn(__dmul_rn(alpha_b, direction[j]), step[j]);
            residual[j] = __dadd_rn(__dmul_rn(nab, h), residual[j]);
        }
        return 1;
    }
    const double na = -alpha;
    double rnew = 0.0;
    for (int64_t j = lo; j < hi; ++j) {
        const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
        const double rj = __dadd_rn(__dmul_rn(na, h), residual[j]);
        residual[j] = rj;
        rnew = fma(rj, rj, rnew);
    }
    const double beta = rnew / r0;
    for (int64_t j = lo; j < hi; ++j) direction[j] = __dadd_rn(__dmul_rn(beta, direction[j]), residual[j]);
    *rtr_e = rnew;
    return 0;
}

// The same for an entity of 64 coefficients and more, by the whole warp: lanes stride the slice (coalesced without any
// staging), dots are per-lane fused multiply-adds closed by the fixed shuffle tree -- segment_dot_kernel's order for
// such entities.  Every lane returns the same flag.
__device__ __forceinline__ int cg_step_entity_warp(int lane, int64_t lo, int64_t hi, double l2, double *direction,
                                                   const double *hd, double *step, double *residual, double *rtr_e, double dl)
{
    const bool reg = l2 != 0.0;
    double dhd = 0.0;
    for (int64_t j = lo + lane; j < hi; j += 32) {
        const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
        dhd = fma(direction[j], h, dhd);
    }
    dhd = warp_sum(dhd);
    const double r0 = *rtr_e;
    const double alpha = r0 / dhd;
    double sts = 0.0;
    for (int64_t j = lo + lane; j < hi; j += 32) {
        const double sj = __dadd_rn(__dmul_rn(alpha, direction[j]), step[j]);
        step[j] = sj;
        sts = fma(sj, sj, sts);
    }
    sts = warp_sum(sts);
    if (sqrt(sts) > dl) {
        double std_ = 0.0, sts2 = 0.0, dtd = 0.0;
        const double na = -alpha;
        for (int64_t j = lo + lane; j < hi; j += 32) {
            const double sj = __dadd_rn(__dmul_rn(na, direction[j]), step[j]);
            step[j] = sj;
            std_ = fma(sj, direction[j], std_);
            sts2 = fma(sj, sj, sts2);
            dtd = fma(direction[j], direction[j], dtd);
        }
        std_ = warp_sum(std_);
        sts2 = warp_sum(sts2);
        dtd = warp_sum(dtd);
        const double dsq = __dmul_rn(dl, dl);
        const double rad = sqrt(__dadd_rn(__dmul_rn(std_, std_), __dmul_rn(dtd, __dsub_rn(dsq, sts2))));
        const double alpha_b = std_ >= 0.0 ? __dsub_rn(dsq, sts2) / __dadd_rn(std_, rad) : __dsub_rn(rad, std_) / dtd;
        const double nab = -alpha_b;
        for (int64_t j = lo + lane; j < hi; j += 32) {
            const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
            step[j] = __dadd_rn(__dmul_rn(alpha_b, direction[j]), step[j]);
            residual[j] = __dadd_rn(__dmul_rn(nab, h), residual[j]);
        }
        return 1;
    }
    const double na = -alpha;
    double rnew = 0.0;
    for (int64_t j = lo + lane; j < hi; j += 32) {
        const double h = reg ? __dadd_rn(__dmul_rn(l2, direction[j]), hd[j]) : hd[j];
        const double rj = __dadd_rn(__dmul_rn(na, h), residual[j]);
        residual[j] = rj;
        rnew = fma(rj, rj, rnew);
    }
    rnew = warp_sum(rnew);
    const double beta = rnew / r0;
    for (int64_t j = lo + lane; j < hi; j += 32) direction[j] = __dadd_rn(__dmul_rn(beta, direction[j]), residual[j]);
    *rtr_e = rnew;
    return 0;
}

// A warp walks 32 consecutive entities.  Their slices are contiguous in the concatenated vectors, so the warp copies
// as many whole entities as fit its staging area to shared memory with coalesced loads, every lane then works on ITS
// entity there, and the three vectors that change go back coalesced; left to walk global memory on its own a lane
// makes every load a wavefront of its own (measured: slower than the unfused loop).  Entities of 64 coefficients and
// more are then done by the whole warp straight from global memory.
#ifndef PML_CG_WARPS
#define PML_CG_WARPS 28
#endif
#ifndef PML_CG_CAP
#define PML_CG_CAP 256
#endif
constexpr int CG_WARPS = PML_CG_WARPS;
constexpr int CG_CAP = PML_CG_CAP;  // doubles per vector and warp: 4 vectors x 2 KB x 28 warps = 224 KB, one CTA per SM.
// (The kernel is latency-bound -- a warp loads, waits, works, stores -- so what counts is warps in flight, not how many
// entities a warp takes at once: 7 warps x 1024 doubles ran 0.36 ms per launch on 10^6 entities, 18 x 384 0.20 ms, 28 x 256 0.17 ms.)
constexpr int CG_BIG = 64;    // segment_dot_kernel's threshold

__global__ void __launch_bounds__(32 * CG_WARPS)
batch_cg_step_kernel(int64_t n_entities, const int64_t *__restrict__ coef_ptr, const uint8_t *__restrict__ run, double l2,
                     double *direction, const double *hd, double *step, double *residual, double *__restrict__ rtr,
                     const double *__restrict__ delta, uint8_t *__restrict__ over)
{
    extern __shared__ __align__(16) double cg_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *s_dir = cg_smem + (size_t)warp * 4 * CG_CAP, *s_hd = s_dir + CG_CAP, *s_step = s_hd + CG_CAP, *s_res = s_step + CG_CAP;
    const int64_t n_groups = (n_entities + 31) >> 5;
    for (int64_t grp = (int64_t)blockIdx.x * CG_WARPS + warp; grp < n_groups; grp += (int64_t)gridDim.x * CG_WARPS) {
        const int64_t e0 = grp << 5;
        const int64_t e_end = e0 + 32 < n_entities ? e0 + 32 : n_entities;
        {
            const int64_t e = e0 + lane;
            const bool mine = e < e_end && run[e] != 0;
            if (e < e_end && !mine) over[e] = 0;
            if (!__any_sync(0xffffffffu, mine)) continue;
        }
        int64_t es = e0;
        while (es < e_end) {                                   // warp-uniform
            const int64_t e = es + lane;
            const bool valid = e < e_end;
            const int64_t p0 = coef_ptr[es];
            const int64_t lo = valid ? coef_ptr[e] : p0, hi = valid ? coef_ptr[e + 1] : p0;
            // how many whole entities from `es` on fit the staging area (a prefix of the lanes)
            const unsigned fits = __ballot_sync(0xffffffffu, valid && hi - p0 <= CG_CAP);
            int take = __ffs(~fits) - 1;                       // leading ones; fits == all ones gives -1 + 0 ...
            if (fits == 0xffffffffu) take = 32;
            const bool mine = valid && run[e] != 0;
            const bool big = hi - lo >= CG_BIG;
            double r = 0.0, dl = 0.0;
            if (mine) {
                r = rtr[e];
                dl = delta[e];
            }
            int took = take;
            if (take > 0) {
                const unsigned small = __ballot_sync(0xffffffffu, lane < take && mine && !big);
                if (small != 0) {
                    const int64_t p1 = __shfl_sync(0xffffffffu, hi, take - 1);
                    const int len = (int)(p1 - p0);
                    for (int k = lane; k < len; k += 32) {
                        s_dir[k] = direction[p0 + k];
                        s_hd[k] = hd[p0 + k];
                        s_step[k] = step[p0 + k];
                        s_res[k] = residual[p0 + k];
                    }
                    __syncwarp();
                    if (lane < take && mine && !big) {
                        over[e] = (uint8_t)cg_step_entity(lo - p0, hi - p0, l2, s_dir, s_hd, s_step, s_res, &r, dl);
                        rtr[e] = r;
                    }
                    __syncwarp();
                    // entities of the range that are masked or left to the warp get their own (unchanged) values back:
                    // harmless, and it keeps the stores coalesced
                    for (int k = lane; k < len; k += 32) {
                        direction[p0 + k] = s_dir[k];
                        step[p0 + k] = s_step[k];
                        residual[p0 + k] = s_res[k];
                    }
                    __syncwarp();
                }
            } else {
                took = 1;                                      // entity `es` alone is larger than the staging area
            }
            unsigned todo = __ballot_sync(0xffffffffu, lane < took && mine && big);
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const int64_t blo = __shfl_sync(0xffffffffu, lo, src), bhi = __shfl_sync(0xffffffffu, hi, src);
                double br = __shfl_sync(0xffffffffu, r, src);
                const double bdl = __shfl_sync(0xffffffffu, dl, src);
                const int flag = cg_step_entity_warp(lane, blo, bhi, l2, direction, hd, step, residual, &br, bdl);
                if (lane == 0) {
                    over[es + src] = (uint8_t)flag;
                    rtr[es + src] = br;
                }
                __syncwarp();
            }
            es += took;
        }
    }
}

int launch_batch_cg_step(cudaStream_t s, int64_t n_entities, const int64_t *coef_ptr, const uint8_t *run, double l2,
                         double *direction, const double *hd, double *step, double *residual, double *rtr,
                         const double *delta, uint8_t *over)
{
    if (n_entities <= 0) return 0;
    const size_t smem = sizeof(double) * 4 * CG_CAP * CG_WARPS;
    if (cudaFuncSetAttribute(batch_cg_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return -1;   // (per device; a few hundred nanoseconds)
    int64_t grid = ((n_entities + 31) / 32 + CG_WARPS - 1) / CG_WARPS;
    if (grid > 148) grid = 148;
    batch_cg_step_kernel<<<(int)grid, 32 * CG_WARPS, smem, s>>>(n_entities, coef_ptr, run, l2, direction, hd, step, residual, rtr,
                                                                 delta, over);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// ------------------------------------------------------------------------------------------
// L-BFGS two-loop recursion for every entity of a batch in ONE kernel (Breeze's LBFGS direction as the lock-step driver
// states it, batched_lbfgs.BatchedLBFGS._direction; lib/optimization/LBFGS.scala:59-107 hands the problem to Breeze):
//   q = -g;  newest pair to oldest: a_i = rho_i (s_i . q), q -= a_i y_i;  q *= (s.y / y.y) of the newest pair;
//   oldest to newest: b = rho_i (y_i . q), q += (a_i - b) s_i.
// The unfused loop is 2m dots and 2m + 1 axpys, each a launch and a pass over every entity's coefficients.  The
// correction pairs live in 2 x m concatenated vectors (slot-major); every entity has its own ring position: head[e] is
// the slot of its newest pair, len[e] how many it holds.  Same rounding as the unfused loop: dots are sequential fused
// multiply-adds in index order below 64 coefficients and lane-strided with the fixed shuffle tree from 64 on
// (segment_dot_kernel), every update is one rounded product plus one rounded sum (segment_axpby_kernel).
// ------------------------------------------------------------------------------------------
#ifndef PML_LB_WARPS
#define PML_LB_WARPS 32
#endif
#ifndef PML_LB_CAP
#define PML_LB_CAP 256
#endif
constexpr int LB_WARPS = PML_LB_WARPS;
constexpr int LB_CAP = PML_LB_CAP;   // doubles per staged vector and warp: 3 x 2 KB + 0.5 KB of ring positions, 32 warps = 208 KB
// (latency-bound like batch_cg_step_kernel, once per pair and loop: 8 warps x 1024 doubles took 1.47 ms per launch on
// 10^6 entities, 32 x 256 0.73 ms; a warp then takes the 32 entities of its group in passes of about a dozen)
constexpr int LB_MAX_M = 16;

struct LbfgsParams {
    int64_t n_entities, d_total;
    const int64_t *coef_ptr;
    const uint8_t *go;
    int m;
    const double *S, *Y, *rho;   // m x d_total, m x d_total, m x n_entities
    const int32_t *head, *len;
    const double *grad;
    double *out;
};

__device__ __forceinline__ int lb_slot(int head, int t, int m)
{
    const int s = head - t;
    return s < 0 ? s + m : s;
}

// an entity of 64 coefficients and more, by the whole warp straight from global memory; q lives in `out`
__device__ __forceinline__ void lbfgs_direction_warp(const LbfgsParams &p, int lane, int64_t e, int64_t lo, int64_t hi)
{
    double *q = p.out;
    for (int64_t j = lo + lane; j < hi; j += 32) q[j] = -p.grad[j];
    const int n = p.len[e], head = p.head[e];
    double alpha[LB_MAX_M];
    double scale = 1.0;
#pragma unroll 1
    for (int t = 0; t < n; ++t) {
        const int slot = lb_slot(head, t, p.m);
        const double *s = p.S + (int64_t)slot * p.d_total, *y = p.Y + (int64_t)slot * p.d_total;
        double sq = 0.0;
        for (int64_t j = lo + lane; j < hi; j += 32) sq = fma(s[j], q[j], sq);
        sq = warp_sum(sq);
        const double a = __dmul_rn(p.rho[(int64_t)slot * p.n_entities + e], sq);
        alpha[t] = a;
        const double na = -a;
        if (t == 0) {
            double sy = 0.0, yy = 0.0;
            for (int64_t j = lo + lane; j < hi; j += 32) {
                sy = fma(s[j], y[j], sy);
                yy = fma(y[j], y[j], yy);
            }
            scale = warp_sum(sy) / warp_sum(yy);
        }
        for (int64_t j = lo + lane; j < hi; j += 32) q[j] = __dadd_rn(q[j], __dmul_rn(na, y[j]));
    }
    for (int64_t j = lo + lane; j < hi; j += 32) q[j] = __dmul_rn(scale, q[j]);
#pragma unroll 1
    for (int t = n - 1; t >= 0; --t) {
        const int slot = lb_slot(head, t, p.m);
        const double *s = p.S + (int64_t)slot * p.d_total, *y = p.Y + (int64_t)slot * p.d_total;
        double yq = 0.0;
        for (int64_t j = lo + lane; j < hi; j += 32) yq = fma(y[j], q[j], yq);
        yq = warp_sum(yq);
        const double c = __dsub_rn(alpha[t], __dmul_rn(p.rho[(int64_t)slot * p.n_entities + e], yq));
        for (int64_t j = lo + lane; j < hi; j += 32) q[j] = __dadd_rn(q[j], __dmul_rn(c, s[j]));
    }
}

// Stage pair t (0 = newest) of the entities of a staged range: element k of the range belongs to an entity whose ring
// position is in pos[k]; elements of entities that hold fewer than t + 1 pairs (or are not worked on) are left alone.
__device__ __forceinline__ void lb_stage(const LbfgsParams &p, int lane, int t, int64_t p0, int n, const uint16_t *pos,
                                         double *sA, double *sB)
{
    for (int k = lane; k < n; k += 32) {
        const unsigned w = pos[k];
        if (t < (int)((w >> 8) & 0x7fu)) {
            const int64_t at = (int64_t)lb_slot((int)(w & 0xffu), t, p.m) * p.d_total + p0 + k;
            sA[k] = p.S[at];
            sB[k] = p.Y[at];
        }
    }
}

// A warp walks 32 consecutive entities as batch_cg_step_kernel does: the gradient slices of as many whole entities as
// fit go to shared memory (q), then pair by pair the s and y slices are staged with coalesced loads and every lane
// works on ITS entity there; q goes back coalesced.  Entities of 64 coefficients and more: the whole warp, from global.
__global__ void __launch_bounds__(32 * LB_WARPS) batch_lbfgs_direction_kernel(const LbfgsParams p)
{
    extern __shared__ __align__(16) double lb_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *sQ = lb_smem + (size_t)warp * 3 * LB_CAP, *sA = sQ + LB_CAP, *sB = sA + LB_CAP;
    uint16_t *sPos = reinterpret_cast<uint16_t *>(lb_smem + (size_t)LB_WARPS * 3 * LB_CAP) + (size_t)warp * LB_CAP;
    const int64_t n_groups = (p.n_entities + 31) >> 5;
    for (int64_t grp = (int64_t)blockIdx.x * LB_WARPS + warp; grp < n_groups; grp += (int64_t)gridDim.x * LB_WARPS) {
        const int64_t e0 = grp << 5;
        const int64_t e_end = e0 + 32 < p.n_entities ? e0 + 32 : p.n_entities;
        {
            const int64_t e = e0 + lane;
            if (!__any_sync(0xffffffffu, e < e_end && (p.go == nullptr || p.go[e] != 0))) continue;
        }
        int64_t es = e0;
        while (es < e_end) {                                   // warp-uniform
            const int64_t e = es + lane;
            const bool valid = e < e_end;
            const int64_t p0 = p.coef_ptr[es];
            const int64_t lo = valid ? p.coef_ptr[e] : p0, hi = valid ? p.coef_ptr[e + 1] : p0;
            const unsigned fits = __ballot_sync(0xffffffffu, valid && hi - p0 <= LB_CAP);
            int take = fits == 0xffffffffu ? 32 : __ffs(~fits) - 1;   // whole entities from `es` on that fit the staging area
            const bool mine = valid && (p.go == nullptr || p.go[e] != 0);
            const bool big = hi - lo >= CG_BIG;
            int took = take;
            if (take > 0) {
                const bool small = lane < take && mine && !big;
                if (__any_sync(0xffffffffu, small)) {
                    const int n = (int)(__shfl_sync(0xffffffffu, hi, take - 1) - p0);
                    const int my_len = small ? p.len[e] : 0, my_head = small ? p.head[e] : 0;
                    const int a0 = (int)(lo - p0), a1 = (int)(hi - p0);
                    for (int k = lane; k < n; k += 32) {
                        sPos[k] = 0;
                        sQ[k] = -p.grad[p0 + k];
                    }
                    __syncwarp();
                    if (small) {
                        const uint16_t w = (uint16_t)(0x8000u | ((unsigned)my_len << 8) | (unsigned)my_head);
                        for (int j = a0; j < a1; ++j) sPos[j] = w;
                    }
                    __syncwarp();
                    const int L = __reduce_max_sync(0xffffffffu, my_len);
                    double alpha[LB_MAX_M];
                    double scale = 1.0;
#pragma unroll 1
                    for (int t = 0; t < L; ++t) {
                        lb_stage(p, lane, t, p0, n, sPos, sA, sB);
                        __syncwarp();
                        if (t < my_len) {
                            double sq = 0.0;
                            for (int j = a0; j < a1; ++j) sq = fma(sA[j], sQ[j], sq);
                            const double a = __dmul_rn(p.rho[(int64_t)lb_slot(my_head, t, p.m) * p.n_entities + e], sq);
                            alpha[t] = a;
                            const double na = -a;
                            if (t == 0) {
                                double sy = 0.0, yy = 0.0;
                                for (int j = a0; j < a1; ++j) {
                                    sy = fma(sA[j], sB[j], sy);
                                    yy = fma(sB[j], sB[j], yy);
                                }
                                scale = sy / yy;
                            }
                            for (int j = a0; j < a1; ++j) sQ[j] = __dadd_rn(sQ[j], __dmul_rn(na, sB[j]));
                        }
                        __syncwarp();
                    }
                    if (small)
                        for (int j = a0; j < a1; ++j) sQ[j] = __dmul_rn(scale, sQ[j]);
#pragma unroll 1
                    for (int t = L - 1; t >= 0; --t) {
                        if (t != L - 1) {                      // pair L-1 is still there from the first loop
                            lb_stage(p, lane, t, p0, n, sPos, sA, sB);
                            __syncwarp();
                        }
                        if (t < my_len) {
                            double yq = 0.0;
                            for (int j = a0; j < a1; ++j) yq = fma(sB[j], sQ[j], yq);
                            const double c = __dsub_rn(alpha[t], __dmul_rn(p.rho[(int64_t)lb_slot(my_head, t, p.m) * p.n_entities + e], yq));
                            for (int j = a0; j < a1; ++j) sQ[j] = __dadd_rn(sQ[j], __dmul_rn(c, sA[j]));
                        }
                        __syncwarp();
                    }
                    for (int k = lane; k < n; k += 32)
                        if (sPos[k] & 0x8000u) p.out[p0 + k] = sQ[k];
                    __syncwarp();
                }
            } else {
                took = 1;                                      // entity `es` alone is larger than the staging area
            }
            unsigned todo = __ballot_sync(0xffffffffu, lane < took && mine && big);
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                lbfgs_direction_warp(p, lane, es + src, __shfl_sync(0xffffffffu, lo, src), __shfl_sync(0xffffffffu, hi, src));
            }
            es += took;
        }
    }
}

int launch_batch_lbfgs_direction(cudaStream_t s, int64_t n_entities, int64_t d_total, const int64_t *coef_ptr, const uint8_t *go,
                                 int m, const double *S, const double *Y, const double *rho, const int32_t *head,
                                 const int32_t *len, const double *grad, double *out)
{
    if (n_entities <= 0) return 0;
    if (m < 1 || m > LB_MAX_M) return -1;
    const size_t smem = sizeof(double) * 3 * LB_CAP * LB_WARPS + sizeof(uint16_t) * LB_CAP * LB_WARPS;
    if (cudaFuncSetAttribute(batch_lbfgs_direction_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return -1;
    LbfgsParams p{n_entities, d_total, coef_ptr, go, m, S, Y, rho, head, len, grad, out};
    int64_t grid = ((n_entities + 31) / 32 + LB_WARPS - 1) / LB_WARPS;
    if (grid > 148) grid = 148;
    batch_lbfgs_direction_kernel<<<(int)grid, 32 * LB_WARPS, smem, s>>>(p);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// Append (s, y) to the ring of every entity with mask[e] != 0: the entity pass moves head / len and stores rho, the
// coefficient pass then copies the two vectors into the entity's new slot.
__global__ void lbfgs_push_entity_kernel(int64_t n_entities, const uint8_t *__restrict__ mask, int m, const double *__restrict__ rho_new,
                                         double *__restrict__ rho, int32_t *__restrict__ head, int32_t *__restrict__ len)
{
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_entities; e += (int64_t)gridDim.x * blockDim.x) {
        if (!mask[e]) continue;
        const int n = len[e];
        const int h = n == 0 ? 0 : (head[e] + 1 == m ? 0 : head[e] + 1);
        head[e] = h;
        len[e] = n < m ? n + 1 : m;
        rho[(int64_t)h * n_entities + e] = rho_new[e];
    }
}

__global__ void lbfgs_push_coef_kernel(int64_t d, const int32_t *__restrict__ ent_of, const uint8_t *__restrict__ mask,
                                       const int32_t *__restrict__ head, const double *__restrict__ s, const double *__restrict__ y,
                                       double *__restrict__ S, double *__restrict__ Y)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x) {
        const int32_t e = ent_of[j];
        if (!mask[e]) continue;
        const int64_t at = (int64_t)head[e] * d + j;
        S[at] = s[j];
        Y[at] = y[j];
    }
}

int launch_batch_lbfgs_push(cudaStream_t st, int64_t n_entities, int64_t d, const int32_t *ent_of, const uint8_t *mask, int m,
                            const double *rho_new, const double *s, const double *y, double *S, double *Y, double *rho,
                            int32_t *head, int32_t *len)
{
    if (n_entities <= 0) return 0;
    if (m < 1 || m > LB_MAX_M) return -1;
    int64_t grid = (n_entities + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    lbfgs_push_entity_kernel<<<(int)grid, 256, 0, st>>>(n_entities, mask, m, rho_new, rho, head, len);
    if (d > 0) {
        grid = (d + 255) / 256;
        if (grid > 148 * 16) grid = 148 * 16;
        lbfgs_push_coef_kernel<<<(int)grid, 256, 0, st>>>(d, ent_of, mask, head, s, y, S, Y);
    }
    return cudaGetLastError() == cudaSuccess ? 2 : -1;
}

// ------------------------------------------------------------------------------------------
// Per-entity NormalizationContext of a batch (every entity of a random-effect coordinate trains in the transformed
// space of ITS projected features: ValueAndGradientAggregator.init / getVector, :90-122 and :199-221, per entity).
// Shifts need an intercept, and the intercept column is 1 in every row, so the per-entity scalars fold into the
// intercept: x_i.ec + marginShift_e = x_i.ec' with ec'[icpt_e] = ec[icpt_e] - sum_j ec_j s_j, and the entity's
// prefactorSum_e = sum_i r_i is vectorSum[icpt_e].  What is left is element-wise.
// ------------------------------------------------------------------------------------------
__global__ void scale_kernel(int64_t d, const double *__restrict__ x, const double *__restrict__ f, double *__restrict__ out)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x)
        out[j] = f ? __dmul_rn(x[j], f[j]) : x[j];
}

int launch_scale(cudaStream_t s, int64_t d, const double *x, const double *f, double *out)
{
    if (d <= 0) return 0;
    int64_t grid = (d + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    scale_kernel<<<(int)grid, 256, 0, s>>>(d, x, f, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// vec[icpt_pos[e]] -= dots[e] for every entity that has an intercept
__global__ void absorb_shift_kernel(int64_t n_entities, const int64_t *__restrict__ icpt_pos, const double *__restrict__ dots,
                                    double *__restrict__ vec)
{
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n_entities; e += (int64_t)gridDim.x * blockDim.x)
        if (icpt_pos[e] >= 0) vec[icpt_pos[e]] -= dots[e];
}

int launch_absorb_shift(cudaStream_t s, int64_t n_entities, const int64_t *icpt_pos, const double *dots, double *vec)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    absorb_shift_kernel<<<(int)grid, 256, 0, s>>>(n_entities, icpt_pos, dots, vec);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// getVector per entity: out_j = (V_j - s_j * P_e) * f_j with P_e = V[icpt_pos[e]]   (out must not alias V)
__global__ void batch_finalize_kernel(int64_t d, const int32_t *__restrict__ ent_of, const double *__restrict__ V,
                                      const double *__restrict__ factors, const double *__restrict__ shifts,
                                      const int64_t *__restrict__ icpt_pos, double *__restrict__ out)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x) {
        double v = V[j];
        if (shifts) {
            const int64_t ip = icpt_pos[ent_of[j]];
            if (ip >= 0) v = __dsub_rn(v, __dmul_rn(shifts[j], V[ip]));
        }
        if (factors) v = __dmul_rn(v, factors[j]);
        out[j] = v;
    }
}

int launch_batch_finalize(cudaStream_t s, int64_t d, const int32_t *ent_of, const double *V, const double *factors,
                          const double *shifts, const int64_t *icpt_pos, double *out)
{
    if (d <= 0) return 0;
    int64_t grid = (d + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    batch_finalize_kernel<<<(int)grid, 256, 0, s>>>(d, ent_of, V, factors, shifts, icpt_pos, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// Counts rows of entities with an intercept that do NOT hold the entry (intercept column, 1.0): folding the shifts
// into the intercept is only valid when there are none.  One warp per entity, lanes stride its rows.
template <typename VT>
__global__ void check_intercepts_kernel(int64_t n_entities, const int64_t *__restrict__ ent_row_ptr,
                                        const uint32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                        const VT *__restrict__ val, const int32_t *__restrict__ rank,
                                        const int64_t *__restrict__ icpt_pos, unsigned long long *__restrict__ bad)
{
    const int lane = threadIdx.x & 31;
    int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    unsigned long long wrong = 0;
    for (; e < n_entities; e += nw) {
        if (icpt_pos[e] < 0) continue;
        const int32_t want = rank[icpt_pos[e]];
        for (int64_t i = ent_row_ptr[e] + lane; i < ent_row_ptr[e + 1]; i += 32) {
            bool found = false;
            for (uint32_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k)
                if ((col[k] & 0x7fffffff) == want && load_value(val, k) == 1.0) found = true;
            if (!found) ++wrong;
        }
    }
    if (wrong) atomicAdd(bad, wrong);  // a count for the error message: order does not matter
}

int launch_check_intercepts(cudaStream_t s, int value_bytes, int64_t n_entities, const int64_t *ent_row_ptr,
                            const uint32_t *row_ptr, const int32_t *col, const void *val, const int32_t *rank,
                            const int64_t *icpt_pos, unsigned long long *bad)
{
    if (n_entities <= 0) return 0;
    int64_t grid = (n_entities * 32 + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    if (value_bytes == 0)
        check_intercepts_kernel<One><<<(int)grid, 256, 0, s>>>(n_entities, ent_row_ptr, row_ptr, col, nullptr, rank, icpt_pos, bad);
    else if (value_bytes == 4)
        check_intercepts_kernel<float><<<(int)grid, 256, 0, s>>>(n_entities, ent_row_ptr, row_ptr, col, (const float *)val, rank,
                                                                 icpt_pos, bad);
    else
        check_intercepts_kernel<double><<<(int)grid, 256, 0, s>>>(n_entities, ent_row_ptr, row_ptr, col, (const double *)val,
                                                                  rank, icpt_pos, bad);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// offsets'_i = base_i + score_{map(i)}: DataSet.addScoresToOffsets (api/data/FixedEffectDataSet.scala:45-54,
// api/data/LocalDataSet.scala:82-89) without the RDD join -- `row_map` carries the row correspondence
// between two coordinates' row orders (e.g. entity-grouped rows -> fixed-effect rows); null = identity.
__global__ void offsets_from_scores_kernel(int64_t n, const double *__restrict__ base, const double *__restrict__ scores,
                                           const int64_t *__restrict__ row_map, double *__restrict__ out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double b = base ? base[i] : 0.0;
        out[i] = b + scores[row_map ? row_map[i] : i];
    }
}

int launch_offsets_from_scores(cudaStream_t s, int64_t n, const double *base, const double *scores, const int64_t *row_map,
                               double *out)
{
    if (n <= 0) return 0;
    int64_t grid = (n + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    offsets_from_scores_kernel<<<(int)grid, 256, 0, s>>>(n, base, scores, row_map, out);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

}  // namespace pml
