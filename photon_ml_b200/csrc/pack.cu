// pack.cu -- one-off packing of LabeledPoint rows (lib/data/LabeledPoint.scala:32-36) into the
// device-resident layout the evaluation kernels read: CSR with columns renumbered by descending
// frequency, plus a CSC copy ordered by (internal column, row) with chunk metadata for
// the column-pass kernels.
//
// This runs once per data set (and never per evaluation); it uses CUB's device radix sort and
// scan for the transposition.  The reference has no counterpart -- it re-reads JVM objects every
// evaluation -- so nothing here mirrors reference code.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

#include <cstdlib>
#include <type_traits>
#include <vector>

#include "internal.h"

namespace pml {

namespace {

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    template <typename T> T *as() { return static_cast<T *>(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};

__global__ void expand_rows_kernel(int64_t n, const uint32_t *__restrict__ row_ptr, int32_t *__restrict__ entry_row)
{
    // one warp per row: entry_row[k] = i for k in [row_ptr[i], row_ptr[i+1])
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += nw) {
        const uint32_t a = row_ptr[w], b = row_ptr[w + 1];
        for (uint32_t k = a + lane; k < b; k += 32) entry_row[k] = (int32_t)w;
    }
}

__global__ void iota_kernel(int64_t n, uint32_t *out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = (uint32_t)i;
}

template <typename VT>
__global__ void permute_kernel(int64_t nnz, const uint32_t *__restrict__ perm, const int32_t *__restrict__ entry_row,
                               const VT *__restrict__ val, int32_t *__restrict__ csc_row, VT *__restrict__ csc_val)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t src = perm[k];
        csc_row[k] = entry_row[src];
        if (val != nullptr) csc_val[k] = val[src];  // null: implicit ones, nothing stored
    }
}

__global__ void count_cols_kernel(int64_t nnz, const int32_t *__restrict__ col, uint32_t *__restrict__ counts)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(counts + col[k], 1u);  // integer counts: order-independent, so packing stays deterministic
}

// rank[orig] = internal id; counts how many columns are non-empty
__global__ void invert_perm_kernel(int32_t d, const int32_t *__restrict__ orig_of, const uint32_t *__restrict__ sorted_counts,
                                   int32_t *__restrict__ rank, unsigned long long *__restrict__ nonempty)
{
    int mine = 0;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < d; c += (int64_t)gridDim.x * blockDim.x) {
        rank[orig_of[c]] = (int32_t)c;
        if (sorted_counts[c] > 0) ++mine;
    }
    if (mine) atomicAdd(nonempty, (unsigned long long)mine);
}

__global__ void clamp_counts_kernel(int32_t d, const uint32_t *__restrict__ counts, uint32_t *__restrict__ keys)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x)
        keys[j] = counts[j] > 0 ? 1u : 0u;
}

__global__ void gather_counts_kernel(int32_t d, const uint32_t *__restrict__ counts, const int32_t *__restrict__ orig_of,
                                     uint32_t *__restrict__ counts_sorted)
{
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < d; c += (int64_t)gridDim.x * blockDim.x)
        counts_sorted[c] = counts[orig_of[c]];
}

__global__ void remap_cols_kernel(int64_t nnz, int32_t *__restrict__ col, const int32_t *__restrict__ rank)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x)
        col[k] = rank[col[k]];
}

__global__ void nonempty_rows_kernel(int64_t n, const uint32_t *__restrict__ row_ptr, int32_t *__restrict__ flag)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        flag[i] = row_ptr[i + 1] > row_ptr[i] ? 1 : 0;
}

__global__ void compact_rows_kernel(int64_t n, const uint32_t *__restrict__ row_ptr, const int32_t *__restrict__ flag,
                                    const int32_t *__restrict__ pos, int64_t nnz, int64_t runs,
                                    uint32_t *__restrict__ run_ptr, int32_t *__restrict__ run_id)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (flag[i]) {
            run_ptr[pos[i]] = row_ptr[i];
            run_id[pos[i]] = (int32_t)i;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) run_ptr[runs] = (uint32_t)nnz;
}

__global__ void mark_starts_kernel(int64_t m, const uint32_t *__restrict__ col_ptr, int32_t *__restrict__ csc_row)
{
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < m; c += (int64_t)gridDim.x * blockDim.x)
        csc_row[col_ptr[c]] |= (int32_t)0x80000000;
}

// chunk_col_lo[c] = number of non-empty columns that start before entry c*chunk_entries;
// head_nonempty[c] = 1 when entry c*chunk_entries is not itself a column start
__global__ void chunk_meta_kernel(int n_chunks, int chunk_entries, int64_t m, const uint32_t *__restrict__ nzcol_ptr,
                                  int32_t *__restrict__ chunk_col_lo, int32_t *__restrict__ head_nonempty,
                                  uint32_t ptr_offset = 0)
{
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= n_chunks; c += gridDim.x * blockDim.x) {
        if (c == n_chunks) {
            chunk_col_lo[c] = (int32_t)m;
            continue;
        }
        const uint32_t base = (uint32_t)((int64_t)c * chunk_entries);
        int64_t lo = 0, hi = m;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (nzcol_ptr[mid] - ptr_offset < base) lo = mid + 1;
            else hi = mid;
        }
        chunk_col_lo[c] = (int32_t)lo;
        head_nonempty[c] = (lo < m && nzcol_ptr[lo] - ptr_offset == base) ? 0 : 1;
    }
}

__global__ void exact_f32_kernel(const double *__restrict__ val, int64_t nnz, int *__restrict__ inexact)
{
    int bad = 0;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        const double v = val[k];
        // NaN compares unequal to itself; a NaN payload survives narrowing well enough for our purposes
        if ((double)(float)v != v && v == v) bad = 1;
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *inexact = 1;
}

template <typename T>
__global__ void all_ones_kernel(const T *__restrict__ val, int64_t nnz, int *__restrict__ other)
{
    int bad = 0;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x)
        if (!(val[k] == (T)1)) bad = 1;  // NaN counts as "not one"
    if (__syncthreads_or(bad) && threadIdx.x == 0) *other = 1;
}

template <typename S, typename D>
__global__ void convert_kernel(const S *__restrict__ src, D *__restrict__ dst, int64_t n)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
        dst[k] = (D)src[k];
}

__global__ void row_ptr_kernel(const int64_t *__restrict__ src, uint32_t *__restrict__ dst, int64_t n1, int64_t limit,
                               unsigned long long *__restrict__ bad)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n1; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = src[i];
        const bool wrong = v < 0 || v > limit || (i > 0 && src[i - 1] > v) || (i == 0 && v != 0);
        if (wrong) atomicAdd(bad, 1ULL);
        dst[i] = (uint32_t)v;
    }
}

__global__ void check_cols_kernel(const int32_t *__restrict__ col, int64_t nnz, int32_t d, unsigned long long *__restrict__ bad)
{
    int wrong = 0;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = col[k];
        if (c < 0 || c >= d) ++wrong;
    }
    if (wrong) atomicAdd(bad, (unsigned long long)wrong);
}

// key of an entry for the tiled layout: hot columns (< K) sort by (row tile, column); cold ones after, by
// (super block of rows, column) -- super blocks keep the slice of r the cold gathers touch inside the L2
__global__ void tile_keys_kernel(int64_t nnz, const int32_t *__restrict__ col, const int32_t *__restrict__ entry_row,
                                 int K, int tile_rows, int n_tiles, int super_rows, uint32_t cold_cols,
                                 uint32_t *__restrict__ keys)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += (int64_t)gridDim.x * blockDim.x) {
        const int32_t c = col[k];
        const int32_t row = entry_row[k];
        keys[k] = c < K ? (uint32_t)(row / tile_rows) * (uint32_t)K + (uint32_t)c
                        : (uint32_t)n_tiles * (uint32_t)K + (uint32_t)(row / super_rows) * cold_cols + (uint32_t)(c - K);
    }
}

// slot of a cold run in cold_partial = its key less the hot key range (keys are unsigned: sort them as such)
__global__ void cold_slots_kernel(int64_t runs, const int32_t *__restrict__ run_keys, uint32_t hot_keys,
                                  int32_t *__restrict__ slot)
{
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < runs; q += (int64_t)gridDim.x * blockDim.x)
        slot[q] = (int32_t)((uint32_t)run_keys[q] - hot_keys);
}

// first sorted position whose key is >= t*K, for t in [0, n_tiles]
__global__ void tile_starts_kernel(int n_tiles, int K, int64_t nnz, const uint32_t *__restrict__ sorted_keys,
                                   uint32_t *__restrict__ hs)
{
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t <= n_tiles; t += gridDim.x * blockDim.x) {
        const uint32_t want = (uint32_t)t * (uint32_t)K;
        int64_t lo = 0, hi = nnz;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (sorted_keys[mid] < want) lo = mid + 1;
            else hi = mid;
        }
        hs[t] = (uint32_t)lo;
    }
}

__global__ void fill_pad_kernel(int64_t len, int32_t pad_idx, int32_t *__restrict__ idx)
{
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < len; k += (int64_t)gridDim.x * blockDim.x)
        idx[k] = pad_idx;
}

// Moves sorted entry i to its place in [hot tiled | cold]; marks run starts; records, per sorted position,
// whether a run starts there and where it landed (for the run lists).
template <typename VT>
__global__ void tile_scatter_kernel(int64_t nnz, const uint32_t *__restrict__ sorted_keys, const uint32_t *__restrict__ perm,
                                    const int32_t *__restrict__ entry_row, const VT *__restrict__ val, int K, int tile_rows,
                                    int n_tiles, const uint32_t *__restrict__ hs, const uint32_t *__restrict__ po,
                                    uint32_t cold_off, int32_t *__restrict__ idx, VT *__restrict__ out_val,
                                    unsigned char *__restrict__ run_flag, uint32_t *__restrict__ run_dest)
{
    const uint32_t hot_keys = (uint32_t)n_tiles * (uint32_t)K;
    const uint32_t n_hot = hs[n_tiles];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t key = sorted_keys[i];
        const uint32_t src = perm[i];
        const bool start = i == 0 || sorted_keys[i - 1] != key;
        int32_t row = entry_row[src];
        uint32_t dest;
        if (key < hot_keys) {
            const uint32_t t = key / (uint32_t)K;
            dest = po[t] + ((uint32_t)i - hs[t]);
            row -= (int32_t)t * tile_rows;
        } else {
            dest = cold_off + ((uint32_t)i - n_hot);
        }
        run_flag[i] = start ? 1 : 0;
        run_dest[i] = dest;
        idx[dest] = start ? (row | (int32_t)0x80000000) : row;
        if (val != nullptr) out_val[dest] = val[src];  // null: implicit ones, nothing stored
    }
}

// Bank-aware order inside the runs of the tiled hot region.  The tiled kernel looks r up in shared memory with 8-byte
// loads, 8 entries per lane: for item i of the 8, the 16 lanes of a half warp are served in one pass only when their
// rows fall into 16 different 8-byte bank pairs (row mod 16 all different); random rows collide about three deep, and a
// dense column read in row order collides eight deep (lane L holds rows 8L..8L+7, so item i of 16 lanes strides by 8).
// A sum does not care about the order of its terms, so within each run the entries are re-dealt: position p of a chunk
// belongs to group (half warp, item) = (p / 128, p % 8), and every group greedily takes an entry whose residue it has
// not seen yet, from the fullest residue class left in the run.  One warp per 256-entry chunk: the lanes copy the
// chunk to shared memory, lane 0 deals (4096 cheap steps, pack time only), the lanes write the new order back.
// Run-start flags stay with their POSITIONS.  Deterministic: the result depends on the chunk's content only.
template <typename VT>
__global__ void __launch_bounds__(256) bank_order_kernel(int64_t n_chunks, int32_t *__restrict__ idx, VT *__restrict__ val,
                                                         int32_t n_table)
{
    // n_table: indices below it are looked up in shared memory (the tiled region: all of them; the CSR stream: the
    // staged prefix of the coefficient vector); the others go through the global path and fit anywhere.  Lanes that
    // read the SAME word share one pass (a broadcast), so an index may join a group that already holds it.
    constexpr int WARPS = 8;
    __shared__ int32_t s_row[WARPS][256];
    __shared__ unsigned char s_perm[WARPS][256], s_sorted[WARPS][256];
    __shared__ int32_t s_owner[WARPS][16][16];   // [group][bank pair] -> the index that occupies it, -1 = free
    __shared__ double s_val[WARPS][(sizeof(VT) > 1) ? 256 : 1];   // wide enough for float and double values
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t c = (int64_t)blockIdx.x * WARPS + warp; c < n_chunks; c += (int64_t)gridDim.x * WARPS) {
        int32_t *ci = idx + c * 256;
        for (int k = lane; k < 256; k += 32) {
            s_row[warp][k] = ci[k];
            s_owner[warp][k >> 4][k & 15] = -1;
            if (val != nullptr) reinterpret_cast<VT *>(s_val[warp])[k] = val[c * 256 + k];
        }
        __syncwarp();
        if (lane == 0) {
            int a = 0;
            while (a < 256) {
                int b = a + 1;
                while (b < 256 && s_row[warp][b] >= 0) ++b;   // [a, b): one run's share of this chunk
                // classes 0..15: bank pair of a table index; class 16: global path, no constraint
                int cnt[17], off[17];
                for (int q = 0; q < 17; ++q) cnt[q] = 0;
                for (int k = a; k < b; ++k) {
                    const int32_t v = s_row[warp][k] & 0x3fffffff;
                    ++cnt[v < n_table ? (v & 15) : 16];
                }
                int acc = a;
                for (int q = 0; q < 17; ++q) { off[q] = acc; acc += cnt[q]; }
                {
                    int fill[17];
                    for (int q = 0; q < 17; ++q) fill[q] = off[q];
                    for (int k = a; k < b; ++k) {
                        const int32_t v = s_row[warp][k] & 0x3fffffff;
                        s_sorted[warp][fill[v < n_table ? (v & 15) : 16]++] = (unsigned char)k;
                    }
                }
                for (int pos = a; pos < b; ++pos) {
                    const int g = (pos >> 7) * 8 + (pos & 7);
                    int best = -1, best_any = -1;
                    for (int q = 0; q < 16; ++q) {
                        if (cnt[q] == 0) continue;
                        if (best_any < 0 || cnt[q] > cnt[best_any]) best_any = q;
                        const int32_t own = s_owner[warp][g][q];
                        const int32_t top = s_row[warp][s_sorted[warp][off[q] + cnt[q] - 1]] & 0x3fffffff;
                        if ((own < 0 || own == top) && (best < 0 || cnt[q] > cnt[best])) best = q;
                    }
                    int q = best;
                    if (q < 0) q = cnt[16] > 0 ? 16 : best_any;   // nothing fits: an unconstrained entry, else a conflict
                    const int src = s_sorted[warp][off[q] + (--cnt[q])];
                    if (q < 16 && s_owner[warp][g][q] < 0) s_owner[warp][g][q] = s_row[warp][src] & 0x3fffffff;
                    s_perm[warp][pos] = (unsigned char)src;
                }
                a = b;
            }
        }
        __syncwarp();
        for (int k = lane; k < 256; k += 32) {
            const int src = s_perm[warp][k];
            const int32_t flag = s_row[warp][k] & (int32_t)0x80000000;
            ci[k] = (s_row[warp][src] & 0x7fffffff) | flag;
            if (val != nullptr) val[c * 256 + k] = reinterpret_cast<const VT *>(s_val[warp])[src];
        }
        __syncwarp();
    }
}

// Final form of the tiled hot region: the 32-bit work copy (flag in bit 31) narrowed to 16-bit words (flag in bit
// 15; local rows are below 2^15), and one byte per chunk that says whether any run starts in it -- chunks without a
// start lie inside one (tile, column) run and take the kernel's no-scan path.  One warp per chunk.
__global__ void __launch_bounds__(256) narrow_tiled_kernel(int64_t n_chunks, const int32_t *__restrict__ idx,
                                                           uint16_t *__restrict__ idx16, unsigned char *__restrict__ uniform,
                                                           int mark_uniform)
{
    const int lane = threadIdx.x & 31;
    int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; c < n_chunks; c += nw) {
        const int4 *ci = reinterpret_cast<const int4 *>(idx + c * 256) + 2 * lane;
        const int4 a = ci[0], b = ci[1];
        const int32_t v[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t lo = ((uint32_t)v[2 * i] & 0x7fffu) | (v[2 * i] < 0 ? 0x8000u : 0u);
            const uint32_t hi = ((uint32_t)v[2 * i + 1] & 0x7fffu) | (v[2 * i + 1] < 0 ? 0x8000u : 0u);
            w[i] = lo | (hi << 16);
        }
        reinterpret_cast<uint4 *>(idx16 + c * 256)[lane] = make_uint4(w[0], w[1], w[2], w[3]);
        const bool start = (a.x | a.y | a.z | a.w | b.x | b.y | b.z | b.w) < 0;
        const bool any = __any_sync(0xffffffffu, start);
        if (lane == 0) uniform[c] = (mark_uniform && !any) ? 1 : 0;
    }
}

inline int grid_for(int64_t n, int block = 256)
{
    int64_t g = (n + block - 1) / block;
    if (g < 1) g = 1;
    if (g > 148 * 16) g = 148 * 16;
    return (int)g;
}

}  // namespace

int narrow_check(cudaStream_t s, const double *val, int64_t nnz, bool *all_exact)
{
    DevBuf flag;
    PML_CUDA(flag.alloc(sizeof(int)));
    PML_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), s));
    if (nnz > 0) exact_f32_kernel<<<grid_for(nnz), 256, 0, s>>>(val, nnz, flag.as<int>());
    PML_CUDA(cudaGetLastError());
    int h = 0;
    PML_CUDA(cudaMemcpyAsync(&h, flag.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    *all_exact = (h == 0);
    return PML_OK;
}

int ones_check(cudaStream_t s, const void *val, int value_bytes, int64_t nnz, bool *all_ones)
{
    DevBuf flag;
    PML_CUDA(flag.alloc(sizeof(int)));
    PML_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), s));
    if (nnz > 0) {
        if (value_bytes == 4) all_ones_kernel<float><<<grid_for(nnz), 256, 0, s>>>((const float *)val, nnz, flag.as<int>());
        else all_ones_kernel<double><<<grid_for(nnz), 256, 0, s>>>((const double *)val, nnz, flag.as<int>());
    }
    PML_CUDA(cudaGetLastError());
    int h = 0;
    PML_CUDA(cudaMemcpyAsync(&h, flag.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    *all_ones = (h == 0);
    return PML_OK;
}

int convert_values(cudaStream_t s, const void *src, int src_bytes, void *dst, int dst_bytes, int64_t nnz)
{
    if (nnz == 0) return PML_OK;
    if (src_bytes == dst_bytes) {
        PML_CUDA(cudaMemcpyAsync(dst, src, (size_t)nnz * src_bytes, cudaMemcpyDeviceToDevice, s));
    } else if (src_bytes == 8) {
        convert_kernel<double, float><<<grid_for(nnz), 256, 0, s>>>((const double *)src, (float *)dst, nnz);
    } else {
        convert_kernel<float, double><<<grid_for(nnz), 256, 0, s>>>((const float *)src, (double *)dst, nnz);
    }
    PML_CUDA(cudaGetLastError());
    return PML_OK;
}

int convert_row_ptr(cudaStream_t s, const int64_t *src, uint32_t *dst, int64_t n_plus_1, int64_t *bad)
{
    DevBuf cnt;
    PML_CUDA(cnt.alloc(sizeof(unsigned long long)));
    PML_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(unsigned long long), s));
    row_ptr_kernel<<<grid_for(n_plus_1), 256, 0, s>>>(src, dst, n_plus_1, 4000000000LL, cnt.as<unsigned long long>());
    PML_CUDA(cudaGetLastError());
    unsigned long long h = 0;
    PML_CUDA(cudaMemcpyAsync(&h, cnt.p, sizeof(h), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    *bad = (int64_t)h;
    return PML_OK;
}

int check_columns(cudaStream_t s, const int32_t *col, int64_t nnz, int32_t d, int64_t *bad)
{
    DevBuf cnt;
    PML_CUDA(cnt.alloc(sizeof(unsigned long long)));
    PML_CUDA(cudaMemsetAsync(cnt.p, 0, sizeof(unsigned long long), s));
    if (nnz > 0) check_cols_kernel<<<grid_for(nnz), 256, 0, s>>>(col, nnz, d, cnt.as<unsigned long long>());
    PML_CUDA(cudaGetLastError());
    unsigned long long h = 0;
    PML_CUDA(cudaMemcpyAsync(&h, cnt.p, sizeof(h), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    *bad = (int64_t)h;
    return PML_OK;
}

// Bank-aware order of `n_chunks` full chunks of an entry stream (see bank_order_kernel)
int bank_order(cudaStream_t s, int64_t n_chunks, int32_t *idx, void *val, int value_bytes, int32_t n_table)
{
    if (n_chunks <= 0) return PML_OK;
    const int grid = (int)(n_chunks / 8 + 1 < 148 * 8 ? n_chunks / 8 + 1 : 148 * 8);
    if (value_bytes == 0) bank_order_kernel<char><<<grid, 256, 0, s>>>(n_chunks, idx, (char *)nullptr, n_table);
    else if (value_bytes == 4) bank_order_kernel<float><<<grid, 256, 0, s>>>(n_chunks, idx, (float *)val, n_table);
    else bank_order_kernel<double><<<grid, 256, 0, s>>>(n_chunks, idx, (double *)val, n_table);
    PML_CUDA(cudaGetLastError());
    return PML_OK;
}

// Fix-up list: runs of consecutive chunks whose heads continue one run (host-side, tiny).
// run_id maps a run ordinal to its output slot (null = identity).
static int build_fix_list(cudaStream_t s, int n_chunks, const int32_t *chunk_lo_dev, const int32_t *head_dev,
                          const int32_t *run_id_dev, int64_t n_runs, FixEntry **fix_out, int *n_fix_out, int *n_long_out,
                          int64_t *bytes)
{
    std::vector<int32_t> h_lo((size_t)n_chunks + 1), h_head((size_t)n_chunks + 1, 0);
    PML_CUDA(cudaMemcpyAsync(h_lo.data(), chunk_lo_dev, sizeof(int32_t) * (size_t)(n_chunks + 1), cudaMemcpyDeviceToHost, s));
    if (n_chunks > 0)
        PML_CUDA(cudaMemcpyAsync(h_head.data(), head_dev, sizeof(int32_t) * (size_t)n_chunks, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    std::vector<int32_t> fix_ord;  // run ordinal per fix entry
    std::vector<FixEntry> fix;
    for (int c = 1; c < n_chunks; ++c) {
        if (!h_head[c]) continue;
        const int32_t ord = h_lo[c] - 1;  // the run that was open when chunk c began
        if (!fix.empty() && fix_ord.back() == ord && fix.back().chunk_end == c) {
            fix.back().chunk_end = c + 1;
        } else {
            FixEntry e;
            e.col = ord;  // replaced by the output slot below when there is a run_id map
            e.chunk_begin = c;
            e.chunk_end = c + 1;
            fix.push_back(e);
            fix_ord.push_back(ord);
        }
    }
    *n_fix_out = (int)fix.size();
    *fix_out = nullptr;
    *n_long_out = 0;
    if (!fix.empty()) {
        if (run_id_dev != nullptr) {
            std::vector<int32_t> h_ids((size_t)n_runs);
            // only a handful of ids are needed, but n_runs*4 bytes once at pack time is simpler than a gather
            PML_CUDA(cudaMemcpyAsync(h_ids.data(), run_id_dev, sizeof(int32_t) * (size_t)n_runs, cudaMemcpyDeviceToHost, s));
            PML_CUDA(cudaStreamSynchronize(s));
            for (size_t q = 0; q < fix.size(); ++q) fix[q].col = h_ids[(size_t)fix_ord[q]];
        }
        // long runs first (stable), so the kernel can give them a warp each
        std::vector<FixEntry> ordered;
        ordered.reserve(fix.size());
        for (const FixEntry &e : fix)
            if (e.chunk_end - e.chunk_begin >= 32) ordered.push_back(e);
        *n_long_out = (int)ordered.size();
        for (const FixEntry &e : fix)
            if (e.chunk_end - e.chunk_begin < 32) ordered.push_back(e);
        fix.swap(ordered);
        PML_CUDA(cudaMalloc(fix_out, sizeof(FixEntry) * fix.size()));
        PML_CUDA(cudaMemcpyAsync(*fix_out, fix.data(), sizeof(FixEntry) * fix.size(), cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaStreamSynchronize(s));
        *bytes += (int64_t)sizeof(FixEntry) * (int64_t)fix.size();
    }
    return PML_OK;
}

// Packing proper.  Columns are renumbered by descending entry count ("internal ids": the m
// non-empty columns are exactly ids [0, m), hottest first; ties keep the original order), the CSR
// column array is rewritten in place to internal ids, and the CSC copy is built in internal-id
// order with the first entry of every column marked in bit 31 of its row index.
static int build_tiled(cudaStream_t s, int sm_count, int64_t n, int64_t nnz, const uint32_t *row_ptr, const int32_t *col,
                       const void *val, int value_bytes, int chunk_entries, const uint32_t *counts_sorted_dev, Packed *out,
                       bool *tiled);

int build_transpose(cudaStream_t s, int sm_count, int64_t n, int32_t d, int64_t nnz, const uint32_t *row_ptr,
                    int32_t *col, const void *val, int value_bytes, int chunk_entries, bool renumber, Packed *out)
{
    out->chunk_entries = chunk_entries;
    out->n_chunks = (int)((nnz + chunk_entries - 1) / chunk_entries);
    int n_chunks = out->n_chunks;
    out->bytes = 0;
    const size_t dd = (size_t)d + 1;

    PML_CUDA(cudaMalloc(&out->rank, sizeof(int32_t) * dd));
    PML_CUDA(cudaMalloc(&out->nzcol_id, sizeof(int32_t) * dd));
    PML_CUDA(cudaMalloc(&out->nzcol_ptr, sizeof(uint32_t) * (dd + 1)));
    out->bytes += (int64_t)dd * 12 + 4;

    DevBuf counts, counts_sorted, ids, entry_row, perm_in, perm_out, keys_out, head, tmp, m_dev;
    PML_CUDA(counts.alloc(sizeof(uint32_t) * dd));
    PML_CUDA(counts_sorted.alloc(sizeof(uint32_t) * dd));
    PML_CUDA(ids.alloc(sizeof(int32_t) * dd));
    PML_CUDA(m_dev.alloc(sizeof(unsigned long long)));

    // 1. entries per column, columns ordered by descending count (stable)
    PML_CUDA(cudaMemsetAsync(counts.p, 0, sizeof(uint32_t) * dd, s));
    PML_CUDA(cudaMemsetAsync(m_dev.p, 0, sizeof(unsigned long long), s));
    if (nnz > 0) count_cols_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, col, counts.as<uint32_t>());
    iota_kernel<<<grid_for(d), 256, 0, s>>>(d, ids.as<uint32_t>());
    PML_CUDA(cudaGetLastError());
    if (d > 0) {
        // Sort key: the entry count (hot columns first), or -- when the caller's column order carries locality of
        // its own, as the block-diagonal batches of per-entity problems do -- just "non-empty first", which the
        // stable sort turns into a plain compaction that keeps the original order.
        DevBuf sort_keys;
        const uint32_t *keys_in = counts.as<uint32_t>();
        if (!renumber) {
            PML_CUDA(sort_keys.alloc(sizeof(uint32_t) * dd));
            clamp_counts_kernel<<<grid_for(d), 256, 0, s>>>(d, counts.as<uint32_t>(), sort_keys.as<uint32_t>());
            keys_in = sort_keys.as<uint32_t>();
        }
        size_t tb = 0;
        PML_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, tb, keys_in, counts_sorted.as<uint32_t>(),
                                                           ids.as<int32_t>(), out->nzcol_id, d, 0, 32, s));
        DevBuf t1;
        PML_CUDA(t1.alloc(tb));
        PML_CUDA(cub::DeviceRadixSort::SortPairsDescending(t1.p, tb, keys_in, counts_sorted.as<uint32_t>(),
                                                           ids.as<int32_t>(), out->nzcol_id, d, 0, 32, s));
        if (!renumber)  // the sorted keys are only 0/1: fetch the real counts in the new order
            gather_counts_kernel<<<grid_for(d), 256, 0, s>>>(d, counts.as<uint32_t>(), out->nzcol_id, counts_sorted.as<uint32_t>());
        invert_perm_kernel<<<grid_for(d), 256, 0, s>>>(d, out->nzcol_id, counts_sorted.as<uint32_t>(), out->rank,
                                                      m_dev.as<unsigned long long>());
        PML_CUDA(cudaGetLastError());
        // 2. column starts over internal ids = exclusive scan of the sorted counts (entry d = nnz)
        PML_CUDA(cudaMemsetAsync(counts_sorted.as<uint32_t>() + d, 0, sizeof(uint32_t), s));
        size_t tb2 = 0;
        PML_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb2, counts_sorted.as<uint32_t>(), out->nzcol_ptr, d + 1, s));
        DevBuf t2;
        PML_CUDA(t2.alloc(tb2));
        PML_CUDA(cub::DeviceScan::ExclusiveSum(t2.p, tb2, counts_sorted.as<uint32_t>(), out->nzcol_ptr, d + 1, s));
    } else {
        PML_CUDA(cudaMemsetAsync(out->nzcol_ptr, 0, sizeof(uint32_t), s));
    }
    unsigned long long m64 = 0;
    PML_CUDA(cudaMemcpyAsync(&m64, m_dev.p, sizeof(m64), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    out->m = (int64_t)m64;
    const int64_t m = out->m;
    out->cold_nnz = nnz;
    out->cold_run_id = out->nzcol_id;

    if (nnz > 0) {
        // CSR columns -> internal ids, in place (both layouts below read them)
        remap_cols_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, col, out->rank);
        PML_CUDA(cudaGetLastError());
        bool tiled = false;
        if (renumber)
        PML_TRY(build_tiled(s, sm_count, n, nnz, row_ptr, col, val, value_bytes, chunk_entries, counts_sorted.as<uint32_t>(),
                            out, &tiled));
        if (tiled) return PML_OK;
    }

    PML_CUDA(cudaMalloc(&out->csc_row, sizeof(int32_t) * (size_t)(nnz + 16)));
    if (value_bytes > 0) PML_CUDA(cudaMalloc(&out->csc_val, (size_t)value_bytes * (size_t)(nnz + 16)));
    out->idx_alloc = out->csc_row;
    out->val_alloc = out->csc_val;
    PML_CUDA(cudaMalloc(&out->chunk_col_lo, sizeof(int32_t) * (size_t)(n_chunks + 1)));
    PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(n_chunks + 1)));
    out->bytes += (int64_t)(sizeof(int32_t) + value_bytes) * (nnz + 16) + (int64_t)sizeof(int32_t) * (n_chunks + 1);

    if (nnz > 0) {
        PML_CUDA(entry_row.alloc(sizeof(int32_t) * (size_t)nnz));
        PML_CUDA(perm_in.alloc(sizeof(uint32_t) * (size_t)nnz));
        PML_CUDA(perm_out.alloc(sizeof(uint32_t) * (size_t)nnz));
        PML_CUDA(keys_out.alloc(sizeof(int32_t) * (size_t)nnz));
        expand_rows_kernel<<<grid_for(n * 32), 256, 0, s>>>(n, row_ptr, entry_row.as<int32_t>());
        iota_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, perm_in.as<uint32_t>());
        PML_CUDA(cudaGetLastError());
        int end_bit = 1;
        while (end_bit < 31 && (1LL << end_bit) < m) ++end_bit;
        size_t tmp_bytes = 0;
        // 4. stable LSD radix sort by internal column keeps rows ascending inside each column
        PML_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const int32_t *)col, keys_out.as<int32_t>(),
                                                 perm_in.as<uint32_t>(), perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
        PML_CUDA(tmp.alloc(tmp_bytes));
        PML_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, (const int32_t *)col, keys_out.as<int32_t>(),
                                                 perm_in.as<uint32_t>(), perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
        if (value_bytes <= 4)
            permute_kernel<float><<<grid_for(nnz), 256, 0, s>>>(nnz, perm_out.as<uint32_t>(), entry_row.as<int32_t>(),
                                                               value_bytes ? (const float *)val : nullptr, out->csc_row,
                                                               (float *)out->csc_val);
        else
            permute_kernel<double><<<grid_for(nnz), 256, 0, s>>>(nnz, perm_out.as<uint32_t>(), entry_row.as<int32_t>(),
                                                                (const double *)val, out->csc_row, (double *)out->csc_val);
        // 5. mark the first entry of every column in bit 31 of its row index
        mark_starts_kernel<<<grid_for(m), 256, 0, s>>>(m, out->nzcol_ptr, out->csc_row);
        PML_CUDA(cudaGetLastError());
    }
    // tail padding is read by the vector loads of the last (partial) chunk's neighbours: keep it benign
    PML_CUDA(cudaMemsetAsync(out->csc_row + nnz, 0, sizeof(int32_t) * 16, s));
    if (value_bytes > 0)
        PML_CUDA(cudaMemsetAsync((char *)out->csc_val + (size_t)nnz * value_bytes, 0, (size_t)value_bytes * 16, s));
    chunk_meta_kernel<<<grid_for((int64_t)n_chunks + 1), 256, 0, s>>>(n_chunks, chunk_entries, m, out->nzcol_ptr,
                                                                     out->chunk_col_lo, head.as<int32_t>());
    PML_CUDA(cudaGetLastError());

    PML_TRY(build_fix_list(s, n_chunks, out->chunk_col_lo, head.as<int32_t>(), out->nzcol_id, m, &out->fix, &out->n_fix,
                           &out->n_fix_long, &out->bytes));
    return PML_OK;
}


// The tiled layout: [hot region | cold region] in one allocation.
//   hot  = entries of the K most frequent columns, ordered by (row tile, column, row); every tile is padded to a
//          multiple of 256 entries with entries that reference the tile's zero slot (index tile_rows, value 0);
//          the row index stored is local to the tile.  One run = one (tile, column) pair -> one slot of `partial`.
//   cold = the remaining columns, ordered by (column, row) as in the plain layout.
// K: columns with at least 2 entries per tile on average, capped by TILE_PARTIAL_BYTES for the partial matrix.
static int build_tiled(cudaStream_t s, int sm_count, int64_t n, int64_t nnz, const uint32_t *row_ptr, const int32_t *col,
                       const void *val, int value_bytes, int chunk_entries, const uint32_t *counts_sorted_dev, Packed *out,
                       bool *tiled)
{
    *tiled = false;
    if (chunk_entries != 256) return PML_OK;  // the tiled kernel works on warp-sized chunks
    int64_t min_rows = TILE_MIN_ROWS_DEFAULT;
    int forced_rows = 0;
    if (const char *e = std::getenv("PML_TILE_MIN_ROWS")) min_rows = std::atoll(e);
    if (const char *e = std::getenv("PML_TILE_ROWS")) forced_rows = std::atoi(e);  // tests: small tiles on small data
    if (std::getenv("PML_NO_TILING") != nullptr || n < min_rows) return PML_OK;
    const int64_t m = out->m;
    int64_t max_rows = TILE_MAX_ROWS, partial_bytes = TILE_PARTIAL_BYTES;
    if (const char *e = std::getenv("PML_TILE_MAX_ROWS")) max_rows = std::atoll(e);        // tuning knobs
    if (const char *e = std::getenv("PML_TILE_PARTIAL_MB")) partial_bytes = std::atoll(e) << 20;
    if (max_rows > 28000) max_rows = 28000;  // 227 KB of shared memory
    // tile geometry: a multiple of the SM count when the data is large enough, tiles of <= max_rows rows
    int64_t n_tiles = (n + max_rows - 1) / max_rows;
    if (n_tiles < sm_count) n_tiles = sm_count;  // one persistent CTA per SM: never fewer tiles than SMs
    n_tiles = (n_tiles + sm_count - 1) / sm_count * sm_count;
    int tile_rows = (int)((n + n_tiles - 1) / n_tiles);
    if (forced_rows > 0) tile_rows = forced_rows < max_rows ? forced_rows : (int)max_rows;
    n_tiles = (n + tile_rows - 1) / tile_rows;
    // K from the sorted counts
    int64_t k_cap = partial_bytes / (8 * n_tiles);
    if (k_cap > m) k_cap = m;
    if (k_cap <= 0) return PML_OK;
    std::vector<uint32_t> h_counts((size_t)k_cap);
    PML_CUDA(cudaMemcpyAsync(h_counts.data(), counts_sorted_dev, sizeof(uint32_t) * (size_t)k_cap, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    int64_t K = 0;
    int64_t min_run = 2;  // average entries per (tile, column) run a column must reach to be tiled
    if (const char *e = std::getenv("PML_TILE_MIN_RUN")) min_run = std::atoll(e);
    while (K < k_cap && (int64_t)h_counts[(size_t)K] >= min_run * n_tiles) ++K;
    if (K < 16 || n_tiles * K >= (1ll << 31)) return PML_OK;
    // cold columns: super blocks of rows whose slice of r (8 bytes a row) stays in the L2 next to the stream
    const int64_t cold_cols = m - K;
    // (r of up to SUPER_MIN_ROWS rows stays there as a whole: measured, no gain from cutting it)
    int64_t super_max = n > SUPER_MIN_ROWS_DEFAULT ? SUPER_ROWS_DEFAULT : n;
    if (const char *e = std::getenv("PML_SUPER_ROWS")) super_max = std::atoll(e);
    if (super_max < 1) super_max = 1;
    int64_t n_super = cold_cols > 0 ? (n + super_max - 1) / super_max : 1;
    if (n_tiles * K + n_super * cold_cols >= (1ll << 32) || n_super * cold_cols >= (1ll << 31)) n_super = 1;
    const int super_rows = (int)((n + n_super - 1) / n_super);
    n_super = (n + super_rows - 1) / super_rows;

    DevBuf entry_row, keys, keys_sorted, perm_in, perm_out, tmp, hs_dev, po_dev, run_flag, hot_dest;
    PML_CUDA(entry_row.alloc(sizeof(int32_t) * (size_t)nnz));
    PML_CUDA(keys.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(keys_sorted.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(perm_in.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(perm_out.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(hs_dev.alloc(sizeof(uint32_t) * (size_t)(n_tiles + 1)));
    PML_CUDA(po_dev.alloc(sizeof(uint32_t) * (size_t)(n_tiles + 1)));
    expand_rows_kernel<<<grid_for(n * 32), 256, 0, s>>>(n, row_ptr, entry_row.as<int32_t>());
    iota_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, perm_in.as<uint32_t>());
    tile_keys_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, col, entry_row.as<int32_t>(), (int)K, tile_rows, (int)n_tiles,
                                                   super_rows, (uint32_t)(cold_cols > 0 ? cold_cols : 1),
                                                   keys.as<uint32_t>());
    PML_CUDA(cudaGetLastError());
    const int64_t max_key = n_tiles * K + n_super * cold_cols;
    int end_bit = 1;
    while (end_bit < 32 && (1LL << end_bit) <= max_key) ++end_bit;
    size_t tmp_bytes = 0;
    PML_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.as<uint32_t>(), keys_sorted.as<uint32_t>(),
                                             perm_in.as<uint32_t>(), perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
    PML_CUDA(tmp.alloc(tmp_bytes));
    PML_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.as<uint32_t>(), keys_sorted.as<uint32_t>(),
                                             perm_in.as<uint32_t>(), perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
    tile_starts_kernel<<<grid_for(n_tiles + 1), 256, 0, s>>>((int)n_tiles, (int)K, nnz, keys_sorted.as<uint32_t>(),
                                                             hs_dev.as<uint32_t>());
    PML_CUDA(cudaGetLastError());
    std::vector<uint32_t> hs((size_t)n_tiles + 1), po((size_t)n_tiles + 1, 0);
    PML_CUDA(cudaMemcpyAsync(hs.data(), hs_dev.p, sizeof(uint32_t) * (size_t)(n_tiles + 1), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    std::vector<int32_t> first_chunk((size_t)n_tiles + 1, 0);
    for (int64_t t = 0; t < n_tiles; ++t) {
        const uint64_t cnt = hs[(size_t)t + 1] - hs[(size_t)t];
        const uint64_t padded = (cnt + 255) / 256 * 256;
        if ((uint64_t)po[(size_t)t] + padded > 4000000000ull) return PML_OK;  // would not fit 32-bit offsets: stay untiled
        po[(size_t)t + 1] = po[(size_t)t] + (uint32_t)padded;
        first_chunk[(size_t)t + 1] = (int32_t)(po[(size_t)t + 1] / 256);
    }
    const int64_t n_hot = hs[(size_t)n_tiles];
    const int64_t hot_len = po[(size_t)n_tiles];
    const int64_t cold_nnz = nnz - n_hot;
    // the run compaction below counts with 32-bit ints: stay untiled
    if (n_hot > 2000000000ll || (n_super > 1 && cold_nnz > 2000000000ll)) return PML_OK;
    const int64_t total = hot_len + cold_nnz;
    PML_CUDA(cudaMemcpyAsync(po_dev.p, po.data(), sizeof(uint32_t) * (size_t)(n_tiles + 1), cudaMemcpyHostToDevice, s));

    TiledSide &h = out->hot;
    // 32-bit work copy of [hot | cold]: scattered, ordered and marked here, then the hot part is narrowed to 16 bits
    // and the cold part moves to an allocation of its own
    DevBuf work;
    PML_CUDA(work.alloc(sizeof(int32_t) * (size_t)(total + 16)));
    int32_t *work_idx = work.as<int32_t>();
    if (value_bytes > 0) {
        PML_CUDA(cudaMalloc(&out->val_alloc, (size_t)value_bytes * (size_t)(total + 16)));
        PML_CUDA(cudaMemsetAsync(out->val_alloc, 0, (size_t)value_bytes * (size_t)(total + 16), s));
    }
    PML_CUDA(cudaMemsetAsync(work_idx + hot_len, 0, sizeof(int32_t) * (size_t)(cold_nnz + 16), s));
    fill_pad_kernel<<<grid_for(hot_len), 256, 0, s>>>(hot_len, tile_rows, work_idx);
    PML_CUDA(run_flag.alloc((size_t)(nnz + 1)));
    PML_CUDA(hot_dest.alloc(sizeof(uint32_t) * (size_t)(nnz + 1)));
    if (value_bytes <= 4)
        tile_scatter_kernel<float><<<grid_for(nnz), 256, 0, s>>>(
            nnz, keys_sorted.as<uint32_t>(), perm_out.as<uint32_t>(), entry_row.as<int32_t>(),
            value_bytes ? (const float *)val : nullptr, (int)K,
            tile_rows, (int)n_tiles, hs_dev.as<uint32_t>(), po_dev.as<uint32_t>(), (uint32_t)hot_len, work_idx,
            (float *)out->val_alloc, run_flag.as<unsigned char>(), hot_dest.as<uint32_t>());
    else
        tile_scatter_kernel<double><<<grid_for(nnz), 256, 0, s>>>(
            nnz, keys_sorted.as<uint32_t>(), perm_out.as<uint32_t>(), entry_row.as<int32_t>(), (const double *)val, (int)K,
            tile_rows, (int)n_tiles, hs_dev.as<uint32_t>(), po_dev.as<uint32_t>(), (uint32_t)hot_len, work_idx,
            (double *)out->val_alloc, run_flag.as<unsigned char>(), hot_dest.as<uint32_t>());
    PML_CUDA(cudaGetLastError());
    if (std::getenv("PML_NO_BANK_ORDER") == nullptr && hot_len > 0)
        PML_TRY(bank_order(s, hot_len / 256, work_idx, out->val_alloc, value_bytes, 0x3fffffff));
    PML_CUDA(cudaMalloc(&h.idx16, sizeof(uint16_t) * (size_t)(hot_len + 16)));
    PML_CUDA(cudaMalloc(&h.uniform, (size_t)(hot_len / 256 + 1)));
    PML_CUDA(cudaMemsetAsync(h.idx16 + hot_len, 0, sizeof(uint16_t) * 16, s));
    narrow_tiled_kernel<<<grid_for(hot_len / 8), 256, 0, s>>>(hot_len / 256, work_idx, h.idx16, h.uniform,
                                                              std::getenv("PML_NO_UNIFORM_FAST") == nullptr ? 1 : 0);
    PML_CUDA(cudaGetLastError());
    PML_CUDA(cudaMalloc(&out->idx_alloc, sizeof(int32_t) * (size_t)(cold_nnz + 16)));
    PML_CUDA(cudaMemcpyAsync(out->idx_alloc, work_idx + hot_len, sizeof(int32_t) * (size_t)(cold_nnz + 16),
                             cudaMemcpyDeviceToDevice, s));
    out->bytes += (int64_t)value_bytes * (total + 16) + (int64_t)sizeof(uint16_t) * (hot_len + 16) + hot_len / 256 + 1 +
                  (int64_t)sizeof(int32_t) * (cold_nnz + 16);

    // hot run list: positions (padded coordinates) and partial slots (= keys) of the run starts
    DevBuf n_sel, run_ptr_hot;
    PML_CUDA(n_sel.alloc(sizeof(int)));
    PML_CUDA(run_ptr_hot.alloc(sizeof(uint32_t) * (size_t)(n_hot + 2)));
    PML_CUDA(cudaMalloc(&h.run_id, sizeof(int32_t) * (size_t)(n_hot + 2)));  // trimmed below would need a copy: keep
    {
        DevBuf t1;
        size_t tb = 0;
        PML_CUDA(cub::DeviceSelect::Flagged(nullptr, tb, hot_dest.as<uint32_t>(), run_flag.as<unsigned char>(),
                                            run_ptr_hot.as<uint32_t>(), n_sel.as<int>(), (int)n_hot, s));
        PML_CUDA(t1.alloc(tb));
        PML_CUDA(cub::DeviceSelect::Flagged(t1.p, tb, hot_dest.as<uint32_t>(), run_flag.as<unsigned char>(),
                                            run_ptr_hot.as<uint32_t>(), n_sel.as<int>(), (int)n_hot, s));
        PML_CUDA(cub::DeviceSelect::Flagged(t1.p, tb, reinterpret_cast<const int32_t *>(keys_sorted.as<uint32_t>()),
                                            run_flag.as<unsigned char>(), h.run_id, n_sel.as<int>(), (int)n_hot, s));
    }
    int runs_hot = 0;
    PML_CUDA(cudaMemcpyAsync(&runs_hot, n_sel.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    const uint32_t hot_len32 = (uint32_t)hot_len;
    PML_CUDA(cudaMemcpyAsync(run_ptr_hot.as<uint32_t>() + runs_hot, &hot_len32, sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    out->bytes += (int64_t)sizeof(int32_t) * (n_hot + 2);

    h.val = out->val_alloc;
    h.len = hot_len;
    h.n_runs = runs_hot;
    h.n_chunks = (int)(hot_len / 256);
    h.tile_rows = tile_rows;
    h.n_tiles = (int)n_tiles;
    h.n_hot_cols = (int)K;
    PML_CUDA(cudaMalloc(&h.chunk_lo, sizeof(int32_t) * (size_t)(h.n_chunks + 1)));
    PML_CUDA(cudaMalloc(&h.tile_first_chunk, sizeof(int32_t) * (size_t)(n_tiles + 1)));
    PML_CUDA(cudaMalloc(&h.partial, sizeof(double) * (size_t)(n_tiles * K)));
    PML_CUDA(cudaMemsetAsync(h.partial, 0, sizeof(double) * (size_t)(n_tiles * K), s));  // slots without a run stay 0
    PML_CUDA(cudaMalloc(&h.carry, sizeof(double) * (size_t)(h.n_chunks + 1)));
    out->bytes += (int64_t)sizeof(int32_t) * (h.n_chunks + 1 + n_tiles + 1) + (int64_t)sizeof(double) * (n_tiles * K + h.n_chunks + 1);
    PML_CUDA(cudaMemcpyAsync(h.tile_first_chunk, first_chunk.data(), sizeof(int32_t) * (size_t)(n_tiles + 1),
                             cudaMemcpyHostToDevice, s));
    {
        DevBuf head;
        PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(h.n_chunks + 1)));
        chunk_meta_kernel<<<grid_for((int64_t)h.n_chunks + 1), 256, 0, s>>>(h.n_chunks, 256, runs_hot, run_ptr_hot.as<uint32_t>(),
                                                                           h.chunk_lo, head.as<int32_t>());
        PML_CUDA(cudaGetLastError());
        PML_TRY(build_fix_list(s, h.n_chunks, h.chunk_lo, head.as<int32_t>(), h.run_id, runs_hot, &h.fix, &h.n_fix,
                               &h.n_fix_long, &out->bytes));
    }

    // cold region: CSC over internal columns [K, m), one copy per super block of rows
    out->csc_row = out->idx_alloc;
    out->csc_val = value_bytes ? (char *)out->val_alloc + (size_t)hot_len * value_bytes : nullptr;
    out->cold_nnz = cold_nnz;
    out->n_chunks = (int)((cold_nnz + 255) / 256);
    PML_CUDA(cudaMalloc(&out->chunk_col_lo, sizeof(int32_t) * (size_t)(out->n_chunks + 1)));
    out->bytes += (int64_t)sizeof(int32_t) * (out->n_chunks + 1);
    if (n_super > 1 && cold_nnz > 0) {
        // one run = one (super block, column) pair -> slot [super block][column - K] of cold_partial, which
        // reduce_partial_kernel adds up over the super blocks; the run list comes from the run-start flags
        DevBuf run_ptr_cold, run_key_cold, t1, head;
        PML_CUDA(run_ptr_cold.alloc(sizeof(uint32_t) * (size_t)(cold_nnz + 2)));
        PML_CUDA(run_key_cold.alloc(sizeof(int32_t) * (size_t)(cold_nnz + 2)));
        size_t tb = 0;
        PML_CUDA(cub::DeviceSelect::Flagged(nullptr, tb, hot_dest.as<uint32_t>() + n_hot, run_flag.as<unsigned char>() + n_hot,
                                            run_ptr_cold.as<uint32_t>(), n_sel.as<int>(), (int)cold_nnz, s));
        PML_CUDA(t1.alloc(tb));
        PML_CUDA(cub::DeviceSelect::Flagged(t1.p, tb, hot_dest.as<uint32_t>() + n_hot, run_flag.as<unsigned char>() + n_hot,
                                            run_ptr_cold.as<uint32_t>(), n_sel.as<int>(), (int)cold_nnz, s));
        PML_CUDA(cub::DeviceSelect::Flagged(t1.p, tb, reinterpret_cast<const int32_t *>(keys_sorted.as<uint32_t>()) + n_hot,
                                            run_flag.as<unsigned char>() + n_hot, run_key_cold.as<int32_t>(),
                                            n_sel.as<int>(), (int)cold_nnz, s));
        int runs_cold = 0;
        PML_CUDA(cudaMemcpyAsync(&runs_cold, n_sel.p, sizeof(int), cudaMemcpyDeviceToHost, s));
        PML_CUDA(cudaStreamSynchronize(s));
        const uint32_t end32 = (uint32_t)(hot_len + cold_nnz);
        PML_CUDA(cudaMemcpyAsync(run_ptr_cold.as<uint32_t>() + runs_cold, &end32, sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaMalloc(&out->cold_slot, sizeof(int32_t) * (size_t)(runs_cold + 1)));
        cold_slots_kernel<<<grid_for(runs_cold), 256, 0, s>>>(runs_cold, run_key_cold.as<int32_t>(),
                                                             (uint32_t)(n_tiles * K), out->cold_slot);
        out->cold_run_id = out->cold_slot;
        out->n_super = (int)n_super;
        out->cold_cols = (int)cold_cols;
        PML_CUDA(cudaMalloc(&out->cold_partial, sizeof(double) * (size_t)(n_super * cold_cols)));
        PML_CUDA(cudaMemsetAsync(out->cold_partial, 0, sizeof(double) * (size_t)(n_super * cold_cols), s));
        out->bytes += (int64_t)sizeof(int32_t) * (runs_cold + 1) + (int64_t)sizeof(double) * n_super * cold_cols;
        PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(out->n_chunks + 1)));
        chunk_meta_kernel<<<grid_for((int64_t)out->n_chunks + 1), 256, 0, s>>>(out->n_chunks, 256, runs_cold,
                                                                              run_ptr_cold.as<uint32_t>(), out->chunk_col_lo,
                                                                              head.as<int32_t>(), (uint32_t)hot_len);
        PML_CUDA(cudaGetLastError());
        PML_TRY(build_fix_list(s, out->n_chunks, out->chunk_col_lo, head.as<int32_t>(), out->cold_run_id, runs_cold, &out->fix,
                               &out->n_fix, &out->n_fix_long, &out->bytes));
    } else {
        out->cold_run_id = out->nzcol_id + K;
        DevBuf head;
        PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(out->n_chunks + 1)));
        chunk_meta_kernel<<<grid_for((int64_t)out->n_chunks + 1), 256, 0, s>>>(out->n_chunks, 256, m - K, out->nzcol_ptr + K,
                                                                              out->chunk_col_lo, head.as<int32_t>(),
                                                                              (uint32_t)n_hot);
        PML_CUDA(cudaGetLastError());
        PML_TRY(build_fix_list(s, out->n_chunks, out->chunk_col_lo, head.as<int32_t>(), out->cold_run_id, m - K, &out->fix,
                               &out->n_fix, &out->n_fix_long, &out->bytes));
    }
    PML_CUDA(cudaStreamSynchronize(s));
    *tiled = true;
    return PML_OK;
}

// Chunk metadata of the CSR stream for row-pass variant 1: marks the first entry of every
// non-empty row in bit 31 of its (internal) column id, lists the non-empty rows when some rows
// are empty, and builds chunk_lo / fix exactly as for the CSC side.
int build_row_side(cudaStream_t s, int64_t n, int64_t nnz, const uint32_t *row_ptr, int32_t *col, int chunk_entries,
                   RowSide *out)
{
    out->n_chunks = (int)((nnz + chunk_entries - 1) / chunk_entries);
    const int n_chunks = out->n_chunks;
    out->bytes = 0;
    PML_CUDA(cudaMalloc(&out->chunk_lo, sizeof(int32_t) * (size_t)(n_chunks + 1)));
    out->bytes += (int64_t)sizeof(int32_t) * (n_chunks + 1);
    DevBuf flag, pos, run_ptr, head;
    PML_CUDA(flag.alloc(sizeof(int32_t) * (size_t)(n + 1)));
    PML_CUDA(pos.alloc(sizeof(int32_t) * (size_t)(n + 1)));
    PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(n_chunks + 1)));
    // flag[i] = row i non-empty (same kernel as for columns: it only compares neighbouring pointers)
    nonempty_rows_kernel<<<grid_for(n), 256, 0, s>>>(n, row_ptr, flag.as<int32_t>());
    PML_CUDA(cudaMemsetAsync(flag.as<int32_t>() + n, 0, sizeof(int32_t), s));
    PML_CUDA(cudaGetLastError());
    {
        DevBuf t;
        size_t tb = 0;
        PML_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, flag.as<int32_t>(), pos.as<int32_t>(), n + 1, s));
        PML_CUDA(t.alloc(tb));
        PML_CUDA(cub::DeviceScan::ExclusiveSum(t.p, tb, flag.as<int32_t>(), pos.as<int32_t>(), n + 1, s));
    }
    int32_t runs32 = 0;
    PML_CUDA(cudaMemcpyAsync(&runs32, pos.as<int32_t>() + n, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    out->n_runs = runs32;
    const int64_t runs = out->n_runs;
    const uint32_t *run_ptr_dev = row_ptr;
    if (runs < n) {
        PML_CUDA(cudaMalloc(&out->nzrow_id, sizeof(int32_t) * (size_t)(runs + 1)));
        out->bytes += (int64_t)sizeof(int32_t) * (runs + 1);
        PML_CUDA(run_ptr.alloc(sizeof(uint32_t) * (size_t)(runs + 1)));
        compact_rows_kernel<<<grid_for(n), 256, 0, s>>>(n, row_ptr, flag.as<int32_t>(), pos.as<int32_t>(), nnz, runs,
                                                       run_ptr.as<uint32_t>(), out->nzrow_id);
        PML_CUDA(cudaGetLastError());
        run_ptr_dev = run_ptr.as<uint32_t>();
    }
    if (nnz > 0) mark_starts_kernel<<<grid_for(runs), 256, 0, s>>>(runs, run_ptr_dev, col);
    chunk_meta_kernel<<<grid_for((int64_t)n_chunks + 1), 256, 0, s>>>(n_chunks, chunk_entries, runs, run_ptr_dev,
                                                                     out->chunk_lo, head.as<int32_t>());
    PML_CUDA(cudaGetLastError());
    PML_TRY(build_fix_list(s, n_chunks, out->chunk_lo, head.as<int32_t>(), out->nzrow_id, runs, &out->fix, &out->n_fix,
                           &out->n_fix_long, &out->bytes));
    return PML_OK;
}

// ------------------------------------------------------------------------------------------
// Sliced-ELL copy of the CSR rows for the one-lane-per-row margin pass (sell_rows_kernel): 32 consecutive rows make a
// slice, padded to the longest of them (rounded up to 4 entries); inside a slice the entries are stored
// [quad of positions][row][4], so a warp's 128-bit loads are coalesced while lane l walks row l.  Rows keep their
// ascending internal column order, which puts the same few hottest coefficients at the same position of most rows: the
// warp's table look-ups at one position are then broadcasts or neighbours, not bank conflicts.  Stored ids are
// column + 1; 0 is padding and selects the zero slot of the table.  Not built when padding would cost more than
// `max_overhead` of the entries (rows of very uneven length): the entry-chunked stream then serves the pass.
// ------------------------------------------------------------------------------------------
__global__ void sell_keys_kernel(int64_t n, int sigma_shift, const uint32_t *__restrict__ row_ptr, uint64_t *__restrict__ keys,
                                 int32_t *__restrict__ rows)
{
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t len = row_ptr[r + 1] - row_ptr[r];
        keys[r] = ((uint64_t)(r >> sigma_shift) << 20) | (len < 0xfffffu ? len : 0xfffffu);
        rows[r] = (int32_t)r;
    }
}

__global__ void sell_width_kernel(int64_t n, int64_t n_slices, const uint32_t *__restrict__ row_ptr, const int32_t *__restrict__ perm,
                                  int64_t *__restrict__ entries)
{
    for (int64_t sl = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; sl <= n_slices; sl += (int64_t)gridDim.x * blockDim.x) {
        if (sl == n_slices) {
            entries[sl] = 0;
            continue;
        }
        const int64_t r0 = sl << 5, r1 = r0 + 32 < n ? r0 + 32 : n;
        uint32_t w = 0;
        for (int64_t i = r0; i < r1; ++i) {
            const int64_t r = perm[i];
            const uint32_t len = row_ptr[r + 1] - row_ptr[r];
            w = len > w ? len : w;
        }
        entries[sl] = 32ll * ((w + 3u) & ~3u);
    }
}

struct NoValue {};   // indicator features: only the indices are stored

// n_hot > 0: the entries of a row are dealt to its positions so that the table look-ups of a warp instruction --
// the same position of 32 rows -- fall into different shared-memory banks: stored id k <= n_hot lives in table slot k,
// 8 bytes wide, i.e. bank pair k mod 16, and a 64-bit shared load is served half a warp at a time.  The row in lane l
// puts at position p an entry of class (l + p) mod 16 if it has one left, else an entry beyond the table (no
// shared-memory access at all); what is left over afterwards fills the holes (those may conflict).  A row's sum is
// taken in the stored order, which is fixed at pack time.
template <typename VT>
__global__ void sell_fill_kernel(int64_t n, int64_t n_slices, const uint32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                 const VT *__restrict__ val, const int32_t *__restrict__ perm, const int64_t *__restrict__ base,
                                 int32_t n_hot, int32_t *__restrict__ idx, VT *__restrict__ out_val)
{
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t sl = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; sl < n_slices; sl += warps) {
        const int64_t b = base[sl];
        const int quads = (int)((base[sl + 1] - b) >> 7);
        const int64_t at_row = (sl << 5) + lane;
        const int64_t row = at_row < n ? perm[at_row] : -1;
        const int64_t start = row >= 0 ? row_ptr[row] : 0;
        const int len = row >= 0 ? (int)(row_ptr[row + 1] - row_ptr[row]) : 0;
        if (n_hot <= 0) {
            for (int q = 0; q < quads; ++q) {
                const int64_t at = b + ((int64_t)q * 32 + lane) * 4;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int pos = 4 * q + j;
                    const bool in = pos < len;
                    idx[at + j] = in ? (col[start + pos] & 0x7fffffff) + 1 : 0;
                    if constexpr (!std::is_same<VT, NoValue>::value) out_val[at + j] = in ? val[start + pos] : (VT)0;
                }
            }
            continue;
        }
        auto klass = [&](int i) {
            const int32_t k = (col[start + i] & 0x7fffffff) + 1;
            return k <= n_hot ? (k & 15) : 16;
        };
        int cur[17];      // first entry of every class that has not been placed yet (len: none left)
        for (int c = 0; c < 17; ++c) cur[c] = len;
        for (int i = len - 1; i >= 0; --i) cur[klass(i)] = i;
        auto advance = [&](int c) {
            int i = cur[c] + 1;
            while (i < len && klass(i) != c) ++i;
            cur[c] = i;
        };
        auto put = [&](int pos, int i) {
            const int64_t at = b + ((int64_t)(pos >> 2) * 32 + lane) * 4 + (pos & 3);
            idx[at] = i < len ? (col[start + i] & 0x7fffffff) + 1 : 0;
            if constexpr (!std::is_same<VT, NoValue>::value) out_val[at] = i < len ? val[start + i] : (VT)0;
        };
        const int width = 4 * quads;
        int holes = 0;
        for (int pos = 0; pos < width; ++pos) {
            int c = (lane + pos) & 15;
            if (cur[c] >= len) c = 16;
            const int i = cur[c];
            put(pos, i);
            if (i < len) advance(c);
            else ++holes;
        }
        if (holes > 0) {
            int c = 0;
            for (int pos = 0; pos < width; ++pos) {
                const int64_t at = b + ((int64_t)(pos >> 2) * 32 + lane) * 4 + (pos & 3);
                if (idx[at] != 0) continue;
                while (c < 17 && cur[c] >= len) ++c;
                if (c == 17) break;
                put(pos, cur[c]);
                advance(c);
            }
        }
    }
}

int build_sell(cudaStream_t s, int64_t n, int64_t nnz, const uint32_t *row_ptr, const int32_t *col, const void *val,
               int value_bytes, double max_overhead, int sigma_shift, int32_t n_hot, SellSide *out)
{
    *out = SellSide();
    if (n <= 0 || nnz <= 0 || n > 0x7fffffffll) return PML_OK;
    const int64_t n_slices = (n + 31) >> 5;
    // rows sorted by length inside windows of 2^sigma_shift rows (stable: rows of one length keep their order), so that
    // the 32 rows of a slice are about equally long and the sums still land in a small window of the output
    int32_t *perm = nullptr;
    {
        DevBuf keys, keys_out, rows, t;
        PML_CUDA(keys.alloc(sizeof(uint64_t) * (size_t)n));
        PML_CUDA(keys_out.alloc(sizeof(uint64_t) * (size_t)n));
        PML_CUDA(rows.alloc(sizeof(int32_t) * (size_t)n));
        PML_CUDA(cudaMalloc(&perm, sizeof(int32_t) * (size_t)n));
        sell_keys_kernel<<<grid_for(n), 256, 0, s>>>(n, sigma_shift, row_ptr, keys.as<uint64_t>(), rows.as<int32_t>());
        size_t tb = 0;
        cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, tb, keys.as<uint64_t>(), keys_out.as<uint64_t>(), rows.as<int32_t>(),
                                                        perm, n, 0, 64, s);
        if (e == cudaSuccess) e = t.alloc(tb);
        if (e == cudaSuccess)
            e = cub::DeviceRadixSort::SortPairs(t.p, tb, keys.as<uint64_t>(), keys_out.as<uint64_t>(), rows.as<int32_t>(), perm, n,
                                                0, 64, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) {
            cudaFree(perm);
            PML_CUDA(e);
        }
    }
    DevBuf entries;
    int64_t *base = nullptr;
    auto drop = [&] {
        cudaFree(base);
        cudaFree(perm);
    };
    if (entries.alloc(sizeof(int64_t) * (size_t)(n_slices + 1)) != cudaSuccess ||
        cudaMalloc(&base, sizeof(int64_t) * (size_t)(n_slices + 1)) != cudaSuccess) {
        drop();
        return fail(PML_ENOMEM, "sliced-ELL slice table");
    }
    sell_width_kernel<<<grid_for(n_slices + 1), 256, 0, s>>>(n, n_slices, row_ptr, perm, entries.as<int64_t>());
    {
        DevBuf t;
        size_t tb = 0;
        cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, tb, entries.as<int64_t>(), base, n_slices + 1, s);
        if (e == cudaSuccess) e = t.alloc(tb);
        if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(t.p, tb, entries.as<int64_t>(), base, n_slices + 1, s);
        int64_t len = 0;
        if (e == cudaSuccess) e = cudaMemcpyAsync(&len, base + n_slices, sizeof(int64_t), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) {
            drop();
            PML_CUDA(e);
        }
        out->len = len;
    }
    if ((double)out->len > (1.0 + max_overhead) * (double)nnz) {   // too much padding: leave the pass to the entry-chunked stream
        drop();
        out->len = 0;
        return PML_OK;
    }
    int32_t *idx = nullptr;
    void *v = nullptr;
    cudaError_t e = cudaMalloc(&idx, sizeof(int32_t) * (size_t)(out->len + 4));
    if (e == cudaSuccess && value_bytes > 0) e = cudaMalloc(&v, (size_t)value_bytes * (size_t)(out->len + 4));
    if (e != cudaSuccess) {                                          // no room for a second copy of the rows: not an error
        cudaGetLastError();
        cudaFree(idx);
        drop();
        out->len = 0;
        return PML_OK;
    }
    const int grid = (int)(n_slices < 148 * 64 ? (n_slices + 7) / 8 : 148 * 8);
    if (value_bytes == 0)
        sell_fill_kernel<NoValue><<<grid, 256, 0, s>>>(n, n_slices, row_ptr, col, nullptr, perm, base, n_hot, idx, nullptr);
    else if (value_bytes == 4)
        sell_fill_kernel<float><<<grid, 256, 0, s>>>(n, n_slices, row_ptr, col, static_cast<const float *>(val), perm, base,
                                                     n_hot, idx, static_cast<float *>(v));
    else
        sell_fill_kernel<double><<<grid, 256, 0, s>>>(n, n_slices, row_ptr, col, static_cast<const double *>(val), perm, base,
                                                      n_hot, idx, static_cast<double *>(v));
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        cudaFree(idx);
        cudaFree(v);
        drop();
        PML_CUDA(e);
    }
    out->idx = idx;
    out->val = v;
    out->base = base;
    out->perm = perm;
    out->n_slices = n_slices;
    out->bytes = (int64_t)sizeof(int64_t) * (n_slices + 1) + (int64_t)(sizeof(int32_t) + value_bytes) * (out->len + 4) +
                 (int64_t)sizeof(int32_t) * n;
    return PML_OK;
}

}  // namespace pml
