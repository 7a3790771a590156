// pack.cu -- one-off packing of LabeledPoint rows (lib/data/LabeledPoint.scala:32-36) into the
// device-resident layout the evaluation kernels read: the CSR arrays as given, plus a CSC copy
// ordered by (column, row) with chunk metadata for col_pass_kernel.
//
// This runs once per data set (and never per evaluation); it uses CUB's device radix sort and
// scan for the transposition.  The reference has no counterpart -- it re-reads JVM objects every
// evaluation -- so nothing here mirrors reference code.
#include <cub/cub.cuh>
#include <cuda_runtime.h>

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
        csc_val[k] = val[src];
    }
}

// col_ptr[j] = first position in the sorted column keys with key >= j   (j in [0, d])
__global__ void col_ptr_kernel(int32_t d, int64_t nnz, const int32_t *__restrict__ sorted_col, uint32_t *__restrict__ col_ptr)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j <= d; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t lo = 0, hi = nnz;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (sorted_col[mid] < (int32_t)j) lo = mid + 1;
            else hi = mid;
        }
        col_ptr[j] = (uint32_t)lo;
    }
}

__global__ void nonempty_flag_kernel(int32_t d, const uint32_t *__restrict__ col_ptr, int32_t *__restrict__ flag)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x)
        flag[j] = col_ptr[j + 1] > col_ptr[j] ? 1 : 0;
}

__global__ void compact_cols_kernel(int32_t d, const uint32_t *__restrict__ col_ptr, const int32_t *__restrict__ flag,
                                    const int32_t *__restrict__ pos, int64_t nnz, int64_t m,
                                    uint32_t *__restrict__ nzcol_ptr, int32_t *__restrict__ nzcol_id)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < d; j += (int64_t)gridDim.x * blockDim.x) {
        if (flag[j]) {
            nzcol_ptr[pos[j]] = col_ptr[j];
            nzcol_id[pos[j]] = (int32_t)j;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) nzcol_ptr[m] = (uint32_t)nnz;
}

// chunk_col_lo[c] = number of non-empty columns that start before entry c*TCHUNK;
// head_nonempty[c] = 1 when entry c*TCHUNK is not itself a column start
__global__ void chunk_meta_kernel(int n_chunks, int64_t m, const uint32_t *__restrict__ nzcol_ptr,
                                  int32_t *__restrict__ chunk_col_lo, int32_t *__restrict__ head_nonempty)
{
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= n_chunks; c += gridDim.x * blockDim.x) {
        if (c == n_chunks) {
            chunk_col_lo[c] = (int32_t)m;
            continue;
        }
        const uint32_t base = (uint32_t)((int64_t)c * TCHUNK);
        int64_t lo = 0, hi = m;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (nzcol_ptr[mid] < base) lo = mid + 1;
            else hi = mid;
        }
        chunk_col_lo[c] = (int32_t)lo;
        head_nonempty[c] = (lo < m && nzcol_ptr[lo] == base) ? 0 : 1;
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

int build_transpose(cudaStream_t s, int64_t n, int32_t d, int64_t nnz, const uint32_t *row_ptr,
                    const int32_t *col, const void *val, int value_bytes, Packed *out)
{
    out->n_chunks = (int)((nnz + TCHUNK - 1) / TCHUNK);
    const int n_chunks = out->n_chunks;
    out->bytes = 0;

    PML_CUDA(cudaMalloc(&out->csc_row, sizeof(int32_t) * (size_t)(nnz + 4)));
    PML_CUDA(cudaMalloc(&out->csc_val, (size_t)value_bytes * (size_t)(nnz + 4)));
    PML_CUDA(cudaMalloc(&out->chunk_col_lo, sizeof(int32_t) * (size_t)(n_chunks + 1)));
    out->bytes += (int64_t)(sizeof(int32_t) + value_bytes) * (nnz + 4) + (int64_t)sizeof(int32_t) * (n_chunks + 1);

    DevBuf entry_row, perm_in, perm_out, keys_out, col_ptr, flag, pos, head, tmp;
    PML_CUDA(entry_row.alloc(sizeof(int32_t) * (size_t)nnz));
    PML_CUDA(perm_in.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(perm_out.alloc(sizeof(uint32_t) * (size_t)nnz));
    PML_CUDA(keys_out.alloc(sizeof(int32_t) * (size_t)nnz));
    PML_CUDA(col_ptr.alloc(sizeof(uint32_t) * (size_t)(d + 1)));
    PML_CUDA(flag.alloc(sizeof(int32_t) * (size_t)(d + 1)));
    PML_CUDA(pos.alloc(sizeof(int32_t) * (size_t)(d + 1)));
    PML_CUDA(head.alloc(sizeof(int32_t) * (size_t)(n_chunks + 1)));

    if (nnz > 0) {
        expand_rows_kernel<<<grid_for(n * 32), 256, 0, s>>>(n, row_ptr, entry_row.as<int32_t>());
        iota_kernel<<<grid_for(nnz), 256, 0, s>>>(nnz, perm_in.as<uint32_t>());
        PML_CUDA(cudaGetLastError());
        int end_bit = 1;
        while (end_bit < 31 && (1LL << end_bit) < (int64_t)d) ++end_bit;
        size_t tmp_bytes = 0;
        // stable LSD radix sort by column keeps rows ascending inside each column
        PML_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, col, keys_out.as<int32_t>(), perm_in.as<uint32_t>(),
                                                 perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
        PML_CUDA(tmp.alloc(tmp_bytes));
        PML_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, col, keys_out.as<int32_t>(), perm_in.as<uint32_t>(),
                                                 perm_out.as<uint32_t>(), nnz, 0, end_bit, s));
        if (value_bytes == 4)
            permute_kernel<float><<<grid_for(nnz), 256, 0, s>>>(nnz, perm_out.as<uint32_t>(), entry_row.as<int32_t>(),
                                                               (const float *)val, out->csc_row, (float *)out->csc_val);
        else
            permute_kernel<double><<<grid_for(nnz), 256, 0, s>>>(nnz, perm_out.as<uint32_t>(), entry_row.as<int32_t>(),
                                                                (const double *)val, out->csc_row, (double *)out->csc_val);
        PML_CUDA(cudaGetLastError());
    }
    col_ptr_kernel<<<grid_for((int64_t)d + 1), 256, 0, s>>>(d, nnz, keys_out.as<int32_t>(), col_ptr.as<uint32_t>());
    nonempty_flag_kernel<<<grid_for(d), 256, 0, s>>>(d, col_ptr.as<uint32_t>(), flag.as<int32_t>());
    PML_CUDA(cudaGetLastError());
    {
        DevBuf tmp2;
        size_t tmp_bytes = 0;
        PML_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag.as<int32_t>(), pos.as<int32_t>(), d + 1, s));
        PML_CUDA(tmp2.alloc(tmp_bytes));
        PML_CUDA(cudaMemsetAsync(flag.as<int32_t>() + d, 0, sizeof(int32_t), s));
        PML_CUDA(cub::DeviceScan::ExclusiveSum(tmp2.p, tmp_bytes, flag.as<int32_t>(), pos.as<int32_t>(), d + 1, s));
        int32_t m32 = 0;
        PML_CUDA(cudaMemcpyAsync(&m32, pos.as<int32_t>() + d, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
        PML_CUDA(cudaStreamSynchronize(s));
        out->m = m32;
    }
    const int64_t m = out->m;
    PML_CUDA(cudaMalloc(&out->nzcol_ptr, sizeof(uint32_t) * (size_t)(m + 1)));
    PML_CUDA(cudaMalloc(&out->nzcol_id, sizeof(int32_t) * (size_t)(m + 1)));
    out->bytes += (int64_t)(sizeof(uint32_t) + sizeof(int32_t)) * (m + 1);
    compact_cols_kernel<<<grid_for(d), 256, 0, s>>>(d, col_ptr.as<uint32_t>(), flag.as<int32_t>(), pos.as<int32_t>(), nnz, m,
                                                   out->nzcol_ptr, out->nzcol_id);
    chunk_meta_kernel<<<grid_for((int64_t)n_chunks + 1), 256, 0, s>>>(n_chunks, m, out->nzcol_ptr, out->chunk_col_lo,
                                                                     head.as<int32_t>());
    PML_CUDA(cudaGetLastError());

    // fix-up list: runs of consecutive chunks whose heads continue one column (host-side, tiny)
    std::vector<int32_t> h_lo((size_t)n_chunks + 1), h_head((size_t)n_chunks + 1, 0);
    PML_CUDA(cudaMemcpyAsync(h_lo.data(), out->chunk_col_lo, sizeof(int32_t) * (size_t)(n_chunks + 1),
                             cudaMemcpyDeviceToHost, s));
    if (n_chunks > 0)
        PML_CUDA(cudaMemcpyAsync(h_head.data(), head.p, sizeof(int32_t) * (size_t)n_chunks, cudaMemcpyDeviceToHost, s));
    PML_CUDA(cudaStreamSynchronize(s));
    std::vector<int32_t> fix_ord;  // column ordinal per fix entry
    std::vector<FixEntry> fix;
    for (int c = 1; c < n_chunks; ++c) {
        if (!h_head[c]) continue;
        const int32_t ord = h_lo[c] - 1;  // the column that was open when chunk c began
        if (!fix.empty() && fix_ord.back() == ord && fix.back().chunk_end == c) {
            fix.back().chunk_end = c + 1;
        } else {
            FixEntry e;
            e.col = ord;  // replaced by the column id below
            e.chunk_begin = c;
            e.chunk_end = c + 1;
            fix.push_back(e);
            fix_ord.push_back(ord);
        }
    }
    out->n_fix = (int)fix.size();
    if (!fix.empty()) {
        std::vector<int32_t> h_ids((size_t)m);
        // only a handful of ids are needed, but m*4 bytes once at pack time is simpler than a gather
        PML_CUDA(cudaMemcpyAsync(h_ids.data(), out->nzcol_id, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, s));
        PML_CUDA(cudaStreamSynchronize(s));
        for (size_t q = 0; q < fix.size(); ++q) fix[q].col = h_ids[(size_t)fix_ord[q]];
        PML_CUDA(cudaMalloc(&out->fix, sizeof(FixEntry) * fix.size()));
        PML_CUDA(cudaMemcpyAsync(out->fix, fix.data(), sizeof(FixEntry) * fix.size(), cudaMemcpyHostToDevice, s));
        PML_CUDA(cudaStreamSynchronize(s));
        out->bytes += (int64_t)sizeof(FixEntry) * (int64_t)fix.size();
    }
    return PML_OK;
}

}  // namespace pml
