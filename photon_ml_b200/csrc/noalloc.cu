// noalloc.cu -- objective VALUE over caller-owned device arrays, without allocating anything.
//
// Who needs this.  Two kinds of Photon ML callers reach the GPU from places where cudaMalloc /
// cudaFree are off the table:
//   * the JNI shim while it holds Java arrays through GetPrimitiveArrayCritical: inside such a
//     critical region the JVM may have paused garbage collection for every thread, and cudaMalloc
//     synchronises the whole device and can block for a long time -- so nothing that may
//     allocate is allowed between Get...Critical and Release...Critical;
//   * Spark executor task threads that evaluate many small per-entity (random-effect) problems
//     concurrently on one GPU: a cudaMalloc/cudaFree pair in one task stalls the kernels of all
//     the others, because both calls serialise the device.
// These callers already keep their rows in device memory; what they cannot do is create a
// pml_dataset (which allocates the CSC copy and scratch).  For them this file evaluates
// Sum_i wt_i * l(z_i, y_i) -- DistributedGLMLossFunction.value for unnormalised data -- straight
// from their CSR arrays, with the only scratch the row kernel needs (one partial sum per CTA and
// a completion ticket) living in a fixed __device__ array that is part of the library image.
//
// The price of a fixed array: a hard row limit, and every call on a device shares the same
// scratch (see the comments on pml_value_noalloc_dev below).
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "internal.h"
#include "pointwise.cuh"

namespace pml {
namespace {

constexpr int NA_THREADS = 256;
constexpr int NA_WARPS = NA_THREADS / 32;
constexpr int NA_ROWS_PER_CTA = NA_WARPS * 32;      // each warp owns 32 rows, one CTA = 256 rows
constexpr int NA_MAX_CTAS = 65536;
constexpr int64_t NA_MAX_ROWS = (int64_t)NA_MAX_CTAS * NA_ROWS_PER_CTA;  // 16,777,216 rows
constexpr int NA_LANES_PER_ROW = 8;
constexpr int NA_MAX_DEVICES = 64;

// Scratch layout (doubles): [0, NA_MAX_CTAS) per-CTA partial sums; slot NA_MAX_CTAS holds the
// completion ticket (read as an unsigned int).  512 KB + 256 B, resident on every GPU that loads
// this library, until the process exits.
constexpr int NA_TICKET_SLOT = NA_MAX_CTAS;
constexpr int NA_SCRATCH_DOUBLES = NA_MAX_CTAS + 32;
static __device__ __align__(256) double g_noalloc_scratch[NA_SCRATCH_DOUBLES];

// One CTA = 256 consecutive rows; sub-warps of 8 lanes reduce one row each (the same shape as
// row_pass_kernel<.,8,.> in kernels.cu, reading int64 row pointers as the caller holds them).
// The CTA's partial goes to scratch[blockIdx.x]; the CTA that draws the last ticket adds the
// partials up in index order, so the result does not depend on scheduling.
template <typename VT>
__global__ void __launch_bounds__(NA_THREADS)
value_rows_noalloc_kernel(int64_t n, const int64_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                          const VT *__restrict__ val, const double *__restrict__ label,
                          const double *__restrict__ offset, const double *__restrict__ weight, int loss,
                          const double *__restrict__ coef, double *scratch, double *__restrict__ out_value)
{
    __shared__ double red[2 * NA_WARPS];
    __shared__ unsigned int ticket;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int RPP = 32 / NA_LANES_PER_ROW;
    const int sub = lane / NA_LANES_PER_ROW, sl = lane % NA_LANES_PER_ROW;
    const int64_t row0 = ((int64_t)blockIdx.x * NA_WARPS + warp) * 32;

    double z_mine = 0.0;
#pragma unroll 1
    for (int pass = 0; pass < NA_LANES_PER_ROW; ++pass) {
        const int64_t row = row0 + pass * RPP + sub;
        double a = 0.0;
        if (row < n) {
            const int64_t end = row_ptr[row + 1];
            for (int64_t k = row_ptr[row] + sl; k < end; k += NA_LANES_PER_ROW)
                a = fma(widen(val[k]), __ldg(coef + col[k]), a);
        }
#pragma unroll
        for (int o = NA_LANES_PER_ROW / 2; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        const double got = __shfl_sync(0xffffffffu, a, (lane % RPP) * NA_LANES_PER_ROW);
        if (lane / RPP == pass) z_mine = got;
    }

    double f = 0.0, unused = 0.0;
    const int64_t my_row = row0 + lane;
    if (my_row < n) {
        const double z = z_mine + (offset ? offset[my_row] : 0.0);
        double l, dl;
        loss_and_dz(loss, z, label[my_row], l, dl);
        f = __dmul_rn(weight ? weight[my_row] : 1.0, l);
    }
    block_sum2<NA_WARPS>(f, unused, red);

    unsigned int *ticket_word = reinterpret_cast<unsigned int *>(scratch + NA_TICKET_SLOT);
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = f;
        __threadfence();
        ticket = atomicAdd(ticket_word, 1u);
    }
    __syncthreads();
    if (ticket != gridDim.x - 1) return;

    // last CTA: every partial is visible (each writer fenced before taking its ticket)
    __threadfence();
    double total = 0.0;
    unused = 0.0;
    const volatile double *partial = scratch;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += NA_THREADS) total += partial[i];
    block_sum2<NA_WARPS>(total, unused, red);
    if (threadIdx.x == 0) *out_value = total;
}

std::mutex g_setup_mu;
double *g_scratch_on_device[NA_MAX_DEVICES] = { nullptr };

// The array's name is only a handle in host code; the address kernels can use has to be asked
// for, once per device, while that device is current.
int scratch_pointer(int device, double **out)
{
    std::lock_guard<std::mutex> lock(g_setup_mu);
    if (g_scratch_on_device[device] == nullptr) {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return fail(PML_ECUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        g_scratch_on_device[device] = p;
    }
    *out = g_scratch_on_device[device];
    return PML_OK;
}

}  // namespace
}  // namespace pml

using namespace pml;

extern "C" int pml_value_noalloc_dev(int device, void *stream, int64_t n, int32_t d, const int64_t *row_ptr_dev,
                                     const int32_t *col_idx_dev, const void *values_dev, int values_dtype,
                                     const double *labels_dev, const double *offsets_dev,
                                     const double *weights_dev, int loss, const double *coef_dev,
                                     double *out_value_dev)
{
    if (n < 0 || d < 0) return fail(PML_EINVAL, "negative size");
    if (n > NA_MAX_ROWS)
        return fail(PML_EINVAL, "pml_value_noalloc_dev handles at most " + std::to_string(NA_MAX_ROWS) +
                                    " rows (fixed scratch); got " + std::to_string(n) +
                                    ". Pack the rows with pml_dataset_create_csr_dev instead.");
    if (loss < PML_LOSS_LOGISTIC || loss > PML_LOSS_SMOOTHED_HINGE) return fail(PML_EINVAL, "unknown loss");
    if (values_dtype != PML_F32 && values_dtype != PML_F64) return fail(PML_EINVAL, "values_dtype must be PML_F32 or PML_F64");
    if (!out_value_dev) return fail(PML_EINVAL, "null out_value_dev");
    if (n > 0 && (!row_ptr_dev || !labels_dev || !coef_dev)) return fail(PML_EINVAL, "null input array");
    if (device < 0 || device >= NA_MAX_DEVICES) return fail(PML_EINVAL, "device index out of range");
    // the calling thread gets its current device back on every way out
    struct DeviceRestore {
        int prev = -1;
        ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
    } restore;
    int current = -1;
    if (cudaGetDevice(&current) == cudaSuccess && current != device) restore.prev = current;
    PML_CUDA(cudaSetDevice(device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);

    double *scratch = nullptr;
    PML_TRY(scratch_pointer(device, &scratch));

    if (n == 0) {
        PML_CUDA(cudaMemsetAsync(out_value_dev, 0, sizeof(double), s));
        return PML_OK;
    }
    // the ticket must start at zero; the partial slots are overwritten before they are read
    PML_CUDA(cudaMemsetAsync(scratch + NA_TICKET_SLOT, 0, sizeof(double), s));
    const int grid = (int)((n + NA_ROWS_PER_CTA - 1) / NA_ROWS_PER_CTA);
    if (values_dtype == PML_F32)
        value_rows_noalloc_kernel<float><<<grid, NA_THREADS, 0, s>>>(n, row_ptr_dev, col_idx_dev,
                                                                     static_cast<const float *>(values_dev), labels_dev,
                                                                     offsets_dev, weights_dev, loss, coef_dev, scratch,
                                                                     out_value_dev);
    else
        value_rows_noalloc_kernel<double><<<grid, NA_THREADS, 0, s>>>(n, row_ptr_dev, col_idx_dev,
                                                                      static_cast<const double *>(values_dev), labels_dev,
                                                                      offsets_dev, weights_dev, loss, coef_dev, scratch,
                                                                      out_value_dev);
    PML_CUDA(cudaGetLastError());
    return PML_OK;
}
