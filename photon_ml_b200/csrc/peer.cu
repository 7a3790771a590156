// peer.cu -- the `merge` step across GPUs (ValueAndGradientAggregator.merge :161-176 as driven by treeAggregate,
// :243-247) fused with getVector (:199-221), over NVLink peer memory, one process per GPU.
//
// Every rank owns a slab [header | raw (d+2, padded) | result (d+1, padded)] allocated with cudaMalloc and opened by all the other
// ranks through CUDA IPC.  After a rank's evaluation kernels have left its raw sums [vectorSum(d), valueSum,
// prefactorSum] in its slab, ONE kernel per rank does the whole exchange:
//   1. barrier: tell every peer "my raw sums are complete", wait until every peer said so;
//   2. rank r adds up slice r of the vector over all ranks' slabs (peer loads, ranks in a fixed order, so every
//      element is summed once, by one rank, in one order: the result is bit-identical on all ranks and from run to
//      run), applies getVector to it (the total prefactorSum is 8 bytes per rank, read by everyone) and stores the
//      finished slice into EVERY rank's result (peer stores);
//   3. barrier: "my stores are out", wait for everyone's -- when the kernel ends, the rank's result is complete and
//      nobody reads its raw sums any more.
// Per rank that is (N-1)/N of the vector in and out over NVLink and two flag round trips -- no staging copies, no
// separate finalize launch, nothing on the host between the evaluation and the finished gradient.  The epoch the
// flags count lives in the slab and is advanced by the kernel itself, so the launch takes no per-call argument and
// can sit in a CUDA graph behind the evaluation kernels.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstring>
#include <string>

#include "internal.h"

namespace pml {
namespace {

constexpr int PEER_THREADS = 512;
constexpr size_t PEER_HEADER_BYTES = 1024;
constexpr long long PEER_TIMEOUT_CYCLES = 60000000000ll;  // ~30 s: a peer that never arrives must not hang the GPU

struct PeerHeader {
    unsigned int ready[PML_MAX_PEERS];   // ready[q] = epoch: rank q's raw sums of that epoch are complete
    unsigned int pad0[32 - PML_MAX_PEERS];
    unsigned int stored[PML_MAX_PEERS];  // stored[q] = epoch: rank q has stored its slice into this rank's result
    unsigned int pad1[32 - PML_MAX_PEERS];
    unsigned int epoch;                  // exchanges completed
    unsigned int arrived;                // CTAs of the running kernel that have finished their stores
};
static_assert(sizeof(PeerHeader) <= PEER_HEADER_BYTES, "header does not fit");

struct PeerParams {
    char *slab[PML_MAX_PEERS];
    int rank, world;
    int64_t d;
    const double *factors, *shifts;
    int *status;  // mapped host word: set to 1 when a barrier times out
};

__device__ __forceinline__ void st_release_sys(unsigned int *p, unsigned int v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int *p)
{
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ double ld_peer(const double *p)
{
    double v;  // straight to L2 / the link: a line of a peer's slab must never be served from a stale L1 copy
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ double2 ld_peer2(const double *p)
{
    double2 v;
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void wait_for(const unsigned int *flag, unsigned int epoch, int *status)
{
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(flag) - epoch) < 0) {
        if (clock64() - t0 > PEER_TIMEOUT_CYCLES) {
            *status = 1;
            break;
        }
        __nanosleep(100);
    }
}

// raw and result are both padded to an even number of doubles: everything moves as 16-byte pairs
__host__ __device__ __forceinline__ int64_t slab_stride(int64_t d) { return (d + 3) & ~(int64_t)1; }
__device__ __forceinline__ double *slab_raw(char *slab) { return reinterpret_cast<double *>(slab + PEER_HEADER_BYTES); }
__device__ __forceinline__ double *slab_result(char *slab, int64_t d)
{
    return reinterpret_cast<double *>(slab + PEER_HEADER_BYTES) + slab_stride(d);
}

template <int WORLD>
__global__ void __launch_bounds__(PEER_THREADS) peer_merge_kernel(const PeerParams p)
{
    __shared__ int is_last;
    PeerHeader *mine = reinterpret_cast<PeerHeader *>(p.slab[p.rank]);
    const unsigned int epoch = mine->epoch + 1;
    const int64_t d = p.d;

    // ---- 1. everyone's raw sums are complete ----
    if (blockIdx.x == 0 && threadIdx.x < WORLD)
        st_release_sys(&reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->ready[p.rank], epoch);
    if (threadIdx.x < WORLD) wait_for(&mine->ready[threadIdx.x], epoch, p.status);
    __syncthreads();

    // ---- 2. this rank's slice: sum over ranks, getVector, store to every rank ----
    double P = 0.0;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) P += ld_peer(slab_raw(p.slab[q]) + d + 1);
    // elements [0, d] (slot d is the value) as pairs; a trailing pair may take the prefactor slot d+1 along (unused)
    const int64_t n_pairs = (d + 2) / 2;
    const int64_t per = (n_pairs + WORLD - 1) / WORLD;
    const int64_t lo = per * p.rank;
    const int64_t hi = lo + per < n_pairs ? lo + per : n_pairs;
    for (int64_t k = lo + (int64_t)blockIdx.x * PEER_THREADS + threadIdx.x; k < hi; k += (int64_t)gridDim.x * PEER_THREADS) {
        const int64_t i = 2 * k;
        double2 part[WORLD];
#pragma unroll
        for (int q = 0; q < WORLD; ++q) part[q] = ld_peer2(slab_raw(p.slab[q]) + i);
        double2 v = make_double2(0.0, 0.0);
#pragma unroll
        for (int q = 0; q < WORLD; ++q) {
            v.x += part[q].x;
            v.y += part[q].y;
        }
        if (i < d) {
            if (p.shifts) v.x = __dsub_rn(v.x, __dmul_rn(p.shifts[i], P));
            if (p.factors) v.x = __dmul_rn(v.x, p.factors[i]);
        }
        if (i + 1 < d) {
            if (p.shifts) v.y = __dsub_rn(v.y, __dmul_rn(p.shifts[i + 1], P));
            if (p.factors) v.y = __dmul_rn(v.y, p.factors[i + 1]);
        }
#pragma unroll
        for (int q = 0; q < WORLD; ++q) *reinterpret_cast<double2 *>(slab_result(p.slab[q], d) + i) = v;
    }

    // ---- 3. everyone's stores are out ----
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&mine->arrived, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!is_last) return;
    __threadfence_system();
    if (threadIdx.x < WORLD)
        st_release_sys(&reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->stored[p.rank], epoch);
    if (threadIdx.x < WORLD) wait_for(&mine->stored[threadIdx.x], epoch, p.status);
    __syncthreads();
    if (threadIdx.x == 0) {
        mine->arrived = 0;
        mine->epoch = epoch;
    }
}

}  // namespace

int launch_peer_merge(cudaStream_t s, int sm_count, const pml_peer *peer, const double *factors, const double *shifts)
{
    PeerParams p;
    std::memset(&p, 0, sizeof(p));
    for (int q = 0; q < peer->world; ++q) p.slab[q] = static_cast<char *>(peer->peer_slab[q]);
    p.rank = peer->rank;
    p.world = peer->world;
    p.d = peer->d;
    p.factors = factors;
    p.shifts = shifts;
    p.status = peer->d_status;
    const int64_t per = ((peer->d + 2) / 2 + peer->world - 1) / peer->world;
    int64_t grid = (per + PEER_THREADS - 1) / PEER_THREADS;
    if (grid > sm_count) grid = sm_count;
    if (grid < 1) grid = 1;
    switch (peer->world) {
    case 1: peer_merge_kernel<1><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 2: peer_merge_kernel<2><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 3: peer_merge_kernel<3><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 4: peer_merge_kernel<4><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 5: peer_merge_kernel<5><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 6: peer_merge_kernel<6><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 7: peer_merge_kernel<7><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    case 8: peer_merge_kernel<8><<<(int)grid, PEER_THREADS, 0, s>>>(p); break;
    default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

size_t peer_slab_bytes(int64_t d) { return PEER_HEADER_BYTES + sizeof(double) * (size_t)(2 * slab_stride(d)); }
double *peer_raw(const pml_peer *peer) { return reinterpret_cast<double *>(static_cast<char *>(peer->slab) + PEER_HEADER_BYTES); }
double *peer_result(const pml_peer *peer) { return peer_raw(peer) + slab_stride(peer->d); }

}  // namespace pml

using namespace pml;

extern "C" {

int pml_peer_create(pml_ctx *ctx, int rank, int world, int32_t d, pml_peer **out)
{
    if (out == nullptr) return fail(PML_EINVAL, "out is null");
    *out = nullptr;
    if (ctx == nullptr) return fail(PML_EINVAL, "null context");
    if (world < 1 || world > PML_MAX_PEERS) return fail(PML_EINVAL, "world size must be between 1 and " + std::to_string(PML_MAX_PEERS));
    if (rank < 0 || rank >= world) return fail(PML_EINVAL, "rank out of range");
    if (d < 0) return fail(PML_EINVAL, "negative dimension");
    PML_CUDA(cudaSetDevice(ctx->device));
    pml_peer *p = new pml_peer();
    p->ctx = ctx;
    p->rank = rank;
    p->world = world;
    p->d = d;
    const size_t bytes = peer_slab_bytes(d);
    cudaError_t e = cudaMalloc(&p->slab, bytes);
    if (e == cudaSuccess) e = cudaMemset(p->slab, 0, bytes);
    if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void **>(&p->h_status), sizeof(int), cudaHostAllocMapped);
    if (e == cudaSuccess) {
        *p->h_status = 0;
        e = cudaHostGetDevicePointer(reinterpret_cast<void **>(&p->d_status), p->h_status, 0);
    }
    if (e != cudaSuccess) {
        pml_peer_destroy(p);
        return fail(e == cudaErrorMemoryAllocation ? PML_ENOMEM : PML_ECUDA, std::string("peer slab: ") + cudaGetErrorString(e));
    }
    p->peer_slab[rank] = p->slab;
    p->connected = world == 1;
    *out = p;
    return PML_OK;
}

int pml_peer_handle(pml_peer *peer, void *out_handle)
{
    if (peer == nullptr || out_handle == nullptr) return fail(PML_EINVAL, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) <= PML_PEER_HANDLE_BYTES, "handle size");
    PML_CUDA(cudaSetDevice(peer->ctx->device));
    cudaIpcMemHandle_t h;
    PML_CUDA(cudaIpcGetMemHandle(&h, peer->slab));
    std::memset(out_handle, 0, PML_PEER_HANDLE_BYTES);
    std::memcpy(out_handle, &h, sizeof(h));
    return PML_OK;
}

int pml_peer_connect(pml_peer *peer, const void *all_handles)
{
    if (peer == nullptr || all_handles == nullptr) return fail(PML_EINVAL, "null argument");
    PML_CUDA(cudaSetDevice(peer->ctx->device));
    const char *src = static_cast<const char *>(all_handles);
    for (int q = 0; q < peer->world; ++q) {
        if (q == peer->rank || peer->peer_slab[q] != nullptr) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, src + (size_t)q * PML_PEER_HANDLE_BYTES, sizeof(h));
        cudaError_t e = cudaIpcOpenMemHandle(&peer->peer_slab[q], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            peer->peer_slab[q] = nullptr;
            return fail(PML_ECUDA, "cannot open the slab of rank " + std::to_string(q) +
                                       " (no peer access between the two GPUs?): " + cudaGetErrorString(e));
        }
        peer->opened[q] = true;
    }
    peer->connected = true;
    return PML_OK;
}

int pml_peer_buffers(pml_peer *peer, double **raw_dev, double **result_dev)
{
    if (peer == nullptr) return fail(PML_EINVAL, "null peer handle");
    if (raw_dev) *raw_dev = peer_raw(peer);
    if (result_dev) *result_dev = peer_result(peer);
    return PML_OK;
}

int pml_peer_destroy(pml_peer *peer)
{
    if (peer == nullptr) return PML_OK;
    if (peer->ctx) {
        cudaSetDevice(peer->ctx->device);
        cudaStreamSynchronize(peer->ctx->stream);
    }
    for (int q = 0; q < PML_MAX_PEERS; ++q)
        if (peer->opened[q] && peer->peer_slab[q]) cudaIpcCloseMemHandle(peer->peer_slab[q]);
    if (peer->slab) cudaFree(peer->slab);
    if (peer->h_status) cudaFreeHost(peer->h_status);
    delete peer;
    return PML_OK;
}

}  // extern "C"
