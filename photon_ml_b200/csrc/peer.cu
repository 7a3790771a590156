// peer.cu -- the `merge` step across GPUs (ValueAndGradientAggregator.merge :161-176 as driven by treeAggregate,
// :243-247) fused with getVector (:199-221), over NVLink peer memory, one process per GPU.
//
// Every rank owns a slab [header | raw (d+2) | result (d+1) | staging (N rows of 1/N of the vector)] allocated with
// cudaMalloc and opened by all the other ranks through CUDA IPC.  After a rank's evaluation kernels have left its raw
// sums [vectorSum(d), valueSum, prefactorSum] in its slab, ONE kernel per rank does the whole exchange, with posted
// stores only (loads over NVLink wait a round trip each and measured 2-3x slower than the same bytes stored):
//   1. push slice q of the raw sums into row [rank] of rank q's staging area, for every peer q;
//   2. barrier: "my pushes are out" to every peer, wait until every peer said so;
//   3. rank r adds up slice r from its own staging rows (local loads, ranks in a fixed order, so every element is
//      summed once, by one rank, in one order: the result is bit-identical on all ranks and from run to run),
//      applies getVector to it and stores the finished slice into EVERY rank's result;
//   4. barrier: "my stores are out", wait for everyone's -- when the kernel ends, the rank's result is complete and
//      nobody writes its staging rows before the next exchange.
// Per rank that is (N-1)/N of the vector out twice over NVLink and two flag round trips -- no separate finalize
// launch, nothing on the host between the evaluation and the finished gradient.  The epoch the
// flags count lives in the slab and is advanced by the kernel itself, so the launch takes no per-call argument and
// can sit in a CUDA graph behind the evaluation kernels.
#include <cuda.h>  // types and prototypes only: the driver entry points are looked up at run time
#include <cuda_runtime.h>
#include <stdint.h>
#include <unistd.h>

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
    unsigned int pushed;                 // CTAs of the running kernel that have pushed their share of the raw sums
    unsigned int pad2[29];
    double pref[PML_MAX_PEERS];          // pref[q] = rank q's prefactorSum of the running exchange
    unsigned int gathered[PML_MAX_PEERS];  // gathered[q] = epoch_g: rank q's slice of the input vectors has arrived here
    unsigned int pad3[32 - PML_MAX_PEERS];
    unsigned int epoch_g;                // all-gathers completed
    unsigned int pushed_g;               // CTAs of the running all-gather that have pushed their share
    unsigned int pad4[30];
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

// false = gave up: the flag never reached `epoch` (a rank that never arrived); the mapped host word tells the caller
__device__ __forceinline__ bool wait_for(const unsigned int *flag, unsigned int epoch, int *status)
{
    const long long t0 = clock64();
    while ((int)(ld_acquire_sys(flag) - epoch) < 0) {
        if (clock64() - t0 > PEER_TIMEOUT_CYCLES) {
            *status = 1;
            return false;
        }
    }
    return true;
}

// raw and result are both padded to an even number of doubles: everything moves as 16-byte pairs
__host__ __device__ __forceinline__ int64_t slab_stride(int64_t d) { return (d + 3) & ~(int64_t)1; }
__device__ __forceinline__ double *slab_raw(char *slab) { return reinterpret_cast<double *>(slab + PEER_HEADER_BYTES); }
__device__ __forceinline__ double *slab_result(char *slab, int64_t d)
{
    return reinterpret_cast<double *>(slab + PEER_HEADER_BYTES) + slab_stride(d);
}
// staging: world rows of ceil(pairs / world) pairs -- row r holds rank r's raw sums of the owner's slice
__device__ __forceinline__ double2 *slab_staging(char *slab, int64_t d)
{
    return reinterpret_cast<double2 *>(reinterpret_cast<double *>(slab + PEER_HEADER_BYTES) + 2 * slab_stride(d));
}

// input vectors of the host-buffer entry points (coefficients, multiply vector): two regions of slab_stride(d)
// doubles behind the staging rows; every rank uploads one slice of each and the all-gather kernel fills in the rest
__host__ __device__ __forceinline__ int64_t staging_doubles(int64_t d, int world)
{
    const int64_t per = ((d + 2) / 2 + world - 1) / world;
    return 2 * per * world;
}
__device__ __forceinline__ double *slab_vec(char *slab, int64_t d, int world, int which)
{
    return reinterpret_cast<double *>(slab + PEER_HEADER_BYTES) + 2 * slab_stride(d) + staging_doubles(d, world) +
           (int64_t)which * slab_stride(d);
}

// The body of the exchange, for rank `rank` as CTA `bid` of `nblk`: the kernel proper passes its own rank and
// grid; the single-device emulation (pml_peer_emulate) runs all ranks' CTAs in one cooperative launch.
template <int WORLD>
__device__ __forceinline__ void peer_merge_body(const PeerParams &p, const int rank, const int bid, const int nblk)
{
    __shared__ int is_last, gave_up;
    if (threadIdx.x == 0) gave_up = 0;
    char *my_slab = p.slab[rank];
    PeerHeader *mine = reinterpret_cast<PeerHeader *>(my_slab);
    const unsigned int epoch = mine->epoch + 1;
    const int64_t d = p.d;
    // elements [0, d] (slot d is the value) as 16-byte pairs; the last pair may take the prefactor slot d+1 along
    const int64_t n_pairs = (d + 2) / 2;
    const int64_t per = (n_pairs + WORLD - 1) / WORLD;
    const int64_t T = (int64_t)nblk * PEER_THREADS;
    const int64_t gtid = (int64_t)bid * PEER_THREADS + threadIdx.x;
    const double *raw = slab_raw(my_slab);
    const double2 *raw2 = reinterpret_cast<const double2 *>(raw);

    // ---- 1. push: slice q of this rank's raw sums goes to row [rank] of rank q's staging area ----
#pragma unroll
    for (int q = 0; q < WORLD; ++q) {
        if (q == rank) continue;
        const int64_t lo_q = per * q;
        const int64_t hi_q = lo_q + per < n_pairs ? lo_q + per : n_pairs;
        double2 *dst = slab_staging(p.slab[q], d) + (int64_t)rank * per - lo_q;
        int64_t k = lo_q + gtid;
        for (; k + 3 * T < hi_q; k += 4 * T) {  // four independent copies in flight per thread
            const double2 a = raw2[k], b = raw2[k + T], c = raw2[k + 2 * T], e = raw2[k + 3 * T];
            dst[k] = a;
            dst[k + T] = b;
            dst[k + 2 * T] = c;
            dst[k + 3 * T] = e;
        }
        for (; k < hi_q; k += T) dst[k] = raw2[k];
    }
    if (bid == 0 && threadIdx.x < WORLD)
        reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->pref[rank] = raw[d + 1];
    __syncthreads();  // the CTA's stores happen-before thread 0's fence, which is cumulative
    if (threadIdx.x == 0) {
        __threadfence_system();
        is_last = atomicAdd(&mine->pushed, 1u) == nblk - 1;
    }
    __syncthreads();
    if (is_last && threadIdx.x < WORLD) {  // every CTA of this rank has pushed: tell everyone
        __threadfence_system();
        if (threadIdx.x == 0) mine->pushed = 0;
        st_release_sys(&reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->ready[rank], epoch);
    }

    // ---- 2. wait for everyone's pushes (all CTAs are resident: the grid never exceeds the SM count) ----
    if (threadIdx.x < WORLD && !wait_for(&mine->ready[threadIdx.x], epoch, p.status)) gave_up = 1;
    __syncthreads();
    // a rank that never arrived: the sums below would be partial -- hand out NaN instead of a plausible number
    const bool poisoned = gave_up != 0;

    // ---- 3. this rank's slice: add the staged rows in rank order, getVector, store to every rank ----
    double P = 0.0;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) P += q == rank ? raw[d + 1] : ld_peer(&mine->pref[q]);
    const int64_t lo = per * rank;
    const int64_t hi = lo + per < n_pairs ? lo + per : n_pairs;
    const double2 *staged = slab_staging(my_slab, d) - lo;
    for (int64_t k = lo + gtid; k < hi; k += T) {
        const int64_t i = 2 * k;
        double2 part[WORLD];
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
            part[q] = q == rank ? raw2[k] : ld_peer2(reinterpret_cast<const double *>(staged + (int64_t)q * per + k));
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
        if (poisoned) v.x = v.y = __longlong_as_double(0x7ff8000000000000ll);
#pragma unroll
        for (int q = 0; q < WORLD; ++q) *reinterpret_cast<double2 *>(slab_result(p.slab[q], d) + i) = v;
    }

    // ---- 4. everyone's stores are out ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        is_last = atomicAdd(&mine->arrived, 1u) == nblk - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence_system();
    if (threadIdx.x < WORLD)
        st_release_sys(&reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->stored[rank], epoch);
    if (threadIdx.x < WORLD) wait_for(&mine->stored[threadIdx.x], epoch, p.status);
    __syncthreads();
    if (threadIdx.x == 0) {
        mine->arrived = 0;
        mine->epoch = epoch;
    }
}

template <int WORLD>
__global__ void __launch_bounds__(PEER_THREADS) peer_merge_kernel(const PeerParams p)
{
    peer_merge_body<WORLD>(p, p.rank, (int)blockIdx.x, (int)gridDim.x);
}

// All ranks of a WORLD-rank exchange as ONE cooperative launch on one device (every CTA resident, so the flag waits
// between "ranks" cannot starve): CTAs [r * per_rank, (r + 1) * per_rank) play rank r.  For tests and for checking a
// build on a box with fewer GPUs than ranks -- waiting kernels of separate launches must not share a GPU.
template <int WORLD>
__global__ void __launch_bounds__(PEER_THREADS) peer_merge_emulated_kernel(const PeerParams p, const int per_rank)
{
    peer_merge_body<WORLD>(p, (int)blockIdx.x / per_rank, (int)blockIdx.x % per_rank, per_rank);
}

// ---- all-gather of the input vectors (host-buffer entry points) ------------------------------------------------
// Every rank has uploaded ITS slice of the coefficients (and of the multiply vector) from host memory over its own
// PCIe link; this kernel stores that slice into every peer's copy (posted NVLink stores) and waits until every
// peer's slice has arrived here: Spark's sc.broadcast (DistributedObjectiveFunction.scala:57-58) with the upload
// spread over all links instead of funnelled through one.
struct GatherParams {
    char *slab[PML_MAX_PEERS];
    int rank, world;
    int64_t d;
    int n_vec;    // 1: region 0 only; 2: regions 0 and 1; -2: region 1 only (cached Hessian-vector: v without coefficients)
    int *status;
};

__host__ __device__ __forceinline__ void vec_slice(int64_t d, int world, int rank, int64_t *lo, int64_t *hi)
{
    const int64_t n_pairs = (d + 1) / 2;
    const int64_t per = (n_pairs + world - 1) / world;
    int64_t a = 2 * per * rank, b = 2 * per * (rank + 1);
    *lo = a < d ? a : d;
    *hi = b < d ? b : d;
}

template <int WORLD>
__device__ __forceinline__ void peer_gather_body(const GatherParams &p, const int rank, const int bid, const int nblk)
{
    __shared__ int is_last;
    char *my_slab = p.slab[rank];
    PeerHeader *mine = reinterpret_cast<PeerHeader *>(my_slab);
    const unsigned int epoch = mine->epoch_g + 1;
    int64_t lo, hi;
    vec_slice(p.d, WORLD, rank, &lo, &hi);
    const int64_t T = (int64_t)nblk * PEER_THREADS;
    const int64_t gtid = (int64_t)bid * PEER_THREADS + threadIdx.x;
    const int first = p.n_vec == -2 ? 1 : 0, last = p.n_vec == 1 ? 0 : 1;
    for (int which = first; which <= last; ++which) {
        const double *src = slab_vec(my_slab, p.d, WORLD, which);
        // slices start at even elements: 16-byte pairs, a trailing odd element on its own
        const int64_t pairs = (hi - lo) / 2;
#pragma unroll
        for (int q = 0; q < WORLD; ++q) {
            if (q == rank) continue;
            double *dst = slab_vec(p.slab[q], p.d, WORLD, which);
            const double2 *s2 = reinterpret_cast<const double2 *>(src + lo);
            double2 *d2 = reinterpret_cast<double2 *>(dst + lo);
            for (int64_t k = gtid; k < pairs; k += T) d2[k] = s2[k];
            if (gtid == 0 && ((hi - lo) & 1)) dst[hi - 1] = src[hi - 1];
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        is_last = atomicAdd(&mine->pushed_g, 1u) == (unsigned int)(nblk - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x < WORLD) {
        __threadfence_system();
        if (threadIdx.x == 0) mine->pushed_g = 0;
        st_release_sys(&reinterpret_cast<PeerHeader *>(p.slab[threadIdx.x])->gathered[rank], epoch);
    }
    if (!is_last) return;
    // the last CTA waits for everyone's slices: when the kernel ends, this rank's copies are complete
    if (threadIdx.x < WORLD) wait_for(&mine->gathered[threadIdx.x], epoch, p.status);
    __syncthreads();
    if (threadIdx.x == 0) mine->epoch_g = epoch;
}

template <int WORLD>
__global__ void __launch_bounds__(PEER_THREADS) peer_gather_kernel(const GatherParams p)
{
    peer_gather_body<WORLD>(p, p.rank, (int)blockIdx.x, (int)gridDim.x);
}

template <int WORLD>
__global__ void __launch_bounds__(PEER_THREADS) peer_gather_emulated_kernel(const GatherParams p, const int per_rank)
{
    peer_gather_body<WORLD>(p, (int)blockIdx.x / per_rank, (int)blockIdx.x % per_rank, per_rank);
}



// ---- the same exchange through an NVSwitch multicast object (NVLS) --------------------------------------------
// When the ranks' slabs are bound to one multicast object (pml_peer_multicast_*), the switch does the adding and
// the fan-out: multimem.ld_reduce returns the sum of all ranks' copies of an address, multimem.st writes all
// copies, multimem.red increments a counter on every rank at once.  Each rank still owns one slice, so every
// element is reduced once and stored everywhere -- identical bits on all ranks -- but it pulls 1/N of the vector
// instead of (N-1)/N and pushes each finished element once instead of N times.
struct McParams {
    char *uc;  // this rank's slab (ordinary mapping)
    char *mc;  // the multicast mapping of all ranks' slabs
    int rank, world;
    int64_t d;
    const double *factors, *shifts;
    int *status;
};

__device__ __forceinline__ double mc_ld_reduce(const double *p)
{
    double v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void mc_st(double *p, double v)
{
    asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

__device__ __forceinline__ void mc_signal(unsigned int *p)
{
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(1u) : "memory");
}

__global__ void __launch_bounds__(PEER_THREADS) peer_merge_mc_kernel(const McParams p)
{
    __shared__ int is_last;
    PeerHeader *mine = reinterpret_cast<PeerHeader *>(p.uc);
    PeerHeader *all = reinterpret_cast<PeerHeader *>(p.mc);
    const unsigned int epoch = mine->epoch + 1;
    const unsigned int target = epoch * (unsigned int)p.world;  // the counters only ever grow: world per exchange
    const int64_t d = p.d;

    // ---- 1. everyone's raw sums are complete ----
    if (blockIdx.x == 0 && threadIdx.x == 0) mc_signal(&all->ready[0]);
    if (threadIdx.x == 0) wait_for(&mine->ready[0], target, p.status);
    __syncthreads();

    // ---- 2. this rank's slice: the switch adds, getVector, one store reaches every rank ----
    const double *raw = slab_raw(p.mc);
    double *result = slab_result(p.mc, d);
    const double P = mc_ld_reduce(raw + d + 1);
    const int64_t per = (d + 1 + p.world - 1) / p.world;  // elements [0, d]; slot d is the value
    const int64_t lo = per * p.rank;
    const int64_t hi = lo + per < d + 1 ? lo + per : d + 1;
    // four independent requests per thread in flight: the switch round trip is what this loop waits for
    const int64_t T = (int64_t)gridDim.x * PEER_THREADS;
    for (int64_t base = lo + (int64_t)blockIdx.x * PEER_THREADS + threadIdx.x; base < hi; base += 4 * T) {
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (base + u * T < hi) v[u] = mc_ld_reduce(raw + base + u * T);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t i = base + u * T;
            if (i >= hi) break;
            double x = v[u];
            if (i < d) {
                if (p.shifts) x = __dsub_rn(x, __dmul_rn(p.shifts[i], P));
                if (p.factors) x = __dmul_rn(x, p.factors[i]);
            }
            mc_st(result + i, x);
        }
    }

    // ---- 3. everyone's stores are out ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        is_last = atomicAdd(&mine->arrived, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    if (threadIdx.x == 0) {
        __threadfence_system();
        mc_signal(&all->stored[0]);
        wait_for(&mine->stored[0], target, p.status);
        mine->arrived = 0;
        mine->epoch = epoch;
    }
}

}  // namespace

int launch_peer_merge(cudaStream_t s, int sm_count, const pml_peer *peer, const double *factors, const double *shifts)
{
    if (peer->mc_va != nullptr) {
        McParams m;
        std::memset(&m, 0, sizeof(m));
        m.uc = static_cast<char *>(peer->uc_va);
        m.mc = static_cast<char *>(peer->mc_va);
        m.rank = peer->rank;
        m.world = peer->world;
        m.d = peer->d;
        m.factors = factors;
        m.shifts = shifts;
        m.status = peer->d_status;
        const int64_t per = (peer->d + 1 + peer->world - 1) / peer->world;
        int64_t grid = (per + PEER_THREADS - 1) / PEER_THREADS;
        if (grid > 4 * sm_count) grid = 4 * sm_count;
        if (grid < 1) grid = 1;
        peer_merge_mc_kernel<<<(int)grid, PEER_THREADS, 0, s>>>(m);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    PeerParams p;
    std::memset(&p, 0, sizeof(p));
    for (int q = 0; q < peer->world; ++q) p.slab[q] = static_cast<char *>(peer->peer_slab[q]);
    p.rank = peer->rank;
    p.world = peer->world;
    p.d = peer->d;
    p.factors = factors;
    p.shifts = shifts;
    p.status = peer->d_status;
    // CTAs wait for the last CTA of every rank: all of them must be resident, one per SM at most
    int64_t grid = ((peer->d + 2) / 2 + PEER_THREADS - 1) / PEER_THREADS;
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

size_t peer_slab_bytes(int64_t d, int world)
{
    return PEER_HEADER_BYTES + sizeof(double) * (size_t)(2 * slab_stride(d) + staging_doubles(d, world) + 2 * slab_stride(d));
}

// this rank's copy of input vector `which` (0 = coefficients, 1 = multiply vector) inside its slab
double *peer_vec(const pml_peer *peer, int which)
{
    char *base = static_cast<char *>(peer->mc_va != nullptr ? peer->uc_va : peer->slab);
    const int world = peer->mc_va != nullptr ? 1 : peer->world;   // the multicast slab is laid out for one staging row
    return reinterpret_cast<double *>(base + PEER_HEADER_BYTES) + 2 * slab_stride(peer->d) + staging_doubles(peer->d, world) +
           (int64_t)which * slab_stride(peer->d);
}

void peer_slice(const pml_peer *peer, int64_t *lo, int64_t *hi) { vec_slice(peer->d, peer->world, peer->rank, lo, hi); }

// n_vec as in GatherParams.  With the multicast transport the slabs are not mapped pairwise: the caller uploads the
// whole vector instead (returns 0 launches).
int launch_peer_gather(cudaStream_t s, int sm_count, const pml_peer *peer, int n_vec)
{
    if (peer->world == 1 || peer->mc_va != nullptr) return 0;
    GatherParams g;
    std::memset(&g, 0, sizeof(g));
    for (int q = 0; q < peer->world; ++q) g.slab[q] = static_cast<char *>(peer->peer_slab[q]);
    g.rank = peer->rank;
    g.world = peer->world;
    g.d = peer->d;
    g.n_vec = n_vec;
    g.status = peer->d_status;
    int64_t lo, hi;
    peer_slice(peer, &lo, &hi);
    int64_t grid = ((hi - lo) / 2 + PEER_THREADS - 1) / PEER_THREADS;
    if (grid > sm_count / 2) grid = sm_count / 2;
    if (grid < 1) grid = 1;
    switch (peer->world) {
    case 2: peer_gather_kernel<2><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 3: peer_gather_kernel<3><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 4: peer_gather_kernel<4><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 5: peer_gather_kernel<5><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 6: peer_gather_kernel<6><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 7: peer_gather_kernel<7><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    case 8: peer_gather_kernel<8><<<(int)grid, PEER_THREADS, 0, s>>>(g); break;
    default: return -1;
    }
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

// The whole exchange of WORLD ranks on ONE device, as cooperative launches: all-gather of `n_vec` input vectors
// (0 = skip), then merge + getVector.  slabs[q] = rank q's slab (peer_slab_bytes(d, world) each, zeroed once).
template <int WORLD>
static int emulate_world(cudaStream_t s, int sm_count, char *const *slabs, int64_t d, int n_vec, const double *factors,
                         const double *shifts, int *status)
{
    const int per_rank = sm_count / WORLD < 4 ? (sm_count / WORLD < 1 ? 1 : sm_count / WORLD) : 4;
    if (n_vec != 0) {
        GatherParams g;
        std::memset(&g, 0, sizeof(g));
        for (int q = 0; q < WORLD; ++q) g.slab[q] = slabs[q];
        g.world = WORLD;
        g.d = d;
        g.n_vec = n_vec;
        g.status = status;
        int pr = per_rank;
        void *args[] = { &g, &pr };
        if (cudaLaunchCooperativeKernel((const void *)peer_gather_emulated_kernel<WORLD>, dim3(per_rank * WORLD),
                                        dim3(PEER_THREADS), args, 0, s) != cudaSuccess)
            return -1;
    }
    PeerParams p;
    std::memset(&p, 0, sizeof(p));
    for (int q = 0; q < WORLD; ++q) p.slab[q] = slabs[q];
    p.world = WORLD;
    p.d = d;
    p.factors = factors;
    p.shifts = shifts;
    p.status = status;
    int pr = per_rank;
    void *args[] = { &p, &pr };
    if (cudaLaunchCooperativeKernel((const void *)peer_merge_emulated_kernel<WORLD>, dim3(per_rank * WORLD),
                                    dim3(PEER_THREADS), args, 0, s) != cudaSuccess)
        return -1;
    return 1;
}

int launch_peer_emulation(cudaStream_t s, int sm_count, int world, char *const *slabs, int64_t d, int n_vec,
                          const double *factors, const double *shifts, int *status)
{
    switch (world) {
    case 1: return emulate_world<1>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 2: return emulate_world<2>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 3: return emulate_world<3>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 4: return emulate_world<4>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 5: return emulate_world<5>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 6: return emulate_world<6>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 7: return emulate_world<7>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    case 8: return emulate_world<8>(s, sm_count, slabs, d, n_vec, factors, shifts, status);
    default: return -1;
    }
}

size_t peer_offset_raw() { return PEER_HEADER_BYTES; }
size_t peer_offset_result(int64_t d) { return PEER_HEADER_BYTES + sizeof(double) * (size_t)slab_stride(d); }
size_t peer_offset_vec(int64_t d, int world, int which)
{
    return PEER_HEADER_BYTES + sizeof(double) * (size_t)(2 * slab_stride(d) + staging_doubles(d, world) + which * slab_stride(d));
}
double *peer_raw(const pml_peer *peer)
{
    char *base = static_cast<char *>(peer->mc_va != nullptr ? peer->uc_va : peer->slab);
    return reinterpret_cast<double *>(base + PEER_HEADER_BYTES);
}
double *peer_result(const pml_peer *peer) { return peer_raw(peer) + slab_stride(peer->d); }

}  // namespace pml

using namespace pml;

// ---- multicast set-up: driver API, looked up at run time (the library links against the runtime only) ----------
namespace {

struct Driver {
    decltype(&cuMulticastCreate) MulticastCreate = nullptr;
    decltype(&cuMulticastAddDevice) MulticastAddDevice = nullptr;
    decltype(&cuMulticastBindMem) MulticastBindMem = nullptr;
    decltype(&cuMulticastUnbind) MulticastUnbind = nullptr;
    decltype(&cuMulticastGetGranularity) MulticastGetGranularity = nullptr;
    decltype(&cuMemCreate) MemCreate = nullptr;
    decltype(&cuMemRelease) MemRelease = nullptr;
    decltype(&cuMemMap) MemMap = nullptr;
    decltype(&cuMemUnmap) MemUnmap = nullptr;
    decltype(&cuMemSetAccess) MemSetAccess = nullptr;
    decltype(&cuMemAddressReserve) MemAddressReserve = nullptr;
    decltype(&cuMemAddressFree) MemAddressFree = nullptr;
    decltype(&cuMemExportToShareableHandle) MemExportToShareableHandle = nullptr;
    decltype(&cuMemImportFromShareableHandle) MemImportFromShareableHandle = nullptr;
    decltype(&cuDeviceGet) DeviceGet = nullptr;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
    decltype(&cuGetErrorString) GetErrorString = nullptr;
    bool ok = false;
};

template <typename F>
bool entry(const char *name, F *out)
{
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        return false;
    }
    *out = reinterpret_cast<F>(fn);
    return true;
}

const Driver &driver()
{
    static const Driver d = [] {
        Driver r;
        r.ok = entry("cuMulticastCreate", &r.MulticastCreate) && entry("cuMulticastAddDevice", &r.MulticastAddDevice) &&
               entry("cuMulticastBindMem", &r.MulticastBindMem) && entry("cuMulticastUnbind", &r.MulticastUnbind) &&
               entry("cuMulticastGetGranularity", &r.MulticastGetGranularity) && entry("cuMemCreate", &r.MemCreate) &&
               entry("cuMemRelease", &r.MemRelease) && entry("cuMemMap", &r.MemMap) && entry("cuMemUnmap", &r.MemUnmap) &&
               entry("cuMemSetAccess", &r.MemSetAccess) && entry("cuMemAddressReserve", &r.MemAddressReserve) &&
               entry("cuMemAddressFree", &r.MemAddressFree) &&
               entry("cuMemExportToShareableHandle", &r.MemExportToShareableHandle) &&
               entry("cuMemImportFromShareableHandle", &r.MemImportFromShareableHandle) &&
               entry("cuDeviceGet", &r.DeviceGet) && entry("cuDeviceGetAttribute", &r.DeviceGetAttribute) &&
               entry("cuGetErrorString", &r.GetErrorString);
        return r;
    }();
    return d;
}

int drv_fail(const char *what, CUresult e)
{
    const char *msg = nullptr;
    if (driver().GetErrorString) driver().GetErrorString(e, &msg);
    return fail(PML_ECUDA, std::string(what) + ": " + (msg ? msg : "driver error ") + " (" + std::to_string((int)e) + ")");
}

#define PML_DRV(call, what)                                   \
    do {                                                      \
        CUresult _e = (call);                                 \
        if (_e != CUDA_SUCCESS) return drv_fail(what, _e);    \
    } while (0)

CUmulticastObjectProp mc_prop(const pml_peer *peer)
{
    CUmulticastObjectProp prop;
    std::memset(&prop, 0, sizeof(prop));
    prop.numDevices = (unsigned int)peer->world;
    prop.size = peer->mc_bytes;
    prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int mc_prepare(pml_peer *peer)
{
    if (!driver().ok) return fail(PML_ECUDA, "this driver has no multicast entry points");
    PML_CUDA(cudaSetDevice(peer->ctx->device));
    PML_CUDA(cudaFree(nullptr));  // the primary context is current from here on
    CUdevice dev;
    PML_DRV(driver().DeviceGet(&dev, peer->ctx->device), "cuDeviceGet");
    int supported = 0;
    PML_DRV(driver().DeviceGetAttribute(&supported, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "cuDeviceGetAttribute");
    if (!supported) return fail(PML_ECUDA, "the device does not support multicast objects (no NVSwitch fabric?)");
    if (peer->mc_bytes == 0) {
        peer->mc_bytes = peer_slab_bytes(peer->d, 1);
        CUmulticastObjectProp prop = mc_prop(peer);
        size_t gran = 0;
        PML_DRV(driver().MulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
        if (gran == 0) gran = 2u << 20;
        peer->mc_bytes = (peer->mc_bytes + gran - 1) / gran * gran;
    }
    return PML_OK;
}

void mc_teardown(pml_peer *peer)
{
    const Driver &d = driver();
    if (!d.ok) return;
    if (peer->mc_va) {
        d.MemUnmap((CUdeviceptr)peer->mc_va, peer->mc_bytes);
        d.MemAddressFree((CUdeviceptr)peer->mc_va, peer->mc_bytes);
        peer->mc_va = nullptr;
    }
    if (peer->uc_va) {
        d.MemUnmap((CUdeviceptr)peer->uc_va, peer->mc_bytes);
        d.MemAddressFree((CUdeviceptr)peer->uc_va, peer->mc_bytes);
        peer->uc_va = nullptr;
    }
    if (peer->mc_bound && peer->mc_handle) {
        CUdevice dev;
        if (d.DeviceGet(&dev, peer->ctx->device) == CUDA_SUCCESS) d.MulticastUnbind(peer->mc_handle, dev, 0, peer->mc_bytes);
        peer->mc_bound = false;
    }
    if (peer->mc_mem) { d.MemRelease(peer->mc_mem); peer->mc_mem = 0; }
    if (peer->mc_handle) { d.MemRelease(peer->mc_handle); peer->mc_handle = 0; }
}

}  // namespace

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
    const size_t bytes = peer_slab_bytes(d, world);
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
    {
        std::lock_guard<std::mutex> lock(ctx->mu);
        ctx->peer_status.push_back(p->h_status);   // pml_ctx_synchronize reports a timed-out exchange
    }
    *out = p;
    return PML_OK;
}

int pml_peer_status(pml_peer *peer, int *out)
{
    if (peer == nullptr || out == nullptr) return fail(PML_EINVAL, "null argument");
    *out = *peer->h_status;
    return PML_OK;
}

int pml_peer_slice(pml_peer *peer, int64_t *out_lo, int64_t *out_hi)
{
    if (peer == nullptr || out_lo == nullptr || out_hi == nullptr) return fail(PML_EINVAL, "null argument");
    peer_slice(peer, out_lo, out_hi);
    return PML_OK;
}

int pml_peer_emulate(pml_ctx *ctx, int world, int32_t d, int repeats, const double *raw, const double *factors,
                     const double *shifts, double *out_result, const double *vec_in, double *vec_out)
{
    if (ctx == nullptr || raw == nullptr || out_result == nullptr) return fail(PML_EINVAL, "null argument");
    if (world < 1 || world > PML_MAX_PEERS) return fail(PML_EINVAL, "world size must be between 1 and " + std::to_string(PML_MAX_PEERS));
    if (d < 0 || repeats < 1) return fail(PML_EINVAL, "bad dimension or repeat count");
    if ((vec_in == nullptr) != (vec_out == nullptr)) return fail(PML_EINVAL, "vec_in and vec_out go together");
    std::lock_guard<std::mutex> lock(ctx->mu);
    PML_CUDA(cudaSetDevice(ctx->device));
    int coop = 0;
    PML_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device));
    if (!coop) return fail(PML_ECUDA, "the device does not support cooperative launches");
    cudaStream_t s = ctx->stream;
    const size_t bytes = peer_slab_bytes(d, world);
    char *slabs[PML_MAX_PEERS] = {};
    double *f_dev = nullptr, *s_dev = nullptr;
    int *h_status = nullptr, *d_status = nullptr;
    int rc = PML_OK;
    auto cleanup = [&] {
        for (int q = 0; q < world; ++q)
            if (slabs[q]) cudaFree(slabs[q]);
        if (f_dev) cudaFree(f_dev);
        if (s_dev) cudaFree(s_dev);
        if (h_status) cudaFreeHost(h_status);
    };
#define EMU_CUDA(expr)                                                                                     \
    do {                                                                                                   \
        cudaError_t _e = (expr);                                                                           \
        if (_e != cudaSuccess) {                                                                           \
            cleanup();                                                                                     \
            return fail(PML_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));                    \
        }                                                                                                  \
    } while (0)
    EMU_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&h_status), sizeof(int), cudaHostAllocMapped));
    *h_status = 0;
    EMU_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void **>(&d_status), h_status, 0));
    for (int q = 0; q < world; ++q) {
        EMU_CUDA(cudaMalloc(&slabs[q], bytes));
        EMU_CUDA(cudaMemsetAsync(slabs[q], 0, bytes, s));
    }
    if (factors) {
        EMU_CUDA(cudaMalloc(&f_dev, sizeof(double) * (size_t)(d + 1)));
        EMU_CUDA(cudaMemcpyAsync(f_dev, factors, sizeof(double) * (size_t)d, cudaMemcpyHostToDevice, s));
    }
    if (shifts) {
        EMU_CUDA(cudaMalloc(&s_dev, sizeof(double) * (size_t)(d + 1)));
        EMU_CUDA(cudaMemcpyAsync(s_dev, shifts, sizeof(double) * (size_t)d, cudaMemcpyHostToDevice, s));
    }
    for (int rep = 0; rep < repeats && rc == PML_OK; ++rep) {
        for (int q = 0; q < world; ++q) {
            EMU_CUDA(cudaMemcpyAsync(slabs[q] + peer_offset_raw(), raw + (size_t)q * (size_t)(d + 2),
                                     sizeof(double) * (size_t)(d + 2), cudaMemcpyHostToDevice, s));
            if (vec_in) {  // every rank uploads ITS slice only: the rest of its copy must come from the all-gather
                int64_t lo, hi;
                vec_slice(d, world, q, &lo, &hi);
                EMU_CUDA(cudaMemsetAsync(slabs[q] + peer_offset_vec(d, world, 0), 0xff, sizeof(double) * (size_t)d, s));
                if (hi > lo)
                    EMU_CUDA(cudaMemcpyAsync(slabs[q] + peer_offset_vec(d, world, 0) + sizeof(double) * (size_t)lo,
                                             vec_in + (size_t)q * (size_t)d + lo, sizeof(double) * (size_t)(hi - lo),
                                             cudaMemcpyHostToDevice, s));
            }
        }
        if (launch_peer_emulation(s, ctx->sm_count, world, slabs, d, vec_in ? 1 : 0, f_dev, s_dev, d_status) < 0) {
            cudaError_t e = cudaGetLastError();
            cleanup();
            return fail(PML_ECUDA, std::string("cooperative launch failed: ") + cudaGetErrorString(e));
        }
        EMU_CUDA(cudaStreamSynchronize(s));
        if (*h_status != 0) rc = fail(PML_ECUDA, "the emulated exchange timed out");
    }
    for (int q = 0; q < world && rc == PML_OK; ++q) {
        EMU_CUDA(cudaMemcpyAsync(out_result + (size_t)q * (size_t)(d + 1), slabs[q] + peer_offset_result(d),
                                 sizeof(double) * (size_t)(d + 1), cudaMemcpyDeviceToHost, s));
        if (vec_out)
            EMU_CUDA(cudaMemcpyAsync(vec_out + (size_t)q * (size_t)d, slabs[q] + peer_offset_vec(d, world, 0),
                                     sizeof(double) * (size_t)d, cudaMemcpyDeviceToHost, s));
    }
    EMU_CUDA(cudaStreamSynchronize(s));
#undef EMU_CUDA
    cleanup();
    return rc;
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

int pml_ctx_pci_bus_id(pml_ctx *ctx, char *out, int len)
{
    if (ctx == nullptr || out == nullptr || len < 16) return fail(PML_EINVAL, "bad argument");
    PML_CUDA(cudaDeviceGetPCIBusId(out, len, ctx->device));
    return PML_OK;
}

int pml_peer_multicast_export(pml_peer *peer, int *out_fd)
{
    if (peer == nullptr || out_fd == nullptr) return fail(PML_EINVAL, "null argument");
    *out_fd = -1;
    if (peer->mc_handle) return fail(PML_ESTATE, "the peer handle already holds a multicast object");
    PML_TRY(mc_prepare(peer));
    CUmulticastObjectProp prop = mc_prop(peer);
    PML_DRV(driver().MulticastCreate(&peer->mc_handle, &prop), "cuMulticastCreate");
    int fd = -1;
    CUresult e = driver().MemExportToShareableHandle(&fd, peer->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (e != CUDA_SUCCESS) {
        driver().MemRelease(peer->mc_handle);
        peer->mc_handle = 0;
        return drv_fail("cuMemExportToShareableHandle", e);
    }
    *out_fd = fd;
    return PML_OK;
}

int pml_peer_multicast_join(pml_peer *peer, int fd)
{
    if (peer == nullptr) return fail(PML_EINVAL, "null peer handle");
    PML_TRY(mc_prepare(peer));
    if (peer->mc_handle == 0) {
        if (fd < 0) return fail(PML_EINVAL, "a rank that did not create the multicast object needs its file descriptor");
        PML_DRV(driver().MemImportFromShareableHandle(&peer->mc_handle, (void *)(uintptr_t)fd,
                                                      CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
                "cuMemImportFromShareableHandle");
    }
    CUdevice dev;
    PML_DRV(driver().DeviceGet(&dev, peer->ctx->device), "cuDeviceGet");
    CUresult e = driver().MulticastAddDevice(peer->mc_handle, dev);
    if (e != CUDA_SUCCESS) {
        mc_teardown(peer);
        return drv_fail("cuMulticastAddDevice", e);
    }
    return PML_OK;
}

int pml_peer_multicast_bind(pml_peer *peer)
{
    if (peer == nullptr) return fail(PML_EINVAL, "null peer handle");
    if (peer->mc_handle == 0) return fail(PML_ESTATE, "pml_peer_multicast_join has not been called");
    if (peer->mc_va) return PML_OK;
    const Driver &d = driver();
    PML_CUDA(cudaSetDevice(peer->ctx->device));
    PML_CUDA(cudaStreamSynchronize(peer->ctx->stream));
    CUmemAllocationProp ap;
    std::memset(&ap, 0, sizeof(ap));
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = peer->ctx->device;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemAccessDesc access;
    std::memset(&access, 0, sizeof(access));
    access.location = ap.location;
    access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    CUdeviceptr uc = 0, mc = 0;
    int rc = PML_OK;
    do {
        CUresult e;
        if ((e = d.MemCreate(&peer->mc_mem, peer->mc_bytes, &ap, 0)) != CUDA_SUCCESS) { rc = drv_fail("cuMemCreate", e); break; }
        // blocks until every rank has added its device
        if ((e = d.MulticastBindMem(peer->mc_handle, 0, peer->mc_mem, 0, peer->mc_bytes, 0)) != CUDA_SUCCESS) { rc = drv_fail("cuMulticastBindMem", e); break; }
        peer->mc_bound = true;
        if ((e = d.MemAddressReserve(&uc, peer->mc_bytes, 0, 0, 0)) != CUDA_SUCCESS) { rc = drv_fail("cuMemAddressReserve", e); break; }
        if ((e = d.MemMap(uc, peer->mc_bytes, 0, peer->mc_mem, 0)) != CUDA_SUCCESS) { d.MemAddressFree(uc, peer->mc_bytes); rc = drv_fail("cuMemMap", e); break; }
        peer->uc_va = (void *)uc;
        if ((e = d.MemSetAccess(uc, peer->mc_bytes, &access, 1)) != CUDA_SUCCESS) { rc = drv_fail("cuMemSetAccess", e); break; }
        if ((e = d.MemAddressReserve(&mc, peer->mc_bytes, 0, 0, 0)) != CUDA_SUCCESS) { rc = drv_fail("cuMemAddressReserve", e); break; }
        if ((e = d.MemMap(mc, peer->mc_bytes, 0, peer->mc_handle, 0)) != CUDA_SUCCESS) { d.MemAddressFree(mc, peer->mc_bytes); rc = drv_fail("cuMemMap (multicast)", e); break; }
        peer->mc_va = (void *)mc;
        if ((e = d.MemSetAccess(mc, peer->mc_bytes, &access, 1)) != CUDA_SUCCESS) { rc = drv_fail("cuMemSetAccess (multicast)", e); break; }
    } while (0);
    if (rc != PML_OK) {
        const std::string keep = pml_last_error();
        mc_teardown(peer);
        return fail(rc, keep);
    }
    PML_CUDA(cudaMemset(peer->uc_va, 0, peer->mc_bytes));
    PML_CUDA(cudaDeviceSynchronize());
    return PML_OK;
}

int pml_peer_multicast_release(pml_peer *peer)
{
    if (peer == nullptr) return fail(PML_EINVAL, "null peer handle");
    PML_CUDA(cudaSetDevice(peer->ctx->device));
    PML_CUDA(cudaStreamSynchronize(peer->ctx->stream));
    mc_teardown(peer);
    return PML_OK;
}

int pml_peer_transport(pml_peer *peer, int *out)
{
    if (peer == nullptr || out == nullptr) return fail(PML_EINVAL, "null argument");
    *out = peer->mc_va != nullptr ? PML_PEER_MULTICAST : PML_PEER_P2P;
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
    mc_teardown(peer);
    if (peer->ctx && peer->h_status) {
        std::lock_guard<std::mutex> lock(peer->ctx->mu);
        auto &v = peer->ctx->peer_status;
        for (size_t i = 0; i < v.size(); ++i)
            if (v[i] == peer->h_status) { v.erase(v.begin() + (long)i); break; }
    }
    for (int q = 0; q < PML_MAX_PEERS; ++q)
        if (peer->opened[q] && peer->peer_slab[q]) cudaIpcCloseMemHandle(peer->peer_slab[q]);
    if (peer->slab) cudaFree(peer->slab);
    if (peer->h_status) cudaFreeHost(peer->h_status);
    delete peer;
    return PML_OK;
}

}  // extern "C"
