// mg_dd.cuh -- device side of the 3D domain decomposition (BASELINE.json configs[2]; the reference has no
// counterpart: it runs one block / one fragment, example_problems/lib/domain_onePatch.knowledge:4-9).
//
// Every rank (= GPU) owns a z-slab of each decomposed level.  Ranks exchange data by storing straight into
// each other's memory over NVLink (peer-mapped pointers: CUDA IPC between processes, plain pointers between
// the virtual ranks of one process) and order those stores with flag words in a small per-rank `DDComm`
// block -- no NCCL call and no host synchronisation inside a cycle, so the whole decomposed cycle of a
// rank is ONE captured CUDA graph.
//
//   halo exchange   k_dd_exchange: "I have reached exchange k" (ack) to both neighbours; wait for theirs
//                   (their kernels no longer read the ghost planes I am about to overwrite); copy my
//                   boundary planes into their ghost planes; fence; "delivered k" (pushed); wait for
//                   theirs.  The kernel ends when my ghost planes are valid: the next kernel of the
//                   stream may read them.  Inside a decomposed cycle the ack round is dropped
//                   (need_ack = 0): "pushed k" from a neighbour means all its kernels before its exchange k
//                   are complete, and the only kernel it runs between its exchanges k and k+1 never reads
//                   the ghost planes exchange k+1 delivers (an out-of-place sweep reads the other buffer; the
//                   restriction / prolongation read other fields) -- mg_engine_dd.inc dd_issue_cycle.
//   agglomeration   k_dd_gather: below the coarsest decomposed level every rank runs the cycle redundantly
//                   on a whole grid; each rank pulls the owned planes of the restricted right-hand side
//                   from all ranks (ready / done flags = a full barrier on either side of the reads).
//   residual norm   k_dd_norm: every rank stores its partial sum of squares into every rank's slot table;
//                   all ranks add the slots in rank order -> bitwise the same norm everywhere.
//
// Counters only grow (exchange / gather / norm rounds since creation), so a replayed graph needs no
// per-launch arguments.  Every wait is bounded: a peer that never arrives sets DDComm::error (reported by
// the host) instead of hanging the GPU.
#pragma once
#include "mg_common.cuh"

namespace mg {

constexpr int DD_MAX_RANKS = 16;
constexpr long long DD_TIMEOUT_CYCLES = 12000000000ll;      // ~6 s at 1.9 GHz

struct DDComm {
    // written by peers
    unsigned long long ack[2];                  // [0]: by the lower neighbour, [1]: by the upper one
    unsigned long long pushed[2];
    unsigned long long ready[DD_MAX_RANKS];     // gather rounds for which rank q's planes are complete
    unsigned long long done[DD_MAX_RANKS];      // gather rounds rank q has finished reading
    unsigned long long nready[DD_MAX_RANKS];    // norm rounds whose partial sum from rank q has arrived
    double sq[2][DD_MAX_RANKS];                 // partial sums of squares, double buffered by round parity
    // local state of this rank's kernels
    unsigned long long epoch_x, epoch_g, epoch_n;
    unsigned int arrivals_x, arrivals_g;
    int error;
    int pad_;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ inline void dd_wait(const unsigned long long* flag, unsigned long long want, int* error) {
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < want) {
        if (clock64() - t0 > DD_TIMEOUT_CYCLES) {
            *error = 1;
            return;
        }
        __nanosleep(64);
    }
}

// grid-stride copy of n2 16-byte words, four independent loads in flight per thread (a pull over NVLink is
// bound by the number of outstanding loads)
__device__ __forceinline__ void dd_copy(double2* __restrict__ dst, const double2* __restrict__ src, long long n2) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n2; i += 4 * stride) {
        const double2 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
        dst[i] = a;
        dst[i + stride] = b;
        dst[i + 2 * stride] = c;
        dst[i + 3 * stride] = d;
    }
    for (; i < n2; i += stride) dst[i] = src[i];
}

// One halo exchange of this rank: src_* = my boundary planes (contiguous), dst_* = the neighbour's ghost
// planes (peer memory), n2 = 16-byte words per direction; lo / hi = the neighbours' DDComm (null: none).
__global__ void __launch_bounds__(256) k_dd_exchange(DDComm* me, DDComm* lo, DDComm* hi, const double2* src_lo,
                                                     double2* dst_lo, const double2* src_hi, double2* dst_hi, long long n2,
                                                     int need_ack) {
    pdl_sync();
    __shared__ unsigned long long k_s;
    if (threadIdx.x == 0) {
        const unsigned long long k = ld_volatile_u64(&me->epoch_x) + 1;
        k_s = k;
        if (need_ack) {
            if (blockIdx.x == 0) {
                __threadfence_system();
                if (lo) st_release_sys(&lo->ack[1], k);        // I am the lower neighbour's upper neighbour
                if (hi) st_release_sys(&hi->ack[0], k);
            }
            if (lo) dd_wait(&me->ack[0], k, &me->error);
            if (hi) dd_wait(&me->ack[1], k, &me->error);
        }
    }
    __syncthreads();
    const unsigned long long k = k_s;
    if (lo) dd_copy(dst_lo, src_lo, n2);
    if (hi) dd_copy(dst_hi, src_hi, n2);
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(&me->arrivals_x, 1u);
        if (prev == gridDim.x - 1) {                       // every CTA's stores are fenced: announce, then wait
            me->arrivals_x = 0;
            __threadfence_system();
            if (lo) st_release_sys(&lo->pushed[1], k);
            if (hi) st_release_sys(&hi->pushed[0], k);
            if (lo) dd_wait(&me->pushed[0], k, &me->error);
            if (hi) dd_wait(&me->pushed[1], k, &me->error);
            me->epoch_x = k;
            __threadfence();
        }
    }
}

struct DDGatherArgs {
    DDComm* comm[DD_MAX_RANKS];
    const double2* src[DD_MAX_RANKS];       // rank q's owned planes of the field (peer memory; own: local)
    long long n2[DD_MAX_RANKS];             // 16-byte words of them
    long long off2[DD_MAX_RANKS];           // where they go in the whole-grid array
    int world, rank;
};

// whole-grid array of the replicated coarse hierarchy <- owned planes of all ranks (pull)
__global__ void __launch_bounds__(256) k_dd_gather(DDGatherArgs a, double2* __restrict__ dst) {
    pdl_sync();
    DDComm* me = a.comm[a.rank];
    __shared__ unsigned long long g_s;
    if (threadIdx.x == 0) {
        const unsigned long long g = ld_volatile_u64(&me->epoch_g) + 1;
        g_s = g;
        if (blockIdx.x == 0) {
            __threadfence_system();
            for (int q = 0; q < a.world; ++q)
                if (q != a.rank) st_release_sys(&a.comm[q]->ready[a.rank], g);
        }
        for (int q = 0; q < a.world; ++q)
            if (q != a.rank) dd_wait(&me->ready[q], g, &me->error);
    }
    __syncthreads();
    const unsigned long long g = g_s;
    for (int q = 0; q < a.world; ++q) dd_copy(dst + a.off2[q], a.src[q], a.n2[q]);
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(&me->arrivals_g, 1u);
        if (prev == gridDim.x - 1) {
            me->arrivals_g = 0;
            __threadfence_system();
            for (int q = 0; q < a.world; ++q)
                if (q != a.rank) st_release_sys(&a.comm[q]->done[a.rank], g);
            for (int q = 0; q < a.world; ++q)              // the sources may be overwritten after this kernel
                if (q != a.rank) dd_wait(&me->done[q], g, &me->error);
            me->epoch_g = g;
            __threadfence();
        }
    }
}

struct DDNormArgs {
    DDComm* comm[DD_MAX_RANKS];
    int world, rank;
};

// *norm holds this rank's sum of squares on entry and sqrt(rank-ordered sum over all ranks) on exit
__global__ void k_dd_norm(DDNormArgs a, double* norm) {
    pdl_sync();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    DDComm* me = a.comm[a.rank];
    const unsigned long long m = ld_volatile_u64(&me->epoch_n) + 1;
    const int b = (int)(m & 1ull);
    const double mine = *norm;
    for (int q = 0; q < a.world; ++q) a.comm[q]->sq[b][a.rank] = mine;
    __threadfence_system();
    for (int q = 0; q < a.world; ++q)
        if (q != a.rank) st_release_sys(&a.comm[q]->nready[a.rank], m);
    for (int q = 0; q < a.world; ++q)
        if (q != a.rank) dd_wait(&me->nready[q], m, &me->error);
    double total = 0.0;
    for (int q = 0; q < a.world; ++q) total = total + ld_relaxed_sys_f64(&me->sq[b][q]);
    *norm = sqrt(total);
    me->epoch_n = m;
}

}  // namespace mg
