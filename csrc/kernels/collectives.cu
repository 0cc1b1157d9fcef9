// Peer-memory collectives for sm_100a: all-reduce (one-shot, two-shot, NVLS),
// broadcast, all-gather, reduce/gather to root, barrier, consensus.
//
// What this replaces in the reference (which has no CUDA kernel at all):
//   * ncclAllReduce/ncclReduce/ncclBroadcast + cudaStreamSynchronize per tensor
//     (/root/reference/srcs/cpp/src/nccl/gpu_collective.cpp:96-127)
//   * the host graph walk with CPU reduction
//     (/root/reference/srcs/go/kungfu/session/session.go:218-286,
//      /root/reference/srcs/go/kungfu/base/op.cpp:57-93)
//   * the separate g/np elementwise op
//     (/root/reference/srcs/python/kungfu/tensorflow/optimizers/sync_sgd.py:103-104)
//
// Design: every rank maps every peer's buffers (symmetric memory over NVLink 5
// / NVSwitch). A kernel is "barrier -> loads/stores on peer pointers ->
// barrier"; the barrier is per CTA index (CTA b of every rank meets CTA b of
// every other rank), so no CTA ever waits for another CTA of its own GPU and
// the launch needs no co-residency guarantee. Reductions add in rank order
// 0..np-1 in fp32, so every rank produces the same bits and a run is
// reproducible, matching the rank-ordered CPU oracle.
#include <algorithm>
#include <cooperative_groups.h>
#include <cstdarg>
#include <cstdio>
#include <type_traits>

#include "common.cuh"

namespace kfb {

static thread_local char g_err[512] = "";
void set_last_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

struct ArArgs {
    PeerPads pads;
    uint32_t *local_epoch;
    float *scratch;
    void *in[KFB_MAX_RANKS];
    void *out[KFB_MAX_RANKS];
    void *mc_in;
    void *mc_out;
    long long nvec;  // number of 16-byte packets
    float scale;
    int rank, np;
    float *sumsq_out;
};

__device__ __forceinline__ int emu_rank(int rank)
{
    return rank < 0 ? (int)blockIdx.y : rank;
}

// Sum of the per-CTA partial norms of all ranks, in a fixed order:
// 1. each CTA pushes its partial to slot [block][rank] of every peer's pad,
// 2. (exit barrier)
// 3. each CTA adds the np partials of its block index in rank order and
//    stores the result in local scratch[block],
// 4. the last CTA to finish adds scratch[0..grid) in block order.
// Every rank evaluates exactly the same expression tree.
__device__ __forceinline__ void
push_partial(const PeerPads &pads, int rank, int np, float partial)
{
    if (threadIdx.x < np)
        st_relaxed_sys_f32(pad_scalar(pads.pad[threadIdx.x], blockIdx.x, rank),
                           partial);
}

__device__ __forceinline__ void
finish_sumsq(const PeerPads &pads, int rank, int np, float *scratch,
             float *sumsq_out)
{
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int r = 0; r < np; ++r)
            t += ld_relaxed_sys_f32(pad_scalar(pads.pad[rank], blockIdx.x, r));
        scratch[blockIdx.x] = t;
        __threadfence();
        unsigned int *ticket =
            reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS);
        const unsigned int n = atomicAdd(ticket, 1u);
        is_last              = (n == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
        __threadfence();
        float t = 0.f;
        for (unsigned b = 0; b < gridDim.x; ++b)
            t += *reinterpret_cast<volatile float *>(scratch + b);
        *sumsq_out = t;
        *reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS) = 0u;
    }
}

// Load one packet from each of np peers (all loads issued before any use so
// np requests are in flight per thread), reduce in rank order.
template <typename T, int OP, int NP>
__device__ __forceinline__ void
load_reduce(void *const *in, int np, long long i,
            typename Pack<T>::acc_t (&acc)[Pack<T>::N])
{
    using P = Pack<T>;
    if (NP > 0) {
        uint4 v[NP > 0 ? NP : 1];
#pragma unroll
        for (int r = 0; r < NP; ++r)
            v[r] = ld_stream16(reinterpret_cast<const uint4 *>(in[r]) + i);
        P::unpack(v[0], acc);
#pragma unroll
        for (int r = 1; r < NP; ++r) {
            typename P::acc_t t[P::N];
            P::unpack(v[r], t);
#pragma unroll
            for (int k = 0; k < P::N; ++k) acc[k] = apply_op<OP>(acc[k], t[k]);
        }
    } else {
        uint4 v0 = ld_stream16(reinterpret_cast<const uint4 *>(in[0]) + i);
        P::unpack(v0, acc);
        for (int r0 = 1; r0 < np; r0 += 4) {
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int r = min(r0 + j, np - 1);
                v[j] = ld_stream16(reinterpret_cast<const uint4 *>(in[r]) + i);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (r0 + j < np) {
                    typename P::acc_t t[P::N];
                    P::unpack(v[j], t);
#pragma unroll
                    for (int k = 0; k < P::N; ++k)
                        acc[k] = apply_op<OP>(acc[k], t[k]);
                }
            }
        }
    }
}

template <typename T, bool NORM>
__device__ __forceinline__ void
scale_and_norm(typename Pack<T>::acc_t (&acc)[Pack<T>::N], float scale,
               float &ss)
{
    using A = typename Pack<T>::acc_t;
    if (std::is_floating_point<A>::value) {
#pragma unroll
        for (int k = 0; k < Pack<T>::N; ++k) {
            acc[k] = (A)(acc[k] * (A)scale);
            if (NORM) ss = fmaf((float)acc[k], (float)acc[k], ss);
        }
    }
}

// ---- two-shot: pull-reduce my shard, push the result to every peer ----------
//
// Element i of shard r is read (from every peer) and then written (to every
// peer) by one thread of rank r only, so in-place operation needs no barrier
// between the two phases: program order of that thread is enough.
template <typename T, int OP, int NP, bool NORM>
__global__ void __launch_bounds__(512)
    k_allreduce_twoshot(const __grid_constant__ ArArgs a)
{
    using P        = Pack<T>;
    const int rank = emu_rank(a.rank);
    const int np   = NP > 0 ? NP : a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    float *scratch = a.scratch + (a.rank < 0 ? rank * KFB_SCRATCH_FLOATS : 0);
    __shared__ float red[32];

    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);

    const long long per = (a.nvec + np - 1) / np;
    const long long lo  = (long long)rank * per;
    const long long hi  = min(lo + per, a.nvec);
    const long long stride = (long long)gridDim.x * blockDim.x;
    float ss               = 0.f;
    constexpr int U        = (NP > 0 && NP <= 2) ? 4 : (NP > 0 && NP <= 4 ? 2 : 1);
    long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < hi; i += U * stride) {
        typename P::acc_t acc[U][P::N];
#pragma unroll
        for (int u = 0; u < U; ++u)
            load_reduce<T, OP, NP>(a.in, np, i + u * stride, acc[u]);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            scale_and_norm<T, NORM>(acc[u], a.scale, ss);
            const uint4 o = P::pack(acc[u]);
            const int n_out = NP > 0 ? NP : np;
#pragma unroll
            for (int k = 0; k < n_out; ++k) {
                const int r = (rank + k) % n_out;
                st_stream16(reinterpret_cast<uint4 *>(a.out[r]) + i + u * stride, o);
            }
        }
    }
    for (; i < hi; i += stride) {
        typename P::acc_t acc[P::N];
        load_reduce<T, OP, NP>(a.in, np, i, acc);
        scale_and_norm<T, NORM>(acc, a.scale, ss);
        const uint4 o = P::pack(acc);
        for (int k = 0; k < np; ++k) {
            const int r = (rank + k) % np;
            st_stream16(reinterpret_cast<uint4 *>(a.out[r]) + i, o);
        }
    }
    if (NORM) {
        const float part = block_sum(ss, red);
        __shared__ float s_part;
        if (threadIdx.x == 0) s_part = part;
        __syncthreads();
        push_partial(a.pads, rank, np, s_part);
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (NORM)
        finish_sumsq(a.pads, rank, np, scratch,
                     a.sumsq_out + (a.rank < 0 ? rank : 0));
}

// ---- one-shot: every rank pulls everything, writes only its own output ------
template <typename T, int OP, int NP, bool NORM>
__global__ void __launch_bounds__(512)
    k_allreduce_oneshot(const __grid_constant__ ArArgs a)
{
    using P        = Pack<T>;
    const int rank = emu_rank(a.rank);
    const int np   = NP > 0 ? NP : a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    float *scratch = a.scratch + (a.rank < 0 ? rank * KFB_SCRATCH_FLOATS : 0);
    __shared__ float red[32];

    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    const long long stride = (long long)gridDim.x * blockDim.x;
    float ss               = 0.f;
    uint4 *out             = reinterpret_cast<uint4 *>(a.out[rank]);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < a.nvec; i += stride) {
        typename P::acc_t acc[P::N];
        load_reduce<T, OP, NP>(a.in, np, i, acc);
        scale_and_norm<T, NORM>(acc, a.scale, ss);
        st_stream16(out + i, P::pack(acc));
    }
    // Inputs must stay untouched until every peer has finished reading them.
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (NORM) {
        // Every rank holds the full result: a purely local fixed-order sum.
        const float part = block_sum(ss, red);
        __shared__ bool is_last;
        if (threadIdx.x == 0) {
            scratch[blockIdx.x] = part;
            __threadfence();
            unsigned int *ticket =
                reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS);
            is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
        }
        __syncthreads();
        if (is_last && threadIdx.x == 0) {
            __threadfence();
            float t = 0.f;
            for (unsigned b = 0; b < gridDim.x; ++b)
                t += *reinterpret_cast<volatile float *>(scratch + b);
            a.sumsq_out[a.rank < 0 ? rank : 0] = t;
            *reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS) = 0u;
        }
    }
}

// ---- NVLS: the switch reduces (multimem.ld_reduce) and multicasts
//      (multimem.st); per-GPU traffic is ~S per direction instead of
//      2(n-1)/n S. SUM of f32 / bf16 / f16 only. ------------------------------
template <typename T> struct Multimem;
template <> struct Multimem<float> {
    __device__ static uint4 ld_reduce(const void *p)
    {
        uint4 r;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 "
                     "{%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "l"(p)
                     : "memory");
        return r;
    }
    __device__ static void st(void *p, const uint4 &v)
    {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], "
                     "{%1,%2,%3,%4};" ::"l"(p),
                     "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                     : "memory");
    }
};
template <> struct Multimem<__nv_bfloat16> {
    __device__ static uint4 ld_reduce(const void *p)
    {
        uint4 r;
        asm volatile(
            "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 "
            "{%0,%1,%2,%3}, [%4];"
            : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
            : "l"(p)
            : "memory");
        return r;
    }
    __device__ static void st(void *p, const uint4 &v)
    {
        asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], "
                     "{%1,%2,%3,%4};" ::"l"(p),
                     "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                     : "memory");
    }
};
template <> struct Multimem<__half> {
    __device__ static uint4 ld_reduce(const void *p)
    {
        uint4 r;
        asm volatile(
            "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 "
            "{%0,%1,%2,%3}, [%4];"
            : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
            : "l"(p)
            : "memory");
        return r;
    }
    __device__ static void st(void *p, const uint4 &v)
    {
        asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], "
                     "{%1,%2,%3,%4};" ::"l"(p),
                     "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                     : "memory");
    }
};

template <typename T, bool NORM>
__global__ void __launch_bounds__(512)
    k_allreduce_nvls(const __grid_constant__ ArArgs a)
{
    using P         = Pack<T>;
    const int rank  = a.rank;
    const int np    = a.np;
    uint32_t *epoch = a.local_epoch;
    __shared__ float red[32];

    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    const long long per    = (a.nvec + np - 1) / np;
    const long long lo     = (long long)rank * per;
    const long long hi     = min(lo + per, a.nvec);
    const long long stride = (long long)gridDim.x * blockDim.x;
    float ss               = 0.f;
    const uint4 *mc_in     = reinterpret_cast<const uint4 *>(a.mc_in);
    uint4 *mc_out          = reinterpret_cast<uint4 *>(a.mc_out);
    constexpr int U        = 4;
    long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < hi; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            v[u] = Multimem<T>::ld_reduce(mc_in + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float acc[P::N];
            P::unpack(v[u], acc);
            scale_and_norm<T, NORM>(acc, a.scale, ss);
            Multimem<T>::st(mc_out + i + u * stride, P::pack(acc));
        }
    }
    for (; i < hi; i += stride) {
        float acc[P::N];
        P::unpack(Multimem<T>::ld_reduce(mc_in + i), acc);
        scale_and_norm<T, NORM>(acc, a.scale, ss);
        Multimem<T>::st(mc_out + i, P::pack(acc));
    }
    if (NORM) {
        const float part = block_sum(ss, red);
        __shared__ float s_part;
        if (threadIdx.x == 0) s_part = part;
        __syncthreads();
        push_partial(a.pads, rank, np, s_part);
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (NORM) finish_sumsq(a.pads, rank, np, a.scratch, a.sumsq_out);
}

// ---- byte movers: broadcast / all-gather / gather ---------------------------
struct MoveArgs {
    PeerPads pads;
    uint32_t *local_epoch;
    const void *in;
    void *in_peers[KFB_MAX_RANKS];
    void *out[KFB_MAX_RANKS];
    void *mc_out;
    long long nvec;
    int rank, np, root;
};

__global__ void __launch_bounds__(512)
    k_broadcast(const __grid_constant__ MoveArgs a)
{
    const int rank = emu_rank(a.rank);
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    // entry: every receiver is done with whatever lived in `out` before
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (rank == a.root) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        const uint4 *src       = reinterpret_cast<const uint4 *>(a.in);
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < a.nvec; i += stride) {
            const uint4 v = ld_stream16(src + i);
            if (a.mc_out) {
                Multimem<float>::st(reinterpret_cast<uint4 *>(a.mc_out) + i, v);
            } else {
                for (int k = 0; k < np; ++k)
                    st_stream16(reinterpret_cast<uint4 *>(a.out[(rank + k) % np]) + i, v);
            }
        }
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

__global__ void __launch_bounds__(512)
    k_allgather(const __grid_constant__ MoveArgs a)
{
    const int rank = emu_rank(a.rank);
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const uint4 *src       = reinterpret_cast<const uint4 *>(
        a.rank < 0 ? a.in_peers[rank] : a.in);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < a.nvec; i += stride) {
        const uint4 v = ld_stream16(src + i);
        for (int k = 0; k < np; ++k) {
            uint4 *dst = reinterpret_cast<uint4 *>(a.out[(rank + k) % np]) +
                         (long long)rank * a.nvec;
            st_stream16(dst + i, v);
        }
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

__global__ void __launch_bounds__(512)
    k_gather(const __grid_constant__ MoveArgs a)
{
    const int rank = emu_rank(a.rank);
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (rank == a.root) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        const long long total  = a.nvec * np;
        uint4 *dst             = reinterpret_cast<uint4 *>(a.out[0]);
        for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
             j < total; j += stride) {
            const int r       = (int)(j / a.nvec);
            const long long i = j - (long long)r * a.nvec;
            st_stream16(dst + j,
                        ld_stream16(reinterpret_cast<const uint4 *>(a.in_peers[r]) + i));
        }
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

// reduce-to-root: only the root pulls; the others just hold their buffers
// still between the two barriers.
template <typename T, int OP>
__global__ void __launch_bounds__(512)
    k_reduce_root(const __grid_constant__ ArArgs a, int root)
{
    using P        = Pack<T>;
    const int rank = emu_rank(a.rank);
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    if (rank == root) {
        const long long stride = (long long)gridDim.x * blockDim.x;
        uint4 *out             = reinterpret_cast<uint4 *>(a.out[0]);
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < a.nvec; i += stride) {
            typename P::acc_t acc[P::N];
            load_reduce<T, OP, 0>(a.in, np, i, acc);
            st_stream16(out + i, P::pack(acc));
        }
    }
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

__global__ void k_barrier(const __grid_constant__ MoveArgs a)
{
    const int rank = emu_rank(a.rank);
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    peer_barrier(a.pads, epoch, rank, a.np, blockIdx.x);
}

// consensus (K12): the reference runs MIN and MAX all-reduces over the bytes
// and compares; here every rank compares its own bytes against every peer's
// in one pass and AND-reduces the verdict.
__global__ void __launch_bounds__(512)
    k_consensus(const __grid_constant__ MoveArgs a, int32_t *agree_out)
{
    const int rank = emu_rank(a.rank);
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    // agree_out holds 1 on entry (set by a launch of its own, kfb_consensus):
    // CTAs only ever clear it, so no CTA's verdict can be overwritten.
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const uint4 *mine = reinterpret_cast<const uint4 *>(a.in_peers[rank]);
    int differ        = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < a.nvec; i += stride) {
        const uint4 m = ld_stream16(mine + i);
        for (int k = 1; k < np; ++k) {
            const uint4 o = ld_stream16(
                reinterpret_cast<const uint4 *>(a.in_peers[(rank + k) % np]) + i);
            differ |= (m.x != o.x) | (m.y != o.y) | (m.z != o.z) | (m.w != o.w);
        }
    }
    if (__syncthreads_or(differ) && threadIdx.x == 0)
        atomicAnd(agree_out + (a.rank < 0 ? rank : 0), 0);
    peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

__global__ void k_set_i32(int32_t *p, int n, int32_t v)
{
    if ((int)threadIdx.x < n) p[threadIdx.x] = v;
}

// ---- single-edge kernels for the strategy-graph executor (K8) ---------------
template <typename T, int OP>
__global__ void __launch_bounds__(512)
    k_edge_reduce(uint4 *dst, const uint4 *src, long long nvec)
{
    using P                = Pack<T>;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < nvec; i += stride) {
        typename P::acc_t x[P::N], y[P::N];
        P::unpack(ld_stream16(dst + i), x);
        P::unpack(ld_stream16(src + i), y);
#pragma unroll
        for (int k = 0; k < P::N; ++k) x[k] = apply_op<OP>(x[k], y[k]);
        st_stream16(dst + i, P::pack(x));
    }
}

__global__ void __launch_bounds__(512)
    k_edge_copy(uint4 *dst, const uint4 *src, long long nvec)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < nvec; i += stride)
        st_stream16(dst + i, ld_stream16(src + i));
}

// ---- a whole strategy walk in ONE launch --------------------------------------------
// The reference walks (reduce graph, broadcast graph) per chunk on the host,
// receiving from `prevs` and sending to `nexts`
// (/root/reference/srcs/go/kungfu/session/session.go:218-286). Here the graphs
// are compiled (csrc/core/graph.cpp:compile_device_plan) into rounds of incoming edges per rank:
// in round r a rank pulls from the sources of its round-r edges, reducing onto
// (reduce phase) or overwriting (broadcast phase) its own chunk; a peer
// barrier separates rounds, because a source's chunk is final only once the
// source finished its own earlier rounds. CTA b of every rank always handles
// the same grid-stride slice of every chunk, so the per-CTA barrier of
// common.cuh is all the ordering the walk needs.
#define KFB_GRAPH_MAX_ROUNDS 64
#define KFB_GRAPH_MAX_EDGES 320
#define KFB_GRAPH_MAX_PLANS 16
struct GraphArgs {
    PeerPads pads;
    uint32_t *local_epoch;
    void *bufs[KFB_MAX_RANKS];
    int rank, np;
    int n_rounds, n_reduce_rounds;
    int round_begin[KFB_GRAPH_MAX_ROUNDS + 1];  // edges of round r: [round_begin[r], round_begin[r+1])
    unsigned char edge_plan[KFB_GRAPH_MAX_EDGES];
    unsigned char edge_src[KFB_GRAPH_MAX_EDGES];
    long long chunk_lo[KFB_GRAPH_MAX_PLANS];  // in 16-byte vectors
    long long chunk_n[KFB_GRAPH_MAX_PLANS];
};

template <typename T, int OP>
__global__ void __launch_bounds__(512)
    k_graph_allreduce(const __grid_constant__ GraphArgs a)
{
    using P                = Pack<T>;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long first  = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    uint4 *mine            = reinterpret_cast<uint4 *>(a.bufs[a.rank]);
    for (int r = 0; r < a.n_rounds; ++r) {
        peer_barrier(a.pads, a.local_epoch, a.rank, a.np, blockIdx.x);
        for (int e = a.round_begin[r]; e < a.round_begin[r + 1]; ++e) {
            const int pl       = a.edge_plan[e];
            const uint4 *src   = reinterpret_cast<const uint4 *>(a.bufs[a.edge_src[e]]) + a.chunk_lo[pl];
            uint4 *dst         = mine + a.chunk_lo[pl];
            const long long nv = a.chunk_n[pl];
            if (r < a.n_reduce_rounds) {
                for (long long i = first; i < nv; i += stride) {
                    typename P::acc_t x[P::N], y[P::N];
                    P::unpack(ld_stream16(dst + i), x);
                    P::unpack(ld_stream16(src + i), y);
#pragma unroll
                    for (int k = 0; k < P::N; ++k) x[k] = apply_op<OP>(x[k], y[k]);
                    st_stream16(dst + i, P::pack(x));
                }
            } else {
                for (long long i = first; i < nv; i += stride) st_stream16(dst + i, ld_stream16(src + i));
            }
        }
    }
    // nobody may overwrite its buffer while a child still pulls from it
    peer_barrier(a.pads, a.local_epoch, a.rank, a.np, blockIdx.x);
}

// ---- host side ------------------------------------------------------------------
static size_t dtype_size(int dt)
{
    switch (dt) {
    case KFB_F32: case KFB_I32: return 4;
    case KFB_BF16: case KFB_F16: return 2;
    case KFB_F64: case KFB_I64: return 8;
    case KFB_U8: return 1;
    }
    return 0;
}

static int check_comm(const kfb_comm *c, int blocks, int threads)
{
    if (!c || c->np < 1 || c->np > KFB_MAX_RANKS || c->rank >= c->np) {
        set_last_error("bad communicator (np=%d rank=%d)", c ? c->np : -1,
                       c ? c->rank : -1);
        return KFB_ERR_ARG;
    }
    if (blocks < 1 || blocks > KFB_MAX_BLOCKS || threads < 32 ||
        threads > 512 || threads % 32) {
        set_last_error("bad launch shape blocks=%d threads=%d (max %d x 512)",
                       blocks, threads, KFB_MAX_BLOCKS);
        return KFB_ERR_ARG;
    }
    return KFB_OK;
}

static void fill_pads(PeerPads &p, const kfb_comm *c)
{
    for (int r = 0; r < KFB_MAX_RANKS; ++r)
        p.pad[r] = reinterpret_cast<uint32_t *>(r < c->np ? c->pads[r] : nullptr);
}

// Emulation needs every (block, rank) CTA co-resident: cooperative launch.
template <typename K, typename... Args>
static int launch(const kfb_comm *c, K kernel, int blocks, int threads,
                  cudaStream_t s, Args... args)
{
    if (c->rank < 0) {
        void *argv[] = {(void *)&args...};
        KFB_CUDA_OK(cudaLaunchCooperativeKernel((const void *)kernel,
                                                dim3(blocks, c->np), dim3(threads),
                                                argv, 0, s));
    } else {
        kernel<<<blocks, threads, 0, s>>>(args...);
        KFB_CUDA_OK(cudaGetLastError());
    }
    return KFB_OK;
}

template <typename T, int OP, bool NORM>
static int launch_ar_np(const kfb_comm *c, int algo, const ArArgs &a, int blocks,
                        int threads, cudaStream_t s)
{
    const bool fast = (OP == KFB_OP_SUM);
#define KFB_AR(KERNEL, NPV)                                                    \
    return launch(c, KERNEL<T, OP, NPV, NORM>, blocks, threads, s, a)
    if (algo == KFB_ALGO_TWOSHOT) {
        if (fast && a.np == 2) KFB_AR(k_allreduce_twoshot, 2);
        if (fast && a.np == 4) KFB_AR(k_allreduce_twoshot, 4);
        if (fast && a.np == 8) KFB_AR(k_allreduce_twoshot, 8);
        KFB_AR(k_allreduce_twoshot, 0);
    } else {
        if (fast && a.np == 2) KFB_AR(k_allreduce_oneshot, 2);
        if (fast && a.np == 4) KFB_AR(k_allreduce_oneshot, 4);
        if (fast && a.np == 8) KFB_AR(k_allreduce_oneshot, 8);
        KFB_AR(k_allreduce_oneshot, 0);
    }
#undef KFB_AR
}

template <typename T, bool FLOATING>
static int launch_ar_op(const kfb_comm *c, int algo, const ArArgs &a, int op,
                        int blocks, int threads, cudaStream_t s)
{
    const bool norm = a.sumsq_out != nullptr;
    switch (op) {
    case KFB_OP_SUM:
        if (FLOATING && norm)
            return launch_ar_np<T, KFB_OP_SUM, FLOATING>(c, algo, a, blocks, threads, s);
        return launch_ar_np<T, KFB_OP_SUM, false>(c, algo, a, blocks, threads, s);
    case KFB_OP_MIN:
        return launch_ar_np<T, KFB_OP_MIN, false>(c, algo, a, blocks, threads, s);
    case KFB_OP_MAX:
        return launch_ar_np<T, KFB_OP_MAX, false>(c, algo, a, blocks, threads, s);
    case KFB_OP_PROD:
        return launch_ar_np<T, KFB_OP_PROD, false>(c, algo, a, blocks, threads, s);
    }
    set_last_error("unknown reduce op %d", op);
    return KFB_ERR_ARG;
}

}  // namespace kfb


using namespace kfb;

extern "C" const char *kfb_last_error(void) { return kfb::g_err; }

extern "C" int kfb_device_info(int device, int *sm_count, int *cc_major,
                               int *cc_minor, int *multicast_supported)
{
    cudaDeviceProp p;
    KFB_CUDA_OK(cudaGetDeviceProperties(&p, device));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    if (multicast_supported) {
        // CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, asked of the driver (symm.cu)
        *multicast_supported = kfb_multicast_supported(device);
    }
    return KFB_OK;
}

static int fill_ar(ArArgs &a, const kfb_comm *c, void *const *in,
                   void *const *out, int64_t count, int dtype)
{
    const size_t es = dtype_size(dtype);
    if (!es) {
        set_last_error("unsupported dtype %d", dtype);
        return KFB_ERR_ARG;
    }
    memset(&a, 0, sizeof(a));
    fill_pads(a.pads, c);
    a.local_epoch = reinterpret_cast<uint32_t *>(c->local_epoch);
    a.scratch     = reinterpret_cast<float *>(c->scratch);
    a.rank        = c->rank;
    a.np          = c->np;
    a.nvec        = (long long)((count * es + 15) / 16);
    a.scale       = 1.f;
    for (int r = 0; r < c->np; ++r) {
        a.in[r]  = in ? in[r] : nullptr;
        a.out[r] = out ? out[r] : nullptr;
        if ((reinterpret_cast<uintptr_t>(a.in[r]) |
             reinterpret_cast<uintptr_t>(a.out[r])) & 15) {
            set_last_error("buffers must be 16-byte aligned");
            return KFB_ERR_ARG;
        }
    }
    return KFB_OK;
}

extern "C" int kfb_allreduce(const kfb_comm *comm, int algo, void *const *in,
                             void *const *out, void *mc_in, void *mc_out,
                             int64_t count, int dtype, int op, float scale,
                             float *sumsq_out, int blocks, int threads,
                             void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    ArArgs a;
    rc = fill_ar(a, comm, in, out, count, dtype);
    if (rc) return rc;
    a.scale       = scale;
    a.sumsq_out   = sumsq_out;
    a.mc_in       = mc_in;
    a.mc_out      = mc_out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (sumsq_out && op != KFB_OP_SUM) {
        set_last_error("sumsq_out is only defined for SUM");
        return KFB_ERR_ARG;
    }
    if (algo == KFB_ALGO_NVLS) {
        if (comm->rank < 0 || !mc_in || !mc_out || op != KFB_OP_SUM) {
            set_last_error("NVLS needs real ranks, multicast pointers and SUM");
            return KFB_ERR_ARG;
        }
        const bool norm = sumsq_out != nullptr;
#define KFB_NVLS(T)                                                            \
    do {                                                                       \
        if (norm)                                                              \
            k_allreduce_nvls<T, true><<<blocks, threads, 0, s>>>(a);           \
        else                                                                   \
            k_allreduce_nvls<T, false><<<blocks, threads, 0, s>>>(a);          \
    } while (0)
        if (dtype == KFB_F32) KFB_NVLS(float);
        else if (dtype == KFB_BF16) KFB_NVLS(__nv_bfloat16);
        else if (dtype == KFB_F16) KFB_NVLS(__half);
        else {
            set_last_error("NVLS supports f32/bf16/f16 only");
            return KFB_ERR_UNSUPP;
        }
#undef KFB_NVLS
        KFB_CUDA_OK(cudaGetLastError());
        return KFB_OK;
    }
    if (algo != KFB_ALGO_ONESHOT && algo != KFB_ALGO_TWOSHOT) {
        set_last_error("unknown algo %d", algo);
        return KFB_ERR_ARG;
    }
    switch (dtype) {
    case KFB_F32: return launch_ar_op<float, true>(comm, algo, a, op, blocks, threads, s);
    case KFB_BF16: return launch_ar_op<__nv_bfloat16, true>(comm, algo, a, op, blocks, threads, s);
    case KFB_F16: return launch_ar_op<__half, true>(comm, algo, a, op, blocks, threads, s);
    case KFB_I32: return launch_ar_op<int32_t, false>(comm, algo, a, op, blocks, threads, s);
    case KFB_F64: return launch_ar_op<double, false>(comm, algo, a, op, blocks, threads, s);
    case KFB_I64: return launch_ar_op<int64_t, false>(comm, algo, a, op, blocks, threads, s);
    case KFB_U8: return launch_ar_op<uint8_t, false>(comm, algo, a, op, blocks, threads, s);
    }
    set_last_error("unsupported dtype %d", dtype);
    return KFB_ERR_ARG;
}

static int fill_move(MoveArgs &m, const kfb_comm *c, int64_t bytes)
{
    memset(&m, 0, sizeof(m));
    fill_pads(m.pads, c);
    m.local_epoch = reinterpret_cast<uint32_t *>(c->local_epoch);
    m.rank        = c->rank;
    m.np          = c->np;
    m.nvec        = (long long)((bytes + 15) / 16);
    return KFB_OK;
}

extern "C" int kfb_broadcast(const kfb_comm *comm, int root, const void *in,
                             void *const *out, void *mc_out, int64_t bytes,
                             int blocks, int threads, void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    if (root < 0 || root >= comm->np) {
        set_last_error("bad root %d", root);
        return KFB_ERR_ARG;
    }
    MoveArgs m;
    fill_move(m, comm, bytes);
    m.in     = in;
    m.root   = root;
    m.mc_out = mc_out;
    for (int r = 0; r < comm->np; ++r) m.out[r] = out[r];
    return launch(comm, k_broadcast, blocks, threads,
                  reinterpret_cast<cudaStream_t>(stream), m);
}

extern "C" int kfb_allgather(const kfb_comm *comm, const void *in,
                             void *const *out, int64_t bytes, int blocks,
                             int threads, void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    if (bytes % 16) {
        set_last_error("allgather block must be a multiple of 16 bytes");
        return KFB_ERR_ARG;
    }
    MoveArgs m;
    fill_move(m, comm, bytes);
    m.in = in;
    for (int r = 0; r < comm->np; ++r) {
        m.out[r] = out[r];
        // emulation: `in` is a host array of np device pointers
        if (comm->rank < 0)
            m.in_peers[r] = reinterpret_cast<void *const *>(in)[r];
    }
    return launch(comm, k_allgather, blocks, threads,
                  reinterpret_cast<cudaStream_t>(stream), m);
}

extern "C" int kfb_gather(const kfb_comm *comm, int root, void *const *in,
                          void *out, int64_t bytes, int blocks, int threads,
                          void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    if (bytes % 16) {
        set_last_error("gather block must be a multiple of 16 bytes");
        return KFB_ERR_ARG;
    }
    MoveArgs m;
    fill_move(m, comm, bytes);
    m.root   = root;
    m.out[0] = out;
    for (int r = 0; r < comm->np; ++r) m.in_peers[r] = in[r];
    return launch(comm, k_gather, blocks, threads,
                  reinterpret_cast<cudaStream_t>(stream), m);
}

extern "C" int kfb_reduce(const kfb_comm *comm, int root, void *const *in,
                          void *out, int64_t count, int dtype, int op,
                          int blocks, int threads, void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    ArArgs a;
    rc = fill_ar(a, comm, in, nullptr, count, dtype);
    if (rc) return rc;
    a.out[0]       = out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define KFB_RED(T)                                                             \
    switch (op) {                                                              \
    case KFB_OP_SUM: return launch(comm, k_reduce_root<T, KFB_OP_SUM>, blocks, threads, s, a, root);   \
    case KFB_OP_MIN: return launch(comm, k_reduce_root<T, KFB_OP_MIN>, blocks, threads, s, a, root);   \
    case KFB_OP_MAX: return launch(comm, k_reduce_root<T, KFB_OP_MAX>, blocks, threads, s, a, root);   \
    case KFB_OP_PROD: return launch(comm, k_reduce_root<T, KFB_OP_PROD>, blocks, threads, s, a, root); \
    }                                                                          \
    break;
    switch (dtype) {
    case KFB_F32: KFB_RED(float)
    case KFB_BF16: KFB_RED(__nv_bfloat16)
    case KFB_F16: KFB_RED(__half)
    case KFB_I32: KFB_RED(int32_t)
    case KFB_F64: KFB_RED(double)
    case KFB_I64: KFB_RED(int64_t)
    case KFB_U8: KFB_RED(uint8_t)
    }
#undef KFB_RED
    set_last_error("unsupported dtype/op %d/%d", dtype, op);
    return KFB_ERR_ARG;
}

extern "C" int kfb_barrier(const kfb_comm *comm, void *stream)
{
    int rc = check_comm(comm, 1, 32);
    if (rc) return rc;
    MoveArgs m;
    fill_move(m, comm, 0);
    return launch(comm, k_barrier, 1, 32,
                  reinterpret_cast<cudaStream_t>(stream), m);
}

extern "C" int kfb_consensus(const kfb_comm *comm, void *const *in,
                             int64_t bytes, int32_t *agree_out, int blocks,
                             int threads, void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    MoveArgs m;
    fill_move(m, comm, bytes);
    for (int r = 0; r < comm->np; ++r) m.in_peers[r] = in[r];
    k_set_i32<<<1, KFB_MAX_RANKS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(agree_out, comm->rank < 0 ? comm->np : 1, 1);
    KFB_CUDA_OK(cudaGetLastError());
    return launch(comm, k_consensus, blocks, threads,
                  reinterpret_cast<cudaStream_t>(stream), m, agree_out);
}

extern "C" int kfb_graph_allreduce(const kfb_comm *comm, void *const *bufs, int dtype, int op, int n_plans,
                                   const int64_t *chunk_lo_bytes, const int64_t *chunk_bytes, int n_rounds,
                                   int n_reduce_rounds, const int32_t *round_begin, const int32_t *edge_plan,
                                   const int32_t *edge_src, int blocks, int threads, void *stream)
{
    int rc = check_comm(comm, blocks, threads);
    if (rc) return rc;
    if (comm->rank < 0) {
        set_last_error("kfb_graph_allreduce: needs real ranks (the edge tables are per rank)");
        return KFB_ERR_ARG;
    }
    if (n_plans < 1 || n_plans > KFB_GRAPH_MAX_PLANS || n_rounds < 0 || n_rounds > KFB_GRAPH_MAX_ROUNDS ||
        n_reduce_rounds < 0 || n_reduce_rounds > n_rounds || round_begin[n_rounds] > KFB_GRAPH_MAX_EDGES) {
        set_last_error("kfb_graph_allreduce: graph too large (plans %d/%d rounds %d/%d edges %d/%d)", n_plans,
                       KFB_GRAPH_MAX_PLANS, n_rounds, KFB_GRAPH_MAX_ROUNDS, n_rounds >= 0 ? round_begin[n_rounds] : -1,
                       KFB_GRAPH_MAX_EDGES);
        return KFB_ERR_ARG;
    }
    if (!dtype_size(dtype)) {
        set_last_error("unsupported dtype %d", dtype);
        return KFB_ERR_ARG;
    }
    GraphArgs a;
    memset(&a, 0, sizeof(a));
    fill_pads(a.pads, comm);
    a.local_epoch = reinterpret_cast<uint32_t *>(comm->local_epoch);
    for (int r = 0; r < comm->np; ++r) a.bufs[r] = bufs[r];
    a.rank            = comm->rank;
    a.np              = comm->np;
    a.n_rounds        = n_rounds;
    a.n_reduce_rounds = n_reduce_rounds;
    for (int r = 0; r <= n_rounds; ++r) a.round_begin[r] = round_begin[r];
    for (int e = 0; e < round_begin[n_rounds]; ++e) {
        if (edge_plan[e] < 0 || edge_plan[e] >= n_plans || edge_src[e] < 0 || edge_src[e] >= comm->np) {
            set_last_error("kfb_graph_allreduce: edge %d out of range", e);
            return KFB_ERR_ARG;
        }
        a.edge_plan[e] = (unsigned char)edge_plan[e];
        a.edge_src[e]  = (unsigned char)edge_src[e];
    }
    for (int i = 0; i < n_plans; ++i) {
        if (chunk_lo_bytes[i] % 16) {
            set_last_error("kfb_graph_allreduce: chunks must start on 16-byte boundaries");
            return KFB_ERR_ARG;
        }
        a.chunk_lo[i] = chunk_lo_bytes[i] / 16;
        a.chunk_n[i]  = (chunk_bytes[i] + 15) / 16;
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define KFB_GRAPH(T)                                                           \
    switch (op) {                                                              \
    case KFB_OP_SUM: k_graph_allreduce<T, KFB_OP_SUM><<<blocks, threads, 0, s>>>(a); break;   \
    case KFB_OP_MIN: k_graph_allreduce<T, KFB_OP_MIN><<<blocks, threads, 0, s>>>(a); break;   \
    case KFB_OP_MAX: k_graph_allreduce<T, KFB_OP_MAX><<<blocks, threads, 0, s>>>(a); break;   \
    case KFB_OP_PROD: k_graph_allreduce<T, KFB_OP_PROD><<<blocks, threads, 0, s>>>(a); break; \
    default: set_last_error("bad op %d", op); return KFB_ERR_ARG;              \
    }                                                                          \
    break;
    switch (dtype) {
    case KFB_F32: KFB_GRAPH(float)
    case KFB_BF16: KFB_GRAPH(__nv_bfloat16)
    case KFB_F16: KFB_GRAPH(__half)
    case KFB_I32: KFB_GRAPH(int32_t)
    case KFB_F64: KFB_GRAPH(double)
    case KFB_I64: KFB_GRAPH(int64_t)
    case KFB_U8: KFB_GRAPH(uint8_t)
    }
#undef KFB_GRAPH
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_edge_reduce(void *dst, const void *src, int64_t count,
                               int dtype, int op, int blocks, int threads,
                               void *stream)
{
    const size_t es = dtype_size(dtype);
    if (!es) {
        set_last_error("unsupported dtype %d", dtype);
        return KFB_ERR_ARG;
    }
    const long long nvec = (long long)((count * es + 15) / 16);
    cudaStream_t s       = reinterpret_cast<cudaStream_t>(stream);
    uint4 *d             = reinterpret_cast<uint4 *>(dst);
    const uint4 *sp      = reinterpret_cast<const uint4 *>(src);
#define KFB_EDGE(T)                                                            \
    switch (op) {                                                              \
    case KFB_OP_SUM: k_edge_reduce<T, KFB_OP_SUM><<<blocks, threads, 0, s>>>(d, sp, nvec); break;   \
    case KFB_OP_MIN: k_edge_reduce<T, KFB_OP_MIN><<<blocks, threads, 0, s>>>(d, sp, nvec); break;   \
    case KFB_OP_MAX: k_edge_reduce<T, KFB_OP_MAX><<<blocks, threads, 0, s>>>(d, sp, nvec); break;   \
    case KFB_OP_PROD: k_edge_reduce<T, KFB_OP_PROD><<<blocks, threads, 0, s>>>(d, sp, nvec); break; \
    default: set_last_error("bad op %d", op); return KFB_ERR_ARG;              \
    }                                                                          \
    break;
    switch (dtype) {
    case KFB_F32: KFB_EDGE(float)
    case KFB_BF16: KFB_EDGE(__nv_bfloat16)
    case KFB_F16: KFB_EDGE(__half)
    case KFB_I32: KFB_EDGE(int32_t)
    case KFB_F64: KFB_EDGE(double)
    case KFB_I64: KFB_EDGE(int64_t)
    case KFB_U8: KFB_EDGE(uint8_t)
    }
#undef KFB_EDGE
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_edge_copy(void *dst, const void *src, int64_t bytes,
                             int blocks, int threads, void *stream)
{
    const long long nvec = (long long)((bytes + 15) / 16);
    k_edge_copy<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<uint4 *>(dst), reinterpret_cast<const uint4 *>(src),
        nvec);
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}
