// Fused data-parallel optimizer steps and multi-tensor helpers for sm_100a.
//
//   kfb_fuse / kfb_defuse        bucket gather/scatter with cast + ||g||^2
//                                (reference: tf.concat fuse/defuse,
//                                 srcs/python/kungfu/tensorflow/ops/__init__.py:29-46)
//   kfb_ssgd_fused_step          reduce-scatter -> SGD on the owned shard ->
//                                all-gather of the new parameters, one kernel
//                                (sync_sgd.py:78-109 did allreduce, g/np and
//                                 the optimizer as separate TF ops)
//   kfb_sma_fused_step           model average + blend + local SGD, one kernel
//                                (sma_sgd.py:50-74)
//   kfb_pair_avg_step            one-sided peer pull + 1/2 blend + SGD + publish
//                                (async_sgd.py:110-142, peer_to_peer.cpp:141-149)
//   kfb_grad_variance            E[g^2]-E[g]^2 in one pass (grad_variance.py:46-59)
//   kfb_gns_finalize             noise-scale estimators + EMA
//                                (ops/monitor.py:6-17, cpu/collective.cpp:223-258)
//
// All flat buffers are processed in packets of 8 elements; the arenas that own
// them pad every tensor to a multiple of 16 elements.
#include <cstring>
#include <mutex>
#include <type_traits>
#include <vector>

#include "common.cuh"
#include "multimem.cuh"

namespace kfb {

// ---- 8-element packets of any storage type as fp32 ---------------------------
template <typename T> struct Pk8;
template <> struct Pk8<float> {
    __device__ static void load(const void *base, long long i, float *v)
    {
        const uint4 *p = reinterpret_cast<const uint4 *>(
                             reinterpret_cast<const float *>(base)) +
                         2 * i;
        const uint4 a = ld_stream16(p), b = ld_stream16(p + 1);
        Pack<float>::unpack(a, v);
        Pack<float>::unpack(b, v + 4);
    }
    __device__ static void store(void *base, long long i, const float *v)
    {
        uint4 *p = reinterpret_cast<uint4 *>(base) + 2 * i;
        st_stream16(p, Pack<float>::pack(v));
        st_stream16(p + 1, Pack<float>::pack(v + 4));
    }
};
template <> struct Pk8<__nv_bfloat16> {
    __device__ static void load(const void *base, long long i, float *v)
    {
        Pack<__nv_bfloat16>::unpack(
            ld_stream16(reinterpret_cast<const uint4 *>(base) + i), v);
    }
    __device__ static void store(void *base, long long i, const float *v)
    {
        st_stream16(reinterpret_cast<uint4 *>(base) + i,
                    Pack<__nv_bfloat16>::pack(v));
    }
};
template <> struct Pk8<__half> {
    __device__ static void load(const void *base, long long i, float *v)
    {
        Pack<__half>::unpack(
            ld_stream16(reinterpret_cast<const uint4 *>(base) + i), v);
    }
    __device__ static void store(void *base, long long i, const float *v)
    {
        st_stream16(reinterpret_cast<uint4 *>(base) + i, Pack<__half>::pack(v));
    }
};

struct SgdH {
    float lr, momentum, dampening, wd;
    int nesterov, first;
    // kind 0: torch.optim.SGD. kind 1: torch.optim.AdamW (decoupled decay),
    // kind 2: torch.optim.Adam (L2 decay folded into the gradient).
    int kind;
    float beta1, beta2, eps;
    float bc1, bc2;         // bias corrections when `step` is NULL
    const long long *step;  // device scalar: completed steps (t = *step + 1)
    float *v;               // Adam second moment, same indexing as the momentum
};

// Adam's bias corrections are a function of the step number. Reading the
// number from device memory keeps a captured CUDA graph valid from one replay
// to the next (the caller advances *step once per training step).
struct Bias {
    float c1, rs2;  // 1 - beta1^t, 1/sqrt(1 - beta2^t)
};
__device__ __forceinline__ Bias opt_bias(const SgdH &h)
{
    Bias b;
    b.c1 = 1.f, b.rs2 = 1.f;
    if (h.kind != 0) {
        float c1 = h.bc1, c2 = h.bc2;
        if (h.step) {
            const float t = (float)(*h.step + 1);
            c1 = 1.f - powf(h.beta1, t);
            c2 = 1.f - powf(h.beta2, t);
        }
        b.c1  = c1;
        b.rs2 = rsqrtf(c2);
    }
    return b;
}

// One update on 8 lanes; p/m/v are the fp32 master weights and optimizer slots.
__device__ __forceinline__ void
opt_update8(float *p, float *m, float *v, const float *g, const SgdH &h, const Bias &b)
{
    if (h.kind == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float d = fmaf(h.wd, p[k], g[k]);
            if (h.momentum != 0.f) {
                m[k] = h.first ? d : fmaf(h.momentum, m[k], (1.f - h.dampening) * d);
                d    = h.nesterov ? fmaf(h.momentum, m[k], d) : m[k];
            }
            p[k] = fmaf(-h.lr, d, p[k]);
        }
    } else {
        const float step_size = h.lr / b.c1;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float gk = g[k];
            if (h.kind == 1) p[k] *= 1.f - h.lr * h.wd;
            else gk = fmaf(h.wd, p[k], gk);
            m[k]            = fmaf(h.beta1, m[k], (1.f - h.beta1) * gk);
            v[k]            = fmaf(h.beta2, v[k], (1.f - h.beta2) * gk * gk);
            const float den = fmaf(sqrtf(v[k]), b.rs2, h.eps);
            p[k]            = fmaf(-step_size, m[k] / den, p[k]);
        }
    }
}

// Which optimizer slots a packet update reads and writes.
__device__ __forceinline__ bool opt_loads_m(const SgdH &h) { return h.kind != 0 || (h.momentum != 0.f && !h.first); }
__device__ __forceinline__ bool opt_stores_m(const SgdH &h) { return h.kind != 0 || h.momentum != 0.f; }

__device__ __forceinline__ void
slots_load(const SgdH &h, const float *mom, const float *v2, long long i, float *m, float *v)
{
    if (opt_loads_m(h)) Pk8<float>::load(mom, i, m);
    if (h.kind != 0) Pk8<float>::load(v2, i, v);
}
__device__ __forceinline__ void
slots_store(const SgdH &h, float *mom, float *v2, long long i, const float *m, const float *v)
{
    if (opt_stores_m(h)) Pk8<float>::store(mom, i, m);
    if (h.kind != 0) Pk8<float>::store(v2, i, v);
}

static SgdH to_h(const kfb_sgd_hyper *h)
{
    SgdH r;
    r.lr        = h->lr;
    r.momentum  = h->momentum;
    r.dampening = h->dampening;
    r.wd        = h->weight_decay;
    r.nesterov  = h->nesterov;
    r.first     = h->first_step;
    r.kind      = h->kind;
    r.beta1     = h->beta1;
    r.beta2     = h->beta2;
    r.eps       = h->eps;
    r.bc1       = h->bias_corr1;
    r.bc2       = h->bias_corr2;
    r.step      = reinterpret_cast<const long long *>(h->step);
    r.v         = h->exp_avg_sq;
    return r;
}

// ---- fuse / defuse -------------------------------------------------------------
#define KFB_FUSE_MAX 64
struct FuseArgs {
    const void *src[KFB_FUSE_MAX];
    long long count[KFB_FUSE_MAX];
    long long off[KFB_FUSE_MAX];
    int n;
};

template <typename S> __device__ __forceinline__ float to_f32(S x);
template <> __device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_f32<__half>(__half x) { return __half2float(x); }
template <typename D> __device__ __forceinline__ D from_f32(float x);
template <> __device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }
template <> __device__ __forceinline__ __half from_f32<__half>(float x) { return __float2half_rn(x); }

// grid = (blocks_x, n tensors). Sources are arbitrary torch tensors (only
// element-aligned), so this kernel is scalar on the source side when the
// source is not 16-byte aligned, vector otherwise. Block partial sums of x^2
// go to scratch[blockIdx.y * gridDim.x + blockIdx.x] -- fixed slots, summed in
// slot order by k_sum_slots, so the norm is reproducible.
template <typename S, typename D, bool DEFUSE>
__global__ void __launch_bounds__(256)
    k_fuse(const __grid_constant__ FuseArgs a, void *flat, float *partials)
{
    const int t            = blockIdx.y;
    const long long n      = a.count[t];
    const long long stride = (long long)gridDim.x * blockDim.x;
    float ss               = 0.f;
    if (!DEFUSE) {
        const S *src = reinterpret_cast<const S *>(a.src[t]);
        D *dst       = reinterpret_cast<D *>(flat) + a.off[t];
        constexpr int V = 16 / sizeof(S);  // source elements per 16 bytes
        const bool vec_ok = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                            ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
        const long long nv = vec_ok ? n / V : 0;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < nv; i += stride) {
            const uint4 v = ld_stream16(reinterpret_cast<const uint4 *>(src) + i);
            const S *e    = reinterpret_cast<const S *>(&v);
            alignas(16) D o[V];
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float f = to_f32<S>(e[k]);
                ss            = fmaf(f, f, ss);
                o[k]          = from_f32<D>(f);
            }
            if (sizeof(D) == sizeof(S)) {
                st_stream16(reinterpret_cast<uint4 *>(dst) + i,
                            *reinterpret_cast<const uint4 *>(o));
            } else if (sizeof(D) < sizeof(S)) {  // f32 -> 16-bit: 8 bytes out
                *reinterpret_cast<uint2 *>(dst + i * V) =
                    *reinterpret_cast<const uint2 *>(o);
            } else {  // 16-bit -> f32: 32 bytes out
                uint4 *q = reinterpret_cast<uint4 *>(dst + i * V);
                st_stream16(q, reinterpret_cast<const uint4 *>(o)[0]);
                st_stream16(q + 1, reinterpret_cast<const uint4 *>(o)[1]);
            }
        }
        for (long long i = nv * V + (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < n; i += stride) {
            const float f = to_f32<S>(src[i]);
            ss            = fmaf(f, f, ss);
            dst[i]        = from_f32<D>(f);
        }
        if (partials) {
            __shared__ float red[32];
            const float part = block_sum(ss, red);
            if (threadIdx.x == 0)
                partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = part;
        }
    } else {
        // flat (S) -> tensors (D)
        const S *src = reinterpret_cast<const S *>(flat) + a.off[t];
        D *dst       = reinterpret_cast<D *>(const_cast<void *>(a.src[t]));
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < n; i += stride)
            dst[i] = from_f32<D>(to_f32<S>(src[i]));
    }
}

// out[0] (+)= sum of partials[0..n) in index order, one warp-free thread per
// 32 slots then a serial tail: deterministic and cheap (n is a few thousand).
__global__ void k_sum_slots(const float *partials, int n, float *out,
                            int accumulate)
{
    __shared__ float part[256];
    // thread t sums slots [t*chunk, (t+1)*chunk) serially; thread 0 then adds
    // the 256 chunk sums serially: the expression tree depends only on n.
    const int chunk = (n + 255) / 256;
    float s         = 0.f;
    for (int i = threadIdx.x * chunk; i < min(n, (int)(threadIdx.x + 1) * chunk); ++i)
        s += partials[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < 256; ++i) t += part[i];
        out[0] = accumulate ? out[0] + t : t;
    }
}

// ---- S-SGD fused step --------------------------------------------------------------
struct StepArgs {
    PeerPads pads;
    uint32_t *local_epoch;
    float *scratch;
    void *grads[KFB_MAX_RANKS];
    void *params[KFB_MAX_RANKS];
    void *stage[KFB_MAX_RANKS];
    const void *grad_local;
    float *master;
    float *mom;
    long long npk;  // number of 8-element packets
    float grad_scale;
    float alpha;
    SgdH h;
    int rank, np;
    float *sumsq_out;
    // KFB_STEP_PRESIGNALED: the arrival of the entry barrier was sent by kfb_peer_signal, only wait.
    // KFB_STEP_NO_EXIT_BARRIER: do not meet the peers again on the way out (a later launch does);
    // sumsq_out then only receives THIS rank's shard partial.
    int flags;
};

template <typename G, typename P, int NP, bool NORM>
__global__ void __launch_bounds__(512)
    k_ssgd_fused(const __grid_constant__ StepArgs a)
{
    const int rank = a.rank < 0 ? (int)blockIdx.y : a.rank;
    const int np   = NP > 0 ? NP : a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    float *scratch = a.scratch + (a.rank < 0 ? rank * KFB_SCRATCH_FLOATS : 0);
    // emulation: each emulated rank has its own master/momentum copy
    float *master = a.master + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0);
    float *mom    = a.mom + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0);
    float *mom2   = a.h.v ? a.h.v + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0) : nullptr;
    const Bias bias = opt_bias(a.h);
    __shared__ float red[32];
    const bool exit_barrier = !(a.flags & KFB_STEP_NO_EXIT_BARRIER);

    if (np > 1) {
        if (a.flags & KFB_STEP_PRESIGNALED) peer_wait_row(a.pads, epoch, rank, np, blockIdx.x);
        else peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    }
    const long long per    = (a.npk + np - 1) / np;
    const long long lo     = (long long)rank * per;
    const long long hi     = min(lo + per, a.npk);
    const long long stride = (long long)gridDim.x * blockDim.x;
    float ss               = 0.f;
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < hi; i += stride) {
        float g[8];
        if (NP > 0) {
            float t[NP > 0 ? NP : 1][8];
#pragma unroll
            for (int r = 0; r < NP; ++r) Pk8<G>::load(a.grads[r], i, t[r]);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = t[0][k];
#pragma unroll
            for (int r = 1; r < NP; ++r)
#pragma unroll
                for (int k = 0; k < 8; ++k) g[k] += t[r][k];
        } else {
            Pk8<G>::load(a.grads[0], i, g);
            for (int r = 1; r < np; ++r) {
                float t[8];
                Pk8<G>::load(a.grads[r], i, t);
#pragma unroll
                for (int k = 0; k < 8; ++k) g[k] += t[k];
            }
        }
        float p[8], m[8], v[8];
        Pk8<float>::load(master, i, p);
        slots_load(a.h, mom, mom2, i, m, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            g[k] *= a.grad_scale;
            if (NORM) ss = fmaf(g[k], g[k], ss);
        }
        opt_update8(p, m, v, g, a.h, bias);
        Pk8<float>::store(master, i, p);
        slots_store(a.h, mom, mom2, i, m, v);
        for (int k = 0; k < np; ++k) Pk8<P>::store(a.params[(rank + k) % np], i, p);
    }
    __shared__ float s_part;
    if (NORM) {
        const float part = block_sum(ss, red);
        if (threadIdx.x == 0) s_part = part;
        __syncthreads();
        if (exit_barrier && threadIdx.x < np)
            st_relaxed_sys_f32(
                pad_scalar(a.pads.pad[threadIdx.x], blockIdx.x, rank), s_part);
    }
    if (np > 1 || NORM) {
        if (np > 1 && exit_barrier) peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
        else __syncthreads();
    }
    if (NORM) {
        __shared__ bool is_last;
        if (threadIdx.x == 0) {
            float t = 0.f;
            if (exit_barrier && !(a.flags & KFB_STEP_PARTIAL_NORM)) {
                for (int r = 0; r < np; ++r)
                    t += ld_relaxed_sys_f32(pad_scalar(a.pads.pad[rank], blockIdx.x, r));
            } else {
                t = s_part;
            }
            scratch[blockIdx.x] = t;
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

// NVLS variant of the fused S-SGD step: the NVSwitch sums the gradient shard
// (multimem.ld_reduce, fp32 accumulation) and multicasts the updated
// parameters (multimem.st), so each GPU moves ~1x the bucket instead of
// 2(n-1)/n x. a.stage[0] / a.stage[1] carry the multicast addresses of the
// gradient and parameter buffers.
template <typename T> struct Mc8;
template <> struct Mc8<float> {
    __device__ static void ld_reduce(const void *mc, long long i, float *v)
    {
        const uint4 *p = reinterpret_cast<const uint4 *>(mc) + 2 * i;
        Pack<float>::unpack(Multimem<float>::ld_reduce(p), v);
        Pack<float>::unpack(Multimem<float>::ld_reduce(p + 1), v + 4);
    }
    __device__ static void st(void *mc, long long i, const float *v)
    {
        uint4 *p = reinterpret_cast<uint4 *>(mc) + 2 * i;
        Multimem<float>::st(p, Pack<float>::pack(v));
        Multimem<float>::st(p + 1, Pack<float>::pack(v + 4));
    }
};
template <> struct Mc8<__nv_bfloat16> {
    __device__ static void ld_reduce(const void *mc, long long i, float *v)
    {
        Pack<__nv_bfloat16>::unpack(Multimem<__nv_bfloat16>::ld_reduce(reinterpret_cast<const uint4 *>(mc) + i), v);
    }
    __device__ static void st(void *mc, long long i, const float *v)
    {
        Multimem<__nv_bfloat16>::st(reinterpret_cast<uint4 *>(mc) + i, Pack<__nv_bfloat16>::pack(v));
    }
};
template <> struct Mc8<__half> {
    __device__ static void ld_reduce(const void *mc, long long i, float *v)
    {
        Pack<__half>::unpack(Multimem<__half>::ld_reduce(reinterpret_cast<const uint4 *>(mc) + i), v);
    }
    __device__ static void st(void *mc, long long i, const float *v)
    {
        Multimem<__half>::st(reinterpret_cast<uint4 *>(mc) + i, Pack<__half>::pack(v));
    }
};

template <typename G, typename P, bool NORM>
__global__ void __launch_bounds__(512)
    k_ssgd_fused_nvls(const __grid_constant__ StepArgs a)
{
    const int rank = a.rank, np = a.np;
    const Bias bias = opt_bias(a.h);
    __shared__ float red[32];
    const bool exit_barrier = !(a.flags & KFB_STEP_NO_EXIT_BARRIER);
    if (a.flags & KFB_STEP_PRESIGNALED) peer_wait_row(a.pads, a.local_epoch, rank, np, blockIdx.x);
    else peer_barrier(a.pads, a.local_epoch, rank, np, blockIdx.x);
    const long long per    = (a.npk + np - 1) / np;
    const long long lo     = (long long)rank * per;
    const long long hi     = min(lo + per, a.npk);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const void *mc_g       = a.stage[0];
    void *mc_p             = a.stage[1];
    float ss               = 0.f;
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
        float g[8], p[8], m[8], v[8];
        Mc8<G>::ld_reduce(mc_g, i, g);
        Pk8<float>::load(a.master, i, p);
        slots_load(a.h, a.mom, a.h.v, i, m, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            g[k] *= a.grad_scale;
            if (NORM) ss = fmaf(g[k], g[k], ss);
        }
        opt_update8(p, m, v, g, a.h, bias);
        Pk8<float>::store(a.master, i, p);
        slots_store(a.h, a.mom, a.h.v, i, m, v);
        Mc8<P>::st(mc_p, i, p);
    }
    __shared__ float s_part;
    if (NORM) {
        const float part = block_sum(ss, red);
        if (threadIdx.x == 0) s_part = part;
        __syncthreads();
        if (exit_barrier && threadIdx.x < np)
            st_relaxed_sys_f32(pad_scalar(a.pads.pad[threadIdx.x], blockIdx.x, rank), s_part);
    }
    if (exit_barrier) peer_barrier(a.pads, a.local_epoch, rank, np, blockIdx.x);
    else {
        // the parameters this rank multicast must be visible before a LATER launch's barrier lets
        // anybody read them: that barrier's release covers this kernel's stores (same stream)
        __threadfence_system();
        __syncthreads();
    }
    if (NORM) {
        __shared__ bool is_last;
        if (threadIdx.x == 0) {
            float t = 0.f;
            if (exit_barrier && !(a.flags & KFB_STEP_PARTIAL_NORM)) {
                for (int r = 0; r < np; ++r)
                    t += ld_relaxed_sys_f32(pad_scalar(a.pads.pad[rank], blockIdx.x, r));
            } else {
                t = s_part;
            }
            a.scratch[blockIdx.x] = t;
            __threadfence();
            unsigned int *ticket = reinterpret_cast<unsigned int *>(a.scratch + KFB_MAX_BLOCKS);
            is_last              = (atomicAdd(ticket, 1u) == gridDim.x - 1);
        }
        __syncthreads();
        if (is_last && threadIdx.x == 0) {
            __threadfence();
            float t = 0.f;
            for (unsigned b = 0; b < gridDim.x; ++b) t += *reinterpret_cast<volatile float *>(a.scratch + b);
            a.sumsq_out[0] = t;
            *reinterpret_cast<unsigned int *>(a.scratch + KFB_MAX_BLOCKS) = 0u;
        }
    }
}

// Arrival of the next collective's entry barrier, sent ahead of the collective
// (see peer_arrive_rows in common.cuh). One CTA.
__global__ void __launch_bounds__(512)
    k_peer_signal(PeerPads pads, uint32_t *local_epoch, int rank, int np, int nrows)
{
    peer_arrive_rows(pads, local_epoch, rank, np, nrows);
}

// Plain local flat SGD on a reduced gradient (any-rank, no peers).
template <typename G, typename P>
__global__ void __launch_bounds__(512)
    k_sgd_flat(const void *grad, void *param, float *master, float *mom,
               long long npk, float grad_scale, SgdH h)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    const Bias bias        = opt_bias(h);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < npk; i += stride) {
        float g[8], p[8], m[8], v[8];
        Pk8<G>::load(grad, i, g);
        Pk8<float>::load(master, i, p);
        slots_load(h, mom, h.v, i, m, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] *= grad_scale;
        opt_update8(p, m, v, g, h, bias);
        Pk8<float>::store(master, i, p);
        slots_store(h, mom, h.v, i, m, v);
        if (param != (void *)master) Pk8<P>::store(param, i, p);
    }
}

// ---- SMA fused step -------------------------------------------------------------------
// NVLS == true: the switch sums the parameter shard (multimem.ld_reduce on the
// multicast alias a.stage[KFB_MAX_RANKS-2]) and multicasts the fp32 average
// into every rank's staging buffer (multimem.st on a.stage[KFB_MAX_RANKS-1]),
// so the NVLink leg moves ~1x the model instead of 2(n-1)/n x.
// PHASE: 0 = everything, 1 = average only (can run while forward/backward of
// this step are still executing: parameters are only read), 2 = blend + update
// only (purely local).
template <typename G, typename P, bool NVLS, int PHASE>
__global__ void __launch_bounds__(512)
    k_sma_fused(const __grid_constant__ StepArgs a)
{
    const int rank = a.rank < 0 ? (int)blockIdx.y : a.rank;
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    float *master = a.master + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0);
    float *mom    = a.mom + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0);
    float *mom2   = a.h.v ? a.h.v + (a.rank < 0 ? (long long)rank * a.npk * 8 : 0) : nullptr;
    const void *grad =
        a.rank < 0 ? a.grads[rank] : a.grad_local;  // LOCAL gradient
    const long long per    = (a.npk + np - 1) / np;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const float inv_np     = 1.f / (float)np;
    const Bias bias        = opt_bias(a.h);

    if (np > 1 && PHASE != 2) {
        // phase 1: average my shard of everybody's parameters, push to all
        peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
        const long long lo = (long long)rank * per;
        const long long hi = min(lo + per, a.npk);
        for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < hi; i += stride) {
            float s[8];
            if (NVLS) {
                Mc8<P>::ld_reduce(a.stage[KFB_MAX_RANKS - 2], i, s);
            } else {
                Pk8<P>::load(a.params[0], i, s);
                for (int r = 1; r < np; ++r) {
                    float t[8];
                    Pk8<P>::load(a.params[r], i, t);
#pragma unroll
                    for (int k = 0; k < 8; ++k) s[k] += t[k];
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] *= inv_np;
            if (NVLS) {
                Mc8<float>::st(a.stage[KFB_MAX_RANKS - 1], i, s);
            } else {
                for (int k = 0; k < np; ++k)
                    Pk8<float>::store(a.stage[(rank + k) % np], i, s);
            }
        }
        // CTA b of every rank has now (1) read the slice of my parameters it
        // needs and (2) delivered its slice of the average to me.
        peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    }
    if (PHASE == 1) return;
    // phase 2: blend + local update. In the single-launch form every CTA only
    // touches the packets CTA b of each rank delivered in phase 1 (same
    // grid-stride pattern inside every shard); as a launch of its own it may
    // have any grid, the averages are complete before it starts.
    for (int s_ = 0; s_ < np; ++s_) {
        const long long lo = (long long)s_ * per;
        const long long hi = min(lo + per, a.npk);
        for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x;
             i < hi; i += stride) {
            float p[8], m[8], v[8], g[8];
            Pk8<float>::load(master, i, p);
            if (np > 1) {
                float avg[8];
                Pk8<float>::load(a.stage[rank], i, avg);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    p[k] = fmaf(a.alpha, avg[k] - p[k], p[k]);
            }
            Pk8<G>::load(grad, i, g);
            slots_load(a.h, mom, mom2, i, m, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] *= a.grad_scale;
            opt_update8(p, m, v, g, a.h, bias);
            Pk8<float>::store(master, i, p);
            slots_store(a.h, mom, mom2, i, m, v);
            Pk8<P>::store(a.params[rank], i, p);
        }
    }
}

// ---- PairAveraging fused step ---------------------------------------------------------
struct PairArgs {
    uint32_t *seq_peer;  // peer's published version word (peer mapped)
    uint32_t *seq_self;
    const void *snap_peer;
    void *snap_self;
    long long slot_bytes;
    void *param;
    const void *grad;
    float *master;
    float *mom;
    float *scratch;
    long long npk;
    float grad_scale;
    SgdH h;
    int32_t *torn_out;
    int has_peer;
    // Device-side peer selection (sel_np > 1): the target of this step is drawn in the kernel from
    // (seed, *sel_step, rank), so nothing in the launch depends on the host's choice and the whole
    // training step can live in one CUDA graph. The caller advances *sel_step once per step.
    int sel_np, sel_rank;
    const long long *sel_step;
    unsigned long long sel_seed;
    uint32_t *seq_all[KFB_MAX_RANKS];
    const void *snap_all[KFB_MAX_RANKS];
};

// Snapshot protocol (per rank, words in its signal pad):
//   seq  number of published versions; version k lives in slot (k & 1)
//   wb   the version the owner is writing right now ("write begin")
// The owner stores wb = seq + 1, fences, writes slot ((seq + 1) & 1), fences,
// stores seq + 1 (release). A reader takes k = seq (acquire), reads slot
// (k & 1), then re-reads wb (acquire): the slot it read is recycled by
// version k + 2, so wb >= k + 2 means the copy may mix two versions. The
// reference tolerates such reads silently (store/blob.go:20-26 copies without
// a write lock); here they are counted in torn_out, without false negatives.
template <typename G, typename P, bool UPDATE>
__global__ void __launch_bounds__(512)
    k_pair_avg(const __grid_constant__ PairArgs a)
{
    __shared__ uint32_t s_k, s_self;
    __shared__ uint32_t *s_seq_peer;
    __shared__ const void *s_snap_peer;
    if (threadIdx.x == 0) {
        uint32_t *seq_peer    = a.seq_peer;
        const void *snap_peer = a.snap_peer;
        if (a.sel_np > 1) {
            // splitmix64 of (seed, step, rank): uniform over the OTHER np-1 ranks (the reference draws
            // U[0, np) and bumps a self-hit to the next rank, async_sgd.py:71-75)
            unsigned long long z =
                a.sel_seed + 0x9E3779B97F4A7C15ull * ((unsigned long long)*a.sel_step * KFB_MAX_RANKS + a.sel_rank + 1);
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            z ^= z >> 31;
            const int target = (a.sel_rank + 1 + (int)(z % (unsigned long long)(a.sel_np - 1))) % a.sel_np;
            seq_peer         = a.seq_all[target];
            snap_peer        = a.snap_all[target];
        }
        s_seq_peer  = seq_peer;
        s_snap_peer = snap_peer;
        s_k         = a.has_peer ? ld_acquire_sys(seq_peer) : 0u;
        s_self      = ld_relaxed_sys(a.seq_self);
        if (UPDATE) {
            // announce the version about to be written before any CTA touches its slot
            st_release_sys(a.seq_self + 1, s_self + 1);
            __threadfence_system();
        }
    }
    __syncthreads();
    const uint32_t k = s_k, kself = s_self;
    const char *peer =
        reinterpret_cast<const char *>(s_snap_peer) + (k & 1) * a.slot_bytes;
    char *mine =
        reinterpret_cast<char *>(a.snap_self) + ((kself + 1) & 1) * a.slot_bytes;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const Bias bias        = opt_bias(a.h);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < a.npk; i += stride) {
        float p[8], m[8], v[8], g[8];
        if (UPDATE) Pk8<float>::load(a.master, i, p);
        else Pk8<P>::load(a.param, i, p);
        if (k != 0) {  // peer has published at least once
            float o[8];
            Pk8<P>::load(peer, i, o);
#pragma unroll
            for (int j = 0; j < 8; ++j) p[j] = 0.5f * (p[j] + o[j]);
        }
        if (UPDATE) {
            Pk8<G>::load(a.grad, i, g);
            slots_load(a.h, a.mom, a.h.v, i, m, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] *= a.grad_scale;
            opt_update8(p, m, v, g, a.h, bias);
            Pk8<float>::store(a.master, i, p);
            slots_store(a.h, a.mom, a.h.v, i, m, v);
            Pk8<P>::store(mine, i, p);
        }
        Pk8<P>::store(a.param, i, p);
    }
    // The last CTA to finish publishes the new version and checks whether the
    // slot we read was being recycled underneath us.
    __shared__ bool is_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned int *ticket =
            reinterpret_cast<unsigned int *>(a.scratch + KFB_MAX_BLOCKS + 1);
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
        if (is_last) {
            *ticket = 0u;
            __threadfence_system();
            if (UPDATE) st_release_sys(a.seq_self, kself + 1);
            if (a.has_peer && a.torn_out && k != 0) {
                const uint32_t wb = ld_acquire_sys(s_seq_peer + 1);
                if ((int32_t)(wb - k) >= 2) atomicAdd(a.torn_out, 1);
            }
        }
    }
}

__global__ void __launch_bounds__(512)
    k_pair_publish(uint32_t *seq_self, const uint4 *param, char *snap_self,
                   long long slot_bytes, long long nvec, float *scratch)
{
    __shared__ uint32_t s_self;
    if (threadIdx.x == 0) {
        s_self = ld_relaxed_sys(seq_self);
        st_release_sys(seq_self + 1, s_self + 1);
        __threadfence_system();
    }
    __syncthreads();
    const uint32_t kself = s_self;
    uint4 *mine = reinterpret_cast<uint4 *>(snap_self + ((kself + 1) & 1) * slot_bytes);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
         i < nvec; i += stride)
        st_stream16(mine + i, ld_stream16(param + i));
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned int *ticket =
            reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS + 1);
        if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
            *ticket = 0u;
            __threadfence_system();
            st_release_sys(seq_self, kself + 1);
        }
    }
}

// Byte copy of the target's current snapshot into a local buffer, with the same
// version check as k_pair_avg (models of several dtypes: pull once, blend per
// dtype afterwards). *got_out := 1 if the peer had published anything.
__global__ void __launch_bounds__(512)
    k_pair_pull(uint32_t *seq_peer, const char *snap_peer, long long slot_bytes, uint4 *dst, long long nvec,
                float *scratch, int32_t *torn_out, int32_t *got_out)
{
    __shared__ uint32_t s_k;
    if (threadIdx.x == 0) s_k = ld_acquire_sys(seq_peer);
    __syncthreads();
    const uint32_t k = s_k;
    if (k != 0) {
        const uint4 *src = reinterpret_cast<const uint4 *>(snap_peer + (k & 1) * slot_bytes);
        const long long stride = (long long)gridDim.x * blockDim.x;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride)
            st_stream16(dst + i, ld_stream16(src + i));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + KFB_MAX_BLOCKS + 1);
        if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
            *ticket = 0u;
            __threadfence_system();
            if (got_out) *got_out = k != 0;
            if (torn_out && k != 0) {
                const uint32_t wb = ld_acquire_sys(seq_peer + 1);
                if ((int32_t)(wb - k) >= 2) atomicAdd(torn_out, 1);
            }
        }
    }
}

// ---- gradient variance -----------------------------------------------------------------
struct VarArgs {
    PeerPads pads;
    uint32_t *local_epoch;
    void *grads[KFB_MAX_RANKS];
    const long long *seg_ends;  // device, cumulative element offsets
    int n_seg;
    float *seg_out;  // device float[n_seg]: this rank's partial sum of v^2
    float *work;     // device: per block {edge value x2, edge segment x2}, then a ticket
    long long npk;
    int rank, np;
};


// No atomics on the sums: the rank's shard is cut into one contiguous range
// per CTA; inside its range a CTA walks the tensors (segments) in order and
// reduces each with the fixed block tree. A tensor that lies inside one CTA's
// range is written directly; a tensor that spans CTAs leaves one partial per
// CTA in `work`, and the last CTA to finish adds those in CTA order. The
// result is therefore bit-identical from run to run.
template <typename G>
__global__ void __launch_bounds__(512)
    k_grad_variance(const __grid_constant__ VarArgs a)
{
    const int rank = a.rank < 0 ? (int)blockIdx.y : a.rank;
    const int np   = a.np;
    uint32_t *epoch =
        a.local_epoch + (a.rank < 0 ? rank * KFB_STATE_EPOCH_WORDS : 0);
    float *seg_out = a.seg_out + (a.rank < 0 ? rank * a.n_seg : 0);
    float *work    = a.work + (a.rank < 0 ? rank * KFB_VAR_WORK_WORDS : 0);
    __shared__ float red[32];
    if (np > 1) peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
    const long long per = (a.npk + np - 1) / np;
    const long long lo  = min((long long)rank * per, a.npk);
    const long long hi  = min(lo + per, a.npk);
    const long long G_  = gridDim.x;
    const long long pb  = (hi - lo + G_ - 1) / G_;  // packets per CTA
    const long long blo = min(lo + (long long)blockIdx.x * pb, hi);
    const long long bhi = min(blo + pb, hi);
    const float inv_np  = 1.f / (float)np;
    float edge_val[2]   = {0.f, 0.f};
    int edge_seg[2]     = {-1, -1};
    if (blo < bhi) {
        // first segment whose end lies beyond the first element of the range
        int sgm = 0;
        {
            const long long e = blo * 8;
            int l = 0, h = a.n_seg - 1;
            while (l < h) {
                const int mid = (l + h) >> 1;
                if (a.seg_ends[mid] > e) h = mid;
                else l = mid + 1;
            }
            sgm = l;
        }
        long long cur = blo;
        while (cur < bhi) {
            const long long seg_lo = sgm == 0 ? 0 : a.seg_ends[sgm - 1] / 8;
            const long long seg_hi = a.seg_ends[sgm] / 8;
            const long long end    = min(bhi, seg_hi);
            float acc              = 0.f;
            for (long long i = cur + threadIdx.x; i < end; i += blockDim.x) {
                float s1[8], s2[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
                for (int r = 0; r < np; ++r) {
                    float t[8];
                    Pk8<G>::load(a.grads[r], i, t);
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        s1[k] += t[k];
                        s2[k] = fmaf(t[k], t[k], s2[k]);
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float mean = s1[k] * inv_np;
                    const float v    = s2[k] * inv_np - mean * mean;
                    acc              = fmaf(v, v, acc);
                }
            }
            const float tot = block_sum(acc, red);
            // does this CTA see all of the tensor's packets that this rank owns?
            const bool whole = max(seg_lo, lo) >= blo && min(seg_hi, hi) <= bhi;
            if (threadIdx.x == 0) {
                if (whole) seg_out[sgm] = tot;
                else {
                    const int slot = (edge_seg[0] < 0) ? 0 : 1;
                    edge_val[slot] = tot;
                    edge_seg[slot] = sgm;
                }
            }
            cur = end;
            if (++sgm >= a.n_seg) break;
        }
    }
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        work[4 * blockIdx.x + 0] = edge_val[0];
        work[4 * blockIdx.x + 1] = edge_val[1];
        reinterpret_cast<int *>(work)[4 * blockIdx.x + 2] = edge_seg[0];
        reinterpret_cast<int *>(work)[4 * blockIdx.x + 3] = edge_seg[1];
        __threadfence();
        unsigned int *ticket = reinterpret_cast<unsigned int *>(work + 4 * KFB_MAX_BLOCKS);
        is_last              = (atomicAdd(ticket, 1u) == gridDim.x - 1);
        if (is_last) {
            *ticket = 0u;
            __threadfence();
            int open_seg   = -1;
            float open_sum = 0.f;
            for (unsigned b = 0; b < gridDim.x; ++b)
                for (int sl = 0; sl < 2; ++sl) {
                    const int sg = *reinterpret_cast<volatile int *>(reinterpret_cast<int *>(work) + 4 * b + 2 + sl);
                    if (sg < 0) continue;
                    const float val = *reinterpret_cast<volatile float *>(work + 4 * b + sl);
                    if (sg != open_seg) {
                        if (open_seg >= 0) seg_out[open_seg] = open_sum;
                        open_seg = sg;
                        open_sum = 0.f;
                    }
                    open_sum += val;
                }
            if (open_seg >= 0) seg_out[open_seg] = open_sum;
        }
    }
    __syncthreads();
    if (np > 1) peer_barrier(a.pads, epoch, rank, np, blockIdx.x);
}

__global__ void k_gns_finalize(const float *ss_local, const float *ss_global,
                               float b_small, float b_big, float alpha,
                               float *state)
{
    // G = (B_big |G_big|^2 - B_small |G_small|^2) / (B_big - B_small)
    // S = (|G_small|^2 - |G_big|^2) / (1/B_small - 1/B_big)
    const float gs = *ss_local, gb = *ss_global;
    const float G  = (b_big * gb - b_small * gs) / (b_big - b_small);
    const float S  = (gs - gb) / (1.f / b_small - 1.f / b_big);
    float g_ema, s_ema;
    if (state[2] != 0.f) {
        g_ema = alpha * state[0] + (1.f - alpha) * G;
        s_ema = alpha * state[1] + (1.f - alpha) * S;
    } else {
        g_ema = G;
        s_ema = S;
    }
    state[0] = g_ema;
    state[1] = s_ema;
    state[2] = 1.f;
    state[3] = s_ema / g_ema;
}

static void fill_pads(PeerPads &p, const kfb_comm *c)
{
    for (int r = 0; r < KFB_MAX_RANKS; ++r)
        p.pad[r] = reinterpret_cast<uint32_t *>(r < c->np ? c->pads[r] : nullptr);
}

static int check_launch(const kfb_comm *c, int blocks, int threads)
{
    if (!c || c->np < 1 || c->np > KFB_MAX_RANKS || c->rank >= c->np) {
        set_last_error("bad communicator");
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

template <typename K, typename A>
static int launch_step(const kfb_comm *c, K kernel, int blocks, int threads,
                       cudaStream_t s, const A &args)
{
    if (c->rank < 0) {
        void *argv[] = {(void *)&args};
        KFB_CUDA_OK(cudaLaunchCooperativeKernel((const void *)kernel,
                                                dim3(blocks, c->np),
                                                dim3(threads), argv, 0, s));
    } else {
        kernel<<<blocks, threads, 0, s>>>(args);
        KFB_CUDA_OK(cudaGetLastError());
    }
    return KFB_OK;
}

}  // namespace kfb

using namespace kfb;

#define KFB_DISPATCH_FLOAT(dt, T, ...)                                         \
    switch (dt) {                                                              \
    case KFB_F32: { using T = float; __VA_ARGS__; } break;                     \
    case KFB_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;            \
    case KFB_F16: { using T = __half; __VA_ARGS__; } break;                    \
    default: set_last_error("dtype %d must be f32/bf16/f16", dt); return KFB_ERR_ARG; \
    }

extern "C" int kfb_fuse(void *const *srcs, const int64_t *counts,
                        const int64_t *dst_offs, int n, int src_dtype,
                        void *dst, int dst_dtype, float *sumsq_out,
                        float *scratch, void *stream)
{
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    // scratch layout: KFB_FUSE_MAX * FUSE_BX partial slots per launch.
    constexpr int BX = 16;
    // Every entry gets one row of BX blocks, so a big tensor is cut into pieces first: 16 blocks
    // cannot stream a 23 M-element embedding gradient at HBM speed. The piece size is a multiple
    // of 16 elements, which keeps the 16-byte alignment of both sides.
    constexpr int64_t PIECE = 1 << 19;
    const size_t esz        = src_dtype == KFB_F32 ? 4 : 2;
    struct Piece {
        void *src;
        int64_t count, off;
    };
    std::vector<Piece> pieces;
    pieces.reserve(n + 64);
    for (int i = 0; i < n; ++i)
        for (int64_t o = 0; o < counts[i]; o += PIECE)
            pieces.push_back({static_cast<char *>(srcs[i]) + o * esz, std::min(PIECE, counts[i] - o), dst_offs[i] + o});
    if (pieces.empty() && sumsq_out) KFB_CUDA_OK(cudaMemsetAsync(sumsq_out, 0, sizeof(float), s));
    n          = static_cast<int>(pieces.size());
    bool first = true;
    for (int base = 0; base < n; base += KFB_FUSE_MAX) {
        FuseArgs a;
        memset(&a, 0, sizeof(a));
        a.n = std::min(KFB_FUSE_MAX, n - base);
        for (int i = 0; i < a.n; ++i) {
            a.src[i]   = pieces[base + i].src;
            a.count[i] = pieces[base + i].count;
            a.off[i]   = pieces[base + i].off;
        }
        float *partials = sumsq_out ? scratch : nullptr;
        if (sumsq_out && !scratch) {
            set_last_error("kfb_fuse: sumsq_out needs scratch of %d floats",
                           KFB_FUSE_MAX * BX);
            return KFB_ERR_ARG;
        }
        dim3 grid(BX, a.n);
        KFB_DISPATCH_FLOAT(src_dtype, S,
            KFB_DISPATCH_FLOAT(dst_dtype, D,
                (k_fuse<S, D, false><<<grid, 256, 0, s>>>(a, dst, partials))));
        KFB_CUDA_OK(cudaGetLastError());
        if (sumsq_out) {
            k_sum_slots<<<1, 256, 0, s>>>(partials, BX * a.n, sumsq_out,
                                          first ? 0 : 1);
            KFB_CUDA_OK(cudaGetLastError());
        }
        first = false;
    }
    return KFB_OK;
}

extern "C" int kfb_sum_slots(const float *slots, int n, float *out, void *stream)
{
    if (n < 0 || !slots || !out) {
        set_last_error("kfb_sum_slots: bad arguments");
        return KFB_ERR_ARG;
    }
    k_sum_slots<<<1, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(slots, n, out, 0);
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

// ---- kfb_fuse for callers that must not allocate ------------------------------------
//
// Some users of this library cannot hand us a scratch buffer because they
// cannot obtain one at the point of the call: a plugin running inside another
// framework's step after that framework has claimed all device memory for its
// own planner, a hook invoked from a latency-critical loop where a
// cudaMalloc (a device-wide synchronising call) is forbidden, or a host
// thread that is not allowed to touch the allocator at all. For them the
// block partials of the norm live in this file-scope device array instead of
// a caller buffer. It is part of the library image: every GPU that loads the
// library holds these 64 KiB until the process exits.
#define KFB_FUSE_NOALLOC_SLOTS 16
static __device__ __align__(256) float g_fuse_noalloc_scratch[KFB_FUSE_NOALLOC_MAX_TENSORS * KFB_FUSE_NOALLOC_SLOTS];

// In host code the array's name is only a handle for the runtime, not a
// device address; the address each device uses is looked up once per device.
static float *g_fuse_noalloc_ptr[64];
static std::mutex g_fuse_noalloc_mu;

static int fuse_noalloc_scratch(float **out)
{
    int dev = 0;
    KFB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_last_error("kfb_fuse_noalloc: device ordinal %d out of range", dev);
        return KFB_ERR_ARG;
    }
    std::lock_guard<std::mutex> lk(g_fuse_noalloc_mu);
    if (!g_fuse_noalloc_ptr[dev]) {
        void *p = nullptr;
        KFB_CUDA_OK(cudaGetSymbolAddress(&p, g_fuse_noalloc_scratch));
        g_fuse_noalloc_ptr[dev] = static_cast<float *>(p);
    }
    *out = g_fuse_noalloc_ptr[dev];
    return KFB_OK;
}

extern "C" int kfb_fuse_noalloc(void *const *srcs, const int64_t *counts,
                                const int64_t *dst_offs, int n, int src_dtype,
                                void *dst, int dst_dtype, float *sumsq_out,
                                void *stream)
{
    if (n < 0 || n > KFB_FUSE_NOALLOC_MAX_TENSORS) {
        set_last_error("kfb_fuse_noalloc: %d tensors exceed the built-in scratch (max %d per call); "
                       "split the call or use kfb_fuse with your own scratch",
                       n, KFB_FUSE_NOALLOC_MAX_TENSORS);
        return KFB_ERR_ARG;
    }
    float *scratch = nullptr;
    if (sumsq_out) {
        // First call on a device resolves the array's address (may load the
        // module): make that call before any stream capture begins.
        int rc = fuse_noalloc_scratch(&scratch);
        if (rc) return rc;
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    constexpr int BX = KFB_FUSE_NOALLOC_SLOTS;
    for (int base = 0; base < n; base += KFB_FUSE_MAX) {
        FuseArgs a;
        memset(&a, 0, sizeof(a));
        a.n = std::min(KFB_FUSE_MAX, n - base);
        for (int i = 0; i < a.n; ++i) {
            a.src[i]   = srcs[base + i];
            a.count[i] = counts[base + i];
            a.off[i]   = dst_offs[base + i];
        }
        // every launched CTA stores its slot (zero included), so nothing in
        // the shared array needs clearing between calls
        float *partials = scratch ? scratch + (size_t)base * BX : nullptr;
        dim3 grid(BX, a.n);
        KFB_DISPATCH_FLOAT(src_dtype, S,
            KFB_DISPATCH_FLOAT(dst_dtype, D,
                (k_fuse<S, D, false><<<grid, 256, 0, s>>>(a, dst, partials))));
        KFB_CUDA_OK(cudaGetLastError());
    }
    if (sumsq_out) {
        if (n == 0) {
            KFB_CUDA_OK(cudaMemsetAsync(sumsq_out, 0, sizeof(float), s));
        } else {
            k_sum_slots<<<1, 256, 0, s>>>(scratch, BX * n, sumsq_out, 0);
            KFB_CUDA_OK(cudaGetLastError());
        }
    }
    return KFB_OK;
}

extern "C" int kfb_defuse(const void *src, int src_dtype, void *const *dsts,
                          const int64_t *counts, const int64_t *src_offs,
                          int n, int dst_dtype, void *stream)
{
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    for (int base = 0; base < n; base += KFB_FUSE_MAX) {
        FuseArgs a;
        memset(&a, 0, sizeof(a));
        a.n = std::min(KFB_FUSE_MAX, n - base);
        for (int i = 0; i < a.n; ++i) {
            a.src[i]   = dsts[base + i];
            a.count[i] = counts[base + i];
            a.off[i]   = src_offs[base + i];
        }
        dim3 grid(16, a.n);
        KFB_DISPATCH_FLOAT(src_dtype, S,
            KFB_DISPATCH_FLOAT(dst_dtype, D,
                (k_fuse<S, D, true><<<grid, 256, 0, s>>>(a, const_cast<void *>(src), nullptr))));
        KFB_CUDA_OK(cudaGetLastError());
    }
    return KFB_OK;
}

static int fill_step(StepArgs &a, const kfb_comm *c, int64_t count)
{
    if (count % 8) {
        set_last_error("flat buffers must hold a multiple of 8 elements (got %lld)",
                       (long long)count);
        return KFB_ERR_ARG;
    }
    memset(&a, 0, sizeof(a));
    fill_pads(a.pads, c);
    a.local_epoch = reinterpret_cast<uint32_t *>(c->local_epoch);
    a.scratch     = reinterpret_cast<float *>(c->scratch);
    a.rank        = c->rank;
    a.np          = c->np;
    a.npk         = count / 8;
    return KFB_OK;
}

extern "C" int kfb_peer_signal(const kfb_comm *comm, int rows, void *stream)
{
    if (!comm || comm->rank < 0 || comm->np < 1 || comm->np > KFB_MAX_RANKS || rows < 1 || rows > KFB_MAX_BLOCKS) {
        set_last_error("kfb_peer_signal: needs a real rank and 1..%d rows", KFB_MAX_BLOCKS);
        return KFB_ERR_ARG;
    }
    if (comm->np == 1) return KFB_OK;
    PeerPads pads;
    fill_pads(pads, comm);
    k_peer_signal<<<1, 512, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        pads, reinterpret_cast<uint32_t *>(comm->local_epoch), comm->rank, comm->np, rows);
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

static thread_local int g_step_flags = 0;
extern "C" void kfb_set_step_flags(int flags) { g_step_flags = flags; }

extern "C" int kfb_ssgd_fused_step(const kfb_comm *comm, void *const *grads,
                                   int grad_dtype, void *const *params,
                                   int param_dtype, float *master, float *mom,
                                   int64_t count, float grad_scale,
                                   const kfb_sgd_hyper *h, float *sumsq_out,
                                   int blocks, int threads, void *stream)
{
    int rc = check_launch(comm, blocks, threads);
    if (rc) return rc;
    StepArgs a;
    rc = fill_step(a, comm, count);
    if (rc) return rc;
    a.flags      = comm->rank < 0 ? 0 : g_step_flags;
    g_step_flags = 0;
    for (int r = 0; r < comm->np; ++r) {
        a.grads[r]  = grads[r];
        a.params[r] = params[r];
    }
    a.master     = master;
    a.mom        = mom;
    a.grad_scale = grad_scale;
    a.h          = to_h(h);
    a.sumsq_out  = sumsq_out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const int np   = comm->np;
#define KFB_SSGD(NPV)                                                          \
    do {                                                                       \
        if (sumsq_out)                                                         \
            return launch_step(comm, k_ssgd_fused<G, P, NPV, true>, blocks, threads, s, a);  \
        return launch_step(comm, k_ssgd_fused<G, P, NPV, false>, blocks, threads, s, a);     \
    } while (0)
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P, {
            if (np == 1) KFB_SSGD(1);
            if (np == 2) KFB_SSGD(2);
            if (np == 4) KFB_SSGD(4);
            if (np == 8) KFB_SSGD(8);
            KFB_SSGD(0);
        }));
#undef KFB_SSGD
    return KFB_OK;
}

extern "C" int kfb_ssgd_fused_step_nvls(const kfb_comm *comm, void *mc_grads, int grad_dtype, void *mc_params,
                                        int param_dtype, float *master, float *mom, int64_t count,
                                        float grad_scale, const kfb_sgd_hyper *h, float *sumsq_out, int blocks,
                                        int threads, void *stream)
{
    int rc = check_launch(comm, blocks, threads);
    if (rc) return rc;
    if (comm->rank < 0 || comm->np < 2 || !mc_grads || !mc_params) {
        set_last_error("NVLS fused step needs real ranks (np >= 2) and multicast addresses");
        return KFB_ERR_ARG;
    }
    StepArgs a;
    rc = fill_step(a, comm, count);
    if (rc) return rc;
    a.flags      = g_step_flags;
    g_step_flags = 0;
    a.stage[0]   = mc_grads;
    a.stage[1]   = mc_params;
    a.master     = master;
    a.mom        = mom;
    a.grad_scale = grad_scale;
    a.h          = to_h(h);
    a.sumsq_out  = sumsq_out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P, {
            if (sumsq_out) k_ssgd_fused_nvls<G, P, true><<<blocks, threads, 0, s>>>(a);
            else k_ssgd_fused_nvls<G, P, false><<<blocks, threads, 0, s>>>(a);
        }));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_sgd_flat(const void *grad, int grad_dtype, void *param,
                            int param_dtype, float *master, float *mom,
                            int64_t count, float grad_scale,
                            const kfb_sgd_hyper *h, int blocks, int threads,
                            void *stream)
{
    if (count % 8) {
        set_last_error("flat buffers must hold a multiple of 8 elements");
        return KFB_ERR_ARG;
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    const SgdH hh  = to_h(h);
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P,
            (k_sgd_flat<G, P><<<blocks, threads, 0, s>>>(grad, param, master, mom, count / 8, grad_scale, hh))));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_sma_fused_step_ex(const kfb_comm *comm, void *const *params,
                                     int param_dtype, void *const *avg_stage,
                                     void *mc_params, void *mc_stage,
                                     const void *grad, int grad_dtype,
                                     float *master, float *mom, int64_t count,
                                     float alpha, float grad_scale,
                                     const kfb_sgd_hyper *h, int phase,
                                     int blocks, int threads, void *stream)
{
    int rc = check_launch(comm, blocks, threads);
    if (rc) return rc;
    if (phase < 0 || phase > 2) {
        set_last_error("kfb_sma_fused_step_ex: phase must be 0 (all), 1 (average) or 2 (update)");
        return KFB_ERR_ARG;
    }
    const bool nvls = mc_params && mc_stage && comm->np > 1 && comm->rank >= 0 && phase != 2;
    StepArgs a;
    rc = fill_step(a, comm, count);
    if (rc) return rc;
    for (int r = 0; r < comm->np; ++r) {
        a.params[r] = params[r];
        a.stage[r]  = avg_stage ? avg_stage[r] : nullptr;
        // emulation: `grad` is a host array of np device pointers
        if (comm->rank < 0 && grad)
            a.grads[r] = reinterpret_cast<void *const *>(grad)[r];
    }
    if (nvls) {
        if (comm->np > KFB_MAX_RANKS - 2) {
            set_last_error("NVLS SMA supports at most %d ranks", KFB_MAX_RANKS - 2);
            return KFB_ERR_ARG;
        }
        a.stage[KFB_MAX_RANKS - 2] = mc_params;
        a.stage[KFB_MAX_RANKS - 1] = mc_stage;
    }
    a.grad_local = grad;
    a.master     = master;
    a.mom        = mom;
    a.alpha      = alpha;
    a.grad_scale = grad_scale;
    a.h          = to_h(h);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define KFB_SMA(NV, PH) return launch_step(comm, k_sma_fused<G, P, NV, PH>, blocks, threads, s, a)
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P, {
            if (phase == 2) KFB_SMA(false, 2);
            if (nvls) {
                if (phase == 0) KFB_SMA(true, 0);
                KFB_SMA(true, 1);
            }
            if (phase == 0) KFB_SMA(false, 0);
            KFB_SMA(false, 1);
        }));
#undef KFB_SMA
    return KFB_OK;
}

extern "C" int kfb_sma_fused_step(const kfb_comm *comm, void *const *params,
                                  int param_dtype, void *const *avg_stage,
                                  const void *grad, int grad_dtype,
                                  float *master, float *mom, int64_t count,
                                  float alpha, float grad_scale,
                                  const kfb_sgd_hyper *h, int blocks,
                                  int threads, void *stream)
{
    return kfb_sma_fused_step_ex(comm, params, param_dtype, avg_stage, nullptr, nullptr, grad, grad_dtype, master,
                                 mom, count, alpha, grad_scale, h, 0, blocks, threads, stream);
}

// Pull + 1/2 blend only (no optimizer update, no publication): the device
// path of PairAveraging around an arbitrary torch optimizer.
extern "C" int kfb_pair_blend(const kfb_comm *comm, int target, const void *snap_peer, int64_t slot_bytes,
                              void *param, int param_dtype, int64_t count, int32_t *torn_out, int blocks,
                              int threads, void *stream)
{
    if (!comm || comm->rank < 0 || comm->rank >= comm->np) {
        set_last_error("pair averaging needs a real rank");
        return KFB_ERR_ARG;
    }
    if (count % 8 || blocks < 1 || blocks > KFB_MAX_BLOCKS) {
        set_last_error("bad count/blocks");
        return KFB_ERR_ARG;
    }
    PairArgs a;
    memset(&a, 0, sizeof(a));
    a.has_peer = (target >= 0 && target < comm->np && target != comm->rank);
    if (!a.has_peer) return KFB_OK;
    a.seq_self   = reinterpret_cast<uint32_t *>(comm->pads[comm->rank]) + KFB_PAD_MISC_WORD;
    a.seq_peer   = reinterpret_cast<uint32_t *>(comm->pads[target]) + KFB_PAD_MISC_WORD;
    a.snap_peer  = snap_peer;
    a.slot_bytes = slot_bytes;
    a.param      = param;
    a.scratch    = reinterpret_cast<float *>(comm->scratch);
    a.npk        = count / 8;
    a.torn_out   = torn_out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    KFB_DISPATCH_FLOAT(param_dtype, P, (k_pair_avg<P, P, false><<<blocks, threads, 0, s>>>(a)));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_pair_avg_step(const kfb_comm *comm, int target,
                                 const void *snap_peer, void *snap_self,
                                 int64_t slot_bytes, void *param,
                                 int param_dtype, const void *grad,
                                 int grad_dtype, float *master, float *mom,
                                 int64_t count, float grad_scale,
                                 const kfb_sgd_hyper *h, int32_t *torn_out,
                                 int blocks, int threads, void *stream)
{
    if (!comm || comm->rank < 0 || comm->rank >= comm->np) {
        set_last_error("pair averaging needs a real rank");
        return KFB_ERR_ARG;
    }
    if (count % 8 || blocks < 1 || blocks > KFB_MAX_BLOCKS) {
        set_last_error("bad count/blocks");
        return KFB_ERR_ARG;
    }
    PairArgs a;
    memset(&a, 0, sizeof(a));
    a.has_peer = (target >= 0 && target < comm->np && target != comm->rank);
    a.seq_self = reinterpret_cast<uint32_t *>(comm->pads[comm->rank]) + KFB_PAD_MISC_WORD;
    a.seq_peer = a.has_peer ? reinterpret_cast<uint32_t *>(comm->pads[target]) + KFB_PAD_MISC_WORD
                            : a.seq_self;
    a.snap_peer  = snap_peer;
    a.snap_self  = snap_self;
    a.slot_bytes = slot_bytes;
    a.param      = param;
    a.grad       = grad;
    a.master     = master;
    a.mom        = mom;
    a.scratch    = reinterpret_cast<float *>(comm->scratch);
    a.npk        = count / 8;
    a.grad_scale = grad_scale;
    a.h          = to_h(h);
    a.torn_out   = torn_out;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P,
            (k_pair_avg<G, P, true><<<blocks, threads, 0, s>>>(a))));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

// Same step with the peer chosen on the device: snaps[r] is rank r's snapshot buffer as mapped here.
extern "C" int kfb_pair_avg_step_dev(const kfb_comm *comm, void *const *snaps, int64_t slot_bytes,
                                     void *param, int param_dtype, const void *grad, int grad_dtype,
                                     float *master, float *mom, int64_t count, float grad_scale,
                                     const kfb_sgd_hyper *h, int32_t *torn_out, const long long *step,
                                     unsigned long long seed, int blocks, int threads, void *stream)
{
    if (!comm || comm->rank < 0 || comm->rank >= comm->np || comm->np > KFB_MAX_RANKS) {
        set_last_error("pair averaging needs a real rank");
        return KFB_ERR_ARG;
    }
    if (count % 8 || blocks < 1 || blocks > KFB_MAX_BLOCKS || !snaps || !step) {
        set_last_error("bad count/blocks/snaps/step");
        return KFB_ERR_ARG;
    }
    PairArgs a;
    memset(&a, 0, sizeof(a));
    a.has_peer = comm->np > 1;
    a.seq_self = reinterpret_cast<uint32_t *>(comm->pads[comm->rank]) + KFB_PAD_MISC_WORD;
    a.seq_peer = a.seq_self;
    a.snap_peer  = snaps[comm->rank];
    a.snap_self  = snaps[comm->rank];
    a.slot_bytes = slot_bytes;
    a.param      = param;
    a.grad       = grad;
    a.master     = master;
    a.mom        = mom;
    a.scratch    = reinterpret_cast<float *>(comm->scratch);
    a.npk        = count / 8;
    a.grad_scale = grad_scale;
    a.h          = to_h(h);
    a.torn_out   = torn_out;
    a.sel_np     = comm->np;
    a.sel_rank   = comm->rank;
    a.sel_step   = step;
    a.sel_seed   = seed;
    for (int r = 0; r < comm->np; ++r) {
        a.seq_all[r]  = reinterpret_cast<uint32_t *>(comm->pads[r]) + KFB_PAD_MISC_WORD;
        a.snap_all[r] = snaps[r];
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    KFB_DISPATCH_FLOAT(grad_dtype, G,
        KFB_DISPATCH_FLOAT(param_dtype, P,
            (k_pair_avg<G, P, true><<<blocks, threads, 0, s>>>(a))));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_pair_pull(const kfb_comm *comm, int target, const void *snap_peer, int64_t slot_bytes, void *dst,
                             int64_t bytes, int32_t *torn_out, int32_t *got_out, int blocks, int threads,
                             void *stream)
{
    if (!comm || comm->rank < 0 || target < 0 || target >= comm->np || target == comm->rank) {
        set_last_error("kfb_pair_pull: needs a real rank and another rank as target");
        return KFB_ERR_ARG;
    }
    if (bytes % 16 || bytes > slot_bytes || blocks < 1 || blocks > KFB_MAX_BLOCKS) {
        set_last_error("kfb_pair_pull: bytes must be a multiple of 16 and fit the slot");
        return KFB_ERR_ARG;
    }
    uint32_t *seq = reinterpret_cast<uint32_t *>(comm->pads[target]) + KFB_PAD_MISC_WORD;
    k_pair_pull<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        seq, reinterpret_cast<const char *>(snap_peer), slot_bytes, reinterpret_cast<uint4 *>(dst), bytes / 16,
        reinterpret_cast<float *>(comm->scratch), torn_out, got_out);
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_pair_publish(const kfb_comm *comm, const void *param,
                                void *snap_self, int64_t slot_bytes,
                                int64_t bytes, int blocks, int threads,
                                void *stream)
{
    if (!comm || comm->rank < 0) {
        set_last_error("pair averaging needs a real rank");
        return KFB_ERR_ARG;
    }
    uint32_t *seq = reinterpret_cast<uint32_t *>(comm->pads[comm->rank]) + KFB_PAD_MISC_WORD;
    k_pair_publish<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        seq, reinterpret_cast<const uint4 *>(param),
        reinterpret_cast<char *>(snap_self), slot_bytes, (bytes + 15) / 16,
        reinterpret_cast<float *>(comm->scratch));
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_grad_variance(const kfb_comm *comm, void *const *grads,
                                 int dtype, int64_t count,
                                 const int64_t *seg_ends_dev, int n_seg,
                                 float *seg_out, float *work, int blocks,
                                 int threads, void *stream)
{
    int rc = check_launch(comm, blocks, threads);
    if (rc) return rc;
    if (count % 8 || n_seg < 1 || !work) {
        set_last_error("bad count / n_seg / work");
        return KFB_ERR_ARG;
    }
    VarArgs a;
    memset(&a, 0, sizeof(a));
    fill_pads(a.pads, comm);
    a.local_epoch = reinterpret_cast<uint32_t *>(comm->local_epoch);
    for (int r = 0; r < comm->np; ++r) a.grads[r] = grads[r];
    a.seg_ends = reinterpret_cast<const long long *>(seg_ends_dev);
    a.n_seg    = n_seg;
    a.seg_out  = seg_out;
    a.work     = work;
    a.npk      = count / 8;
    a.rank     = comm->rank;
    a.np       = comm->np;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    KFB_DISPATCH_FLOAT(dtype, G,
        return launch_step(comm, k_grad_variance<G>, blocks, threads, s, a));
    return KFB_OK;
}

extern "C" int kfb_gns_finalize(const float *sumsq_local,
                                const float *sumsq_global, float batch_small,
                                float batch_big, float alpha, float *state,
                                void *stream)
{
    k_gns_finalize<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        sumsq_local, sumsq_global, batch_small, batch_big, alpha, state);
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}
