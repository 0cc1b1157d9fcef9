// Shared device helpers for the kungfu_b200 sm_100a kernels.
//
// Everything here is written for one target only (sm_100a, B200): 16-byte
// vector accesses, system-scope release/acquire flags for NVLink peers, fp32
// accumulation for 16-bit types.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>

#include "../../include/kungfu_b200.h"

namespace kfb {

// ---- error plumbing (host) -------------------------------------------------
void set_last_error(const char *fmt, ...);
#define KFB_CUDA_OK(expr)                                                      \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            kfb::set_last_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,   \
                                cudaGetErrorString(_e));                       \
            return KFB_ERR_CUDA;                                               \
        }                                                                      \
    } while (0)

// ---- memory-model helpers --------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v)
                 : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];"
                 : "=r"(v)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];"
                 : "=r"(v)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_f32(float *p, float v)
{
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v)
                 : "memory");
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float *p)
{
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];"
                 : "=f"(v)
                 : "l"(p)
                 : "memory");
    return v;
}

// 16-byte streaming load/store. Peer (NVLink) addresses bypass the local L2
// anyway; for local data we do not want these one-touch streams to evict the
// model's working set from L1.
__device__ __forceinline__ uint4 ld_stream16(const void *p)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ void st_stream16(void *p, const uint4 &v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(
                     p),
                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// ---- 16-byte packet <-> fp32 lanes ----------------------------------------
template <typename T> struct Pack;  // 16 bytes of T <-> accumulator lanes

template <> struct Pack<float> {
    static constexpr int N = 4;
    using acc_t            = float;
    __device__ static void unpack(const uint4 &v, float *a)
    {
        a[0] = __uint_as_float(v.x);
        a[1] = __uint_as_float(v.y);
        a[2] = __uint_as_float(v.z);
        a[3] = __uint_as_float(v.w);
    }
    __device__ static uint4 pack(const float *a)
    {
        return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]),
                          __float_as_uint(a[2]), __float_as_uint(a[3]));
    }
};

template <> struct Pack<int32_t> {
    static constexpr int N = 4;
    using acc_t            = int32_t;
    __device__ static void unpack(const uint4 &v, int32_t *a)
    {
        a[0] = (int32_t)v.x;
        a[1] = (int32_t)v.y;
        a[2] = (int32_t)v.z;
        a[3] = (int32_t)v.w;
    }
    __device__ static uint4 pack(const int32_t *a)
    {
        return make_uint4((uint32_t)a[0], (uint32_t)a[1], (uint32_t)a[2],
                          (uint32_t)a[3]);
    }
};

template <> struct Pack<__nv_bfloat16> {
    static constexpr int N = 8;
    using acc_t            = float;
    __device__ static void unpack(const uint4 &v, float *a)
    {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // bf16 -> fp32 is a 16-bit shift: no conversion instruction.
            a[2 * i]     = __uint_as_float(w[i] << 16);
            a[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ static uint4 pack(const float *a)
    {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
            w[i]             = *reinterpret_cast<uint32_t *>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

template <> struct Pack<__half> {
    static constexpr int N = 8;
    using acc_t            = float;
    __device__ static void unpack(const uint4 &v, float *a)
    {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h    = *reinterpret_cast<const __half2 *>(&w[i]);
            float2 f     = __half22float2(h);
            a[2 * i]     = f.x;
            a[2 * i + 1] = f.y;
        }
    }
    __device__ static uint4 pack(const float *a)
    {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = __floats2half2_rn(a[2 * i], a[2 * i + 1]);
            w[i]      = *reinterpret_cast<uint32_t *>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

template <> struct Pack<double> {
    static constexpr int N = 2;
    using acc_t            = double;
    __device__ static void unpack(const uint4 &v, double *a)
    {
        a[0] = __hiloint2double(v.y, v.x);
        a[1] = __hiloint2double(v.w, v.z);
    }
    __device__ static uint4 pack(const double *a)
    {
        return make_uint4(__double2loint(a[0]), __double2hiint(a[0]),
                          __double2loint(a[1]), __double2hiint(a[1]));
    }
};
template <> struct Pack<int64_t> {
    static constexpr int N = 2;
    using acc_t            = int64_t;
    __device__ static void unpack(const uint4 &v, int64_t *a)
    {
        a[0] = (int64_t)(((uint64_t)v.y << 32) | v.x);
        a[1] = (int64_t)(((uint64_t)v.w << 32) | v.z);
    }
    __device__ static uint4 pack(const int64_t *a)
    {
        return make_uint4((uint32_t)a[0], (uint32_t)((uint64_t)a[0] >> 32),
                          (uint32_t)a[1], (uint32_t)((uint64_t)a[1] >> 32));
    }
};
template <> struct Pack<uint8_t> {
    static constexpr int N = 16;
    using acc_t            = int32_t;  // widened lanes; packed back modulo 256
    __device__ static void unpack(const uint4 &v, int32_t *a)
    {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = (w[i >> 2] >> (8 * (i & 3))) & 0xff;
    }
    __device__ static uint4 pack(const int32_t *a)
    {
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 16; ++i)
            w[i >> 2] |= ((uint32_t)a[i] & 0xffu) << (8 * (i & 3));
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// ---- reduce ops ------------------------------------------------------------
template <int OP, typename A> __device__ __forceinline__ A apply_op(A x, A y)
{
    if (OP == KFB_OP_SUM) return x + y;
    if (OP == KFB_OP_MIN) return x < y ? x : y;
    if (OP == KFB_OP_MAX) return x > y ? x : y;
    return x * y;  // KFB_OP_PROD
}

// ---- block reduction of one float (deterministic tree) ---------------------
__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Returns the block total in every thread of warp 0 (valid in thread 0).
__device__ __forceinline__ float block_sum(float v, float *smem32)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) smem32[wid] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float t      = (threadIdx.x < nw) ? smem32[threadIdx.x] : 0.f;
    if (wid == 0) t = warp_sum(t);
    __syncthreads();
    return t;
}

// ---- cross-GPU block barrier -------------------------------------------------
//
// Block b of every rank meets block b of every other rank. Each rank owns a
// "pad" in symmetric (peer-mapped) memory: flags[block][src_rank]. A rank
// announces arrival by storing its new epoch into its own column of every
// peer's pad (release, system scope), then spins on its own pad until every
// column reached that epoch (acquire, system scope). The epoch lives in
// *local* memory (one counter per block) and only ever grows, so the flags
// never need clearing and the kernel stays CUDA-graph replayable. After a
// resize the pads and epochs are re-created from zero (new membership = new
// epoch space), which is what the reference's connection token did
// (/root/reference/srcs/go/rchannel/connection/connection.go:77-87).
#ifndef KFB_BARRIER_TIMEOUT_NS
#define KFB_BARRIER_TIMEOUT_NS 120000000000ull  // 120 s: far above any legitimate skew between ranks
#endif

struct PeerPads {
    uint32_t *pad[KFB_MAX_RANKS];  // pad[r] = rank r's pad (peer mapped)
};

__device__ __forceinline__ uint32_t *pad_flag(uint32_t *pad, int block, int src)
{
    return pad + (size_t)block * KFB_MAX_RANKS + src;
}
// Floats exchanged between ranks (norm partials): [block][src_rank]
__device__ __forceinline__ float *pad_scalar(uint32_t *pad, int block, int src)
{
    return reinterpret_cast<float *>(pad + KFB_PAD_FLAG_WORDS) +
           (size_t)block * KFB_MAX_RANKS + src;
}

__device__ __forceinline__ void
peer_barrier(const PeerPads &pads, uint32_t *local_epoch, int rank, int np,
             int block)
{
    __syncthreads();
    if (threadIdx.x < np) {
        const uint32_t e = local_epoch[block] + 1;
        __threadfence_system();
        st_release_sys(pad_flag(pads.pad[threadIdx.x], block, rank), e);
        const uint32_t *mine = pad_flag(pads.pad[rank], block, threadIdx.x);
        // A peer that died (or a collective issued by only some ranks) must not
        // wedge the GPU forever: after KFB_BARRIER_TIMEOUT_NS the kernel traps,
        // which surfaces as a CUDA error in the host process instead of a hang.
        unsigned long long t0 = 0;
        unsigned spins        = 0;
        while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
            __nanosleep(20);
            if ((++spins & 0x3fffu) == 0) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > KFB_BARRIER_TIMEOUT_NS) {
                    printf("kungfu_b200: rank %d gave up waiting for rank %d at a peer barrier (block %d, epoch %u)\n",
                           rank, (int)threadIdx.x, block, e);
                    __trap();
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) local_epoch[block] += 1;
    // No trailing barrier needed: the next reader of local_epoch[block] sits
    // behind the leading __syncthreads() of the next peer_barrier call.
}

// Split-phase form of the same barrier, for kernels that overlap other work:
// waiting inside a resident CTA takes an SM slot away from whatever else is
// running, so the arrival is sent early by a tiny launch of its own
// (peer_arrive_rows: "my inputs of the next collective are final") and the
// collective kernel, launched later, only checks that every peer's arrival is
// there (peer_wait_row), which by then it normally is.
__device__ __forceinline__ void
peer_arrive_rows(const PeerPads &pads, uint32_t *local_epoch, int rank, int np, int nrows)
{
    // one CTA; thread t handles (row, peer) = (t / np, t % np)
    for (int t = threadIdx.x; t < nrows * np; t += blockDim.x) {
        const int row = t / np, peer = t % np;
        const uint32_t e = local_epoch[row] + 1;
        __threadfence_system();
        st_release_sys(pad_flag(pads.pad[peer], row, rank), e);
    }
    __syncthreads();
    for (int row = threadIdx.x; row < nrows; row += blockDim.x) local_epoch[row] += 1;
}

__device__ __forceinline__ void
peer_wait_row(const PeerPads &pads, const uint32_t *local_epoch, int rank, int np, int block)
{
    if (threadIdx.x < np) {
        const uint32_t e     = local_epoch[block];  // already advanced by this rank's arrival
        const uint32_t *mine = pad_flag(pads.pad[rank], block, threadIdx.x);
        unsigned long long t0 = 0;
        unsigned spins        = 0;
        while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
            __nanosleep(20);
            if ((++spins & 0x3fffu) == 0) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > KFB_BARRIER_TIMEOUT_NS) {
                    printf("kungfu_b200: rank %d gave up waiting for rank %d's arrival (block %d, epoch %u)\n",
                           rank, (int)threadIdx.x, block, e);
                    __trap();
                }
            }
        }
    }
    __syncthreads();
}

}  // namespace kfb
