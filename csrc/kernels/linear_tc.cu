// Y = act(X . W^T + bias) on the 5th-generation tensor cores (sm_100a).
//
// X [M,K] and W [N,K] are bf16 with K contiguous (torch.nn.Linear's layout), so
// both operands are "K-major" and no transpose is ever materialised.
//
//   warp 0      TMA producer: cp.async.bulk.tensor (128B swizzle) of X and W
//               tiles, 64 K-elements (one swizzle row) per stage
//   warp 1      MMA issuer: one thread issues tcgen05.mma (bf16 -> fp32) with
//               the accumulator in tensor memory; tcgen05.commit hands smem
//               stages back to the producer and accumulators to the epilogue
//               through mbarriers
//   warp 2      allocates / frees the 512 TMEM columns (two 128x256 accumulators)
//   warps 4-19  epilogue: tcgen05.ld the accumulator, + bias, activation, bf16
//               pack, 16-byte stores; runs while the tensor cores are already
//               on the next tile (double-buffered accumulators)
//
// Two tilings of the same kernel:
//   PAIR = false  one CTA per 128x256 tile (cta_group::1), 4 stages of 48 KiB
//   PAIR = true   a cluster of two CTAs (one TPC) per 256x256 tile: each CTA
//                 stages its own 128 rows of X and HALF of the W tile, the leader
//                 issues tcgen05.mma.cta_group::2 (M = 256) which reads both
//                 CTAs' shared memory and writes both CTAs' tensor memory. A
//                 stage is 32 KiB per CTA, so 6 stages fit and L2->SM traffic
//                 per flop drops by a third.
//
// The kernel is persistent (one CTA per SM, static tile schedule). The fused
// epilogue removes the separate bias / GELU passes over the [M,N] activation;
// with `z` given the pre-activation is stored too (what the backward needs).
//
// The reference has no counterpart: its models come from TensorFlow/Keras and
// its own native code is communication only. This kernel serves the BERT
// benchmark model of this repo (kungfu_b200/models/bert.py).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "kungfu_b200.h"

namespace kfb
{
void set_last_error(const char *fmt, ...);  // collectives.cu

constexpr int LT_BM     = 128;  // rows per CTA
constexpr int LT_BN     = 256;
constexpr int LT_BK     = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int LT_UMMA_K = 16;
constexpr int LT_A_BYTES        = LT_BM * LT_BK * 2;
constexpr int LT_EPI_WARPS      = 16;  // 4 TMEM lane quarters x 4 column parts
constexpr int LT_FIRST_EPI_WARP = 4;
constexpr int LT_EPI_COLS       = LT_BN / (LT_EPI_WARPS / 4);  // accumulator columns per epilogue warp
constexpr int LT_THREADS        = (LT_FIRST_EPI_WARP + LT_EPI_WARPS) * 32;
constexpr int LT_TMEM_COLS      = 2 * LT_BN;

template <bool PAIR> struct LtCfg {
    static constexpr int STAGES      = PAIR ? 6 : 4;
    static constexpr int B_ROWS      = PAIR ? LT_BN / 2 : LT_BN;  // W rows this CTA stages
    static constexpr int B_BYTES     = B_ROWS * LT_BK * 2;
    static constexpr int STAGE_BYTES = LT_A_BYTES + B_BYTES;
    static constexpr int NUM_BARS    = 2 * STAGES + 4;
    static constexpr int SMEM_BYTES  = 1024 + STAGES * STAGE_BYTES + NUM_BARS * 8 + 16;
    // bytes that land per stage on the barrier the MMA thread waits on
    static constexpr uint32_t TX_BYTES = (PAIR ? 2u : 1u) * STAGE_BYTES;
    // instruction descriptor: D fp32 (bit 4), A/B bf16 (bits 7, 10), both K-major, N>>3 at 17, M>>4 at 24
    static constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(LT_BN >> 3) << 17) |
                                      (static_cast<uint32_t>((PAIR ? 2 * LT_BM : LT_BM) >> 4) << 24);
};

// ---- PTX wrappers ----------------------------------------------------------------

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

// shared::cluster address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}

__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// arrive on a barrier that may live in the other CTA of the cluster (shared::cluster address)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar)
{
    // default (.release.cta) semantics: the consumer only needs this warp's TMEM reads to be done, which
    // tcgen05.wait::ld + fence::before_thread_sync order; a cluster-scope release would drain the global stores
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}

__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}

// Bounded wait: a pipeline bug must end in a trap, not in a wedged GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) {
            printf("[kungfu_b200] linear_tc: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x,
                   threadIdx.x);
            __trap();
        }
    }
}

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

// CTA-pair form: the data lands in this CTA's shared memory, the byte count is
// signalled on `cluster_bar`, which may be the leader CTA's barrier.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, uint32_t cluster_bar, int c0,
                                                 int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
        "%4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(cluster_bar), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// tcgen05.commit: the mbarrier gets one arrival once every MMA issued so far by
// this thread has finished (implies fence::before_thread_sync). The pair form
// arrives on the barrier at the same offset in both CTAs of the pair.
template <bool PAIR> __device__ __forceinline__ void tc_commit(uint32_t bar)
{
    if (PAIR) {
        const uint16_t mask = 3;
        asm volatile(
            "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                bar),
            "h"(mask)
            : "memory");
    } else {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                     : "memory");
    }
}

// D[tmem] (+)= A[smem] . B[smem]^T, (128 | 256) x 256 x 16, bf16 inputs, fp32 accumulate.
template <bool PAIR>
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate)
{
    if (PAIR) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
            "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
            "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp gets row (lane base + t).
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for a K-major bf16 tile whose rows are 128
// bytes (64 elements) under the 128-byte swizzle TMA wrote: 8-row groups are
// 1024 bytes apart (stride byte offset), the leading byte offset is unused for
// swizzled K-major tiles (set to 1), descriptor version 1, layout type 2.
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t smem_addr)
{
    uint64_t d = 0;
    d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>(1) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

__device__ __forceinline__ float gelu_erf(float z)
{
    // 0.5 z (1 + erf(z / sqrt 2)); erf by Abramowitz-Stegun 7.1.25 (|error| < 2.5e-5,
    // two orders below the bf16 output's resolution): 2 MUFU + ~12 FP32 ops per
    // element, no slow-path branches, so the epilogue stays under the MMA time.
    const float x = fabsf(z) * 0.70710678f;
    float t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.47047f, x, 1.0f)));
    float p = fmaf(0.7478556f, t, -0.0958798f);
    p       = fmaf(p, t, 0.3480242f) * t;
    float ex;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(x * x * -1.44269504f));
    const float e  = fmaf(-p, ex, 1.0f);
    const float hz = 0.5f * z;
    return fmaf(hz, copysignf(e, z), hz);
}

template <int ACT> __device__ __forceinline__ float apply_act(float z)
{
    if (ACT == KFB_ACT_GELU) return gelu_erf(z);
    if (ACT == KFB_ACT_RELU) return fmaxf(z, 0.0f);
    return z;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t *>(&h);
}

// Tile order: groups of LT_GROUP_M row-blocks sweep all column-blocks before the
// next group starts, so the ~148 tiles in flight cover a near-square patch of
// the output and re-read a small set of X / W panels out of L2 (instead of one
// wave streaming all of X from HBM).
constexpr int LT_GROUP_M = 8;

__device__ __forceinline__ void tile_to_blocks(int tile, int num_m, int num_n, int &m_blk, int &n_blk)
{
    const int per_group = LT_GROUP_M * num_n;
    const int group     = tile / per_group;
    const int first_m   = group * LT_GROUP_M;
    const int rows      = min(num_m - first_m, LT_GROUP_M);
    const int in_group  = tile - group * per_group;
    m_blk               = first_m + in_group % rows;
    n_blk               = in_group / rows;
}

// ---- the kernel ------------------------------------------------------------------

template <int ACT, bool SAVE_Z, bool PAIR>
__global__ void __launch_bounds__(LT_THREADS, 1)
    k_linear_act(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
                 const void *__restrict__ bias, int bias_bf16, __nv_bfloat16 *__restrict__ y,
                 __nv_bfloat16 *__restrict__ z, int M, int N, int K)
{
    using Cfg = LtCfg<PAIR>;
    extern __shared__ uint8_t lt_smem_raw[];
    // the 128-byte swizzle pattern is a function of address bits [4,10): tiles need 1024-byte alignment
    // (the dynamic shared window starts at the same offset in both CTAs of a pair, so offsets match)
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(lt_smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t *bars      = reinterpret_cast<uint64_t *>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + Cfg::NUM_BARS);

    const uint32_t smem_base = smem_u32(smem);
    const uint32_t bar_base  = smem_u32(bars);
    auto full_bar            = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar           = [&](int s) { return bar_base + 8u * (Cfg::STAGES + s); };
    auto tmem_full_bar       = [&](int a) { return bar_base + 8u * (2 * Cfg::STAGES + a); };
    auto tmem_empty_bar      = [&](int a) { return bar_base + 8u * (2 * Cfg::STAGES + 2 + a); };

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;  // 0 = leader (issues the MMAs)

    const int tile_rows   = PAIR ? 2 * LT_BM : LT_BM;
    const int num_m       = (M + tile_rows - 1) / tile_rows;
    const int num_n       = (N + LT_BN - 1) / LT_BN;
    const int num_tiles   = num_m * num_n;
    const int num_kb      = (K + LT_BK - 1) / LT_BK;
    const int worker      = PAIR ? blockIdx.x >> 1 : blockIdx.x;
    const int num_workers = PAIR ? gridDim.x >> 1 : gridDim.x;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tm_x)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tm_w)) : "memory");
    } else if (warp == 1 && lane == 0) {
        for (int s = 0; s < Cfg::STAGES; ++s) {
            mbar_init(full_bar(s), 1);   // the (leader's) producer's arrive.expect_tx
            mbar_init(empty_bar(s), 1);  // one tcgen05.commit arrival
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tmem_full_bar(a), 1);
            mbar_init(tmem_empty_bar(a), LT_EPI_WARPS * (PAIR ? 2 : 1));  // both CTAs' epilogue warps
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    } else if (warp == 2) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"(LT_TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                             smem_u32(tmem_slot)),
                         "r"(LT_TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    if (PAIR)
        cluster_sync_all();  // the peer's barriers must be initialised before anything signals them
    else
        __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ===== TMA producer (every CTA stages its own operands) =====
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = worker; tile < num_tiles; tile += num_workers) {
            int m_blk, n_blk;
            tile_to_blocks(tile, num_m, num_n, m_blk, n_blk);
            const int m0 = m_blk * tile_rows + static_cast<int>(cta_rank) * LT_BM;
            const int n0 = n_blk * LT_BN + (PAIR ? static_cast<int>(cta_rank) * Cfg::B_ROWS : 0);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(empty_bar(stage), phase ^ 1);  // this CTA's slot has been consumed
                const uint32_t a_smem = smem_base + stage * Cfg::STAGE_BYTES;
                if (PAIR) {
                    // both CTAs' bytes are counted on the leader's barrier: that is the one the MMA thread waits on
                    const uint32_t sig = map_to_cta(full_bar(stage), 0);
                    if (cta_rank == 0) mbar_arrive_expect_tx(full_bar(stage), Cfg::TX_BYTES);
                    tma_load_2d_pair(a_smem, &tm_x, sig, kb * LT_BK, m0);
                    tma_load_2d_pair(a_smem + LT_A_BYTES, &tm_w, sig, kb * LT_BK, n0);
                } else {
                    mbar_arrive_expect_tx(full_bar(stage), Cfg::TX_BYTES);
                    tma_load_2d(a_smem, &tm_x, full_bar(stage), kb * LT_BK, m0);
                    tma_load_2d(a_smem + LT_A_BYTES, &tm_w, full_bar(stage), kb * LT_BK, n0);
                }
                if (++stage == Cfg::STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1 && lane == 0 && cta_rank == 0) {
        // ===== MMA issuer (a single thread; of the leader CTA for a pair) =====
        int stage = 0, accum = 0;
        uint32_t phase = 0, accum_phase = 0;
        for (int tile = worker; tile < num_tiles; tile += num_workers) {
            mbar_wait(tmem_empty_bar(accum), accum_phase ^ 1);  // the epilogue(s) drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + accum * LT_BN;
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(full_bar(stage), phase);  // TMA has landed this stage (in both CTAs)
                tc_fence_after();
                const uint32_t a_smem = smem_base + stage * Cfg::STAGE_BYTES;
                const uint64_t a_desc = make_kmajor_sw128_desc(a_smem);
                const uint64_t b_desc = make_kmajor_sw128_desc(a_smem + LT_A_BYTES);
#pragma unroll
                for (int k = 0; k < LT_BK / LT_UMMA_K; ++k) {
                    // advancing 16 elements (32 bytes) along K inside the swizzle row: +2 in the >>4 address field
                    tc_mma_bf16<PAIR>(d_tmem, a_desc + 2 * k, b_desc + 2 * k, Cfg::IDESC, (kb | k) != 0);
                }
                tc_commit<PAIR>(empty_bar(stage));  // stage reusable once these MMAs have read it
                if (++stage == Cfg::STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            tc_commit<PAIR>(tmem_full_bar(accum));  // accumulator complete
            if (++accum == 2) {
                accum = 0;
                accum_phase ^= 1;
            }
        }
    } else if (warp >= LT_FIRST_EPI_WARP) {
        // ===== epilogue: TMEM -> registers -> bias/activation -> global =====
        const int quarter = warp & 3;                        // the TMEM lanes this warp may read: 32*(warp%4)..+31
        const int part    = (warp - LT_FIRST_EPI_WARP) >> 2;  // which LT_EPI_COLS of the 256 accumulator columns
        int accum = 0;
        uint32_t accum_phase = 0;
        const float *bias_f         = reinterpret_cast<const float *>(bias);
        const __nv_bfloat16 *bias_h = reinterpret_cast<const __nv_bfloat16 *>(bias);
        for (int tile = worker; tile < num_tiles; tile += num_workers) {
            int m_blk, n_blk;
            tile_to_blocks(tile, num_m, num_n, m_blk, n_blk);
            const int m0  = m_blk * tile_rows + static_cast<int>(cta_rank) * LT_BM;
            const int n0  = n_blk * LT_BN;
            const int row = m0 + quarter * 32 + lane;
            // bias for this warp's columns (lane j holds column j of the 32-column chunk), fetched one
            // chunk ahead - the first one while the tensor cores are still working on the tile
            auto load_bias = [&](int nb) -> float {
                if (bias == nullptr || nb >= N) return 0.0f;
                return bias_bf16 ? __bfloat162float(bias_h[nb]) : bias_f[nb];
            };
            float b_next = load_bias(n0 + part * LT_EPI_COLS + lane);
            mbar_wait(tmem_full_bar(accum), accum_phase);
            tc_fence_after();
#pragma unroll 1  // one copy of the epilogue body: the unrolled form overflows the instruction cache
            for (int c = 0; c < LT_EPI_COLS / 32; ++c) {
                const int col0 = part * LT_EPI_COLS + c * 32;
                const int n    = n0 + col0;
                if (n >= N) break;  // warp-uniform
                uint32_t v[32];
                tc_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + accum * LT_BN + col0, v);
                const float b_lane = b_next;
                if (c + 1 < LT_EPI_COLS / 32) b_next = load_bias(n + 32 + lane);
                tc_ld_wait();
                float f[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + __shfl_sync(0xffffffffu, b_lane, j);
                if (row < M) {
                    const size_t off = static_cast<size_t>(row) * N + n;
                    if (n + 32 <= N) {
                        if (SAVE_Z) {
                            uint4 *zp = reinterpret_cast<uint4 *>(z + off);
#pragma unroll
                            for (int q = 0; q < 4; ++q)
                                zp[q] = make_uint4(pack_bf16(f[8 * q], f[8 * q + 1]), pack_bf16(f[8 * q + 2], f[8 * q + 3]),
                                                   pack_bf16(f[8 * q + 4], f[8 * q + 5]),
                                                   pack_bf16(f[8 * q + 6], f[8 * q + 7]));
                        }
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = apply_act<ACT>(f[j]);
                        uint4 *yp = reinterpret_cast<uint4 *>(y + off);
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            yp[q] = make_uint4(pack_bf16(f[8 * q], f[8 * q + 1]), pack_bf16(f[8 * q + 2], f[8 * q + 3]),
                                               pack_bf16(f[8 * q + 4], f[8 * q + 5]),
                                               pack_bf16(f[8 * q + 6], f[8 * q + 7]));
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (n + j < N) {
                                if (SAVE_Z) z[off + j] = __float2bfloat16(f[j]);
                                y[off + j] = __float2bfloat16(apply_act<ACT>(f[j]));
                            }
                        }
                    }
                }
            }
            // hand the accumulator back to the MMA thread (which lives in the leader CTA)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR)
                    mbar_arrive_cluster(map_to_cta(tmem_empty_bar(accum), 0));
                else
                    mbar_arrive(tmem_empty_bar(accum));
            }
            if (++accum == 2) {
                accum = 0;
                accum_phase ^= 1;
            }
        }
    }

    __syncwarp();  // the single-thread roles rejoin their warps before the aligned barrier
    tc_fence_before();
    if (PAIR)
        cluster_sync_all();  // neither CTA may leave while the other can still touch its smem / TMEM / barriers
    else
        __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        if (PAIR)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(LT_TMEM_COLS)
                         : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(LT_TMEM_COLS)
                         : "memory");
    }
}

// dz = dy * act'(z), one pass (the backward of the fused epilogue)
template <int ACT>
__global__ void __launch_bounds__(256) k_act_bwd(const uint4 *__restrict__ dy, const uint4 *__restrict__ z,
                                                  uint4 *__restrict__ dz, size_t n8)
{
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n8;
         i += static_cast<size_t>(gridDim.x) * blockDim.x) {
        uint4 a = dy[i], b = z[i], o;
        const __nv_bfloat162 *ap = reinterpret_cast<const __nv_bfloat162 *>(&a);
        const __nv_bfloat162 *bp = reinterpret_cast<const __nv_bfloat162 *>(&b);
        __nv_bfloat162 *op       = reinterpret_cast<__nv_bfloat162 *>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float2 g = __bfloat1622float2(ap[j]), v = __bfloat1622float2(bp[j]), r;
            if (ACT == KFB_ACT_GELU) {
                // d/dz [z Phi(z)] = Phi(z) + z phi(z)
                const float px = 0.5f * (1.0f + erff(v.x * 0.70710678f)), py = 0.5f * (1.0f + erff(v.y * 0.70710678f));
                r.x = g.x * (px + v.x * 0.39894228f * __expf(-0.5f * v.x * v.x));
                r.y = g.y * (py + v.y * 0.39894228f * __expf(-0.5f * v.y * v.y));
            } else {
                r.x = v.x > 0.0f ? g.x : 0.0f;
                r.y = v.y > 0.0f ? g.y : 0.0f;
            }
            op[j] = __floats2bfloat162_rn(r.x, r.y);
        }
        dz[i] = o;
    }
}

// ---- host side ---------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// [rows, K] bf16, K contiguous; box = 64 (K) x box_rows, 128-byte swizzle, zero fill out of bounds
static bool make_kmajor_map(CUtensorMap *map, const void *base, int rows, int K, int box_rows)
{
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    cuuint64_t dims[2]    = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(K) * 2};
    cuuint32_t box[2]     = {static_cast<cuuint32_t>(LT_BK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2]    = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// -1: choose per shape, 0: one CTA per tile, 1: CTA pairs (kfb_linear_act_set_mode / KUNGFU_LINEAR_TC_PAIR)
static int g_pair_mode = -2;

static int pair_mode()
{
    if (g_pair_mode == -2) {
        const char *e = getenv("KUNGFU_LINEAR_TC_PAIR");
        g_pair_mode   = e ? atoi(e) : -1;
    }
    return g_pair_mode;
}

template <int ACT, bool SAVE_Z, bool PAIR>
static int launch_linear(const void *x, const void *w, const void *bias, int bias_bf16, void *y, void *z, int M, int N,
                         int K, int sms, cudaStream_t stream)
{
    using Cfg = LtCfg<PAIR>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [] {
        attr_err = cudaFuncSetAttribute(k_linear_act<ACT, SAVE_Z, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::SMEM_BYTES);
    });
    if (attr_err != cudaSuccess) {
        set_last_error("linear_act: cannot opt in to %d bytes of shared memory: %s", Cfg::SMEM_BYTES,
                       cudaGetErrorString(attr_err));
        return static_cast<int>(attr_err);
    }
    CUtensorMap tx, tw;
    if (!make_kmajor_map(&tx, x, M, K, LT_BM) || !make_kmajor_map(&tw, w, N, K, Cfg::B_ROWS)) {
        set_last_error("linear_act: cuTensorMapEncodeTiled failed (M=%d N=%d K=%d)", M, N, K);
        return static_cast<int>(cudaErrorNotSupported);
    }
    const int tile_rows = PAIR ? 2 * LT_BM : LT_BM;
    const int tiles     = ((M + tile_rows - 1) / tile_rows) * ((N + LT_BN - 1) / LT_BN);
    const int workers   = PAIR ? sms / 2 : sms;
    const int grid      = (tiles < workers ? tiles : workers) * (PAIR ? 2 : 1);

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim            = dim3(grid);
    cfg.blockDim           = dim3(LT_THREADS);
    cfg.dynamicSmemBytes   = Cfg::SMEM_BYTES;
    cfg.stream             = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id               = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs                = attr;
    cfg.numAttrs             = 1;
    const cudaError_t e =
        cudaLaunchKernelEx(&cfg, k_linear_act<ACT, SAVE_Z, PAIR>, tx, tw, bias, bias_bf16,
                           static_cast<__nv_bfloat16 *>(y), static_cast<__nv_bfloat16 *>(z), M, N, K);
    if (e != cudaSuccess) set_last_error("linear_act launch: %s", cudaGetErrorString(e));
    return static_cast<int>(e);
}

template <bool PAIR>
static int dispatch_linear(const void *x, const void *w, const void *bias, int bb, void *y, void *z, int M, int N, int K,
                           int act, int sms, cudaStream_t s)
{
    const bool sz = z != nullptr;
    switch (act) {
    case KFB_ACT_NONE: return launch_linear<KFB_ACT_NONE, false, PAIR>(x, w, bias, bb, y, nullptr, M, N, K, sms, s);
    case KFB_ACT_GELU:
        return sz ? launch_linear<KFB_ACT_GELU, true, PAIR>(x, w, bias, bb, y, z, M, N, K, sms, s)
                  : launch_linear<KFB_ACT_GELU, false, PAIR>(x, w, bias, bb, y, z, M, N, K, sms, s);
    case KFB_ACT_RELU:
        return sz ? launch_linear<KFB_ACT_RELU, true, PAIR>(x, w, bias, bb, y, z, M, N, K, sms, s)
                  : launch_linear<KFB_ACT_RELU, false, PAIR>(x, w, bias, bb, y, z, M, N, K, sms, s);
    }
    set_last_error("linear_act: unknown activation %d", act);
    return static_cast<int>(cudaErrorInvalidValue);
}
}  // namespace kfb

extern "C" void kfb_linear_act_set_mode(int mode) { kfb::g_pair_mode = mode < -1 || mode > 1 ? -1 : mode; }

extern "C" int kfb_linear_act(const void *x, const void *w, const void *bias, int bias_dtype, void *y, void *z, int M,
                              int N, int K, int act, void *stream_)
{
    using namespace kfb;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    // TMA needs 16-byte global strides; the epilogue's vector stores need 16-byte rows
    if ((K % 8) != 0 || (N % 8) != 0) {
        set_last_error("linear_act: K and N must be multiples of 8 (got K=%d N=%d)", K, N);
        return static_cast<int>(cudaErrorInvalidValue);
    }
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(y) |
         reinterpret_cast<uintptr_t>(z)) & 15) {
        set_last_error("linear_act: x, w, y, z must be 16-byte aligned");
        return static_cast<int>(cudaErrorMisalignedAddress);
    }
    if (bias && bias_dtype != KFB_F32 && bias_dtype != KFB_BF16) {
        set_last_error("linear_act: bias must be f32 or bf16");
        return static_cast<int>(cudaErrorInvalidValue);
    }
    static int sms_of[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (sms_of[dev] == 0) cudaDeviceGetAttribute(&sms_of[dev], cudaDevAttrMultiProcessorCount, dev);
    const int sms = sms_of[dev] > 0 ? sms_of[dev] : 148;
    const int bb  = bias_dtype == KFB_BF16;
    int mode      = pair_mode();
    if (mode < 0) mode = M > LT_BM ? 1 : 0;  // a pair needs two row blocks to be worth its second SM
    return mode ? dispatch_linear<true>(x, w, bias, bb, y, z, M, N, K, act, sms, stream)
                : dispatch_linear<false>(x, w, bias, bb, y, z, M, N, K, act, sms, stream);
}

extern "C" int kfb_act_bwd(const void *dy, const void *z, void *dz, size_t n, int act, void *stream_)
{
    using namespace kfb;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n == 0) return 0;
    if (n % 8) {
        set_last_error("act_bwd: element count must be a multiple of 8");
        return static_cast<int>(cudaErrorInvalidValue);
    }
    const size_t n8 = n / 8;
    size_t want     = (n8 + 255) / 256;
    const int grid  = static_cast<int>(want < 148 * 8 ? want : 148 * 8);
    if (act == KFB_ACT_GELU)
        k_act_bwd<KFB_ACT_GELU><<<grid, 256, 0, stream>>>(static_cast<const uint4 *>(dy), static_cast<const uint4 *>(z),
                                                          static_cast<uint4 *>(dz), n8);
    else if (act == KFB_ACT_RELU)
        k_act_bwd<KFB_ACT_RELU><<<grid, 256, 0, stream>>>(static_cast<const uint4 *>(dy), static_cast<const uint4 *>(z),
                                                          static_cast<uint4 *>(dz), n8);
    else {
        set_last_error("act_bwd: unknown activation %d", act);
        return static_cast<int>(cudaErrorInvalidValue);
    }
    return static_cast<int>(cudaGetLastError());
}
