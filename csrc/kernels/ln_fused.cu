// y = LayerNorm(x + dropout(h)) * gamma + beta, forward and backward, bf16.
//
// The transformer block's "add & norm" (BERT: after the attention projection
// and after the FFN) as ONE pass each way instead of dropout + add + layer-norm
// (3 kernels, 4 reads + 3 writes of the [tokens, hidden] activation) and, on the
// way back, layer-norm input grad + gamma/beta grad + dropout grad (3-4 kernels):
//
//   forward   reads h, x; writes y and s = x + dropout(h) (what backward needs),
//             mean / rstd per row
//   backward  reads dy, s; writes dx (= ds), dh (= ds * mask / (1-p)) and
//             per-block partial sums of dgamma / dbeta, summed by a small second
//             kernel in a fixed order (deterministic, no atomics)
//
// No dropout mask is stored: it is regenerated from a counter-based generator
// (Philox 4x32-10) keyed by (seed, stream id of the call site) and indexed by
// (training step, element), so forward and backward of the same step agree and
// a CUDA graph replays with fresh masks as long as the step counter (a device
// scalar the caller bumps once per step) advances.
//
// One warp per row, hidden size = VPL * 256 (256, 512, 768, 1024): lane l owns
// 8 consecutive elements of every 256-element slice, so all global accesses are
// 16-byte and coalesced, the row lives in registers, statistics are two-pass.
//
// The reference has no counterpart (its models are TensorFlow's); this serves
// the BERT benchmark model (kungfu_b200/models/bert.py).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "kungfu_b200.h"

namespace kfb
{
void set_last_error(const char *fmt, ...);  // collectives.cu

constexpr int LN_THREADS = 256;
constexpr int LN_WARPS   = LN_THREADS / 32;

// Philox 4x32-10 (Salmon et al., SC'11): 128 random bits per (counter, key)
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

struct DropoutCfg {
    float p;                   // drop probability; 0 disables
    uint64_t seed;
    const long long *step;     // device scalar: training step (may be NULL = 0)
    uint32_t stream;           // distinguishes call sites within a step
};

// keep-mask of the 8 elements of vector `vec` (global index of the 16-byte group): bit j = keep element j
__device__ __forceinline__ uint32_t keep_mask8(const DropoutCfg &d, unsigned long long step, unsigned long long vec)
{
    const uint4 r = philox4x32_10(make_uint4((uint32_t)vec, (uint32_t)(vec >> 32), (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)d.seed ^ d.stream, (uint32_t)(d.seed >> 32)));
    const uint32_t thr = (uint32_t)(d.p * 65536.0f);  // drop when the 16-bit draw is below p * 2^16
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        m |= ((w[j] & 0xFFFFu) >= thr ? 1u : 0u) << (2 * j);
        m |= ((w[j] >> 16) >= thr ? 1u : 0u) << (2 * j + 1);
    }
    return m;
}

__device__ __forceinline__ void unpack8(const uint4 &v, float (&f)[8])
{
    const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h[j]);
        f[2 * j]       = t.x;
        f[2 * j + 1]   = t.y;
    }
}

__device__ __forceinline__ uint4 pack8(const float (&f)[8])
{
    uint4 v;
    __nv_bfloat162 *h = reinterpret_cast<__nv_bfloat162 *>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    return v;
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int VPL>
__global__ void __launch_bounds__(LN_THREADS)
    k_add_dropout_ln_fwd(const uint4 *__restrict__ h, const uint4 *__restrict__ x, const uint4 *__restrict__ gamma,
                         const uint4 *__restrict__ beta, uint4 *__restrict__ y, uint4 *__restrict__ s_out,
                         float *__restrict__ mean_out, float *__restrict__ rstd_out, long long M, float eps,
                         DropoutCfg d)
{
    constexpr int H   = VPL * 256;
    constexpr int VEC = H / 8;  // 16-byte vectors per row
    const int lane    = threadIdx.x & 31;
    const long long warp0  = (long long)blockIdx.x * LN_WARPS + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * LN_WARPS;
    const unsigned long long step = d.step ? (unsigned long long)*d.step : 0ull;
    const float keep_scale        = d.p > 0.f ? 1.0f / (1.0f - d.p) : 1.0f;

    float g[VPL][8], b[VPL][8];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        unpack8(gamma[v * 32 + lane], g[v]);
        unpack8(beta[v * 32 + lane], b[v]);
    }
    for (long long row = warp0; row < M; row += nwarps) {
        float s[VPL][8];
        float sum = 0.f;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const long long vi = row * VEC + v * 32 + lane;
            float hv[8];
            unpack8(h[vi], hv);
            if (d.p > 0.f) {
                const uint32_t keep = keep_mask8(d, step, (unsigned long long)vi);
#pragma unroll
                for (int j = 0; j < 8; ++j) hv[j] = (keep >> j) & 1u ? hv[j] * keep_scale : 0.f;
            }
            if (x != nullptr) {
                float xv[8];
                unpack8(x[vi], xv);
#pragma unroll
                for (int j = 0; j < 8; ++j) hv[j] += xv[j];
            }
            // statistics are taken over the bf16-rounded sum: exactly what backward will read back
            const uint4 packed = pack8(hv);
            if (s_out != nullptr) s_out[vi] = packed;
            unpack8(packed, s[v]);
#pragma unroll
            for (int j = 0; j < 8; ++j) sum += s[v][j];
        }
        const float mean = warp_sum(sum) * (1.0f / H);
        float sq         = 0.f;
#pragma unroll
        for (int v = 0; v < VPL; ++v)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float c = s[v][j] - mean;
                sq            = fmaf(c, c, sq);
            }
        const float rstd = rsqrtf(warp_sum(sq) * (1.0f / H) + eps);
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = fmaf((s[v][j] - mean) * rstd, g[v][j], b[v][j]);
            y[row * VEC + v * 32 + lane] = pack8(o);
        }
        if (lane == 0 && mean_out != nullptr) {
            mean_out[row] = mean;
            rstd_out[row] = rstd;
        }
    }
}

// partial: [gridDim.x][2][H] floats (dgamma row then dbeta row per block)
template <int VPL>
__global__ void __launch_bounds__(LN_THREADS)
    k_add_dropout_ln_bwd(const uint4 *__restrict__ dy, const uint4 *__restrict__ s, const float *__restrict__ mean,
                         const float *__restrict__ rstd, const uint4 *__restrict__ gamma, uint4 *__restrict__ dx,
                         uint4 *__restrict__ dh, float *__restrict__ partial, long long M, DropoutCfg d)
{
    constexpr int H   = VPL * 256;
    constexpr int VEC = H / 8;
    const int lane    = threadIdx.x & 31;
    const int warp    = threadIdx.x >> 5;
    const long long warp0  = (long long)blockIdx.x * LN_WARPS + warp;
    const long long nwarps = (long long)gridDim.x * LN_WARPS;
    const unsigned long long step = d.step ? (unsigned long long)*d.step : 0ull;
    const float keep_scale        = d.p > 0.f ? 1.0f / (1.0f - d.p) : 1.0f;

    float g[VPL][8], dg[VPL][8], db[VPL][8];
#pragma unroll
    for (int v = 0; v < VPL; ++v) {
        unpack8(gamma[v * 32 + lane], g[v]);
#pragma unroll
        for (int j = 0; j < 8; ++j) dg[v][j] = db[v][j] = 0.f;
    }
    for (long long row = warp0; row < M; row += nwarps) {
        const float mu = mean[row], rs = rstd[row];
        float xh[VPL][8], gy[VPL][8];
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const long long vi = row * VEC + v * 32 + lane;
            float dyv[8], sv[8];
            unpack8(dy[vi], dyv);
            unpack8(s[vi], sv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                xh[v][j] = (sv[j] - mu) * rs;
                gy[v][j] = dyv[j] * g[v][j];
                c1 += gy[v][j];
                c2 = fmaf(gy[v][j], xh[v][j], c2);
                dg[v][j] = fmaf(dyv[j], xh[v][j], dg[v][j]);
                db[v][j] += dyv[j];
            }
        }
        c1 = warp_sum(c1) * (1.0f / H);
        c2 = warp_sum(c2) * (1.0f / H);
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
            const long long vi = row * VEC + v * 32 + lane;
            float ds[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) ds[j] = rs * (gy[v][j] - c1 - xh[v][j] * c2);
            if (dx != nullptr) dx[vi] = pack8(ds);
            if (dh != nullptr) {
                if (d.p > 0.f) {
                    const uint32_t keep = keep_mask8(d, step, (unsigned long long)vi);
#pragma unroll
                    for (int j = 0; j < 8; ++j) ds[j] = (keep >> j) & 1u ? ds[j] * keep_scale : 0.f;
                }
                dh[vi] = pack8(ds);
            }
        }
    }
    // block-level sum of the per-warp column accumulators, one quantity at a time (24 KiB of smem for H = 768)
    __shared__ float red[LN_WARPS][H];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        __syncthreads();
#pragma unroll
        for (int v = 0; v < VPL; ++v)
#pragma unroll
            for (int j = 0; j < 8; ++j) red[warp][(v * 32 + lane) * 8 + j] = q == 0 ? dg[v][j] : db[v][j];
        __syncthreads();
        for (int c = threadIdx.x; c < H; c += LN_THREADS) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < LN_WARPS; ++w) t += red[w][c];
            partial[((size_t)blockIdx.x * 2 + q) * H + c] = t;
        }
    }
}

// dgamma[c] = sum_b partial[b][0][c], dbeta[c] = sum_b partial[b][1][c]. 32 x 32 threads per block: 32
// consecutive columns (coalesced), 32 threads per column each summing every 32nd block, then a fixed-order
// combine through shared memory (deterministic).
__global__ void __launch_bounds__(1024)
    k_ln_bwd_finalize(const float *__restrict__ partial, int blocks, int H, __nv_bfloat16 *__restrict__ dgamma,
                      __nv_bfloat16 *__restrict__ dbeta)
{
    __shared__ float red[32][33];
    const int c = blockIdx.x * 32 + threadIdx.x;  // column in [0, 2H)
    float t     = 0.f;
    if (c < 2 * H) {
        const int q = c / H, col = c % H;
        for (int b = threadIdx.y; b < blocks; b += 32) t += partial[((size_t)b * 2 + q) * H + col];
    }
    red[threadIdx.y][threadIdx.x] = t;
    __syncthreads();
    if (threadIdx.y == 0 && c < 2 * H) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) s += red[k][threadIdx.x];
        (c < H ? dgamma : dbeta)[c % H] = __float2bfloat16(s);
    }
}

static int sm_count()
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    static int cached[64];
    if (dev >= 0 && dev < 64) {
        if (cached[dev] == 0) cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev);
        if (cached[dev] > 0) sms = cached[dev];
    }
    return sms;
}

// forward: up to 4 CTAs per SM; backward: its register footprint allows one CTA per SM, so one wave
static int ln_grid(long long M, int per_sm)
{
    const long long want = (M + LN_WARPS - 1) / LN_WARPS;
    const long long cap  = (long long)sm_count() * per_sm;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

// ---- column sum: out[N] = sum over rows of a[M, N] (bias gradients) -------------------------
//
// grid (N / 64, R): a block owns 64 columns (one bf16 pair per lane) of one row split; its 8 warps
// stride over the rows, combine through shared memory and write one partial row. The last block
// of a column group to finish (ticket counter, reset for the next call) adds the R partials in
// order - deterministic, one launch, no atomics on the data.
// V = bf16 elements per lane: 8 (16-byte loads, a warp spans 256 columns) when N % 256 == 0, else 2
template <int V>
__global__ void __launch_bounds__(256)
    k_colsum(const __nv_bfloat16 *__restrict__ a, long long M, int N, __nv_bfloat16 *__restrict__ out,
             float *__restrict__ partial, unsigned int *__restrict__ tickets)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int R        = gridDim.y;
    const int col      = (blockIdx.x * 32 + lane) * V;  // first column of this lane
    const long long r0 = (M * blockIdx.y) / R, r1 = (M * (blockIdx.y + 1)) / R;
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
#pragma unroll 4
    for (long long r = r0 + warp; r < r1; r += 8) {
        if (V == 8) {
            float f[8];
            unpack8(*reinterpret_cast<const uint4 *>(a + r * N + col), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        } else {
            const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(a + r * N + col));
            acc[0] += v.x;
            acc[1] += v.y;
        }
    }
    __shared__ float red[8][32 * V + 1];
    __shared__ bool is_last;
#pragma unroll
    for (int j = 0; j < V; ++j) red[warp][lane * V + j] = acc[j];
    __syncthreads();
    // all 256 threads combine the 8 warps' rows: thread t owns columns t, t + 256, ...
    for (int c = threadIdx.x; c < 32 * V; c += 256) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][c];
        partial[(size_t)blockIdx.y * N + blockIdx.x * 32 * V + c] = t;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = atomicAdd(&tickets[blockIdx.x], 1u) == (unsigned)R - 1;
    __syncthreads();
    if (is_last) {
        __threadfence();
        for (int c = threadIdx.x; c < 32 * V; c += 256) {
            float s = 0.f;
#pragma unroll 8  // independent loads in flight: the chain of R L2 round trips was the kernel's critical path
            for (int y = 0; y < R; ++y) s += __ldcg(partial + (size_t)y * N + blockIdx.x * 32 * V + c);
            out[blockIdx.x * 32 * V + c] = __float2bfloat16(s);
        }
        if (threadIdx.x == 0) tickets[blockIdx.x] = 0u;
    }
}

// ---- gradient of "split a packed qkv into heads": (dq, dk, dv) [B, h, S, d] -> dqkv [B, S, 3, h, d] ----
//
// Autograd does this as a stack (one pass) followed by a permuting copy (another pass); this is one
// 16-byte-vectorised gather. Sources may be strided in (b, h, s); d is contiguous.
struct PackQkvArgs {
    const uint4 *src[3];
    long long sb[3], sh[3], ss[3];  // strides in elements
    int B, h, S, d8;                // d8 = d / 8
};

__global__ void __launch_bounds__(256) k_pack_qkv_grad(const __grid_constant__ PackQkvArgs a, uint4 *__restrict__ out)
{
    const long long total = (long long)a.B * a.S * 3 * a.h * a.d8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        long long t  = i;
        const int v  = (int)(t % a.d8);
        t /= a.d8;
        const int hh = (int)(t % a.h);
        t /= a.h;
        const int w = (int)(t % 3);
        t /= 3;
        const int s = (int)(t % a.S);
        const int b = (int)(t / a.S);
        const long long off = b * a.sb[w] + hh * a.sh[w] + s * a.ss[w];  // elements; multiple of 8 by contract
        out[i]              = a.src[w][off / 8 + v];
    }
}
}  // namespace kfb

extern "C" int kfb_ln_partial_floats(long long M, int H)
{
    return kfb::ln_grid(M, 1) * 2 * H;
}

extern "C" int kfb_colsum_scratch_floats(int N)
{
    return 64 * N + N / 64 + 64;  // up to 64 row splits of partials, then the tickets
}

extern "C" int kfb_colsum(const void *a, long long M, int N, void *out, float *scratch, void *stream)
{
    using namespace kfb;
    if (N <= 0 || (N % 64) != 0 || M <= 0) {
        set_last_error("colsum: N must be a positive multiple of 64 and M > 0 (got M=%lld N=%d)", M, N);
        return (int)cudaErrorInvalidValue;
    }
    const bool wide  = (N % 256) == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0;
    const int groups = wide ? N / 256 : N / 64;
    int R            = (3 * sm_count() + groups - 1) / groups;
    if (R > 64) R = 64;
    if (R > M) R = (int)M;
    if (R < 1) R = 1;
    unsigned int *tickets = reinterpret_cast<unsigned int *>(scratch + (size_t)64 * N);
    cudaStream_t st       = static_cast<cudaStream_t>(stream);
    if (wide)
        k_colsum<8><<<dim3(groups, R), 256, 0, st>>>(static_cast<const __nv_bfloat16 *>(a), M, N,
                                                      static_cast<__nv_bfloat16 *>(out), scratch, tickets);
    else
        k_colsum<2><<<dim3(groups, R), 256, 0, st>>>(static_cast<const __nv_bfloat16 *>(a), M, N,
                                                      static_cast<__nv_bfloat16 *>(out), scratch, tickets);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) set_last_error("colsum launch: %s", cudaGetErrorString(e));
    return (int)e;
}

extern "C" int kfb_pack_qkv_grad(const void *dq, const void *dk, const void *dv, const long long *strides9, int B,
                                 int h, int S, int d, void *out, void *stream)
{
    using namespace kfb;
    if (d % 8) {
        set_last_error("pack_qkv_grad: head dim %d must be a multiple of 8", d);
        return (int)cudaErrorInvalidValue;
    }
    PackQkvArgs a;
    const void *src[3] = {dq, dk, dv};
    for (int w = 0; w < 3; ++w) {
        a.src[w] = static_cast<const uint4 *>(src[w]);
        a.sb[w]  = strides9[3 * w];
        a.sh[w]  = strides9[3 * w + 1];
        a.ss[w]  = strides9[3 * w + 2];
        if ((a.sb[w] | a.sh[w] | a.ss[w]) % 8 || (reinterpret_cast<uintptr_t>(src[w]) & 15)) {
            set_last_error("pack_qkv_grad: sources must be 16-byte aligned with strides that are multiples of 8");
            return (int)cudaErrorInvalidValue;
        }
    }
    a.B = B, a.h = h, a.S = S, a.d8 = d / 8;
    const long long total = (long long)B * S * 3 * h * a.d8;
    if (total == 0) return 0;
    long long want = (total + 255) / 256;
    const int grid = (int)(want < (long long)sm_count() * 8 ? want : (long long)sm_count() * 8);
    k_pack_qkv_grad<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a, static_cast<uint4 *>(out));
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) set_last_error("pack_qkv_grad launch: %s", cudaGetErrorString(e));
    return (int)e;
}

#define KFB_LN_DISPATCH(H, ...)                                                                   \
    switch (H) {                                                                                  \
    case 256: { constexpr int VPL = 1; __VA_ARGS__; } break;                                      \
    case 512: { constexpr int VPL = 2; __VA_ARGS__; } break;                                      \
    case 768: { constexpr int VPL = 3; __VA_ARGS__; } break;                                      \
    case 1024: { constexpr int VPL = 4; __VA_ARGS__; } break;                                     \
    default:                                                                                      \
        kfb::set_last_error("add_dropout_ln: hidden size %d not in {256, 512, 768, 1024}", H);    \
        return (int)cudaErrorInvalidValue;                                                        \
    }

extern "C" int kfb_add_dropout_ln_fwd(const void *h, const void *x, const void *gamma, const void *beta, void *y,
                                      void *s_out, float *mean, float *rstd, long long M, int H, float eps, float p,
                                      unsigned long long seed, const long long *step, unsigned int stream_id,
                                      void *stream)
{
    using namespace kfb;
    if (M <= 0) return 0;
    if (p < 0.f || p >= 1.f) {
        set_last_error("add_dropout_ln: dropout probability %f outside [0, 1)", p);
        return (int)cudaErrorInvalidValue;
    }
    const DropoutCfg d{p, seed, step, stream_id};
    const int grid = ln_grid(M, 4);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KFB_LN_DISPATCH(H, (k_add_dropout_ln_fwd<VPL><<<grid, LN_THREADS, 0, st>>>(
                           static_cast<const uint4 *>(h), static_cast<const uint4 *>(x),
                           static_cast<const uint4 *>(gamma), static_cast<const uint4 *>(beta),
                           static_cast<uint4 *>(y), static_cast<uint4 *>(s_out), mean, rstd, M, eps, d)));
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) set_last_error("add_dropout_ln fwd launch: %s", cudaGetErrorString(e));
    return (int)e;
}

extern "C" int kfb_add_dropout_ln_bwd(const void *dy, const void *s, const float *mean, const float *rstd,
                                      const void *gamma, void *dx, void *dh, void *dgamma, void *dbeta,
                                      float *partial, long long M, int H, float p, unsigned long long seed,
                                      const long long *step, unsigned int stream_id, void *stream)
{
    using namespace kfb;
    if (M <= 0) return 0;
    const DropoutCfg d{p, seed, step, stream_id};
    const int grid = ln_grid(M, 1);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    KFB_LN_DISPATCH(H, (k_add_dropout_ln_bwd<VPL><<<grid, LN_THREADS, 0, st>>>(
                           static_cast<const uint4 *>(dy), static_cast<const uint4 *>(s), mean, rstd,
                           static_cast<const uint4 *>(gamma), static_cast<uint4 *>(dx), static_cast<uint4 *>(dh),
                           partial, M, d)));
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) {
        k_ln_bwd_finalize<<<(2 * H + 31) / 32, dim3(32, 32), 0, st>>>(
            partial, grid, H, static_cast<__nv_bfloat16 *>(dgamma), static_cast<__nv_bfloat16 *>(dbeta));
        e = cudaGetLastError();
    }
    if (e != cudaSuccess) set_last_error("add_dropout_ln bwd launch: %s", cudaGetErrorString(e));
    return (int)e;
}
