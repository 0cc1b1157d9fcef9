// Fused batch-norm (+ residual add) (+ ReLU) for channels-last 16-bit
// activations with fp32 statistics: forward and backward, sm_100a.
//
// Why it is here: the library's headline benchmark is ResNet-50 throughput
// (/root/reference/benchmarks/system/benchmark_kungfu_torch.py:61-135,
// BASELINE.md 2.1) and on a B200 that model is bound by HBM traffic of its
// normalisation / activation / residual layers, not by tensor-core math.
// PyTorch runs BN, add and ReLU as separate passes (8 tensor passes forward,
// 8 backward per block output); these kernels need 4 forward (x twice,
// residual, y) and 7 backward (dy, x, y twice each + dx; the residual gradient
// is written by the same pass).
//
// Layout: x[rows][C] with C % 8 == 0; a thread owns 8 consecutive channels
// (one 16-byte packet) and walks rows, so its per-channel scale/shift stay in
// registers. Channel sums are reduced block-wise in shared memory, written as
// per-block partials and combined in fp64 in a fixed order (deterministic).
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>

#include "common.cuh"

namespace kfb {

constexpr int BN_THREADS   = 256;
// CTAs per SM the backward kernels are compiled for (4 => at most 64 registers)
#ifndef KFB_BN_BWD_OCC
#define KFB_BN_BWD_OCC 4
#endif
constexpr int BN_MAX_BLOCKS = 592;  // 4 CTAs per SM

template <typename T> struct Ld8 {
    __device__ static void load(const void *p, float *v) { Pack<T>::unpack(ld_stream16(p), v); }
    __device__ static void store(void *p, const float *v) { st_stream16(p, Pack<T>::pack(v)); }
};

// Block-level reduction of per-thread channel accumulators.
// acc[A][8]: A arrays of 8 channel lanes per thread. Result (sum over the
// RY row-slots of the block) goes to out[a * C + c] for this block.
template <int A>
__device__ __forceinline__ void
block_channel_reduce(float (&acc)[A][8], int tx, int ty, int C8, int RY, bool active, float *smem, float *out, int C)
{
    // smem layout: [A][RY][C] floats, processed in passes if it would not fit
    for (int a = 0; a < A; ++a) {
        __syncthreads();
        if (active) {
#pragma unroll
            for (int k = 0; k < 8; ++k) smem[(size_t)ty * C + tx * 8 + k] = acc[a][k];
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            float s = 0.f;
            for (int r = 0; r < RY; ++r) s += smem[(size_t)r * C + c];
            out[(size_t)a * C + c] = s;
        }
    }
}

// ---- forward -------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void
bn_stats_body(const uint4 *__restrict__ x, long long rows, int C, float *__restrict__ partial, float *smem)
{
    const int C8 = C >> 3, RY = BN_THREADS / C8;
    const int tx = threadIdx.x % C8, ty = threadIdx.x / C8;
    const bool active = ty < RY;
    float acc[2][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[0][k] = acc[1][k] = 0.f;
    if (active) {
        const long long stride = (long long)gridDim.x * RY;
        long long row = (long long)blockIdx.x * RY + ty;
        for (; row + 3 * stride < rows; row += 4 * stride) {
            float v[4][8];
#pragma unroll
            for (int u = 0; u < 4; ++u) Ld8<T>::load(x + (row + u * stride) * C8 + tx, v[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    acc[0][k] += v[u][k];
                    acc[1][k] = fmaf(v[u][k], v[u][k], acc[1][k]);
                }
        }
        for (; row < rows; row += stride) {
            float v[8];
            Ld8<T>::load(x + row * C8 + tx, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                acc[0][k] += v[k];
                acc[1][k] = fmaf(v[k], v[k], acc[1][k]);
            }
        }
    }
    block_channel_reduce<2>(acc, tx, ty, C8, RY, active, smem, partial + (size_t)blockIdx.x * 2 * C, C);
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS)
    k_bn_stats(const uint4 *__restrict__ x, long long rows, int C, float *__restrict__ partial)
{
    extern __shared__ float smem[];
    bn_stats_body<T>(x, rows, C, partial, smem);
}

// Partials are combined by CTAs of 32 channels x 8 slices: coalesced 128-byte
// reads, 8-way parallel over the partial blocks, fp64, fixed order.
constexpr int FIN_TX = 32, FIN_TY = 32;
__device__ __forceinline__ bool
reduce_partials(const float *__restrict__ partial, int nblocks, int C, double *s_out, double *q_out)
{
    __shared__ double sh[2][FIN_TY][FIN_TX];
    const int tx = threadIdx.x % FIN_TX, ty = threadIdx.x / FIN_TX;
    const int c  = blockIdx.x * FIN_TX + tx;
    double s = 0.0, q = 0.0;
    if (c < C) {
#pragma unroll 4
        for (int b = ty; b < nblocks; b += FIN_TY) {
            s += (double)partial[(size_t)b * 2 * C + c];
            q += (double)partial[(size_t)b * 2 * C + C + c];
        }
    }
    sh[0][ty][tx] = s;
    sh[1][ty][tx] = q;
    __syncthreads();
    if (ty != 0 || c >= C) return false;
    s = q = 0.0;
#pragma unroll
    for (int k = 0; k < FIN_TY; ++k) {
        s += sh[0][k][tx];
        q += sh[1][k][tx];
    }
    *s_out = s;
    *q_out = q;
    return true;
}

__global__ void __launch_bounds__(FIN_TX *FIN_TY)
    k_bn_finalize(const float *__restrict__ partial, int nblocks, int C, long long rows, float eps, float momentum,
                  float *running_mean, float *running_var, float *save_mean, float *save_invstd)
{
    double s, q;
    if (!reduce_partials(partial, nblocks, C, &s, &q)) return;
    const int c       = blockIdx.x * FIN_TX + threadIdx.x % FIN_TX;
    const double mean = s / (double)rows;
    double var        = q / (double)rows - mean * mean;
    if (var < 0.0) var = 0.0;
    save_mean[c]   = (float)mean;
    save_invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) {
        const double unbiased = rows > 1 ? var * (double)rows / (double)(rows - 1) : var;
        running_mean[c] = (float)((1.0 - momentum) * running_mean[c] + momentum * mean);
        running_var[c]  = (float)((1.0 - momentum) * running_var[c] + momentum * unbiased);
    }
}

__global__ void k_bn_eval_prep(const float *running_mean, const float *running_var, int C, float eps,
                               float *save_mean, float *save_invstd)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    save_mean[c]   = running_mean[c];
    save_invstd[c] = rsqrtf(running_var[c] + eps);
}

// The apply passes walk the rows from the END: the statistics pass that ran
// just before swept them front to back, so the tail of the tensor is what the
// 126 MB L2 still holds.
template <typename T, bool RELU, bool RES>
__device__ __forceinline__ void
bn_apply_body(const uint4 *__restrict__ x, const uint4 *__restrict__ res, uint4 *__restrict__ y,
              const float *__restrict__ weight, const float *__restrict__ bias,
              const float *mean, const float *invstd, long long rows, int C,
              uint8_t *__restrict__ mask)
{
    const int C8 = C >> 3, RY = BN_THREADS / C8;
    const int tx = threadIdx.x % C8, ty = threadIdx.x / C8;
    if (ty >= RY) return;
    float scale[8], shift[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c   = tx * 8 + k;
        const float w = weight ? weight[c] : 1.f, b = bias ? bias[c] : 0.f;
        scale[k]      = w * invstd[c];
        shift[k]      = fmaf(-mean[c], scale[k], b);
    }
    const long long stride = (long long)gridDim.x * RY;
    long long row          = (long long)blockIdx.x * RY + ty;
    for (; row + 3 * stride < rows; row += 4 * stride) {
        float v[4][8], r[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            Ld8<T>::load(x + (rows - 1 - (row + u * stride)) * C8 + tx, v[u]);
            if (RES) Ld8<T>::load(res + (rows - 1 - (row + u * stride)) * C8 + tx, r[u]);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            unsigned bits = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float o = fmaf(v[u][k], scale[k], shift[k]);
                if (RES) o += r[u][k];
                if (RELU) {
                    bits |= (o > 0.f ? 1u : 0u) << k;
                    o = fmaxf(o, 0.f);
                }
                v[u][k] = o;
            }
            Ld8<T>::store(y + (rows - 1 - (row + u * stride)) * C8 + tx, v[u]);
            if (RELU && mask) mask[(rows - 1 - (row + u * stride)) * C8 + tx] = (uint8_t)bits;
        }
    }
    for (; row < rows; row += stride) {
        float v[8], r[8];
        Ld8<T>::load(x + (rows - 1 - row) * C8 + tx, v);
        if (RES) Ld8<T>::load(res + (rows - 1 - row) * C8 + tx, r);
        unsigned bits = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float o = fmaf(v[k], scale[k], shift[k]);
            if (RES) o += r[k];
            if (RELU) {
                bits |= (o > 0.f ? 1u : 0u) << k;
                o = fmaxf(o, 0.f);
            }
            v[k] = o;
        }
        Ld8<T>::store(y + (rows - 1 - row) * C8 + tx, v);
        if (RELU && mask) mask[(rows - 1 - row) * C8 + tx] = (uint8_t)bits;
    }
}

template <typename T, bool RELU, bool RES>
__global__ void __launch_bounds__(BN_THREADS)
    k_bn_apply(const uint4 *__restrict__ x, const uint4 *__restrict__ res, uint4 *__restrict__ y,
               const float *__restrict__ weight, const float *__restrict__ bias,
               const float *__restrict__ mean, const float *__restrict__ invstd, long long rows, int C,
               uint8_t *__restrict__ mask)
{
    bn_apply_body<T, RELU, RES>(x, res, y, weight, bias, mean, invstd, rows, C, mask);
}

// ---- backward --------------------------------------------------------------------------
// partial[b][0][c] = sum dz, partial[b][1][c] = sum dz * xhat
template <typename T, bool RELU>
__device__ __forceinline__ void
bn_bwd_stats_body(const uint4 *__restrict__ dy, const uint4 *__restrict__ x, const uint8_t *__restrict__ mask,
                  const float *__restrict__ mean, const float *__restrict__ invstd, long long rows, int C,
                  float *__restrict__ partial, float *smem)
{
    const int C8 = C >> 3, RY = BN_THREADS / C8;
    const int tx = threadIdx.x % C8, ty = threadIdx.x / C8;
    const bool active = ty < RY;
    float acc[2][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[0][k] = acc[1][k] = 0.f;
    if (active) {
        float mu[8], is[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            mu[k] = mean[tx * 8 + k];
            is[k] = invstd[tx * 8 + k];
        }
        const long long stride = (long long)gridDim.x * RY;
        long long row          = (long long)blockIdx.x * RY + ty;
        for (; row + stride < rows; row += 2 * stride) {
            float g[2][8], v[2][8];
            unsigned o[2] = {0xffu, 0xffu};
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                Ld8<T>::load(dy + (row + u * stride) * C8 + tx, g[u]);
                Ld8<T>::load(x + (row + u * stride) * C8 + tx, v[u]);
                if (RELU) o[u] = mask[(row + u * stride) * C8 + tx];
            }
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float dz = (RELU && !((o[u] >> k) & 1u)) ? 0.f : g[u][k];
                    acc[0][k] += dz;
                    acc[1][k] = fmaf(dz, (v[u][k] - mu[k]) * is[k], acc[1][k]);
                }
        }
        for (; row < rows; row += stride) {
            float g[8], v[8];
            unsigned o = 0xffu;
            Ld8<T>::load(dy + row * C8 + tx, g);
            Ld8<T>::load(x + row * C8 + tx, v);
            if (RELU) o = mask[row * C8 + tx];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float dz = (RELU && !((o >> k) & 1u)) ? 0.f : g[k];
                acc[0][k] += dz;
                acc[1][k] = fmaf(dz, (v[k] - mu[k]) * is[k], acc[1][k]);
            }
        }
    }
    block_channel_reduce<2>(acc, tx, ty, C8, RY, active, smem, partial + (size_t)blockIdx.x * 2 * C, C);
}

template <typename T, bool RELU>
__global__ void __launch_bounds__(BN_THREADS, KFB_BN_BWD_OCC)
    k_bn_bwd_stats(const uint4 *__restrict__ dy, const uint4 *__restrict__ x, const uint8_t *__restrict__ mask,
                   const float *__restrict__ mean, const float *__restrict__ invstd, long long rows, int C,
                   float *__restrict__ partial)
{
    extern __shared__ float smem[];
    bn_bwd_stats_body<T, RELU>(dy, x, mask, mean, invstd, rows, C, partial, smem);
}

__global__ void __launch_bounds__(FIN_TX *FIN_TY)
    k_bn_bwd_finalize(const float *__restrict__ partial, int nblocks, int C, float *dweight, float *dbias,
                      float *sums /* [2][C] */)
{
    double s, q;
    if (!reduce_partials(partial, nblocks, C, &s, &q)) return;
    const int c = blockIdx.x * FIN_TX + threadIdx.x % FIN_TX;
    sums[c]     = (float)s;
    sums[C + c] = (float)q;
    if (dbias) dbias[c] = (float)s;
    if (dweight) dweight[c] = (float)q;
}

template <typename T, bool RELU, bool RES>
__device__ __forceinline__ void
bn_bwd_apply_body(const uint4 *__restrict__ dy, const uint4 *__restrict__ x, const uint8_t *__restrict__ mask,
                  const float *__restrict__ weight, const float *__restrict__ mean,
                  const float *__restrict__ invstd, const float *sums, uint4 *__restrict__ dx,
                  uint4 *__restrict__ dres, long long rows, int C)
{
    const int C8 = C >> 3, RY = BN_THREADS / C8;
    const int tx = threadIdx.x % C8, ty = threadIdx.x / C8;
    if (ty >= RY) return;
    const float inv_m = 1.f / (float)rows;
    float mu[8], is[8], sc[8], a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = tx * 8 + k;
        mu[k]       = mean[c];
        is[k]       = invstd[c];
        sc[k]       = (weight ? weight[c] : 1.f) * is[k];
        a[k]        = sums[c] * inv_m;      // mean of dz
        b[k]        = sums[C + c] * inv_m;  // mean of dz * xhat
    }
    const long long stride = (long long)gridDim.x * RY;
    long long row          = (long long)blockIdx.x * RY + ty;
    for (; row + stride < rows; row += 2 * stride) {
        float g[2][8], v[2][8];
            unsigned o[2] = {0xffu, 0xffu};
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            Ld8<T>::load(dy + (rows - 1 - (row + u * stride)) * C8 + tx, g[u]);
            Ld8<T>::load(x + (rows - 1 - (row + u * stride)) * C8 + tx, v[u]);
            if (RELU) o[u] = mask[(rows - 1 - (row + u * stride)) * C8 + tx];
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float dz   = (RELU && !((o[u] >> k) & 1u)) ? 0.f : g[u][k];
                const float xhat = (v[u][k] - mu[k]) * is[k];
                g[u][k]          = dz;
                v[u][k]          = sc[k] * (dz - a[k] - xhat * b[k]);
            }
            Ld8<T>::store(dx + (rows - 1 - (row + u * stride)) * C8 + tx, v[u]);
            if (RES) Ld8<T>::store(dres + (rows - 1 - (row + u * stride)) * C8 + tx, g[u]);
        }
    }
    for (; row < rows; row += stride) {
        float g[8], v[8];
            unsigned o = 0xffu;
        Ld8<T>::load(dy + (rows - 1 - row) * C8 + tx, g);
        Ld8<T>::load(x + (rows - 1 - row) * C8 + tx, v);
        if (RELU) o = mask[(rows - 1 - row) * C8 + tx];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float dz   = (RELU && !((o >> k) & 1u)) ? 0.f : g[k];
            const float xhat = (v[k] - mu[k]) * is[k];
            g[k]             = dz;
            v[k]             = sc[k] * (dz - a[k] - xhat * b[k]);
        }
        Ld8<T>::store(dx + (rows - 1 - row) * C8 + tx, v);
        if (RES) Ld8<T>::store(dres + (rows - 1 - row) * C8 + tx, g);
    }
}

template <typename T, bool RELU, bool RES>
// (the residual variant needs ~75 registers: capping it at 64 spills and measures slower)
__global__ void __launch_bounds__(BN_THREADS, RES ? 3 : KFB_BN_BWD_OCC)
    k_bn_bwd_apply(const uint4 *__restrict__ dy, const uint4 *__restrict__ x, const uint8_t *__restrict__ mask,
                   const float *__restrict__ weight, const float *__restrict__ mean,
                   const float *__restrict__ invstd, const float *__restrict__ sums, uint4 *__restrict__ dx,
                   uint4 *__restrict__ dres, long long rows, int C)
{
    bn_bwd_apply_body<T, RELU, RES>(dy, x, mask, weight, mean, invstd, sums, dx, dres, rows, C);
}

static int bn_grid(long long rows, int C)
{
    const int RY      = BN_THREADS / (C >> 3);
    long long blocks  = (rows + (long long)RY * 4 - 1) / ((long long)RY * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > BN_MAX_BLOCKS) blocks = BN_MAX_BLOCKS;
    return (int)blocks;
}

// Grid = at most one resident wave (SMs x CTAs/SM): rows are walked grid-stride,
// so a partial second wave would only add a tail.
template <typename Kern> static int one_wave(Kern kern, size_t smem)
{
    static std::mutex mu;
    static std::map<const void *, int> cache;
    const void *key = reinterpret_cast<const void *>(kern);
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    int dev = 0, sms = 148, occ = 1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, BN_THREADS, smem) != cudaSuccess || occ < 1) occ = 1;
    if (occ > 4) occ = 4;
    int g = sms * occ;
    if (g > BN_MAX_BLOCKS) g = BN_MAX_BLOCKS;
    cache[key] = g;
    return g;
}

static int bn_check(long long rows, int C)
{
    if (C % 8 || C < 8 || (C >> 3) > BN_THREADS || rows < 1) {
        set_last_error("bn_act: channels must be a multiple of 8 in [8, %d] (got %d)", BN_THREADS * 8, C);
        return KFB_ERR_UNSUPP;
    }
    return KFB_OK;
}

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_bn_scratch_floats(long long rows, int channels)
{
    // partial sums [blocks][2][C] + reduced sums [2][C]
    if (channels % 8 || channels < 8 || (channels >> 3) > BN_THREADS) return 0;
    return (bn_grid(rows, channels) + 1) * 2 * channels;
}

template <typename T>
static int bn_fwd_t(const void *x, const void *residual, void *y, const float *weight, const float *bias,
                    float *running_mean, float *running_var, float *save_mean, float *save_invstd,
                    long long rows, int C, float eps, float momentum, int relu, int training, float *scratch,
                    uint8_t *mask, cudaStream_t s)
{
    const int grid = bn_grid(rows, C);
    const int RY   = BN_THREADS / (C >> 3);
    if (training) {
        const size_t smem = (size_t)RY * C * sizeof(float);
        const int gs      = std::min(grid, one_wave(k_bn_stats<T>, smem));
        k_bn_stats<T><<<gs, BN_THREADS, smem, s>>>(reinterpret_cast<const uint4 *>(x), rows, C, scratch);
        KFB_CUDA_OK(cudaGetLastError());
        k_bn_finalize<<<(C + FIN_TX - 1) / FIN_TX, FIN_TX * FIN_TY, 0, s>>>(scratch, gs, C, rows, eps, momentum, running_mean,
                                                       running_var, save_mean, save_invstd);
    } else {
        k_bn_eval_prep<<<(C + 127) / 128, 128, 0, s>>>(running_mean, running_var, C, eps, save_mean, save_invstd);
    }
    KFB_CUDA_OK(cudaGetLastError());
    const uint4 *xp = reinterpret_cast<const uint4 *>(x), *rp = reinterpret_cast<const uint4 *>(residual);
    uint4 *yp       = reinterpret_cast<uint4 *>(y);
#define KFB_BN_APPLY(R, S)                                                                                  \
    k_bn_apply<T, R, S><<<std::min(grid, one_wave(k_bn_apply<T, R, S>, 0)), BN_THREADS, 0, s>>>(xp, rp, yp, weight, bias, save_mean, save_invstd, rows, C, mask)
    if (relu && residual) KFB_BN_APPLY(true, true);
    else if (relu) KFB_BN_APPLY(true, false);
    else if (residual) KFB_BN_APPLY(false, true);
    else KFB_BN_APPLY(false, false);
#undef KFB_BN_APPLY
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_bn_act_fwd(const void *x, const void *residual, void *y, const float *weight,
                              const float *bias, float *running_mean, float *running_var, float *save_mean,
                              float *save_invstd, int64_t rows, int channels, float eps, float momentum, int relu,
                              int training, int dtype, float *scratch, void *relu_mask, void *stream)
{
    int rc = bn_check(rows, channels);
    if (rc) return rc;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16)
        return bn_fwd_t<__nv_bfloat16>(x, residual, y, weight, bias, running_mean, running_var, save_mean,
                                       save_invstd, rows, channels, eps, momentum, relu, training, scratch,
                                       reinterpret_cast<uint8_t *>(relu_mask), s);
    if (dtype == KFB_F16)
        return bn_fwd_t<__half>(x, residual, y, weight, bias, running_mean, running_var, save_mean, save_invstd,
                                rows, channels, eps, momentum, relu, training, scratch, reinterpret_cast<uint8_t *>(relu_mask), s);
    set_last_error("bn_act supports bf16 / f16 activations");
    return KFB_ERR_UNSUPP;
}

template <typename T>
static int bn_bwd_t(const void *dy, const void *x, const void *mask, const float *weight, const float *save_mean,
                    const float *save_invstd, void *dx, void *dres, float *dweight, float *dbias, long long rows,
                    int C, int relu, float *scratch, cudaStream_t s)
{
    const int grid    = bn_grid(rows, C);
    const int RY      = BN_THREADS / (C >> 3);
    const size_t smem = (size_t)RY * C * sizeof(float);
    float *sums       = scratch + (size_t)grid * 2 * C;
    const uint4 *gp = reinterpret_cast<const uint4 *>(dy), *xp = reinterpret_cast<const uint4 *>(x)
                ;
    const uint8_t *yp = reinterpret_cast<const uint8_t *>(mask);
    const int gs = std::min(grid, relu ? one_wave(k_bn_bwd_stats<T, true>, smem) : one_wave(k_bn_bwd_stats<T, false>, smem));
    if (relu)
        k_bn_bwd_stats<T, true><<<gs, BN_THREADS, smem, s>>>(gp, xp, yp, save_mean, save_invstd, rows, C, scratch);
    else
        k_bn_bwd_stats<T, false><<<gs, BN_THREADS, smem, s>>>(gp, xp, yp, save_mean, save_invstd, rows, C, scratch);
    KFB_CUDA_OK(cudaGetLastError());
    k_bn_bwd_finalize<<<(C + FIN_TX - 1) / FIN_TX, FIN_TX * FIN_TY, 0, s>>>(scratch, gs, C, dweight, dbias, sums);
    KFB_CUDA_OK(cudaGetLastError());
    uint4 *dxp = reinterpret_cast<uint4 *>(dx), *drp = reinterpret_cast<uint4 *>(dres);
#define KFB_BN_BWD(R, S)                                                                                    \
    k_bn_bwd_apply<T, R, S><<<std::min(grid, one_wave(k_bn_bwd_apply<T, R, S>, 0)), BN_THREADS, 0, s>>>(gp, xp, yp, weight, save_mean, save_invstd, sums, dxp, drp, rows, C)
    if (relu && dres) KFB_BN_BWD(true, true);
    else if (relu) KFB_BN_BWD(true, false);
    else if (dres) KFB_BN_BWD(false, true);
    else KFB_BN_BWD(false, false);
#undef KFB_BN_BWD
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_bn_act_bwd(const void *dy, const void *x, const void *relu_mask, const float *weight,
                              const float *save_mean, const float *save_invstd, void *dx, void *dresidual,
                              float *dweight, float *dbias, int64_t rows, int channels, int relu, int dtype,
                              float *scratch, void *stream)
{
    int rc = bn_check(rows, channels);
    if (rc) return rc;
    if (relu && !relu_mask) {
        set_last_error("bn_act backward with relu needs the mask written by the forward");
        return KFB_ERR_ARG;
    }
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16)
        return bn_bwd_t<__nv_bfloat16>(dy, x, relu_mask, weight, save_mean, save_invstd, dx, dresidual, dweight, dbias, rows,
                                       channels, relu, scratch, s);
    if (dtype == KFB_F16)
        return bn_bwd_t<__half>(dy, x, relu_mask, weight, save_mean, save_invstd, dx, dresidual, dweight, dbias, rows,
                                channels, relu, scratch, s);
    set_last_error("bn_act supports bf16 / f16 activations");
    return KFB_ERR_UNSUPP;
}
