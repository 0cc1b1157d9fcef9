// 3x3 / stride 2 / pad 1 max pooling for channels-last 16-bit activations
// (the ResNet stem), forward and backward, sm_100a.
//
// PyTorch's NHWC max-pool keeps int64 argmax indices (4x the activation bytes)
// and its backward runs at a small fraction of HBM speed on B200; here the
// argmax is one byte per element (window slot 0..8), the backward is a gather
// (no atomics, deterministic): every input packet looks at the <= 4 windows
// that contain it. Part of the ResNet-50 benchmark path
// (/root/reference/benchmarks/system/benchmark_kungfu_torch.py:61).
#include "common.cuh"

namespace kfb {

template <typename T>
__global__ void __launch_bounds__(256)
    k_maxpool_fwd(const uint4 *__restrict__ x, uint4 *__restrict__ y, uint2 *__restrict__ idx, int N, int H, int W,
                  int C8, int OH, int OW)
{
    const long long total  = (long long)N * OH * OW * C8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int c  = (int)(i % C8);
        long long p  = i / C8;
        const int ow = (int)(p % OW);
        p /= OW;
        const int oh = (int)(p % OH);
        const int n  = (int)(p / OH);
        float best[8];
        unsigned slot[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            best[k] = -INFINITY;
            slot[k] = 0;
        }
#pragma unroll
        for (int kh = 0; kh < 3; ++kh) {
            const int ih = oh * 2 - 1 + kh;
            if (ih < 0 || ih >= H) continue;
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const int iw = ow * 2 - 1 + kw;
                if (iw < 0 || iw >= W) continue;
                float v[8];
                Pack<T>::unpack(__ldg(x + (((long long)n * H + ih) * W + iw) * C8 + c), v);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    if (v[k] > best[k]) {
                        best[k] = v[k];
                        slot[k] = kh * 3 + kw;
                    }
                }
            }
        }
        st_stream16(y + i, Pack<T>::pack(best));
        uint2 s;
        s.x = slot[0] | (slot[1] << 8) | (slot[2] << 16) | (slot[3] << 24);
        s.y = slot[4] | (slot[5] << 8) | (slot[6] << 16) | (slot[7] << 24);
        idx[i] = s;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_maxpool_bwd(const uint4 *__restrict__ dy, const uint2 *__restrict__ idx, uint4 *__restrict__ dx, int N, int H,
                  int W, int C8, int OH, int OW)
{
    const long long total  = (long long)N * H * W * C8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int c  = (int)(i % C8);
        long long p  = i / C8;
        const int iw = (int)(p % W);
        p /= W;
        const int ih = (int)(p % H);
        const int n  = (int)(p / H);
        float g[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = 0.f;
        const int oh0 = ih >> 1, oh1 = (ih + 1) >> 1;  // equal when ih is even
        const int ow0 = iw >> 1, ow1 = (iw + 1) >> 1;
        for (int oh = oh0; oh <= oh1; ++oh) {
            if (oh >= OH) break;
            const int kh = ih - (oh * 2 - 1);
            for (int ow = ow0; ow <= ow1; ++ow) {
                if (ow >= OW) break;
                const unsigned want = (unsigned)(kh * 3 + (iw - (ow * 2 - 1)));
                const long long o   = (((long long)n * OH + oh) * OW + ow) * C8 + c;
                const uint2 s       = __ldg(idx + o);
                float d[8];
                Pack<T>::unpack(__ldg(dy + o), d);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const unsigned sl = ((k < 4 ? s.x : s.y) >> (8 * (k & 3))) & 0xffu;
                    if (sl == want) g[k] += d[k];
                }
            }
        }
        st_stream16(dx + i, Pack<T>::pack(g));
    }
}

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_maxpool3x3s2_fwd(const void *x, void *y, void *idx, int N, int H, int W, int C, int dtype,
                                    void *stream)
{
    if (C % 8) {
        set_last_error("maxpool: channels must be a multiple of 8");
        return KFB_ERR_UNSUPP;
    }
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1, C8 = C / 8;
    const long long total = (long long)N * OH * OW * C8;
    const int blocks      = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    cudaStream_t s        = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16)
        k_maxpool_fwd<__nv_bfloat16><<<blocks, 256, 0, s>>>((const uint4 *)x, (uint4 *)y, (uint2 *)idx, N, H, W, C8, OH, OW);
    else if (dtype == KFB_F16)
        k_maxpool_fwd<__half><<<blocks, 256, 0, s>>>((const uint4 *)x, (uint4 *)y, (uint2 *)idx, N, H, W, C8, OH, OW);
    else {
        set_last_error("maxpool supports bf16 / f16");
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_maxpool3x3s2_bwd(const void *dy, const void *idx, void *dx, int N, int H, int W, int C,
                                    int dtype, void *stream)
{
    if (C % 8) {
        set_last_error("maxpool: channels must be a multiple of 8");
        return KFB_ERR_UNSUPP;
    }
    const int OH = (H + 2 - 3) / 2 + 1, OW = (W + 2 - 3) / 2 + 1, C8 = C / 8;
    const long long total = (long long)N * H * W * C8;
    const int blocks      = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    cudaStream_t s        = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16)
        k_maxpool_bwd<__nv_bfloat16><<<blocks, 256, 0, s>>>((const uint4 *)dy, (const uint2 *)idx, (uint4 *)dx, N, H, W, C8, OH, OW);
    else if (dtype == KFB_F16)
        k_maxpool_bwd<__half><<<blocks, 256, 0, s>>>((const uint4 *)dy, (const uint2 *)idx, (uint4 *)dx, N, H, W, C8, OH, OW);
    else {
        set_last_error("maxpool supports bf16 / f16");
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

// ---- global average pooling (N,H,W,C) -> (N,C), and its broadcast backward ----
namespace kfb {

template <typename T>
__global__ void __launch_bounds__(256)
    k_gap_fwd(const uint4 *__restrict__ x, uint4 *__restrict__ y, int N, int HW, int C8)
{
    // one thread per (n, c8): HW is small (49 for ResNet), consecutive threads read consecutive packets
    const long long total = (long long)N * C8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C8);
        const long long n = i / C8;
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.f;
        for (int p = 0; p < HW; ++p) {
            float v[8];
            Pack<T>::unpack(ld_stream16(x + (n * HW + p) * C8 + c), v);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
        const float inv = 1.f / (float)HW;
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] *= inv;
        st_stream16(y + i, Pack<T>::pack(acc));
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
    k_gap_bwd(const uint4 *__restrict__ dy, uint4 *__restrict__ dx, int N, int HW, int C8)
{
    const long long total = (long long)N * HW * C8;
    const float inv       = 1.f / (float)HW;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c       = (int)(i % C8);
        const long long n = i / ((long long)HW * C8);
        float v[8];
        Pack<T>::unpack(__ldg(dy + n * C8 + c), v);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] *= inv;
        st_stream16(dx + i, Pack<T>::pack(v));
    }
}

}  // namespace kfb

extern "C" int kfb_gap_fwd(const void *x, void *y, int N, int HW, int C, int dtype, void *stream)
{
    if (C % 8) {
        kfb::set_last_error("global avg pool: channels must be a multiple of 8");
        return KFB_ERR_UNSUPP;
    }
    const long long total = (long long)N * (C / 8);
    const int blocks      = (int)std::min<long long>((total + 255) / 256, 148LL * 8);
    cudaStream_t s        = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16) kfb::k_gap_fwd<__nv_bfloat16><<<blocks, 256, 0, s>>>((const uint4 *)x, (uint4 *)y, N, HW, C / 8);
    else if (dtype == KFB_F16) kfb::k_gap_fwd<__half><<<blocks, 256, 0, s>>>((const uint4 *)x, (uint4 *)y, N, HW, C / 8);
    else {
        kfb::set_last_error("global avg pool supports bf16 / f16");
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}

extern "C" int kfb_gap_bwd(const void *dy, void *dx, int N, int HW, int C, int dtype, void *stream)
{
    if (C % 8) {
        kfb::set_last_error("global avg pool: channels must be a multiple of 8");
        return KFB_ERR_UNSUPP;
    }
    const long long total = (long long)N * HW * (C / 8);
    const int blocks      = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    cudaStream_t s        = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == KFB_BF16) kfb::k_gap_bwd<__nv_bfloat16><<<blocks, 256, 0, s>>>((const uint4 *)dy, (uint4 *)dx, N, HW, C / 8);
    else if (dtype == KFB_F16) kfb::k_gap_bwd<__half><<<blocks, 256, 0, s>>>((const uint4 *)dy, (uint4 *)dx, N, HW, C / 8);
    else {
        kfb::set_last_error("global avg pool supports bf16 / f16");
        return KFB_ERR_UNSUPP;
    }
    KFB_CUDA_OK(cudaGetLastError());
    return KFB_OK;
}
