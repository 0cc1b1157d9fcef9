#include "core/base.hpp"

#include <immintrin.h>

#include <algorithm>
#include <cstring>

namespace kf {

bool dtype_valid(uint32_t dt)
{
    switch (dt) {
    case U8: case U16: case U32: case U64: case I8: case I16: case I32: case I64:
    case F16: case F32: case F64: case BOOL: case BF16:
        return true;
    }
    return false;
}

const char *dtype_name(uint32_t dt)
{
    switch (dt) {
    case U8: return "uint8";
    case U16: return "uint16";
    case U32: return "uint32";
    case U64: return "uint64";
    case I8: return "int8";
    case I16: return "int16";
    case I32: return "int32";
    case I64: return "int64";
    case F16: return "float16";
    case F32: return "float32";
    case F64: return "float64";
    case BOOL: return "bool";
    case BF16: return "bfloat16";
    }
    return "?";
}

static const char *kStrategyNames[] = {
    "TREE", "BINARY_TREE", "RING", "STAR", "MULTI_STAR", "CLIQUE",
    "BINARY_TREE_STAR", "MULTI_BINARY_TREE_STAR", "AUTO",
};

const char *strategy_name(Strategy s) { return kStrategyNames[(int)s]; }

bool parse_strategy(const std::string &name, Strategy *out)
{
    for (int i = 0; i <= (int)Auto; ++i) {
        if (name == kStrategyNames[i]) {
            *out = (Strategy)i;
            return true;
        }
    }
    return false;
}

std::vector<std::string> strategy_names()
{
    return std::vector<std::string>(kStrategyNames, kStrategyNames + 9);
}

float half_to_float(uint16_t h) { return _cvtsh_ss(h); }
uint16_t float_to_half(float f) { return _cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT); }

uint16_t float_to_bf16(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // quiet NaN
    u += 0x7fffu + ((u >> 16) & 1);  // round to nearest even
    return (uint16_t)(u >> 16);
}

namespace {

template <typename T> struct OpSum { T operator()(T a, T b) const { return a + b; } };
template <typename T> struct OpMin { T operator()(T a, T b) const { return std::min(a, b); } };
template <typename T> struct OpMax { T operator()(T a, T b) const { return std::max(a, b); } };
template <typename T> struct OpProd { T operator()(T a, T b) const { return a * b; } };

// The element-wise loop of every reduce in the graph walk. The library is built with -O2, where GCC's
// vectoriser gives up on loops that need a run-time overlap check, so this one is compiled at -O3 and told
// that iterations are independent (z may BE x or y, the in-place case, but never overlaps them shifted):
// 8-lane AVX2 adds / mins / maxes instead of one scalar per cycle.
template <typename T, typename F>
__attribute__((optimize("O3", "tree-vectorize"))) void apply(const void *x, const void *y, void *z, size_t n, F f)
{
    const T *px = static_cast<const T *>(x);
    const T *py = static_cast<const T *>(y);
    T *pz       = static_cast<T *>(z);
#pragma GCC ivdep
    for (size_t i = 0; i < n; ++i) pz[i] = f(px[i], py[i]);
}

template <typename T> void apply_op(const void *x, const void *y, void *z, size_t n, Op op)
{
    switch (op) {
    case SUM: apply<T>(x, y, z, n, OpSum<T>()); break;
    case MIN: apply<T>(x, y, z, n, OpMin<T>()); break;
    case MAX: apply<T>(x, y, z, n, OpMax<T>()); break;
    case PROD: apply<T>(x, y, z, n, OpProd<T>()); break;
    }
}

// 8 lanes of fp32 through one of the four ops. min / max take their operands swapped so that the
// result for unordered inputs is the one std::min(a, b) / std::max(a, b) give in the scalar tail.
template <Op OP> inline __m256 vec_op(__m256 a, __m256 b)
{
    if (OP == SUM) return _mm256_add_ps(a, b);
    if (OP == MIN) return _mm256_min_ps(b, a);
    if (OP == MAX) return _mm256_max_ps(b, a);
    return _mm256_mul_ps(a, b);
}

inline float scalar_op(Op op, float a, float b)
{
    switch (op) {
    case SUM: return a + b;
    case MIN: return std::min(a, b);
    case MAX: return std::max(a, b);
    default: return a * b;
    }
}

template <Op OP> size_t half_lanes(const uint16_t *px, const uint16_t *py, uint16_t *pz, size_t n)
{
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {
        __m256 a = _mm256_cvtph_ps(_mm_loadu_si128((const __m128i *)(px + i)));
        __m256 b = _mm256_cvtph_ps(_mm_loadu_si128((const __m128i *)(py + i)));
        _mm_storeu_si128((__m128i *)(pz + i), _mm256_cvtps_ph(vec_op<OP>(a, b), _MM_FROUND_TO_NEAREST_INT));
    }
    return i;
}

void apply_half(const void *x, const void *y, void *z, size_t n, Op op)
{
    const uint16_t *px = static_cast<const uint16_t *>(x);
    const uint16_t *py = static_cast<const uint16_t *>(y);
    uint16_t *pz       = static_cast<uint16_t *>(z);
    size_t i           = 0;
    switch (op) {
    case SUM: i = half_lanes<SUM>(px, py, pz, n); break;
    case MIN: i = half_lanes<MIN>(px, py, pz, n); break;
    case MAX: i = half_lanes<MAX>(px, py, pz, n); break;
    case PROD: i = half_lanes<PROD>(px, py, pz, n); break;
    }
    for (; i < n; ++i) pz[i] = float_to_half(scalar_op(op, half_to_float(px[i]), half_to_float(py[i])));
}

template <Op OP> size_t bf16_lanes(const uint16_t *px, const uint16_t *py, uint16_t *pz, size_t n)
{
    const __m256i bias = _mm256_set1_epi32(0x7fff);
    const __m256i one  = _mm256_set1_epi32(1);
    size_t i           = 0;
    for (; i + 8 <= n; i += 8) {
        __m256i ai = _mm256_slli_epi32(_mm256_cvtepu16_epi32(_mm_loadu_si128((const __m128i *)(px + i))), 16);
        __m256i bi = _mm256_slli_epi32(_mm256_cvtepu16_epi32(_mm_loadu_si128((const __m128i *)(py + i))), 16);
        __m256 s   = vec_op<OP>(_mm256_castsi256_ps(ai), _mm256_castsi256_ps(bi));
        __m256i u  = _mm256_castps_si256(s);
        // round to nearest even (NaN payloads are not expected in gradients)
        __m256i lsb = _mm256_and_si256(_mm256_srli_epi32(u, 16), one);
        u           = _mm256_srli_epi32(_mm256_add_epi32(u, _mm256_add_epi32(bias, lsb)), 16);
        __m128i lo  = _mm256_castsi256_si128(u);
        __m128i hi  = _mm256_extracti128_si256(u, 1);
        _mm_storeu_si128((__m128i *)(pz + i), _mm_packus_epi32(lo, hi));
    }
    return i;
}

void apply_bf16(const void *x, const void *y, void *z, size_t n, Op op)
{
    const uint16_t *px = static_cast<const uint16_t *>(x);
    const uint16_t *py = static_cast<const uint16_t *>(y);
    uint16_t *pz       = static_cast<uint16_t *>(z);
    size_t i           = 0;
    switch (op) {
    case SUM: i = bf16_lanes<SUM>(px, py, pz, n); break;
    case MIN: i = bf16_lanes<MIN>(px, py, pz, n); break;
    case MAX: i = bf16_lanes<MAX>(px, py, pz, n); break;
    case PROD: i = bf16_lanes<PROD>(px, py, pz, n); break;
    }
    for (; i < n; ++i) pz[i] = float_to_bf16(scalar_op(op, bf16_to_float(px[i]), bf16_to_float(py[i])));
}

}  // namespace

void transform2(const void *x, const void *y, void *z, size_t n, uint32_t dt, Op op)
{
    switch (dt) {
    case U8: case BOOL: apply_op<uint8_t>(x, y, z, n, op); break;
    case U16: apply_op<uint16_t>(x, y, z, n, op); break;
    case U32: apply_op<uint32_t>(x, y, z, n, op); break;
    case U64: apply_op<uint64_t>(x, y, z, n, op); break;
    case I8: apply_op<int8_t>(x, y, z, n, op); break;
    case I16: apply_op<int16_t>(x, y, z, n, op); break;
    case I32: apply_op<int32_t>(x, y, z, n, op); break;
    case I64: apply_op<int64_t>(x, y, z, n, op); break;
    case F32: apply_op<float>(x, y, z, n, op); break;
    case F64: apply_op<double>(x, y, z, n, op); break;
    case F16: apply_half(x, y, z, n, op); break;
    case BF16: apply_bf16(x, y, z, n, op); break;
    }
}

Workspace Workspace::slice(size_t b, size_t e) const
{
    Workspace w;
    const size_t es = dtype_size(dtype);
    w.send  = static_cast<const char *>(send) + b * es;
    w.recv  = static_cast<char *>(recv) + b * es;
    w.count = e - b;
    w.dtype = dtype;
    w.op    = op;
    w.name  = "part::" + name + "[" + std::to_string(b) + ":" + std::to_string(e) + "]";
    return w;
}

void Workspace::forward() const
{
    if (!inplace()) memcpy(recv, send, bytes());
}

std::vector<Interval> even_partition(size_t n, size_t k)
{
    std::vector<Interval> out;
    if (k == 0) return out;
    const size_t q = n / k, r = n % k;
    size_t off = 0;
    for (size_t i = 0; i < k; ++i) {
        const size_t len = q + (i < r ? 1 : 0);
        out.push_back({off, off + len});
        off += len;
    }
    return out;
}

}  // namespace kf
