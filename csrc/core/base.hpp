// Data types, reduce ops, strategy names, CPU reduction kernel.
//
// Parity: /root/reference/srcs/cpp/include/kungfu/{dtype.h,op.h,strategy.h},
// /root/reference/srcs/go/kungfu/base/{op.cpp,f16.c,strategy.go}. The dtype
// codes keep the reference's bit layout (category<<16 | bytes<<8 | 8) so values
// interchange; bf16 is an added category.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace kf {

constexpr uint32_t type_code(uint32_t cat, uint32_t bytes) { return (cat << 16) | (bytes << 8) | 8; }

enum DType : uint32_t {
    U8   = type_code(0, 1),
    U16  = type_code(0, 2),
    U32  = type_code(0, 4),
    U64  = type_code(0, 8),
    I8   = type_code(1, 1),
    I16  = type_code(1, 2),
    I32  = type_code(1, 4),
    I64  = type_code(1, 8),
    F16  = type_code(2, 2),
    F32  = type_code(2, 4),
    F64  = type_code(2, 8),
    BOOL = type_code(3, 1),
    BF16 = type_code(4, 2),
};

inline size_t dtype_size(uint32_t dt) { return (dt >> 8) & 0xff; }
bool dtype_valid(uint32_t dt);
const char *dtype_name(uint32_t dt);

enum Op : int { SUM = 0, MIN = 1, MAX = 2, PROD = 3 };

// Same ordinal values as KungFu_Strategy (strategy.h:7-17).
enum Strategy : int {
    Tree = 0,
    BinaryTree,
    Ring,
    Star,
    MultiStar,
    Clique,
    BinaryTreeStar,
    MultiBinaryTreeStar,
    Auto,
};
const char *strategy_name(Strategy s);
bool parse_strategy(const std::string &name, Strategy *out);
std::vector<std::string> strategy_names();

// z[i] = x[i] (op) y[i]; x/y/z may alias. fp16/bf16 accumulate in fp32 and
// round once (F16C/AVX2 fast path for SUM).
void transform2(const void *x, const void *y, void *z, size_t n, uint32_t dt, Op op);

float half_to_float(uint16_t h);
uint16_t float_to_half(float f);
inline float bf16_to_float(uint16_t b)
{
    union { uint32_t u; float f; } v;
    v.u = (uint32_t)b << 16;
    return v.f;
}
uint16_t float_to_bf16(float f);

// Exponential moving average with first-sample passthrough
// (/root/reference/srcs/cpp/include/kungfu/utils/ema.hpp:19-28).
template <typename T> class EMA
{
    T alpha_;
    bool has_ = false;
    T value_  = 0;

  public:
    explicit EMA(T alpha) : alpha_(alpha) {}
    T update(T x)
    {
        if (has_) {
            value_ = alpha_ * value_ + (1 - alpha_) * x;
        } else {
            has_   = true;
            value_ = x;
        }
        return value_;
    }
    T get() const { return value_; }
};

// A typed, named pair of buffers: the unit every collective works on
// (/root/reference/srcs/go/kungfu/base/workspace.go:10-50).
struct Workspace {
    const void *send = nullptr;
    void *recv       = nullptr;
    size_t count     = 0;
    uint32_t dtype   = U8;
    Op op            = SUM;
    std::string name;

    size_t bytes() const { return count * dtype_size(dtype); }
    bool inplace() const { return send == recv; }
    // [b, e) element slice named "part::<name>[b:e]"
    Workspace slice(size_t b, size_t e) const;
    void forward() const;  // recv := send unless in place
};

// Split [0, n) into k near-equal intervals (/root/reference/srcs/go/plan/interval.go:12-27).
struct Interval {
    size_t begin, end;
    size_t len() const { return end - begin; }
};
std::vector<Interval> even_partition(size_t n, size_t k);

}  // namespace kf
