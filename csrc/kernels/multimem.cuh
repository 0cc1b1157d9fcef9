// multimem (NVLS) load-reduce / store wrappers shared by kernels outside
// collectives.cu (which carries the same definitions for its own use).
// multimem.ld_reduce: the NVSwitch reads the same offset of every GPU bound to
// the multicast object and returns the sum (fp32 accumulation for 16-bit
// types); multimem.st: the switch writes the value to every GPU.
#pragma once
#include "common.cuh"

namespace kfb {

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

}  // namespace kfb
