// Two fp32 values that travel together: on sm_100a the CUDA cores have packed FMUL2 / FADD2 / FFMA2 (one instruction,
// two independent IEEE fp32 results), which the compiler only emits for the explicit .f32x2 PTX forms.  The worker warps
// of the tensor-core kernels push their element-wise jet math through this type two points at a time (jet_math.cuh is
// templated on the value type), which halves the arithmetic instructions of that chain; each half is bit-identical to
// the scalar operation (round-to-nearest fp32, fma fused once).  On the host the same type is two plain floats, so the
// CPU harness (tests/harness/jet_host.cu) can run the packed instantiation of the formulas.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace tdq {

struct F2 {
    float x, y;
    __host__ __device__ __forceinline__ F2() {}
    __host__ __device__ __forceinline__ F2(float f) : x(f), y(f) {}
    __host__ __device__ __forceinline__ F2(float a, float b) : x(a), y(b) {}
};

#ifdef __CUDA_ARCH__
#define TDQ_F2_OP2(NAME, PTX)                                                                                              \
    __device__ __forceinline__ F2 NAME(F2 a, F2 b) {                                                                       \
        F2 d;                                                                                                              \
        asm("{\n\t.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t" PTX                          \
            " rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;\n\t}"                                                                   \
            : "=f"(d.x), "=f"(d.y)                                                                                         \
            : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));                                                                     \
        return d;                                                                                                          \
    }
TDQ_F2_OP2(operator*, "mul.rn.f32x2")
TDQ_F2_OP2(operator+, "add.rn.f32x2")
TDQ_F2_OP2(operator-, "sub.rn.f32x2")
#undef TDQ_F2_OP2
// a * b + c, fused
__device__ __forceinline__ F2 jm_fma(F2 a, F2 b, F2 c) {
    F2 d;
    asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
        : "=f"(d.x), "=f"(d.y)
        : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
#else
__host__ __device__ __forceinline__ F2 operator*(F2 a, F2 b) { return F2(a.x * b.x, a.y * b.y); }
__host__ __device__ __forceinline__ F2 operator+(F2 a, F2 b) { return F2(a.x + b.x, a.y + b.y); }
__host__ __device__ __forceinline__ F2 operator-(F2 a, F2 b) { return F2(a.x - b.x, a.y - b.y); }
__host__ __device__ __forceinline__ F2 jm_fma(F2 a, F2 b, F2 c) { return F2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
#endif
__host__ __device__ __forceinline__ F2& operator+=(F2& a, F2 b) { a = a + b; return a; }

}  // namespace tdq
