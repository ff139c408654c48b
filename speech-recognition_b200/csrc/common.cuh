// common.cuh -- shared helpers of libsrb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/srb200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libsrb200 is written for sm_100a (B200) only"
#endif

namespace srb {

void set_error(const char *fmt, ...);
int check_cuda(cudaError_t e, const char *what);
int sm_count();
int smem_optin();

#define SRB_CUDA(call)                                   \
    do {                                                 \
        int _rc = ::srb::check_cuda((call), #call);      \
        if (_rc) return _rc;                             \
    } while (0)

#define SRB_REQUIRE(cond, code, ...)                     \
    do {                                                 \
        if (!(cond)) {                                   \
            ::srb::set_error(__VA_ARGS__);               \
            return (code);                               \
        }                                                \
    } while (0)

constexpr float LOG2E_F = 1.4426950408889634f;
constexpr float LN2_F = 0.6931471805599453f;

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lg2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// packed fp32 FMA (fma.rn.f32x2, sm_100+): d = a * b + c on both halves in one issue slot
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }

__device__ __forceinline__ int lower_bound_i32(const int32_t *a, int n, int64_t v) {
    // first index i in [0, n] with a[i] >= v
    int lo = 0, hi = n;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if ((int64_t)a[mid] < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

}  // namespace srb
