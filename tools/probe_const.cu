// probe: 20 Gaussians x 39 dims per frame, one frame per thread; parameters (a) as kernel-parameter constants (uniform
// datapath, LDCU), (b) from shared memory with broadcast LDS.128 (the K1 core's form).  Prints ms per 10 M frames.
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
struct WP { float2 ab[20][39]; float c[20]; };
__global__ void __launch_bounds__(128, 4) kc(const __grid_constant__ WP p, const float* __restrict__ x, float* __restrict__ out, int n) {
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        float2 xx[39];
#pragma unroll
        for (int d = 0; d < 39; ++d) { float v = x[(size_t)(t & 0xfffff) * 39 + d]; xx[d] = make_float2(v * v, v); }
        float best = 1e30f;
#pragma unroll
        for (int g = 0; g < 20; ++g) {
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int d = 0; d < 39; ++d) acc = __ffma2_rn(xx[d], p.ab[g][d], acc);
            best = fminf(best, acc.x + acc.y + p.c[g]);
        }
        out[t] = best;
    }
}
__global__ void __launch_bounds__(128, 4) ks(const WP* __restrict__ pg, const float* __restrict__ x, float* __restrict__ out, int n) {
    __shared__ float4 ab[5][39][2];   // blocks of 4 Gaussians: (A0,B0,A1,B1),(A2,B2,A3,B3)
    __shared__ float cs[20];
    for (int i = threadIdx.x; i < 20 * 39; i += blockDim.x) {
        int g = i / 39, d = i % 39; float2 v = pg->ab[g][d];
        float* q = reinterpret_cast<float*>(&ab[g / 4][d][(g % 4) / 2]) + 2 * (g % 2); q[0] = v.x; q[1] = v.y;
    }
    if (threadIdx.x < 20) cs[threadIdx.x] = pg->c[threadIdx.x];
    __syncthreads();
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        float2 xx[39];
#pragma unroll
        for (int d = 0; d < 39; ++d) { float v = x[(size_t)(t & 0xfffff) * 39 + d]; xx[d] = make_float2(v * v, v); }
        float best = 1e30f;
#pragma unroll 1
        for (int b = 0; b < 5; ++b) {
            float2 a0 = make_float2(0.f, 0.f), a1 = a0, a2 = a0, a3 = a0;
#pragma unroll
            for (int d = 0; d < 39; ++d) {
                const float4 p01 = ab[b][d][0], p23 = ab[b][d][1];
                a0 = __ffma2_rn(xx[d], make_float2(p01.x, p01.y), a0);
                a1 = __ffma2_rn(xx[d], make_float2(p01.z, p01.w), a1);
                a2 = __ffma2_rn(xx[d], make_float2(p23.x, p23.y), a2);
                a3 = __ffma2_rn(xx[d], make_float2(p23.z, p23.w), a3);
            }
            best = fminf(best, fminf(fminf(a0.x + a0.y + cs[4 * b], a1.x + a1.y + cs[4 * b + 1]),
                                     fminf(a2.x + a2.y + cs[4 * b + 2], a3.x + a3.y + cs[4 * b + 3])));
        }
        out[t] = best;
    }
}
int main() {
    const int n = 10000000;
    float *x, *out; WP* pg;
    cudaMalloc(&x, (size_t)(1 << 20) * 39 * 4); cudaMalloc(&out, (size_t)n * 4); cudaMalloc(&pg, sizeof(WP));
    std::vector<float> hx((size_t)(1 << 20) * 39);
    for (size_t i = 0; i < hx.size(); ++i) hx[i] = (float)((i * 2654435761u) % 1000) / 500.f - 1.f;
    cudaMemcpy(x, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice);
    WP hp;
    for (int g = 0; g < 20; ++g) { for (int d = 0; d < 39; ++d) hp.ab[g][d] = make_float2(-0.5f - 0.01f * g, 0.1f * d - 0.02f * g); hp.c[g] = 0.3f * g; }
    cudaMemcpy(pg, &hp, sizeof(WP), cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 4;
    for (int v = 0; v < 2; ++v) {
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            if (v == 0) kc<<<grid, 128>>>(hp, x, out, n); else ks<<<grid, 128>>>(pg, x, out, n);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            printf("%s rep %d: %.3f ms per 10 M frames (%s)\n", v == 0 ? "const-param" : "smem-bcast", rep, ms, cudaGetErrorString(cudaGetLastError()));
        }
        std::vector<float> ho(8); cudaMemcpy(ho.data(), out + 12345, 32, cudaMemcpyDeviceToHost); printf("  out %g %g\n", ho[0], ho[7]);
    }
    return 0;
}
