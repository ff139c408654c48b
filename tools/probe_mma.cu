// probe: issue rate of legacy mma.sync.m16n8k8 TF32 on sm_100a (decides whether a 3xTF32 scorer is worth building)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k(int iters, float *sink) {
    float d[8][4];
    unsigned a[4] = {0x3f800000u + threadIdx.x, 0x3f800001u, 0x3f800002u, 0x3f800003u}, b[2] = {0x3f000000u, 0x3f000001u};
#pragma unroll
    for (int i = 0; i < 8; ++i) d[i][0] = d[i][1] = d[i][2] = d[i][3] = (float)i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
    if (s == 123.456f) sink[0] = s;
}
int main() {
    float *sink; cudaMalloc(&sink, 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int wps = 1; wps <= 4; wps *= 2) {          // CTAs of 256 threads per SM
        const int iters = 20000, grid = sms * wps;
        k<<<grid, 256>>>(100, sink);
        cudaEventRecord(e0); k<<<grid, 256>>>(iters, sink); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double mmas = (double)grid * 8 * iters * 8;   // warps * iters * 8
        printf("ctas/SM %d: %.3f ms, %.2f T FMA/s tf32 (m16n8k8 = 1024 FMA), %.2f mma/clk/SM @1.965GHz, err=%d\n", wps, ms,
               mmas * 1024 / (ms * 1e-3) / 1e12, mmas / (ms * 1e-3) / sms / 1.965e9, (int)cudaGetLastError());
    }
    return 0;
}
