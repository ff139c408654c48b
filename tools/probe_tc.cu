// probe_tc.cu -- stand-alone check of the tcgen05 TF32x3 building blocks used by score_tc.cu:
//   (1) operands in shared memory (K-major, no swizzle, chunk-plane layout of tc05.cuh), (2) A operand in TMEM written
//   with tcgen05.st (thread = row), (3) accuracy of the 3-product TF32 split against FP64, (4) cycles per MMA.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe_tc tools/probe_tc.cu ; run on a B200.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../speech-recognition_b200/csrc/tc05.cuh"

using namespace srb::tc;

constexpr int M = 128, N = 224, K = 80;
constexpr int TM_D = 0, TM_AH = 256, TM_AL = 256 + K;

__global__ void __launch_bounds__(128, 1)
probe(const float *A, const float *B, float *Dss, float *Dts, long long *clk) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint32_t *Ah = reinterpret_cast<uint32_t *>(sm), *Al = Ah + M * K, *Bh = Al + M * K, *Bl = Bh + N * K;
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase_s;
    const int t = threadIdx.x, w = t >> 5;
    if (w == 0) tmem_alloc<512>(&tbase_s);
    if (t == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tb = tbase_s;
    const uint32_t lane_base = tb + ((uint32_t)(w * 32) << 16);
    // A: thread = row
    for (int k0 = 0; k0 < K; k0 += 8) {
        uint32_t h[8], l[8];
        for (int j = 0; j < 8; ++j) split_tf32(A[t * K + k0 + j], h[j], l[j]);
        for (int j = 0; j < 8; ++j) {
            const int k = k0 + j;
            Ah[(k >> 2) * M * 4 + t * 4 + (k & 3)] = h[j];
            Al[(k >> 2) * M * 4 + t * 4 + (k & 3)] = l[j];
        }
        tmem_st8(lane_base + TM_AH + k0, h);
        tmem_st8(lane_base + TM_AL + k0, l);
    }
    tmem_st_wait();
    for (int i = t; i < N * K; i += 128) {
        const int n = i / K, k = i % K;
        uint32_t h, l;
        split_tf32(B[i], h, l);
        Bh[(k >> 2) * N * 4 + n * 4 + (k & 3)] = h;
        Bl[(k >> 2) * N * 4 + n * 4 + (k & 3)] = l;
    }
    fence_async_smem();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(M, N);
    const uint32_t aH = smem_u32(Ah), aL = smem_u32(Al), bH = smem_u32(Bh), bL = smem_u32(Bl);
    uint32_t phase = 0;
    // ---- (1) SS
    if (t == 0) {
        for (int kk = 0; kk < K / 8; ++kk) {
            const uint64_t dah = smem_desc(aH + kk * 2 * M * 16, M * 16, 128), dal = smem_desc(aL + kk * 2 * M * 16, M * 16, 128);
            const uint64_t dbh = smem_desc(bH + kk * 2 * N * 16, N * 16, 128), dbl = smem_desc(bL + kk * 2 * N * 16, N * 16, 128);
            mma_tf32_ss(tb + TM_D, dal, dbh, idesc, kk > 0);
            mma_tf32_ss(tb + TM_D, dah, dbl, idesc, true);
            mma_tf32_ss(tb + TM_D, dah, dbh, idesc, true);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, phase);
    phase ^= 1;
    fence_after_sync();
    for (int c0 = 0; c0 < N; c0 += 16) {
        float v[16];
        tmem_ld16(lane_base + TM_D + c0, v);
        for (int j = 0; j < 16; ++j) Dss[t * N + c0 + j] = v[j];
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    // ---- (2) TS
    if (t == 0) {
        for (int kk = 0; kk < K / 8; ++kk) {
            const uint64_t dbh = smem_desc(bH + kk * 2 * N * 16, N * 16, 128), dbl = smem_desc(bL + kk * 2 * N * 16, N * 16, 128);
            mma_tf32_ts(tb + TM_D, tb + TM_AL + kk * 8, dbh, idesc, kk > 0);
            mma_tf32_ts(tb + TM_D, tb + TM_AH + kk * 8, dbl, idesc, true);
            mma_tf32_ts(tb + TM_D, tb + TM_AH + kk * 8, dbh, idesc, true);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, phase);
    phase ^= 1;
    fence_after_sync();
    for (int c0 = 0; c0 < N; c0 += 16) {
        float v[16];
        tmem_ld16(lane_base + TM_D + c0, v);
        for (int j = 0; j < 16; ++j) Dts[t * N + c0 + j] = v[j];
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    // ---- (4) cycles per MMA, 240 back to back.  mode: 0 SS N=224, 1 TS N=224, 2-7 TS at narrower N (B rows stay 224
    //      apart), 8-10 SS at N = 112 / 64 / 32, 11 TS N=32 alternating two accumulators, 12 SS N=32 alternating two
    for (int mode = 0; mode < 13; ++mode) {
        long long t0 = 0;
        const int nn_tab[13] = {N, N, 112, 64, 32, 160, 128, 16, 112, 64, 32, 32, 32};
        const int nn = nn_tab[mode];
        const bool ss = mode == 0 || (mode >= 8 && mode <= 10) || mode == 12;
        const bool two = mode >= 11;
        const uint32_t idn = idesc_tf32(M, nn);
        if (t == 0) {
            t0 = clock64();
            int flip = 0;
            for (int rep = 0; rep < 8; ++rep)
                for (int kk = 0; kk < K / 8; ++kk) {
                    const uint64_t dah = smem_desc(aH + kk * 2 * M * 16, M * 16, 128);
                    const uint64_t dbh = smem_desc(bH + kk * 2 * N * 16, N * 16, 128);
                    for (int j = 0; j < 3; ++j) {
                        const uint32_t dcol = tb + TM_D + (two ? 64 * (flip ^= 1) : 0);
                        if (ss) mma_tf32_ss(dcol, dah, dbh, idn, true);
                        else mma_tf32_ts(dcol, tb + TM_AH + kk * 8, dbh, idn, true);
                    }
                }
            mma_commit(&bar);
        }
        mbar_wait(&bar, phase);
        phase ^= 1;
        fence_after_sync();
        if (t == 0) clk[mode] = clock64() - t0;
        __syncthreads();
    }
    fence_before_sync();
    __syncthreads();
    if (w == 0) tmem_free<512>(tb);
}

static void run(const char *name, const std::vector<float> &A, const std::vector<float> &B) {
    float *dA, *dB, *dS, *dT;
    long long *dC;
    cudaMalloc(&dA, A.size() * 4);
    cudaMalloc(&dB, B.size() * 4);
    cudaMalloc(&dS, M * N * 4);
    cudaMalloc(&dT, M * N * 4);
    cudaMalloc(&dC, 128);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(dS, 0xff, M * N * 4);
    cudaMemset(dT, 0xff, M * N * 4);
    const size_t bytes = (size_t)(2 * M * K + 2 * N * K) * 4;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    probe<<<1, 128, bytes>>>(dA, dB, dS, dT, dC);
    cudaError_t e = cudaDeviceSynchronize();
    printf("[%s] kernel: %s\n", name, cudaGetErrorString(e));
    if (e != cudaSuccess) exit(1);
    std::vector<float> S(M * N), T(M * N);
    long long clk[16];
    cudaMemcpy(S.data(), dS, M * N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(T.data(), dT, M * N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(clk, dC, 104, cudaMemcpyDeviceToHost);
    double es = 0, et = 0, mag = 0, e32 = 0;
    int bad_s = 0, bad_t = 0;
    for (int r = 0; r < M; ++r)
        for (int n = 0; n < N; ++n) {
            double ref = 0, am = 0;
            float f32 = 0.f;
            for (int k = 0; k < K; ++k) {
                ref += (double)A[r * K + k] * B[n * K + k];
                am += fabs((double)A[r * K + k] * B[n * K + k]);
                f32 = fmaf(A[r * K + k], B[n * K + k], f32);
            }
            const double ds = fabs(S[r * N + n] - ref) / (am + 1e-30), dt = fabs(T[r * N + n] - ref) / (am + 1e-30);
            if (!(ds < 1e-5)) ++bad_s;
            if (!(dt < 1e-5)) ++bad_t;
            if (ds > es) es = ds;
            if (dt > et) et = dt;
            e32 = fmax(e32, fabs(f32 - ref) / (am + 1e-30));
            mag = fmax(mag, am);
        }
    printf("[%s] max |err| / sum|a*b|:  SS %.3e (bad %d)   TS %.3e (bad %d)   fp32 fmaf chain %.3e\n", name, es, bad_s, et, bad_t, e32);
    printf("[%s] D[5][7] ss=%g ts=%g ; 240 MMAs (M128 N224 K8): SS %lld clk (%.1f/MMA)  TS %lld clk (%.1f/MMA)\n", name,
           S[5 * N + 7], T[5 * N + 7], clk[0], clk[0] / 240.0, clk[1], clk[1] / 240.0);
    printf("[%s] TS clk/MMA at N=112: %.1f  N=64: %.1f  N=32: %.1f  N=160: %.1f  N=128: %.1f  N=16: %.1f\n", name, clk[2] / 240.0,
           clk[3] / 240.0, clk[4] / 240.0, clk[5] / 240.0, clk[6] / 240.0, clk[7] / 240.0);
    printf("[%s] SS clk/MMA at N=112: %.1f  N=64: %.1f  N=32: %.1f ; two accumulators alternating at N=32: TS %.1f  SS %.1f\n", name,
           clk[8] / 240.0, clk[9] / 240.0, clk[10] / 240.0, clk[11] / 240.0, clk[12] / 240.0);
    cudaFree(dA); cudaFree(dB); cudaFree(dS); cudaFree(dT); cudaFree(dC);
}

int main() {
    std::vector<float> A(M * K), B(N * K);
    for (int r = 0; r < M; ++r)
        for (int k = 0; k < K; ++k) A[r * K + k] = ((r * 3 + k * 5) % 17 - 8) * 0.25f;
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) B[n * K + k] = ((n * 7 + k * 3) % 13 - 6) * 0.5f;
    run("exact", A, B);
    srand(7);
    for (auto &v : A) v = (float)((rand() / (double)RAND_MAX - 0.5) * 20.0);
    for (auto &v : B) v = (float)((rand() / (double)RAND_MAX - 0.5) * 3.0);
    run("random", A, B);
    return 0;
}
