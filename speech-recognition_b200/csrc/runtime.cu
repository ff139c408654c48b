// runtime.cu -- error reporting, device queries, model packing, FP32 FMA peak probe.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace srb {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return 0;
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return SRB_ECUDA;
}

static int g_sm[64], g_smem[64];
static bool g_have[64];

static int query(int *sm, int *smem) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return SRB_ENODEVICE;
    if (!g_have[dev]) {
        if (cudaDeviceGetAttribute(&g_sm[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return SRB_ENODEVICE;
        if (cudaDeviceGetAttribute(&g_smem[dev], cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
            return SRB_ENODEVICE;
        g_have[dev] = true;
    }
    *sm = g_sm[dev];
    *smem = g_smem[dev];
    return 0;
}

int sm_count() {
    int a = 0, b = 0;
    return query(&a, &b) ? 0 : a;
}
int smem_optin() {
    int a = 0, b = 0;
    return query(&a, &b) ? 0 : b;
}

// ---- model packing (replaces the per-update inverse of hmm_state.py:13-17 and the per-call constants of :39-40)
__global__ void pack_model_kernel(const double *__restrict__ mu, const double *__restrict__ var,
                                  const double *__restrict__ w, const double *__restrict__ shift, int R, int M, int Mp,
                                  int D, float *__restrict__ ab, float *__restrict__ c, float *__restrict__ shift_f) {
    const int gp = blockIdx.x * blockDim.x + threadIdx.x;   // padded Gaussian index
    if (gp < D && shift_f) shift_f[gp] = (float)shift[gp];
    if (gp >= R * Mp) return;
    const int r = gp / Mp, m = gp % Mp;
    float2 *abp = reinterpret_cast<float2 *>(ab) + ((size_t)(gp >> 2) * D) * 4 + (gp & 3);
    if (m >= M) {
        for (int d = 0; d < D; ++d) abp[(size_t)d * 4] = make_float2(0.f, 0.f);
        c[gp] = -CUDART_INF_F;
        return;
    }
    const size_t g = (size_t)r * M + m;
    const double L2E = 1.4426950408889634074;
    double logdet = 0.0, q = 0.0;
    for (int d = 0; d < D; ++d) {
        const double v = var[g * D + d], mm = mu[g * D + d] - shift[d];
        logdet += log(v);
        q += mm * mm / v;
        abp[(size_t)d * 4] = make_float2((float)(-0.5 / v * L2E), (float)(mm / v * L2E));
    }
    const double lw = w[g] > 0.0 ? log(w[g]) : -CUDART_INF;
    c[gp] = (float)((lw - 0.5 * ((double)D * 1.8378770664093454836 + logdet) - 0.5 * q) * L2E);
}

// ---- FP32 FMA peak probe: 16 independent accumulator chains per thread, no memory traffic.
template <bool PACKED>
__global__ void __launch_bounds__(256) fma_peak_kernel(int iters, float *sink, float seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(seed + i, seed - i);
    const float2 m = make_float2(1.0000001f, 0.9999999f), b = make_float2(seed * 1e-7f, -seed * 1e-7f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (PACKED) {
                    a[i] = ffma2(a[i], m, b);
                } else {
                    a[i].x = fmaf(a[i].x, m.x, b.x);
                    a[i].y = fmaf(a[i].y, m.y, b.y);
                }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    if (s == 123.456f) sink[0] = s;   // never true: keeps the chains alive
}

}  // namespace srb

using namespace srb;

extern "C" {

int srb_version(void) { return 100; }

const char *srb_last_error_string(void) { return g_err; }

int srb_device_info(int *sm, int *smem, int *cc_major, int *cc_minor) {
    int a = 0, b = 0, dev = 0;
    if (query(&a, &b)) {
        set_error("no CUDA device");
        return SRB_ENODEVICE;
    }
    if (sm) *sm = a;
    if (smem) *smem = b;
    cudaGetDevice(&dev);
    if (cc_major) cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    if (cc_minor) cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev);
    return SRB_OK;
}

int64_t srb_packed_ab_floats(int R, int M, int D) {
    const int64_t Mp = (M + 3) / 4 * 4;
    return (int64_t)R * Mp * D * 2;
}
int64_t srb_packed_c_floats(int R, int M) { return (int64_t)R * ((M + 3) / 4 * 4); }

int srb_pack_model(const double *mu, const double *var, const double *w, const double *shift, int R, int M, int D,
                   float *ab, float *c, float *shift_f32, void *stream) {
    SRB_REQUIRE(mu && var && w && shift && ab && c, SRB_EINVAL, "srb_pack_model: null pointer");
    SRB_REQUIRE(R > 0 && M > 0 && D > 0 && D <= SRB_MAX_DIM, SRB_EINVAL, "srb_pack_model: bad R/M/D (%d,%d,%d)", R, M, D);
    const int Mp = (M + 3) / 4 * 4;
    int n = R * Mp;
    if (n < D) n = D;
    pack_model_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(mu, var, w, shift, R, M, Mp, D, ab, c, shift_f32);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_fma_peak(int packed, int iters, float *ms_out, double *fma_count_out, void *stream) {
    SRB_REQUIRE(ms_out && fma_count_out && iters > 0, SRB_EINVAL, "srb_fma_peak: bad argument");
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    cudaStream_t st = (cudaStream_t)stream;
    float *sink = nullptr;
    SRB_CUDA(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    SRB_CUDA(cudaEventCreate(&e0));
    SRB_CUDA(cudaEventCreate(&e1));
    const int grid = sms * 8;   // 8 CTAs of 256 threads per SM = 64 warps/SM
    for (int rep = 0; rep < 2; ++rep) {   // first pass warms up
        SRB_CUDA(cudaEventRecord(e0, st));
        if (packed) fma_peak_kernel<true><<<grid, 256, 0, st>>>(iters, sink, 1.5f);
        else fma_peak_kernel<false><<<grid, 256, 0, st>>>(iters, sink, 1.5f);
        SRB_CUDA(cudaEventRecord(e1, st));
        SRB_CUDA(cudaEventSynchronize(e1));
    }
    SRB_CUDA(cudaGetLastError());
    SRB_CUDA(cudaEventElapsedTime(ms_out, e0, e1));
    *fma_count_out = (double)grid * 256.0 * (double)iters * 8.0 * 8.0 * 2.0;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    return SRB_OK;
}

}  // extern "C"
