// mfcc.cu -- MFCC front-end, the step right before the scoring path.
// Replaces sr/feature/feature.py:7-88 (segment, zero_padding, mfcc_features, standardize) and core.py:13-31
// (delta_feature, load_wav_as_mfcc1) for a batch of utterances:
//   srb_mfcc               pre-emphasis 0.97 -> 400-sample frames every 160 (zero padded past the end) -> centred in 512
//                          -> Hamming(512) -> |rFFT-512|^2 / 512 -> 40 triangular mel filters (host-built, the reference's
//                          floor()ed bin edges) -> eps guard -> log10 -> DCT-II ortho, coefficients 1..13
//   srb_delta_standardize  [mfcc, delta, delta-delta] (simple differences, one-sided at the ends) -> per-utterance
//                          (x - mean) / std per dimension (population std, like np.std)
// One warp per frame: the 512-point FFT lives in the warp's 4 KB of shared memory (radix-2, bit-reversed load, 9
// stages of 8 butterflies per lane); twiddles, window and the DCT table are built per CTA with sincospi/cospi.
#include "common.cuh"

namespace srb {

constexpr int MF_NFFT = 512, MF_FLEN = 400, MF_FSTEP = 160, MF_NFILT = 40, MF_NCEP = 13, MF_NBIN = MF_NFFT / 2 + 1;
constexpr int MF_WARPS = 8;
constexpr float MF_PREEMPH = 0.97f;

__global__ void __launch_bounds__(MF_WARPS * 32)
mfcc_kernel(const float *__restrict__ signal, const int64_t *__restrict__ sig_off, const int64_t *__restrict__ frame_off,
            int B, int64_t F, const float *__restrict__ fb_w, const int32_t *__restrict__ fb_range,
            float *__restrict__ fbank_out, float *__restrict__ mfcc_out) {
    __shared__ float2 tw[MF_NFFT];                    // per stage st, contiguous: tw[2^st - 1 + k] = e^{-2 pi i k / 2^(st+1)}
    __shared__ float win[MF_NFFT];                    // np.hamming(512)
    __shared__ float dct[MF_NCEP][MF_NFILT];          // sqrt(2/40) cos(pi (2m+1) c / 80), c = 1..13
    __shared__ float2 buf[MF_WARPS][MF_NFFT + MF_NFFT / 16];   // index i lives at i + i/16 (conflict-free bit reversal)
    __shared__ float fbl[MF_WARPS][MF_NFILT];
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    for (int i = tid; i < MF_NFFT - 1; i += blockDim.x) {
        const int st = 31 - __clz(i + 1), k = i + 1 - (1 << st);
        float s, c;
        sincospif(-(float)k / (float)(1 << st), &s, &c);
        tw[i] = make_float2(c, s);
    }
    for (int n = tid; n < MF_NFFT; n += blockDim.x) win[n] = 0.54f - 0.46f * cospif(2.f * (float)n / (MF_NFFT - 1));
    for (int i = tid; i < MF_NCEP * MF_NFILT; i += blockDim.x) {
        const int c = i / MF_NFILT + 1, m = i % MF_NFILT;
        dct[c - 1][m] = sqrtf(2.f / MF_NFILT) * cospif((float)((2 * m + 1) * c) / (2.f * MF_NFILT));
    }
    __syncthreads();
    float2 *x = buf[wq];
    constexpr int PADL = (MF_NFFT - MF_FLEN) / 2;      // feature.py:34-40: 56 zeros left, 56 right

    for (int64_t f = (int64_t)blockIdx.x * MF_WARPS + wq; f < F; f += (int64_t)gridDim.x * MF_WARPS) {
        // utterance of frame f (warp uniform)
        int lo = 0, hi = B;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (frame_off[mid] <= f) lo = mid; else hi = mid;
        }
        const int64_t s0 = sig_off[lo], slen = sig_off[lo + 1] - s0;
        const int64_t first = (f - frame_off[lo]) * MF_FSTEP;
        // pre-emphasis + framing + centring + window, written in bit-reversed order
        for (int n = lane; n < MF_NFFT; n += 32) {
            float v = 0.f;
            const int j = n - PADL;
            if (j >= 0 && j < MF_FLEN) {
                const int64_t t = first + j;
                if (t < slen) {
                    const float cur = signal[s0 + t];
                    v = t > 0 ? cur - MF_PREEMPH * signal[s0 + t - 1] : cur;       // feature.py:46-47
                }
                v *= win[n];
            }
            const int r = (int)(__brev((unsigned)n) >> (32 - 9));
            x[r + (r >> 4)] = make_float2(v, 0.f);
        }
        __syncwarp();
#pragma unroll 1
        for (int st = 0; st < 9; ++st) {
            const int half = 1 << st;
#pragma unroll
            for (int q = 0; q < MF_NFFT / 2 / 32; ++q) {
                const int b = lane + 32 * q;           // butterfly index 0..255
                const int k = b & (half - 1);
                const int j0 = ((b >> st) << (st + 1)) | k, j1 = j0 + half;
                const int i0 = j0 + (j0 >> 4), i1 = j1 + (j1 >> 4);
                const float2 w = tw[half - 1 + k];
                const float2 a = x[i0], c = x[i1];
                const float2 t = make_float2(c.x * w.x - c.y * w.y, c.x * w.y + c.y * w.x);
                x[i0] = make_float2(a.x + t.x, a.y + t.y);
                x[i1] = make_float2(a.x - t.x, a.y - t.y);
            }
            __syncwarp();
        }
        // power spectrum in place (x[k].x), k = 0..256
        for (int k = lane; k < MF_NBIN; k += 32) {
            const float2 v = x[k + (k >> 4)];
            x[k + (k >> 4)].x = (v.x * v.x + v.y * v.y) * (1.f / MF_NFFT);
        }
        __syncwarp();
        // mel filter bank (feature.py:57-76) + log10
        for (int m = lane; m < MF_NFILT; m += 32) {
            const int k0 = fb_range[2 * m], k1 = fb_range[2 * m + 1];
            float e = 0.f;
            for (int k = k0; k < k1; ++k) e = fmaf(x[k + (k >> 4)].x, fb_w[m * MF_NBIN + k], e);
            if (e == 0.f) e = 2.220446049250313e-16f;                              // np.finfo(float).eps
            const float l = log10f(e);
            fbl[wq][m] = l;
            if (fbank_out) fbank_out[f * MF_NFILT + m] = l;
        }
        __syncwarp();
        if (lane < MF_NCEP) {
            float acc = 0.f;
#pragma unroll 8
            for (int m = 0; m < MF_NFILT; ++m) acc = fmaf(fbl[wq][m], dct[lane][m], acc);
            mfcc_out[f * MF_NCEP + lane] = acc;
        }
        __syncwarp();
    }
}

// One CTA per utterance.  Pass 1 writes [c, delta c, delta delta c] and accumulates the per-dimension mean; pass 2
// the variance about that mean; pass 3 standardises in place.
__global__ void __launch_bounds__(256)
delta_standardize_kernel(const float *__restrict__ mfcc, const int64_t *__restrict__ frame_off, int B,
                         float *__restrict__ out, int standardize) {
    constexpr int C = MF_NCEP, D = 3 * MF_NCEP;
    __shared__ double acc[D];
    __shared__ float mean_s[D], istd_s[D];
    const int u = blockIdx.x;
    if (u >= B) return;
    const int64_t f0 = frame_off[u];
    const int T = (int)(frame_off[u + 1] - f0);
    const float *c = mfcc + f0 * C;
    float *o = out + f0 * D;
    auto cep = [&](int t, int k) { return c[(int64_t)t * C + k]; };
    auto d1 = [&](int t, int k) {                       // core.py:13-22
        if (t == 0) return cep(1, k) - cep(0, k);
        if (t == T - 1) return cep(T - 1, k) - cep(T - 2, k);
        return cep(t + 1, k) - cep(t - 1, k);
    };
    auto d2 = [&](int t, int k) {
        if (t == 0) return d1(1, k) - d1(0, k);
        if (t == T - 1) return d1(T - 1, k) - d1(T - 2, k);
        return d1(t + 1, k) - d1(t - 1, k);
    };
    for (int i = threadIdx.x; i < T * C; i += blockDim.x) {
        const int t = i / C, k = i % C;
        o[(int64_t)t * D + k] = cep(t, k);
        o[(int64_t)t * D + C + k] = d1(t, k);
        o[(int64_t)t * D + 2 * C + k] = d2(t, k);
    }
    if (!standardize) return;
    __syncthreads();
    // per-dimension moments: thread group of 6 rows x 39 dims, double accumulation through shared atomics per pass
    for (int pass = 0; pass < 2; ++pass) {
        if (threadIdx.x < D) acc[threadIdx.x] = 0.0;
        __syncthreads();
        const int d = threadIdx.x % D, r0 = threadIdx.x / D, rows = blockDim.x / D;   // 256 / 39 = 6 row groups
        if (r0 < rows) {
            double s = 0.0;
            for (int t = r0; t < T; t += rows) {
                const float v = o[(int64_t)t * D + d];
                if (pass == 0) s += v;
                else s += (double)(v - mean_s[d]) * (double)(v - mean_s[d]);
            }
            atomicAdd(&acc[d], s);
        }
        __syncthreads();
        if (threadIdx.x < D) {
            if (pass == 0) mean_s[threadIdx.x] = (float)(acc[threadIdx.x] / T);
            else istd_s[threadIdx.x] = (float)(1.0 / sqrt(acc[threadIdx.x] / T));   // feature.py:85-88, np.std (ddof 0)
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < T * D; i += blockDim.x) {
        const int d = i % D;
        o[i] = (o[i] - mean_s[d]) * istd_s[d];
    }
}

}  // namespace srb

using namespace srb;

extern "C" {

int64_t srb_mfcc_num_frames(int64_t nsamples, int sample_rate) {
    // feature.py:7-16: `slen > frame_len` compares samples with SECONDS, so it is true for any non-empty signal
    if (nsamples <= 0 || sample_rate != 16000) return 0;
    return (nsamples + MF_FSTEP - 1) / MF_FSTEP;
}

int srb_mfcc(const float *signal, const int64_t *sig_off, const int64_t *frame_off, int32_t B, int64_t F, int sample_rate,
             const float *fb_w, const int32_t *fb_range, float *fbank, float *mfcc, void *stream) {
    SRB_REQUIRE(sample_rate == 16000, SRB_EUNSUPPORTED,
                "srb_mfcc: built for 16 kHz (400-sample frames every 160, NFFT 512), got %d", sample_rate);
    SRB_REQUIRE(B >= 0 && F >= 0, SRB_EINVAL, "srb_mfcc: bad sizes");
    if (B == 0 || F == 0) return SRB_OK;
    SRB_REQUIRE(signal && sig_off && frame_off && fb_w && fb_range && mfcc, SRB_EINVAL, "srb_mfcc: null pointer");
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int64_t want = (F + MF_WARPS - 1) / MF_WARPS;
    const int grid = (int)(want < (int64_t)sms * 4 ? want : (int64_t)sms * 4);
    mfcc_kernel<<<grid, MF_WARPS * 32, 0, (cudaStream_t)stream>>>(signal, sig_off, frame_off, B, F, fb_w, fb_range, fbank, mfcc);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_delta_standardize(const float *mfcc, const int64_t *frame_off, int32_t B, float *features, int standardize,
                          void *stream) {
    SRB_REQUIRE(B >= 0, SRB_EINVAL, "srb_delta_standardize: bad sizes");
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(mfcc && frame_off && features, SRB_EINVAL, "srb_delta_standardize: null pointer");
    delta_standardize_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(mfcc, frame_off, B, features, standardize);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
