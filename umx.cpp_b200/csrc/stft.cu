// STFT / iSTFT kernels (reference: src/dsp.cpp:103-258, src/dsp.hpp:21-103).
//
// One CTA of 256 threads transforms one frame.  Both stereo channels ride in ONE complex 4096-point
// FFT (z = L + iR); see fft4096.cuh.  Fused into the forward kernel: segment gather from the shifted
// track (umx.cpp:120-122,217-218), symmetric edge padding (dsp.cpp:109-128), Hann window, magnitude,
// the 1487-bin crop + channel stacking of the network input as fp16 hi/lo planes
// (inference.cpp:29,58-68) and the per-track max|X| the Wiener stage needs (wiener.cpp:36-52).
// Fused into the inverse kernel: Hermitian merge of the two channels, window, 1/4096,
// 1/(window-sum-square + 1e-8), overlap-add in shared memory (dsp.cpp:231-258) and either the plain
// chunk store or the triangular-weighted accumulation into the track (umx.cpp:234-249).
#include "common.cuh"
#include "fft4096.cuh"
#include "kernels.cuh"

#include <cmath>
#include <mutex>
#include <vector>

namespace umx
{

__device__ float2 g_tw[4096];  // exp(-2*pi*i*j/4096)
__device__ float g_win[4096];  // periodic Hann, src/dsp.hpp:61-77
__device__ float g_istft_scale[4096]; // win[n] / 4096 / (wss + 1e-8) of a frame fully inside the signal (dsp.cpp:247-256)

// ------------------------------------------------------------------------------------------------
// tables (once per device)
// ------------------------------------------------------------------------------------------------
int init_fft_tables(int device)
{
    static std::mutex mu;
    static bool done[64] = {};
    std::lock_guard<std::mutex> lk(mu);
    if (device < 0 || device >= 64)
    {
        set_error("init_fft_tables: bad device");
        return -2;
    }
    if (done[device])
        return 0;
    std::vector<float2> tw(4096);
    std::vector<float> win(4096);
    static constexpr float PI = 3.14159265359F; // src/dsp.hpp:63
    const float floatN = (float)(kNfft + 1);
    for (int j = 0; j < 4096; ++j)
    {
        const double a = 2.0 * M_PI * (double)j / 4096.0;
        tw[j] = make_float2((float)std::cos(a), (float)-std::sin(a));
        // same expression, same float steps and the same libm cosf as src/dsp.hpp:71-75
        win[j] = 0.5F * (1.0F - cosf(2.0F * PI * (float)j / (floatN - 1)));
    }
    UMX_CUDA_OK(cudaMemcpyToSymbol(g_tw, tw.data(), sizeof(float2) * 4096));
    UMX_CUDA_OK(cudaMemcpyToSymbol(g_win, win.data(), sizeof(float) * 4096));
    // interior frames: every output position is covered by 4 frames, added in increasing frame order, i.e.
    // window offsets nl + 3072, + 2048, + 1024, + 0 (init_const_normalized_window, src/dsp.hpp:79-102)
    std::vector<float> scale(4096);
    for (int n = 0; n < 4096; ++n)
    {
        const int nl = n & 1023;
        float wss = 0.f;
        for (int j = 3; j >= 0; --j)
            wss += win[nl + 1024 * j] * win[nl + 1024 * j];
        scale[n] = win[n] * (1.0f / (float)kNfft) / (wss + 1e-8f);
    }
    UMX_CUDA_OK(cudaMemcpyToSymbol(g_istft_scale, scale.data(), sizeof(float) * 4096));
    done[device] = true;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) stft_kernel(const StftTrack *__restrict__ tracks, int n_buf, int F,
                                                   float2 *__restrict__ spec, __half *__restrict__ xh,
                                                   __half *__restrict__ xl, float *__restrict__ maxabs, int skip_zero)
{
    __shared__ float2 s[kFftSmemFloat2];
    __shared__ float s_red[8];
    const int t = threadIdx.x;
    const int f = blockIdx.x;
    const int trk = blockIdx.y;
    const StftTrack tk = tracks[trk];
    // frames beyond the chunk are all zero (the window is the zero-extended chunk); the first of them is computed and stands
    // for the rest: nothing downstream reads the others (see frames_reaching and LstmLayerArgs::chunks)
    if (skip_zero && f > frames_reaching(tk.n_valid, F))
        return;

    float2 r[16];
#pragma unroll
    for (int a = 0; a < 16; ++a)
    {
        const int n = 256 * a + t;
        const int smp = padded_to_sample(kHop * f + n, n_buf);
        float2 v = make_float2(0.f, 0.f);
        if (smp < tk.n_valid)
        {
            const long o = (long)smp + tk.src_shift;
            if (o >= 0 && o < tk.n_src)
                v = __ldg(&tk.audio[o]);
        }
        const float w = g_win[n];
        r[a] = make_float2(v.x * w, v.y * w);
    }
    fft_pass1<false>(t, r, s, g_tw);
    __syncthreads();
    fft_pass2_load(t, r, s);
    __syncthreads();
    fft_pass2_store<false>(t, r, s, g_tw);
    __syncthreads();
    fft_pass3<false>(t, r, s);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k)
        s[t + 256 * k] = r[k];
    __syncthreads();

    // separate the two real spectra: XL = (Z[b] + conj Z[-b])/2, XR = (Z[b] - conj Z[-b])/(2i)
    const long row = (long)trk * F + f;
    float2 *specL = spec + (row * 2 + 0) * kBins;
    float2 *specR = spec + (row * 2 + 1) * kBins;
    float mx = 0.f;
    auto emit = [&](int b, float2 zb) {
        const float2 zm = s[(4096 - b) & 4095];
        const float2 xL = make_float2(0.5f * (zb.x + zm.x), 0.5f * (zb.y - zm.y));
        const float2 xR = make_float2(0.5f * (zb.y + zm.y), 0.5f * (zm.x - zb.x));
        specL[b] = xL;
        specR[b] = xR;
        const float mL = sqrtf(xL.x * xL.x + xL.y * xL.y);
        const float mR = sqrtf(xR.x * xR.x + xR.y * xR.y);
        mx = fmaxf(mx, fmaxf(mL, mR));
        if (xh != nullptr && b < kBinsCropped)
        {
            __half h, l;
            split_half(mL, h, l);
            xh[row * kFc1InPad + b] = h;
            xl[row * kFc1InPad + b] = l;
            split_half(mR, h, l);
            xh[row * kFc1InPad + kBinsCropped + b] = h;
            xl[row * kFc1InPad + kBinsCropped + b] = l;
        }
    };
#pragma unroll
    for (int j = 0; j < 8; ++j)
        emit(t + 256 * j, r[j]);
    if (t == 0)
        emit(2048, r[8]);

    if (maxabs != nullptr)
    {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((t & 31) == 0)
            s_red[t >> 5] = mx;
        __syncthreads();
        if (t == 0)
        {
            float m = s_red[0];
#pragma unroll
            for (int i = 1; i < 8; ++i)
                m = fmaxf(m, s_red[i]);
            atomicMax(reinterpret_cast<int *>(maxabs) + trk, __float_as_int(m)); // m >= 0
        }
    }
}

// ------------------------------------------------------------------------------------------------
// inverse
// ------------------------------------------------------------------------------------------------
// window-sum-square at padded position p = 1024*f + n, frames added in increasing order
// (init_const_normalized_window, src/dsp.hpp:79-102)
__device__ __forceinline__ float wss_at(int f, int n, int F, const float *win)
{
    float acc = 0.f;
    const int fb = f + (n >> 10);
    const int nl = n & 1023;
#pragma unroll
    for (int j = 3; j >= 0; --j)
    {
        const int fp = fb - j;
        if (fp >= 0 && fp < F)
        {
            const float w = __ldg(&win[nl + 1024 * j]);
            acc += w * w;
        }
    }
    return acc;
}

__device__ __forceinline__ float tri_weight(int k, int n_buf)
{
    // umx.cpp:197-206: weight(i) = weight(seg-1-i) = i+1, divided by the maximum (seg/2), pow(.,1)
    const int tri = min(k + 1, n_buf - k);
    return (float)tri / (float)(n_buf / 2);
}

// cos / sin of k pi / 8
__device__ constexpr float kCos8[16] = {1.0f, 0.923879533f, 0.707106781f, 0.382683432f, 0.0f, -0.382683432f, -0.707106781f, -0.923879533f, -1.0f, -0.923879533f, -0.707106781f, -0.382683432f, 0.0f, 0.382683432f, 0.707106781f, 0.923879533f};
__device__ constexpr float kSin8[16] = {0.0f, 0.382683432f, 0.707106781f, 0.923879533f, 1.0f, 0.923879533f, 0.707106781f, 0.382683432f, 0.0f, -0.382683432f, -0.707106781f, -0.923879533f, -1.0f, -0.923879533f, -0.707106781f, -0.382683432f};

__global__ void __launch_bounds__(256, 3) istft_kernel(const float2 *__restrict__ stems, long stem_stride, int F,
                                                    int frames_per_cta, const IstftTrack *__restrict__ tracks,
                                                    int n_buf, int stem_begin)
{
    // 66 KB of SMEM and <= 85 registers: 3 CTAs per SM.  (A cp.async-staged variant that prefetched the next
    // frame at 2 CTAs per SM measured slower: the kernel is bound by FFT instruction issue, not by load latency.)
    extern __shared__ float2 sm[];
    float2 *s = sm;                     // FFT workspace
    float2 *ring = sm + kFftSmemFloat2; // overlap-add accumulators (L,R), 4096 positions

    const int t = threadIdx.x;
    const int stem = stem_begin + blockIdx.y;
    const int trk = blockIdx.z;
    const IstftTrack tk = tracks[trk];
    const int f0 = blockIdx.x * frames_per_cta;
    // frames that reach an emitted sample: sample i sits at padded position i + 2048, frame f covers [1024 f, 1024 f + 4096).
    // The reference always transforms the whole 60 s window and drops what lies beyond the chunk (src/dsp.cpp:203-205); a
    // short last chunk (or a track with no chunk in this segment) needs only its first few frames.
    const int f_need = frames_reaching(tk.n_wave, F);
    if (f0 >= f_need)
        return;
    const int f1 = min(f_need, f0 + frames_per_cta);
    const bool last_cta = (f1 == F);
    const float2 *base = stems + (long)stem * stem_stride;
    const int f_first = max(f0 - 3, 0);

    for (int i = t; i < 4096; i += 256)
        ring[i] = make_float2(0.f, 0.f);

    // interior frames (every sample covered by 4 frames): the window-sum-square of the periodic Hann at hop N/4 is 1.5, so
    // the scale win[n] / 4096 / (wss + 1e-8) is KS * (1 - cos(theta_t + k pi/8)), n = t + 256 k: two FMAs on this
    // thread's (cos, sin)(theta_t) instead of a table load per sample (agrees with the table to a few ulp)
    const float2 tw_t = g_tw[t]; // (cos, -sin)(2 pi t / 4096)
    constexpr float KS = 0.5f / (4096.0f * (1.5f + 1e-8f));

    float2 r[16];
    for (int f = f_first; f < f1; ++f)
    {
        const float2 *yL = base + ((long)trk * F + f) * 2 * kBins;
        const float2 *yR = yL + kBins;
        if (f + 1 < f1)
        { // the next frame's 257 lines on their way into L2 while this one is transformed
            const char *nx = reinterpret_cast<const char *>(yL + 2 * kBins);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 128 * t));
            if (t == 0)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + 128 * 256));
        }
#pragma unroll
        for (int a = 0; a < 16; ++a)
        {
            const int k = 256 * a + t;
            const int kk = k <= 2048 ? k : 4096 - k;
            float2 L = __ldg(&yL[kk]), R = __ldg(&yR[kk]);
            if (k > 2048)
            { // Hermitian extension
                L.y = -L.y;
                R.y = -R.y;
            }
            else if (k == 0 || k == 2048)
            { // the half-spectrum inverse ignores the imaginary part of DC and Nyquist
                L.y = 0.f;
                R.y = 0.f;
            }
            r[a] = make_float2(L.x - R.y, L.y + R.x); // Z = YL + i*YR
        }
        __syncthreads(); // the previous frame's emission (ring reads) and FFT workspace reads are finished
        fft_pass1<true>(t, r, s, g_tw);
        __syncthreads();
        fft_pass2_load(t, r, s);
        __syncthreads();
        fft_pass2_store<true>(t, r, s, g_tw);
        __syncthreads();
        fft_pass3<true>(t, r, s);
        // r[k] = (yL, yR)[n], n = t + 256k ; window, 1/4096, 1/(wss+1e-8), overlap-add
        const bool interior = (f >= 3 && f + 3 < F); // every sample of the frame is covered by 4 frames
#pragma unroll
        for (int k = 0; k < 16; ++k)
        {
            const int n = t + 256 * k;
            float sc;
            if (interior)
            {
                sc = fmaf(tw_t.x, -KS * kCos8[k], fmaf(tw_t.y, -KS * kSin8[k], KS)); // KS (1 - (cos t cos k - sin t sin k)), tw_t.y = -sin
            }
            else
                sc = __ldg(&g_win[n]) * (1.0f / (float)kNfft) / (wss_at(f, n, F, g_win) + 1e-8f);
            const float cL = r[k].x * sc;
            const float cR = r[k].y * sc;
            const int idx = (kHop * f + n) & 4095;
            if (n >= 3072)
                ring[idx] = make_float2(cL, cR);
            else
            {
                float2 o = ring[idx];
                ring[idx] = make_float2(o.x + cL, o.y + cR);
            }
        }
        __syncthreads();
        if (f >= f0)
        {
            const int hops = (last_cta && f == F - 1) ? 4 : 1; // the final frame also completes hops F..F+2
            for (int h = 0; h < hops; ++h)
            {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    const int p = kHop * (f + h) + t + 256 * j;
                    const int i = p - kPad;
                    if (i >= 0 && i < tk.n_wave)
                    {
                        const float2 v = ring[p & 4095];
                        if (tk.accumulate)
                        {
                            float2 *dst = tk.out + (long)stem * tk.out_stem_stride + tk.out_offset + i;
                            const float wgt = tri_weight(i, n_buf);
                            float2 o = *dst;
                            *dst = make_float2(o.x + wgt * v.x, o.y + wgt * v.y);
                        }
                        else
                            tk.out[(long)stem * tk.out_stem_stride + tk.out_offset + i] = v;
                    }
                }
            }
        }
        // (the __syncthreads at the top of the next iteration orders this emission before the next overlap-add)
    }
}

// ------------------------------------------------------------------------------------------------
// out = acc / sum_weight, trimmed to the original track (umx.cpp:251-273,136-147); output samples
// [i_begin, i_end) of each track (a range is final as soon as the last segment touching it has been added)
// ------------------------------------------------------------------------------------------------
__global__ void finalize_kernel(const FinalizeTrack *__restrict__ tracks, int shift, int n_buf, int stride, int n_stems,
                                int stem_begin)
{
    const FinalizeTrack tk = tracks[blockIdx.y];
    const long i = tk.i_begin + (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tk.i_end)
        return;
    const long p = i + shift;
    // segments s with s*stride <= p < s*stride + chunk_len(s), added in increasing s
    float sw = 0.f;
    // one 64-bit division per block (uniform), not per thread: the block's 256 positions span at most two segments' starts
    const long p0 = tk.i_begin + (long)blockIdx.x * blockDim.x + shift;
    const long s0 = p0 / stride;
    const long rem = p - s0 * stride;
    const long s_hi = rem < stride ? s0 : (rem < 2L * stride ? s0 + 1 : p / stride);
    for (long sgm = (s_hi > 0 ? s_hi - 1 : 0); sgm <= s_hi; ++sgm)
    {
        const long off = sgm * stride;
        if (off >= tk.L)
            continue;
        const long clen = min((long)n_buf, tk.L - off);
        const long k = p - off;
        if (k >= 0 && k < clen)
            sw += tri_weight((int)k, n_buf);
    }
    for (int st = stem_begin; st < stem_begin + n_stems; ++st)
    {
        const float2 a = tk.acc[(long)st * tk.L + p];
        tk.out[(long)st * tk.n + i] = make_float2(a.x / sw, a.y / sw);
    }
}

// ------------------------------------------------------------------------------------------------
// layout conversion between the internal [frame][channel][bin] order and the reference's Eigen
// column-major Tensor3d(2, F, 2049) == [bin][frame][channel] (only used by the kernel-level API)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void to_ref_layout_kernel(const T *__restrict__ in, T *__restrict__ out, int F)
{
    __shared__ T tile[32][33];
    const int c = blockIdx.z;
    const int b0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y)
    {
        const int f = f0 + j, b = b0 + threadIdx.x;
        if (f < F && b < kBins)
            tile[j][threadIdx.x] = in[((long)f * 2 + c) * kBins + b];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y)
    {
        const int b = b0 + j, f = f0 + threadIdx.x;
        if (f < F && b < kBins)
            out[((long)b * F + f) * 2 + c] = tile[threadIdx.x][j];
    }
}

template <typename T>
__global__ void from_ref_layout_kernel(const T *__restrict__ in, T *__restrict__ out, int F)
{
    __shared__ T tile[32][33];
    const int c = blockIdx.z;
    const int b0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y)
    {
        const int b = b0 + j, f = f0 + threadIdx.x;
        if (f < F && b < kBins)
            tile[j][threadIdx.x] = in[((long)b * F + f) * 2 + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y)
    {
        const int f = f0 + j, b = b0 + threadIdx.x;
        if (f < F && b < kBins)
            out[((long)f * 2 + c) * kBins + b] = tile[threadIdx.x][j];
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
int launch_stft(const StftTrack *d_tracks, int n_tracks, int n_buf, float2 *spec, __half *xh, __half *xl,
                float *maxabs, cudaStream_t st, int skip_zero)
{
    const int F = n_buf / kHop + 1;
    dim3 grid(F, n_tracks);
    stft_kernel<<<grid, 256, 0, st>>>(d_tracks, n_buf, F, spec, xh, xl, maxabs, skip_zero);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_istft(const float2 *stems, long stem_stride, int n_stems, const IstftTrack *d_tracks, int n_tracks,
                 int n_buf, cudaStream_t st, int stem_begin)
{
    const int F = n_buf / kHop + 1;
    const int G = kIstftFramesPerCta;
    const size_t smem = sizeof(float2) * (kFftSmemFloat2 + 4096);
    static PerDeviceOnce once;
    if (int rc = once.run([&](int) {
            UMX_CUDA_OK(cudaFuncSetAttribute(istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            return 0;
        }))
        return rc;
    dim3 grid(ceil_div(F, G), n_stems, n_tracks);
    istft_kernel<<<grid, 256, smem, st>>>(stems, stem_stride, F, G, d_tracks, n_buf, stem_begin);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_finalize(const FinalizeTrack *d_tracks, int n_tracks, long max_range, int shift, int n_buf, int stride, int n_stems,
                    cudaStream_t st, int stem_begin)
{
    if (max_range <= 0 || n_tracks <= 0)
        return 0;
    finalize_kernel<<<dim3(ceil_div(max_range, 256), n_tracks), 256, 0, st>>>(d_tracks, shift, n_buf, stride, n_stems, stem_begin);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_to_ref_layout_c64(const float2 *in, float2 *out, int F, cudaStream_t st)
{
    dim3 grid(ceil_div(kBins, 32), ceil_div(F, 32), 2), block(32, 8);
    to_ref_layout_kernel<float2><<<grid, block, 0, st>>>(in, out, F);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}
int launch_from_ref_layout_c64(const float2 *in, float2 *out, int F, cudaStream_t st)
{
    dim3 grid(ceil_div(kBins, 32), ceil_div(F, 32), 2), block(32, 8);
    from_ref_layout_kernel<float2><<<grid, block, 0, st>>>(in, out, F);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}
int launch_to_ref_layout_f32(const float *in, float *out, int F, cudaStream_t st)
{
    dim3 grid(ceil_div(kBins, 32), ceil_div(F, 32), 2), block(32, 8);
    to_ref_layout_kernel<float><<<grid, block, 0, st>>>(in, out, F);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}
int launch_from_ref_layout_f32(const float *in, float *out, int F, cudaStream_t st)
{
    dim3 grid(ceil_div(kBins, 32), ceil_div(F, 32), 2), block(32, 8);
    from_ref_layout_kernel<float><<<grid, block, 0, st>>>(in, out, F);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx
