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
#include "wiener_math.cuh"

#include <cmath>
#include <cstdlib>
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
                                                   __half *__restrict__ xl, float *__restrict__ maxabs, int skip_zero,
                                                   int reflect_pad, int win_len, int stat_windows)
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
        const int smp = reflect_pad ? padded_to_sample_reflect(kHop * f + n, n_buf) : padded_to_sample(kHop * f + n, n_buf);
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
            atomicMax(reinterpret_cast<int *>(maxabs) + trk * stat_windows + (win_len > 0 ? f / win_len : 0), __float_as_int(m)); // m >= 0
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
// Wiener filter pass + inverse STFT + overlap-add in one kernel (reference: src/wiener.cpp:276-422 feeding src/dsp.cpp:178-258,
// then umx.cpp:234-249).  The four filtered stem spectrograms are never written to memory: a CTA takes a run of frames of one
// track and, frame by frame,
//   A. all 512 threads, thread per bin: mixture (L, R) and the 4 x 2 masks from HBM, the four sources' spatial covariances of
//      the bin from L2, the PSDs, Cxx^-1 — ONCE for the four stems — then every stem's y_j = G_j x, merged as Z = Y_L + i Y_R
//      with its Hermitian extension straight into that stem's FFT workspace in shared memory;
//   B. each half of the CTA (256 threads) runs the complex inverse FFT (fft4096.cuh) of two stems, one after the other;
//   C. window, 1/4096, 1/(window-sum-square + 1e-8) and the overlap-add, which needs no shared memory: sample n = t + 256 k of a
//      frame belongs to thread t for every frame, so the three partial hops a stem carries from frame to frame live in that
//      thread's registers (added in the reference's frame order); complete hops go to the chunk / track accumulator.
// DRAM traffic per (frame, bin): 48 bytes in (16 mixture + 32 masks) instead of 48 in + 64 out + 64 in of the two-kernel form.
// ------------------------------------------------------------------------------------------------
constexpr int kWiThreads = 512;
constexpr int kWiFrames = 32; // frames per CTA (+ 3 halo frames whose hops the previous CTA emits)
// shared memory: four FFT workspaces, then the staged inputs of one frame (mixture 2 x 2049 complex, 4 x (2 x 2049) masks)
constexpr int kWiWsBytes = 4 * kFftSmemFloat2 * 8;            // 131,584
constexpr int kWiSpecBytes = 2 * kBins * 8;                   // 32,784 (multiple of 16)
constexpr int kWiMaskBytes = ((kFc3Out * 4 + 15) / 16) * 16;  // 16,400 per target
constexpr int kWiSmemBytes = kWiWsBytes + kWiSpecBytes + 4 * kWiMaskBytes + 16 + 256 * 8; // + mbarrier + 256-entry twiddle table
static_assert(kWiSmemBytes <= 232448, "fused Wiener/iSTFT kernel: shared memory budget exceeded");
static_assert(kMaskLd * 4 % 16 == 0 && kWiSpecBytes % 16 == 0, "bulk copies need 16-byte aligned rows");

__device__ __forceinline__ float istft_scale(int f, int n, int k, int F, float2 tw_t)
{
    // interior frames (every sample covered by 4 frames): the window-sum-square of the periodic Hann at hop N/4 is 1.5, so the
    // scale win[n] / 4096 / (wss + 1e-8) is KS * (1 - cos(theta_t + k pi/8)), n = t + 256 k (see istft_kernel)
    constexpr float KS = 0.5f / (4096.0f * (1.5f + 1e-8f));
    if (f >= 3 && f + 3 < F)
        return fmaf(tw_t.x, -KS * kCos8[k], fmaf(tw_t.y, -KS * kSin8[k], KS));
    return __ldg(&g_win[n]) * (1.0f / (float)kNfft) / (wss_at(f, n, F, g_win) + 1e-8f);
}

namespace wi
{
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    if (mbar_try_wait(bar, parity))
        return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 4000000000LL) // a protocol bug must abort the kernel, never hang the GPU
            __trap();
}
// global -> shared bulk copy (TMA, 1-D), bytes counted on the mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void bar_half(int half) { asm volatile("bar.sync %0, 256;" ::"r"(1 + half) : "memory"); }
// mem += v, no return value: the accumulators are touched by one thread per launch, launches are stream-ordered
__device__ __forceinline__ void red_add2(float2 *p, float2 v)
{
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(v.x), "f"(v.y) : "memory");
}
} // namespace wi

__global__ void __launch_bounds__(kWiThreads, 1)
    wiener_istft_kernel(const float2 *__restrict__ spec, const float *__restrict__ mask, long mask_target_stride, int mask_ld,
                        const float *__restrict__ maxabs, const float *__restrict__ stats, int F,
                        const IstftTrack *__restrict__ tracks, int n_buf, int stem_begin, int stem_count, int frame_begin,
                        int frame_end, int stat_windows, int win_len, int flags, int frames_per_cta)
{
    const int psd_abs2 = flags & 1, reg_once = (flags >> 1) & 1, no_em = (flags >> 2) & 1; // WienerArgs: upstream-faithful variants
    extern __shared__ __align__(128) uint8_t smem_raw[];
    float2 *sm = reinterpret_cast<float2 *>(smem_raw);                                   // four FFT workspaces
    const float2 *in_spec = reinterpret_cast<const float2 *>(smem_raw + kWiWsBytes);     // [2][2049] mixture of the frame
    const uint8_t *in_mask = smem_raw + kWiWsBytes + kWiSpecBytes;                       // 4 x [2 x 2049] masks
    const uint32_t bar = wi::smem_u32(smem_raw + kWiWsBytes + kWiSpecBytes + 4 * kWiMaskBytes);
    float2 *t256 = reinterpret_cast<float2 *>(smem_raw + kWiWsBytes + kWiSpecBytes + 4 * kWiMaskBytes + 16); // tw[16 m]
    const int tid = threadIdx.x, half = tid >> 8, t = tid & 255;
    const int trk = blockIdx.z;
    const int n_wave = tracks[trk].n_wave;
    const int f0 = frame_begin + blockIdx.x * frames_per_cta; // this launch emits the hops of frames [frame_begin, frame_end)
    const int f_need = min(frame_end, frames_reaching(n_wave, F)); // frames that reach an emitted sample (see istft_kernel)
    if (f0 >= f_need)
        return;
    const int f1 = min(f_need, f0 + frames_per_cta);
    const bool last_cta = (f1 == F);
    const int f_first = max(f0 - 3, 0);

    // the frame's inputs arrive by bulk copies issued one frame ahead (they land while the previous frame's FFTs run)
    auto stage = [&](int f) {
        const long row = (long)trk * F + f;
        wi::mbar_expect_tx(bar, kWiSpecBytes + 4 * kWiMaskBytes);
        wi::bulk_load(wi::smem_u32(in_spec), spec + row * 2 * kBins, kWiSpecBytes, bar);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            wi::bulk_load(wi::smem_u32(in_mask + j * kWiMaskBytes), mask + (long)j * mask_target_stride + row * mask_ld, kWiMaskBytes, bar);
    };
    if (tid == 0)
    {
        wi::mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        stage(f_first);
    }
    if (tid < 256)
        t256[tid] = g_tw[16 * tid];
    __syncthreads();

    // this half's stems: slot 0 -> stem index `half`, slot 1 -> `half + 2` (relative to stem_begin); workspace = stem index
    float2 acc[2][3][4]; // [slot][hops still missing 1, 2, 3 frames][j]: partial sums of positions 1024 (f + h) + t + 256 j
#pragma unroll
    for (int sl = 0; sl < 2; ++sl)
#pragma unroll
        for (int h = 0; h < 3; ++h)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                acc[sl][h][j] = make_float2(0.f, 0.f);

    uint32_t parity = 0;
    for (int f = f_first; f < f1; ++f)
    {
        // ---- A: filter; Z_j = Y_jL + i Y_jR and its Hermitian extension into workspace j
        // statistics and max|X| of this frame's window (one window = the whole segment unless win_len > 0)
        const int slot = trk * stat_windows + wiener_window(f, win_len);
        const float ma = wiener_max_abs(maxabs, slot);
        const float inv_ma = 1.0f / ma;
        const float *st_base = stats + (long)slot * 16 * kBins;
        wi::mbar_wait(bar, parity);
        parity ^= 1;
#pragma unroll 2
        for (int b = tid; b < kBins; b += kWiThreads)
        {
            float R[16];
#pragma unroll
            for (int i = 0; i < 16; ++i)
                R[i] = no_em ? 0.f : __ldg(st_base + (long)i * kBins + b);
            WienerIn in;
            in.XL = in_spec[b];
            in.XR = in_spec[kBins + b];
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const float *mk = reinterpret_cast<const float *>(in_mask + j * kWiMaskBytes);
                in.mL[j] = mk[b];
                in.mR[j] = mk[kBins + b];
            }
            WienerCommon c;
            wiener_common(in, R, inv_ma, 0, c, psd_abs2, reg_once);
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const int idx = j - stem_begin;
                if (idx < 0 || idx >= stem_count)
                    continue;
                float2 yL, yR;
                if (no_em)
                { // no EM iteration: y_j = mask_j X
                    yL = make_float2(in.mL[j] * in.XL.x, in.mL[j] * in.XL.y);
                    yR = make_float2(in.mR[j] * in.XR.x, in.mR[j] * in.XR.y);
                }
                else
                    wiener_source(c, R, j, ma, yL, yR);
                float2 *w = sm + idx * kFftSmemFloat2;
                if (b == 0 || b == kBins - 1)
                    w[b == 0 ? 0 : 2048] = make_float2(yL.x, yR.x); // the half-spectrum inverse ignores Im of DC and Nyquist
                else
                {
                    w[b] = make_float2(yL.x - yR.y, yL.y + yR.x);
                    w[4096 - b] = make_float2(yL.x + yR.y, yR.x - yL.y); // conj(Y_L) + i conj(Y_R)
                }
            }
        }
        __syncthreads(); // workspaces complete; the staged inputs have been consumed
        if (tid == 0 && f + 1 < f1)
            stage(f + 1);
        // ---- B + C: two stems per half, one after the other; the halves only meet again at the end of the frame
#pragma unroll
        for (int sl = 0; sl < 2; ++sl)
        {
            const int idx = half + 2 * sl;
            if (idx >= stem_count) // uniform per half
                continue;
            float2 *s = sm + idx * kFftSmemFloat2;
            float2 r[16];
#pragma unroll
            for (int a = 0; a < 16; ++a)
                r[a] = s[256 * a + t];
            // two per-thread constants, fetched per transform rather than held in registers across the frame loop (the kernel
            // has none to spare): the constant factor of the pass-2 twiddles and (cos, -sin)(2 pi t / 4096) for the window
            const float2 tw_base = __ldg(&g_tw[(t >> 4) * (t & 15)]);
            const float2 tw_now = __ldg(&g_tw[t]);
            wi::bar_half(half);
            fft_pass1_t256<true>(t, r, s, t256);
            wi::bar_half(half);
            fft_pass2_load(t, r, s);
            wi::bar_half(half);
            fft_pass2_store_t256<true>(t, r, s, t256, tw_base);
            wi::bar_half(half);
            fft_pass3<true>(t, r, s);
            // r[k] = (y_L, y_R)[n], n = t + 256 k: scale (in place), then overlap-add in registers.
            // (The 16 interior scales depend on (t, k) only, but hoisted out of the frame loop they cost 16 registers the kernel
            // does not have: they were spilled, and with 227 KB of shared memory there is no L1 left to catch the reloads.  Two
            // FMAs each from tw_now, which is loaded per transform, so the compiler cannot hoist them.)
#pragma unroll
            for (int k = 0; k < 16; ++k)
            {
                const float sc = istft_scale(f, t + 256 * k, k, F, tw_now);
                r[k].x *= sc;
                r[k].y *= sc;
            }
            // (the track descriptor is read again here rather than kept in registers across the frame loop)
            const IstftTrack *tkp = tracks + trk;
            float2 *out = tkp->out + (long)(stem_begin + idx) * tkp->out_stem_stride + tkp->out_offset;
            const int accumulate = tkp->accumulate;
            const int ex_lo = tkp->ex_lo, ex_hi = tkp->final_out ? tkp->ex_hi : 0;
            auto emit = [&](int hop, int j, float2 v) {
                const int i = kHop * hop + t + 256 * j - kPad;
                if (i >= 0 && i < n_wave)
                {
                    if (accumulate)
                    { // out[off + k] += w[k] * y[k] (umx.cpp:234-249)
                        const float wgt = tri_weight(i, n_buf);
                        const float2 c = make_float2(wgt * v.x, wgt * v.y);
                        if (i >= ex_lo && i < ex_hi)
                        { // no other segment reaches this sample: (0 + w y) / (0 + w), the accumulator is never touched
                            const long io = tkp->final_off + i;
                            if (io >= 0 && io < tkp->final_n)
                                tkp->final_out[(long)(stem_begin + idx) * tkp->final_n + io] = make_float2(__fdiv_rn(c.x, wgt), __fdiv_rn(c.y, wgt));
                        }
                        else
                            wi::red_add2(out + i, c);
                    }
                    else
                        out[i] = v;
                }
            };
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                // hop f is complete: contributions were added in frame order f-3, f-2, f-1, f (as the reference does)
                const float2 done = make_float2(acc[sl][0][j].x + r[j].x, acc[sl][0][j].y + r[j].y);
                acc[sl][0][j] = make_float2(acc[sl][1][j].x + r[4 + j].x, acc[sl][1][j].y + r[4 + j].y);
                acc[sl][1][j] = make_float2(acc[sl][2][j].x + r[8 + j].x, acc[sl][2][j].y + r[8 + j].y);
                acc[sl][2][j] = r[12 + j];
                if (f >= f0)
                    emit(f, j, done);
            }
            if (last_cta && f == F - 1)
            { // the final frame also completes hops F .. F+2 (nothing follows it)
#pragma unroll
                for (int h = 0; h < 3; ++h)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        emit(F + h, j, acc[sl][h][j]);
            }
        }
        __syncthreads(); // every pass-3 read is done before the next frame's filter pass overwrites the workspaces
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
                float *maxabs, cudaStream_t st, int skip_zero, int reflect_pad, int win_len, int stat_windows)
{
    const int F = n_buf / kHop + 1;
    UMX_CHECK(win_len <= 0 || ceil_div(F, win_len) <= stat_windows, "stft: more Wiener windows than max|X| slots");
    dim3 grid(F, n_tracks);
    stft_kernel<<<grid, 256, 0, st>>>(d_tracks, n_buf, F, spec, xh, xl, maxabs, skip_zero, reflect_pad, win_len, stat_windows);
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

__global__ void zero_f32_kernel(float *p, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        p[i] = 0.f;
}
int launch_zero_f32(float *p, int n, cudaStream_t st)
{
    zero_f32_kernel<<<ceil_div(n, 256), 256, 0, st>>>(p, n);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

// p[track * stride + off + i] = 0 for i < count, track < n_tracks
__global__ void zero_strided_kernel(float *p, long stride, long off, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count)
        p[(long)blockIdx.y * stride + off + i] = 0.f;
}
int launch_zero_strided(float *p, long stride, long off, int count, int n_tracks, cudaStream_t st)
{
    zero_strided_kernel<<<dim3(ceil_div(count, 256), n_tracks), 256, 0, st>>>(p, stride, off, count);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

int launch_wiener_istft(const WienerArgs &w, const IstftTrack *d_tracks, int n_buf, cudaStream_t st, int frame_begin, int frame_end)
{
    if (frame_end < 0)
        frame_end = w.F;
    UMX_CHECK(frame_begin >= 0 && frame_begin < frame_end && frame_end <= w.F, "wiener_istft: bad frame range");
    UMX_CHECK(!w.mask_is_mag && w.stats && w.maxabs && w.stem_count >= 1 && w.stem_count <= 4 && w.stat_windows >= 1,
              "wiener_istft: bad arguments");
    UMX_CHECK(w.win_len <= 0 || ceil_div(w.F, w.win_len) <= w.stat_windows, "wiener_istft: more windows than statistics slots");
    const size_t smem = kWiSmemBytes;
    static PerDeviceOnce once;
    if (int rc = once.run([&](int) {
            UMX_CUDA_OK(cudaFuncSetAttribute(wiener_istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            return 0;
        }))
        return rc;
    static const int frames_per_cta = [] {
        const char *e = getenv("UMX_WI_FRAMES"); // frames per CTA (+ 3 halo frames): tuning knob, default 32
        const int v = e ? atoi(e) : kWiFrames;
        return v >= 4 && v <= 1024 ? v : kWiFrames;
    }();
    dim3 grid(ceil_div(frame_end - frame_begin, frames_per_cta), 1, w.n_tracks);
    wiener_istft_kernel<<<grid, kWiThreads, smem, st>>>(w.spec, w.mask, w.mask_target_stride, w.mask_ld, w.maxabs, w.stats, w.F,
                                                        d_tracks, n_buf, w.stem_begin, w.stem_count, frame_begin, frame_end,
                                                        w.stat_windows, w.win_len,
                                                        (w.psd_abs2 ? 1 : 0) | (w.reg_once ? 2 : 0) | (w.no_em ? 4 : 0), frames_per_cta);
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
