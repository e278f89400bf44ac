// Wiener-EM (reference: src/wiener.cpp:92-425, one EM iteration, whole-segment statistics).
//
// Three launches per segment batch, all bandwidth-bound, thread-per-bin so every global access is
// coalesced along the bin axis of the internal [frame][channel][bin] layout:
//   wiener_reduce    per (bin tile, frame chunk, track): partial sums over frames of y y^H (2x2, Hermitian:
//                    R00, R11, Re/Im R01) and of the PSD v for each of the 4 sources  (wiener.cpp:181-258)
//   wiener_finalize  fixed-order sum of the chunk partials, R = S / (eps + sum v)     (wiener.cpp:260-271)
//   wiener_apply     per (frame, bin): v_j, Cxx = sum_j (sqrt(eps) I + v_j R_j) (regulariser added once PER
//                    SOURCE, wiener.cpp:301-325), 2x2 inverse (:54-84), gain_j = v_j R_j Cxx^-1 (:335-375),
//                    y_j = gain_j x (:380-401), times max_abs (:408-422)
// The initial estimate is y_j = polar(mag_j, arg X)/max_abs (wiener.cpp:99-146).  In the pipeline
// mag_j = mask_j |X| (inference.cpp:173-185), so y_j = mask_j X / max_abs and neither the magnitudes nor
// the phase are materialised.  PSD quirk kept: v = sum_ch (re+im)^2 / 2 (wiener.cpp:187-202).
#include "common.cuh"
#include "kernels.cuh"
#include "wiener_math.cuh"

namespace umx
{

constexpr int kWienerQ = 20;         // per source: S00, S11, Re S01, Im S01, V
constexpr int kWienerApplyFrames = 16;

constexpr int kWienerUnroll = 4; // frames whose loads are in flight together (the kernels are latency-bound otherwise)

// YIN: the estimates come from WienerArgs::y_in (second and later EM iterations); compile-time so the first-iteration path, the
// one every segment takes, keeps its prefetched loads
template <bool YIN>
__global__ void __launch_bounds__(256) wiener_reduce_kernel(const WienerArgs a)
{
    const int b = blockIdx.x * 256 + threadIdx.x;
    const int chunk = blockIdx.y, trk = blockIdx.z;
    if (b >= kBins)
        return;
    // chunk = a run of frames whose partial sums are kept apart: F/19 frames each for whole-segment statistics, one window each
    // when the statistics are per window (a.win_len > 0)
    const int fc = a.win_len > 0 ? a.win_len : (a.F + kWienerChunks - 1) / kWienerChunks;
    // frames beyond the chunk have an all-zero mixture spectrum: exactly zero terms in every sum (and the masks there are not computed)
    const int f_lim = a.chunks ? frames_reaching(a.chunks[trk].n_wave, a.F) : a.F;
    const int f_begin = chunk * fc, f_end = min(f_lim, f_begin + fc);
    const float ma = 1.0f / wiener_max_abs(a.maxabs, trk * a.stat_windows + (a.win_len > 0 ? chunk : 0)); // reciprocal: see initial_estimate
    float acc[kWienerQ];
#pragma unroll
    for (int i = 0; i < kWienerQ; ++i)
        acc[i] = 0.f;
    for (int f0 = f_begin; f0 < f_end; f0 += kWienerUnroll)
    {
        WienerIn in[kWienerUnroll];
#pragma unroll
        for (int u = 0; u < kWienerUnroll; ++u)
            wiener_load(a, (long)trk * a.F + min(f0 + u, f_end - 1), b, in[u]);
#pragma unroll
        for (int u = 0; u < kWienerUnroll; ++u)
        {
            if (f0 + u >= f_end)
                break;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                float2 yL, yR;
                if (YIN)
                    wiener_estimate(a, in[u], (long)trk * a.F + f0 + u, b, j, ma, yL, yR);
                else
                {
                    yL = initial_estimate(in[u].mL[j], in[u].XL, ma, a.mask_is_mag);
                    yR = initial_estimate(in[u].mR[j], in[u].XR, ma, a.mask_is_mag);
                }
                acc[5 * j + 0] += yL.x * yL.x + yL.y * yL.y;
                acc[5 * j + 1] += yR.x * yR.x + yR.y * yR.y;
                acc[5 * j + 2] += yL.x * yR.x + yL.y * yR.y; // yL * conj(yR)
                acc[5 * j + 3] += yL.y * yR.x - yL.x * yR.y;
                acc[5 * j + 4] += wiener_psd(yL, yR, a.psd_abs2);
            }
        }
    }
    float *p = a.partials + ((long)trk * kWienerChunks + chunk) * kWienerQ * kBins + b;
#pragma unroll
    for (int i = 0; i < kWienerQ; ++i)
        p[(long)i * kBins] = acc[i];
}

__global__ void __launch_bounds__(256) wiener_finalize_kernel(const WienerArgs a)
{
    const int b = blockIdx.x * 256 + threadIdx.x;
    const int trk = blockIdx.y, win = blockIdx.z; // one window: all chunks (whole-segment statistics) or chunk `win`
    if (b >= kBins)
        return;
    const float *p = a.partials + (long)trk * kWienerChunks * kWienerQ * kBins + b;
    float *s = a.stats + ((long)trk * a.stat_windows + win) * 16 * kBins + b;
    const int c_begin = a.win_len > 0 ? win : 0, c_end = a.win_len > 0 ? win + 1 : kWienerChunks;
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        float q[5] = {0.f, 0.f, 0.f, 0.f, kWienerEps}; // weight starts at eps (wiener.cpp:210)
        for (int c = c_begin; c < c_end; ++c)
#pragma unroll
            for (int i = 0; i < 5; ++i)
                q[i] += p[((long)c * kWienerQ + 5 * j + i) * kBins];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            s[(long)(4 * j + i) * kBins] = q[i] / q[4];
    }
}

__global__ void __launch_bounds__(256) wiener_apply_kernel(const WienerArgs a)
{
    const int b = blockIdx.x * 256 + threadIdx.x;
    const int trk = blockIdx.z;
    if (b >= kBins)
        return;
    // the stems are only read by the iSTFT of this track's chunk: frames beyond it are never used (see istft_kernel)
    const int f_lim = a.chunks ? frames_reaching(a.chunks[trk].n_wave, a.F) : a.F;
    const int f_begin = blockIdx.y * kWienerApplyFrames, f_end = min(f_lim, f_begin + kWienerApplyFrames);
    if (f_begin >= f_end)
        return;
    int win = -1;
    float ma = 1.f, inv_ma = 1.f;
    float R[16];
    for (int f0 = f_begin; f0 < f_end; f0 += 2)
    {
        WienerIn in[2];
#pragma unroll
        for (int u = 0; u < 2; ++u)
            wiener_load(a, (long)trk * a.F + min(f0 + u, f_end - 1), b, in[u]);
#pragma unroll
        for (int u = 0; u < 2; ++u)
        {
            const int f = f0 + u;
            if (f >= f_end)
                break;
            if (wiener_window(f, a.win_len) != win)
            { // statistics and max|X| of this frame's window (one window unless a.win_len > 0)
                win = wiener_window(f, a.win_len);
                ma = wiener_max_abs(a.maxabs, trk * a.stat_windows + win);
                inv_ma = 1.0f / ma;
                const float *s = a.stats + ((long)trk * a.stat_windows + win) * 16 * kBins + b;
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    R[i] = s[(long)i * kBins];
            }
            const long row = (long)trk * a.F + f;
            WienerCommon c;
            wiener_common_est(in[u].XL, in[u].XR, R, inv_ma, c, a.psd_abs2, a.reg_once,
                              [&](int j, float2 &yL, float2 &yR) { wiener_estimate(a, in[u], row, b, j, inv_ma, yL, yR); });
            const float ma_out = a.out_normalized ? 1.0f : ma; // an iteration that feeds another leaves its outputs divided by max_abs
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (j >= a.stem_begin && j < a.stem_begin + a.stem_count) // target-parallel ranks only write their own stems
                {
                    float2 y0, y1;
                    if (a.no_em)
                    { // no EM iteration: the initial estimate itself
                        y0 = initial_estimate(in[u].mL[j], in[u].XL, 1.f, a.mask_is_mag);
                        y1 = initial_estimate(in[u].mR[j], in[u].XR, 1.f, a.mask_is_mag);
                    }
                    else
                        wiener_source(c, R, j, ma_out, y0, y1);
                    float2 *o = a.out + (long)j * a.out_stem_stride + (row * 2) * kBins + b;
                    o[0] = y0;
                    o[kBins] = y1;
                }
            if (a.mix_out)
            {
                a.mix_out[(row * 2 + 0) * kBins + b] = c.x0;
                a.mix_out[(row * 2 + 1) * kBins + b] = c.x1;
            }
        }
    }
}

int launch_wiener(const WienerArgs &a, cudaStream_t st)
{
    const int tiles = ceil_div(kBins, 256);
    const int n_win = a.win_len > 0 ? ceil_div(a.F, a.win_len) : 1;
    UMX_CHECK(n_win <= kWienerChunks && n_win <= a.stat_windows, "wiener: at most 19 windows per segment (win_len too small for this window length)");
    if (!a.no_em)
    {
        const dim3 grid(tiles, a.win_len > 0 ? n_win : kWienerChunks, a.n_tracks);
        if (a.y_in)
            wiener_reduce_kernel<true><<<grid, 256, 0, st>>>(a);
        else
            wiener_reduce_kernel<false><<<grid, 256, 0, st>>>(a);
        UMX_CUDA_OK(cudaGetLastError());
        wiener_finalize_kernel<<<dim3(tiles, a.n_tracks, n_win), 256, 0, st>>>(a);
        UMX_CUDA_OK(cudaGetLastError());
    }
    if (!a.out)
        return 0; // statistics only
    wiener_apply_kernel<<<dim3(tiles, ceil_div(a.F, kWienerApplyFrames), a.n_tracks), 256, 0, st>>>(a);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx

// =====================================================================================================
// Allocation-free Wiener entry point (no counterpart in the reference; declared in include/umx_cuda.h).
//
// Some hosts of this library may not allocate device memory at the time they need a Wiener pass: a
// real-time audio thread cannot afford cudaMalloc's implicit device-wide synchronisation and unbounded
// latency, and a plugin living inside a process whose GPU memory is entirely owned by another
// allocator (a fixed start-up budget, no driver allocations afterwards) is simply not allowed to.
// For them the scratch of the three Wiener kernels (per-chunk partial sums, the normalised spatial
// covariances and the per-track max|X|) lives in the file-scope device array below instead of in
// cudaMalloc'ed buffers.  Costs and rules, all visible to the caller:
//   * the array (UMX_WIENER_NOALLOC_MAX_TRACKS tracks, about 6.5 MB) is part of the library image:
//     every GPU that loads the library carries it until the process exits, used or not;
//   * the FIRST call on each device looks the array's device address up (cudaGetSymbolAddress with that
//     device current).  Make that first call outside any stream capture;
//   * all calls share the one array and it keeps the previous call's contents: the max|X| slots are
//     cleared at the start of every call, everything else is fully overwritten before it is read;
//   * two calls whose kernels overlap on the GPU (different streams, different host threads) race on
//     the array.  Serialise them.
// =====================================================================================================
#include "../../include/umx_cuda.h"

#include <mutex>

namespace umx
{
namespace
{
constexpr int kNoAllocMaxTracks = UMX_WIENER_NOALLOC_MAX_TRACKS;
constexpr size_t kNaPartials = (size_t)kWienerChunks * kWienerQ * kBins; // floats per track
constexpr size_t kNaStats = (size_t)16 * kBins;
constexpr size_t kNaMaxabsOff = (size_t)kNoAllocMaxTracks * (kNaPartials + kNaStats);
constexpr size_t kNaTotal = kNaMaxabsOff + 64;

// In HOST code this name is only a stand-in for the symbol, never a device address.
static __device__ __align__(256) float g_wiener_noalloc_scratch[kNaTotal];

// per-track max |X| over a [n_tracks*F][2][2049] spectrogram (non-negative floats order like ints)
__global__ void __launch_bounds__(256) spec_maxabs_kernel(const float2 *__restrict__ spec, long per_track,
                                                          float *__restrict__ maxabs)
{
    __shared__ float red[8];
    const int trk = blockIdx.y;
    const float2 *p = spec + (long)trk * per_track;
    float mx = 0.f;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < per_track; i += (long)gridDim.x * blockDim.x)
    {
        const float2 x = p[i];
        mx = fmaxf(mx, sqrtf(x.x * x.x + x.y * x.y));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0)
        red[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        for (int i = 1; i < 8; ++i)
            mx = fmaxf(mx, red[i]);
        atomicMax(reinterpret_cast<int *>(maxabs) + trk, __float_as_int(mx));
    }
}

// device address of the scratch array on the current device, looked up once per device
int noalloc_scratch(float **out)
{
    static std::mutex mu;
    static float *ptr[64] = {};
    int dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&dev));
    UMX_CHECK(dev >= 0 && dev < 64, "umx_cuda_wiener_noalloc: device index out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev])
    {
        float *p = nullptr;
        UMX_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_wiener_noalloc_scratch));
        UMX_CHECK(p != nullptr, "umx_cuda_wiener_noalloc: cudaGetSymbolAddress returned a null address");
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return 0;
}
} // namespace
} // namespace umx

extern "C" int umx_cuda_wiener_noalloc(const float *d_mix, const float *d_mags, int F, int n_tracks, float *d_out,
                                       void *stream)
{
    using namespace umx;
    UMX_CHECK(d_mix && d_mags && d_out, "umx_cuda_wiener_noalloc: null pointer");
    UMX_CHECK(F >= 1, "umx_cuda_wiener_noalloc: F must be >= 1");
    // the filter pass walks frames in blocks of 16 along grid.y (<= 65535): reject up front, before anything is queued
    UMX_CHECK(F <= 16 * 65535, "umx_cuda_wiener_noalloc: F must be <= 1,048,560 frames");
    if (n_tracks < 1 || n_tracks > kNoAllocMaxTracks)
    {
        set_error("umx_cuda_wiener_noalloc: n_tracks = " + std::to_string(n_tracks) + " is outside [1, " +
                  std::to_string(kNoAllocMaxTracks) + "]: the built-in scratch array holds at most " +
                  std::to_string(kNoAllocMaxTracks) + " tracks; use umx_cuda_wiener or a session for more");
        return -2;
    }
    float *scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch))
        return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // The scratch array used is the CURRENT device's copy: the caller's buffers and stream must live on that device.
    // (Queries only; nothing here allocates.)
    int cur_dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&cur_dev));
    for (const void *p : {static_cast<const void *>(d_mix), static_cast<const void *>(d_mags), static_cast<const void *>(d_out)})
    {
        cudaPointerAttributes pa{};
        if (cudaPointerGetAttributes(&pa, p) != cudaSuccess || (pa.type != cudaMemoryTypeDevice && pa.type != cudaMemoryTypeManaged) ||
            (pa.type == cudaMemoryTypeDevice && pa.device != cur_dev))
        {
            cudaGetLastError();
            set_error("umx_cuda_wiener_noalloc: d_mix, d_mags and d_out must be device memory of the current device (" +
                      std::to_string(cur_dev) + ")");
            return -2;
        }
    }
#if CUDART_VERSION >= 12080
    if (st != nullptr && st != cudaStreamLegacy && st != cudaStreamPerThread)
    {
        int st_dev = -1;
        if (cudaStreamGetDevice(st, &st_dev) == cudaSuccess && st_dev != cur_dev)
        {
            set_error("umx_cuda_wiener_noalloc: the stream belongs to device " + std::to_string(st_dev) +
                      " but device " + std::to_string(cur_dev) + " is current");
            return -2;
        }
        cudaGetLastError();
    }
#endif
    const long e = (long)n_tracks * F * 2 * kBins; // elements of one [n_tracks*F][2][2049] tensor
    float *maxabs = scratch + kNaMaxabsOff;
    // the only state that must start at zero: the atomicMax targets
    UMX_CUDA_OK(cudaMemsetAsync(maxabs, 0, sizeof(float) * kNoAllocMaxTracks, st));
    const long per_track = (long)F * 2 * kBins;
    spec_maxabs_kernel<<<dim3((unsigned)std::min<long>(ceil_div(per_track, 256), 592), n_tracks), 256, 0, st>>>(
        reinterpret_cast<const float2 *>(d_mix), per_track, maxabs);
    UMX_CUDA_OK(cudaGetLastError());
    WienerArgs w{};
    w.spec = reinterpret_cast<const float2 *>(d_mix);
    w.mask = d_mags;
    w.mask_target_stride = e;
    w.mask_ld = 2 * kBins;
    w.mask_is_mag = 1;
    w.maxabs = maxabs;
    w.stat_windows = 1;
    w.partials = scratch;
    w.stats = scratch + (size_t)kNoAllocMaxTracks * kNaPartials;
    w.out = reinterpret_cast<float2 *>(d_out);
    w.out_stem_stride = e;
    w.mix_out = nullptr; // the caller's mix is left untouched
    w.F = F;
    w.n_tracks = n_tracks;
    w.stem_begin = 0;
    w.stem_count = 4;
    return launch_wiener(w, st);
}
