// Per-(frame, bin) arithmetic of the Wiener-EM filter pass (reference: src/wiener.cpp:276-422), shared by the stand-alone
// filter kernel (wiener.cu) and the kernel that filters while it loads the inverse STFT's input (stft.cu).
#pragma once
#include "common.cuh"
#include "kernels.cuh"

namespace umx
{

// max_abs of (track, window): max(1, max|X| / 10) (wiener.cpp:36-52); `slot` = track * kWienerChunks + window
__device__ __forceinline__ float wiener_max_abs(const float *maxabs, int slot)
{
    return fmaxf(1.0f, maxabs[slot] / kWienerScale);
}
__device__ __forceinline__ int wiener_window(int f, int win_len) { return win_len > 0 ? f / win_len : 0; }
// PSD of one source from its two channels' estimates: the reference's ((re+im)_L^2 + (re+im)_R^2) / 2 (wiener.cpp:187-202, sic)
// or upstream's mean_ch |y|^2
__device__ __forceinline__ float wiener_psd(float2 yL, float2 yR, int psd_abs2)
{
    if (psd_abs2)
        return ((yL.x * yL.x + yL.y * yL.y) + (yR.x * yR.x + yR.y * yR.y)) * 0.5f;
    const float sL = yL.x + yL.y, sR = yR.x + yR.y;
    return (sL * sL + sR * sR) * 0.5f;
}

// y_jc / max_abs for one channel.  inv_ma = 1 / max_abs: the reference's own build uses -freciprocal-math, and an
// IEEE division per element (with its divergent slow path) made these kernels issue-bound instead of HBM-bound.
__device__ __forceinline__ float2 initial_estimate(float m, float2 X, float inv_ma, int is_mag)
{
    if (is_mag)
    { // polar(mag, arg X): X == 0 has phase 0
        const float a2 = X.x * X.x + X.y * X.y;
        if (a2 > 0.f)
        {
            const float sc = m * rsqrtf(a2) * inv_ma;
            return make_float2(sc * X.x, sc * X.y);
        }
        return make_float2(m * inv_ma, 0.f);
    }
    const float sc = m * inv_ma;
    return make_float2(sc * X.x, sc * X.y);
}

// one frame of inputs for one bin: mix (L, R) and the 4 sources' (L, R) masks / magnitudes
struct WienerIn
{
    float2 XL, XR;
    float mL[4], mR[4];
};
__device__ __forceinline__ void wiener_load(const float2 *__restrict__ spec, const float *__restrict__ mask, long mask_target_stride,
                                            int mask_ld, long row, int b, WienerIn &w)
{
    w.XL = __ldg(&spec[(row * 2 + 0) * kBins + b]);
    w.XR = __ldg(&spec[(row * 2 + 1) * kBins + b]);
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        const float *mk = mask + (long)j * mask_target_stride + row * mask_ld;
        w.mL[j] = mask ? __ldg(mk + b) : 0.f; // (no masks in the second and later EM iterations: WienerArgs::y_in)
        w.mR[j] = mask ? __ldg(mk + kBins + b) : 0.f;
    }
}
__device__ __forceinline__ void wiener_load(const WienerArgs &a, long row, int b, WienerIn &w)
{
    wiener_load(a.spec, a.mask, a.mask_target_stride, a.mask_ld, row, b, w);
}

// what all four sources' outputs of a (frame, bin) share: the PSDs and u = Cxx^-1 x
struct WienerCommon
{
    float2 x0, x1; // mix / max_abs (wiener.cpp:118-129)
    float v[4];    // v_j = ((re+im)_L^2 + (re+im)_R^2) / 2 of the initial estimate (wiener.cpp:187-202, sic)
    float2 u0, u1; // Cxx^-1 (x0, x1)
};
// R: the four sources' normalised spatial covariances of this bin, R[4j .. 4j+3] = R00, R11, Re R01, Im R01.
// est(j, yL, yR): the current estimate of source j (both channels), divided by max_abs.
template <class Est>
__device__ __forceinline__ void wiener_common_est(float2 XL, float2 XR, const float (&R)[16], float inv_ma, WienerCommon &c,
                                                  int psd_abs2, int reg_once, Est est)
{
    const float reg = sqrtf(kWienerEps); // wiener.cpp:155-163
    c.x0 = make_float2(XL.x * inv_ma, XL.y * inv_ma);
    c.x1 = make_float2(XR.x * inv_ma, XR.y * inv_ma);
    // Cxx = sum_j (sqrt(eps) I + v_j R_j): the reference adds the regulariser once PER SOURCE (wiener.cpp:301-325); upstream once
    const float reg_j = reg_once ? 0.f : reg;
    float c00 = reg_once ? reg : 0.f, c11 = c00, c01r = 0.f, c01i = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        float2 yL, yR;
        est(j, yL, yR);
        c.v[j] = wiener_psd(yL, yR, psd_abs2);
        c00 += reg_j + c.v[j] * R[4 * j + 0];
        c11 += reg_j + c.v[j] * R[4 * j + 1];
        c01r += c.v[j] * R[4 * j + 2];
        c01i += c.v[j] * R[4 * j + 3];
    }
    // inverse of the Hermitian 2x2 [[c00, c01],[conj c01, c11]] (invert4D, wiener.cpp:54-84): inv = [[c11, -c01],[-conj c01, c00]] / det
    const float det = c00 * c11 - (c01r * c01r + c01i * c01i);
    const float inv_det = 1.0f / det; // conj(det)/norm(det) with det real (one division per bin-frame)
    // u = Cxx^-1 x.  The reference forms the gain matrix G_j = v_j R_j Cxx^-1 per source and applies it to x (wiener.cpp:335-401);
    // G_j x = v_j R_j (Cxx^-1 x) is the same product with the part all sources share taken out of the loop over sources.
    c.u0 = make_float2(inv_det * (c11 * c.x0.x - (c01r * c.x1.x - c01i * c.x1.y)), inv_det * (c11 * c.x0.y - (c01r * c.x1.y + c01i * c.x1.x)));
    c.u1 = make_float2(inv_det * (c00 * c.x1.x - (c01r * c.x0.x + c01i * c.x0.y)), inv_det * (c00 * c.x1.y - (c01r * c.x0.y - c01i * c.x0.x)));
}
// first iteration: the estimates are mask (or magnitude with the mixture's phase) times the mixture
__device__ __forceinline__ void wiener_common(const WienerIn &in, const float (&R)[16], float inv_ma, int is_mag, WienerCommon &c,
                                              int psd_abs2 = 0, int reg_once = 0)
{
    wiener_common_est(in.XL, in.XR, R, inv_ma, c, psd_abs2, reg_once, [&](int j, float2 &yL, float2 &yR) {
        yL = initial_estimate(in.mL[j], in.XL, inv_ma, is_mag);
        yR = initial_estimate(in.mR[j], in.XR, inv_ma, is_mag);
    });
}
// y_j = v_j R_j u, times max_abs (wiener.cpp:335-422) for source j; R10 = conj(R01)
__device__ __forceinline__ void wiener_source(const WienerCommon &c, const float (&R)[16], int j, float ma, float2 &y0, float2 &y1)
{
    const float r00 = R[4 * j + 0], r11 = R[4 * j + 1], r01r = R[4 * j + 2], r01i = R[4 * j + 3];
    const float g = c.v[j] * ma;
    y0.x = g * (r00 * c.u0.x + (r01r * c.u1.x - r01i * c.u1.y));
    y0.y = g * (r00 * c.u0.y + (r01r * c.u1.y + r01i * c.u1.x));
    y1.x = g * ((r01r * c.u0.x + r01i * c.u0.y) + r11 * c.u1.x);
    y1.y = g * ((r01r * c.u0.y - r01i * c.u0.x) + r11 * c.u1.y);
}

// the current estimate of source j at (row, bin b), divided by max_abs: the previous iteration's output, or the initial estimate
__device__ __forceinline__ void wiener_estimate(const WienerArgs &a, const WienerIn &in, long row, int b, int j, float inv_ma, float2 &yL,
                                                float2 &yR)
{
    if (a.y_in)
    {
        const float2 *p = a.y_in + (long)j * a.y_in_stem_stride + (row * 2) * kBins + b;
        yL = __ldg(p);
        yR = __ldg(p + kBins);
    }
    else
    {
        yL = initial_estimate(in.mL[j], in.XL, inv_ma, a.mask_is_mag);
        yR = initial_estimate(in.mR[j], in.XR, inv_ma, a.mask_is_mag);
    }
}

} // namespace umx
