// 4096-point complex FFT as three radix-16 passes for a 256-thread CTA.
//
// Replaces the reference's per-frame Eigen::FFT (kissfft) calls in stft_inner / istft_inner
// (src/dsp.cpp:209-258).  One complex transform serves BOTH stereo channels: z = L + iR, and the
// two real spectra are separated afterwards (forward) or merged beforehand (inverse).
//
// Index algebra (n = 256a + 16b + c, k = p + 16q + 256r, W = exp(-/+ 2*pi*i/4096)):
//   pass 1  thread t=(b,c)=16b+c : 16-pt DFT over a of x[256a+t]  -> A1[p], times W^(16*b*p), stored at 256p+t
//   pass 2  thread t=(p,c)=16p+c : 16-pt DFT over b of A1[256p+16b+c] -> A2[q], times W^(c*(p+16q)),
//                                  stored TRANSPOSED and padded at c*257 + 16q + p   (bank-conflict free)
//   pass 3  thread t=p+16q       : 16-pt DFT over c of A2[c*257+t] -> X[t + 256r], r = 0..15, in registers
// The functions are host/device so tests/host_fft_check.cpp can run the exact same code on the CPU.
#pragma once

#if defined(__CUDACC__)
#define UMX_HD __host__ __device__ __forceinline__
#include <cuda_runtime.h>
#else
#define UMX_HD inline
struct float2
{
    float x, y;
};
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
#endif

namespace umx
{

constexpr int kFftSmemFloat2 = 16 * 257; // padded pass-2 layout (>= 4096)

UMX_HD float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
UMX_HD float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
UMX_HD float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
// multiply by the conjugate of b
UMX_HD float2 cmulc(float2 a, float2 b) { return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }

// 4-point DFT, in place. INV=false: exp(-i...), INV=true: exp(+i...)
template <bool INV>
UMX_HD void dft4(float2 &a0, float2 &a1, float2 &a2, float2 &a3)
{
    float2 s02 = cadd(a0, a2), d02 = csub(a0, a2);
    float2 s13 = cadd(a1, a3), d13 = csub(a1, a3);
    // -i*d13 (forward) or +i*d13 (inverse)
    float2 j13 = INV ? make_float2(-d13.y, d13.x) : make_float2(d13.y, -d13.x);
    a0 = cadd(s02, s13);
    a2 = csub(s02, s13);
    a1 = cadd(d02, j13);
    a3 = csub(d02, j13);
}

// multiply by W16^m (cos t, -/+ sin t), t = 2*pi*m/16, for the m that occur in a 4x4 split
template <bool INV, int M>
UMX_HD float2 mul_w16(float2 v)
{
    constexpr float c1 = 0.92387953251128674f, s1 = 0.38268343236508977f, c2 = 0.70710678118654752f;
    constexpr float cs = (M == 0) ? 1.f : (M == 1) ? c1 : (M == 2) ? c2 : (M == 3) ? s1 : (M == 4) ? 0.f : (M == 6) ? -c2 : -c1;
    constexpr float sn = (M == 0) ? 0.f : (M == 1) ? s1 : (M == 2) ? c2 : (M == 3) ? c1 : (M == 4) ? 1.f : (M == 6) ? c2 : -s1;
    static_assert(M == 0 || M == 1 || M == 2 || M == 3 || M == 4 || M == 6 || M == 9, "unexpected twiddle");
    if (M == 0)
        return v;
    const float2 w = make_float2(cs, INV ? sn : -sn);
    return cmul(v, w);
}

// 16-point DFT in registers, natural order in and out
template <bool INV>
UMX_HD void dft16(float2 (&v)[16])
{
    // step A: DFT4 over n1 for each n2 (inputs v[4*n1+n2]) -> v[4*k1+n2], then twiddle W16^(n2*k1)
    dft4<INV>(v[0], v[4], v[8], v[12]);
    dft4<INV>(v[1], v[5], v[9], v[13]);
    dft4<INV>(v[2], v[6], v[10], v[14]);
    dft4<INV>(v[3], v[7], v[11], v[15]);
    v[5] = mul_w16<INV, 1>(v[5]);
    v[6] = mul_w16<INV, 2>(v[6]);
    v[7] = mul_w16<INV, 3>(v[7]);
    v[9] = mul_w16<INV, 2>(v[9]);
    v[10] = mul_w16<INV, 4>(v[10]);
    v[11] = mul_w16<INV, 6>(v[11]);
    v[13] = mul_w16<INV, 3>(v[13]);
    v[14] = mul_w16<INV, 6>(v[14]);
    v[15] = mul_w16<INV, 9>(v[15]);
    // step B: DFT4 over n2 for each k1 (inputs v[4*k1+n2]) -> v[4*k1+k2] = X[k1+4*k2]
    dft4<INV>(v[0], v[1], v[2], v[3]);
    dft4<INV>(v[4], v[5], v[6], v[7]);
    dft4<INV>(v[8], v[9], v[10], v[11]);
    dft4<INV>(v[12], v[13], v[14], v[15]);
    // transpose the 4x4 index so that v[k] = X[k]
    float2 t;
#define UMX_SWAP(i, j) \
    t = v[i];          \
    v[i] = v[j];       \
    v[j] = t;
    UMX_SWAP(1, 4)
    UMX_SWAP(2, 8)
    UMX_SWAP(3, 12)
    UMX_SWAP(6, 9)
    UMX_SWAP(7, 13)
    UMX_SWAP(11, 14)
#undef UMX_SWAP
}

// tw[j] = (cos(2*pi*j/4096), -sin(2*pi*j/4096)), j in [0,4096): forward twiddle; inverse uses the conjugate
template <bool INV>
UMX_HD float2 twmul(float2 v, const float2 *tw, int j)
{
    const float2 w = tw[j & 4095];
    return INV ? cmulc(v, w) : cmul(v, w);
}

// pass 1: r[a] = x[256a+t] in, writes s[256p+t]
template <bool INV>
UMX_HD void fft_pass1(int t, float2 (&r)[16], float2 *s, const float2 *tw)
{
    dft16<INV>(r);
    const int b = t >> 4;
#pragma unroll
    for (int p = 0; p < 16; ++p)
    {
        float2 v = r[p];
        if (p != 0)
            v = twmul<INV>(v, tw, 16 * b * p);
        s[256 * p + t] = v;
    }
}

// The same two passes with a 256-entry twiddle table t256[m] = tw[16 m] (2 KB: small enough for shared memory next to four
// workspaces).  Pass 1 needs tw[16 b p] = t256[(b p) & 255]; pass 2 needs tw[c p + 16 c q] = tw[c p] * t256[(c q) & 255], the
// first factor being a per-thread constant (`base`, loaded once per kernel).
template <bool INV>
UMX_HD void fft_pass1_t256(int t, float2 (&r)[16], float2 *s, const float2 *t256)
{
    dft16<INV>(r);
    const int b = t >> 4;
#pragma unroll
    for (int p = 0; p < 16; ++p)
    {
        float2 v = r[p];
        if (p != 0)
        {
            const float2 w = t256[(b * p) & 255];
            v = INV ? cmulc(v, w) : cmul(v, w);
        }
        s[256 * p + t] = v;
    }
}

template <bool INV>
UMX_HD void fft_pass2_store_t256(int t, float2 (&r)[16], float2 *s, const float2 *t256, float2 base)
{
    const int p = t >> 4, c = t & 15;
    dft16<INV>(r);
#pragma unroll
    for (int q = 0; q < 16; ++q)
    {
        const float2 w = q == 0 ? base : cmul(base, t256[(c * q) & 255]);
        s[c * 257 + 16 * q + p] = INV ? cmulc(r[q], w) : cmul(r[q], w);
    }
}

// pass 2a: load ; (barrier) ; pass 2b: transform + transposed store
UMX_HD void fft_pass2_load(int t, float2 (&r)[16], const float2 *s)
{
    const int p = t >> 4, c = t & 15;
#pragma unroll
    for (int b = 0; b < 16; ++b)
        r[b] = s[256 * p + 16 * b + c];
}

template <bool INV>
UMX_HD void fft_pass2_store(int t, float2 (&r)[16], float2 *s, const float2 *tw)
{
    const int p = t >> 4, c = t & 15;
    dft16<INV>(r);
#pragma unroll
    for (int q = 0; q < 16; ++q)
    {
        float2 v = r[q];
        v = twmul<INV>(v, tw, c * (p + 16 * q));
        s[c * 257 + 16 * q + p] = v;
    }
}

// pass 3: r[k] = X[t + 256k] out
template <bool INV>
UMX_HD void fft_pass3(int t, float2 (&r)[16], const float2 *s)
{
#pragma unroll
    for (int c = 0; c < 16; ++c)
        r[c] = s[c * 257 + t];
    dft16<INV>(r);
}

// ---- padded-signal index math (src/dsp.cpp:109-128,152-155) ------------------------------------
// Maps a position P of the (n_buf + 4096)-long padded buffer to the sample index s in [0, n_buf) of
// the un-padded buffer it takes its value from (symmetric edge copy, edge sample repeated).
UMX_HD int padded_to_sample(int P, int n_buf)
{
    if (P < 2048)
        P = 4095 - P;
    else if (P >= n_buf + 2048)
        P = 2 * n_buf + 4095 - P;
    return P - 2048;
}
// The same map for true reflect padding (torch.stft(center=True, pad_mode="reflect"), what upstream Open-Unmix does: the edge
// sample is NOT repeated): option "stft_pad" = 1.
UMX_HD int padded_to_sample_reflect(int P, int n_buf)
{
    if (P < 2048)
        return 2048 - P;
    if (P >= n_buf + 2048)
        return 2 * n_buf + 2046 - P;
    return P - 2048;
}

} // namespace umx
