// Small elementwise kernels: weight dequantisation (reference: src/model.cpp:578-665), GEMM operand
// preparation, and layout conversions used by the kernel-level API.
#include "common.cuh"
#include "kernels.cuh"

namespace umx
{

// matrix(i,j) = q*scale + offset, compiled by the reference build to ONE fused multiply-add
// (src/model.cpp:614,660) -> __fmaf_rn is bit-exact with it.
template <typename Q>
__global__ void dequant_kernel(const Q *__restrict__ q, float scale, float offset, float *__restrict__ out, long n)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
        out[i] = __fmaf_rn((float)q[i], scale, offset);
}

int launch_dequant_u8(const uint8_t *q, float scale, float offset, float *out, long n, cudaStream_t st)
{
    if (n <= 0)
        return 0;
    dequant_kernel<uint8_t><<<ceil_div(n, 256), 256, 0, st>>>(q, scale, offset, out, n);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}
int launch_dequant_u16(const uint16_t *q, float scale, float offset, float *out, long n, cudaStream_t st)
{
    if (n <= 0)
        return 0;
    dequant_kernel<uint16_t><<<ceil_div(n, 256), 256, 0, st>>>(q, scale, offset, out, n);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void make_planes_kernel(const float *__restrict__ w, int N, int K, const float *__restrict__ kscale,
                                   int kscale_mod, __half *__restrict__ hi, __half *__restrict__ lo, int n_pad, int ld)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)n_pad * ld)
        return;
    const int n = (int)(i / ld), k = (int)(i % ld);
    float v = 0.f;
    if (n < N && k < K)
    {
        v = w[(long)n * K + k];
        if (kscale)
            v *= kscale[k % kscale_mod];
    }
    __half h, l;
    split_half(v, h, l);
    hi[i] = h;
    lo[i] = l;
}

int launch_make_planes(const float *w, int N, int K, const float *kscale, int kscale_mod, __half *hi, __half *lo,
                       int n_pad, int ld, cudaStream_t st)
{
    const long total = (long)n_pad * ld;
    make_planes_kernel<<<ceil_div(total, 256), 256, 0, st>>>(w, N, K, kscale, kscale_mod, hi, lo, n_pad, ld);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

// one warp per output row
__global__ void row_dot_kernel(const float *__restrict__ w, int N, int K, const float *__restrict__ v, int vmod,
                               float *__restrict__ out)
{
    const int n = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const int lane = threadIdx.x & 31;
    if (n >= N)
        return;
    double acc = 0.0;
    for (int k = lane; k < K; k += 32)
        acc += (double)v[k % vmod] * (double)w[(long)n * K + k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0)
        out[n] = (float)acc;
}

int launch_row_dot(const float *w, int N, int K, const float *v, int vmod, float *out, cudaStream_t st)
{
    row_dot_kernel<<<ceil_div(N, 8), 256, 0, st>>>(w, N, K, v, vmod, out);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void pack_whh_kernel(const uint8_t *__restrict__ q, int Hh, uint8_t *__restrict__ out)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x; // over (k, unit)
    if (i >= (long)Hh * Hh)
        return;
    const int k = (int)(i / Hh), u = (int)(i % Hh);
#pragma unroll
    for (int g = 0; g < 4; ++g)
        out[i * 4 + g] = q[((long)g * Hh + u) * Hh + k];
}

int launch_pack_whh(const uint8_t *q, int Hh, uint8_t *out, cudaStream_t st)
{
    pack_whh_kernel<<<ceil_div((long)Hh * Hh, 256), 256, 0, st>>>(q, Hh, out);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void colmajor_to_planes_kernel(const float *__restrict__ in, int F, int H, __half *__restrict__ hi,
                                          __half *__restrict__ lo, int ld)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)F * H)
        return;
    const int f = (int)(i / H), h = (int)(i % H);
    __half a, b;
    split_half(in[(long)h * F + f], a, b);
    hi[(long)f * ld + h] = a;
    lo[(long)f * ld + h] = b;
}

int launch_colmajor_to_planes(const float *in, int F, int H, __half *hi, __half *lo, int ld, cudaStream_t st)
{
    colmajor_to_planes_kernel<<<ceil_div((long)F * H, 256), 256, 0, st>>>(in, F, H, hi, lo, ld);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void rowmajor_to_colmajor_kernel(const float *__restrict__ in, int F, int H, int ld, float *__restrict__ out)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)F * H)
        return;
    const int h = (int)(i / F), f = (int)(i % F);
    out[i] = in[(long)f * ld + h];
}

int launch_rowmajor_to_colmajor(const float *in, int F, int H, int ld, float *out, cudaStream_t st)
{
    rowmajor_to_colmajor_kernel<<<ceil_div((long)F * H, 256), 256, 0, st>>>(in, F, H, ld, out);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void mask_to_mag_kernel(const float *__restrict__ mask, int ld, const float2 *__restrict__ spec,
                                   float *__restrict__ mag, long M)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x; // over (row, c, b)
    if (i >= M * 2 * kBins)
        return;
    const long row = i / (2 * kBins);
    const int cb = (int)(i % (2 * kBins));
    const float2 x = spec[i];
    mag[i] = mask[row * ld + cb] * sqrtf(x.x * x.x + x.y * x.y);
}

int launch_mask_to_mag(const float *mask, int ld, const float2 *spec, float *mag, long M, cudaStream_t st)
{
    mask_to_mag_kernel<<<ceil_div(M * 2 * kBins, 256), 256, 0, st>>>(mask, ld, spec, mag, M);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

__global__ void rowsum_kernel(const __half *__restrict__ hi, const __half *__restrict__ lo, long rows, int K, int ld,
                              float *__restrict__ out)
{
    const long r = (long)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const int lane = threadIdx.x & 31;
    if (r >= rows)
        return;
    const __half2 *h2 = reinterpret_cast<const __half2 *>(hi + r * ld);
    const __half2 *l2 = reinterpret_cast<const __half2 *>(lo + r * ld);
    float acc = 0.f;
    for (int k = lane; k < K / 2; k += 32)
    {
        const float2 a = __half22float2(h2[k]), b = __half22float2(l2[k]);
        acc += (a.x + b.x) + (a.y + b.y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0)
        out[r] = acc;
}

int launch_rowsum(const __half *hi, const __half *lo, long rows, int K, int ld, float *out, cudaStream_t st)
{
    UMX_CHECK(K % 2 == 0 && ld % 2 == 0, "rowsum: K and pitch must be even");
    rowsum_kernel<<<ceil_div(rows, 8), 256, 0, st>>>(hi, lo, rows, K, ld, out);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

// W_ih as the centred quantised integers in fp16, rows re-ordered gate-interleaved: out row 4*unit + gate =
// in row gate*Hh + unit (the order the recurrence kernel reads the input projection in)
__global__ void q_to_half_kernel(const uint8_t *__restrict__ q, int Hh, int K, __half *__restrict__ out)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)4 * Hh * K)
        return;
    const int row = (int)(i / K), c = (int)(i % K);
    const int unit = row >> 2, gate = row & 3;
    out[i] = __int2half_rn((int)q[((long)gate * Hh + unit) * K + c] - 127); // centred: keeps x.q small next to offset*rowsum(x)
}

int launch_q_to_half(const uint8_t *q, int Hh, int K, __half *out, cudaStream_t st)
{
    q_to_half_kernel<<<ceil_div((long)4 * Hh * K, 256), 256, 0, st>>>(q, Hh, K, out);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

// x[f][c*1487 + b] = |spec[f][c][b]|, b < 1487, as fp16 (hi, lo) planes (src/inference.cpp:29,58-68)
__global__ void spec_to_x_kernel(const float2 *__restrict__ spec, long M, __half *__restrict__ xh, __half *__restrict__ xl)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M * kFc1In)
        return;
    const long row = i / kFc1In;
    const int j = (int)(i % kFc1In);
    const int c = j / kBinsCropped, b = j % kBinsCropped;
    const float2 x = spec[(row * 2 + c) * kBins + b];
    __half h, l;
    split_half(sqrtf(x.x * x.x + x.y * x.y), h, l);
    xh[row * kFc1InPad + j] = h;
    xl[row * kFc1InPad + j] = l;
}

int launch_spec_to_x(const float2 *spec, long M, __half *xh, __half *xl, cudaStream_t st)
{
    spec_to_x_kernel<<<ceil_div(M * kFc1In, 256), 256, 0, st>>>(spec, M, xh, xl);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx
