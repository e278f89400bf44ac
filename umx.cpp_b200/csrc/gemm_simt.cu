// Plain fp32 SIMT GEMM with the same operands and fused epilogues as the tensor-core kernel.
//
// This is the VALIDATION kernel for gemm_tc.cu (tests compare the two) and a bring-up aid
// (session option "gemm_impl" = 1); it is not the product path.  It reconstructs A = A_hi + A_lo and
// B = B_hi + B_lo (22+ significant bits) and accumulates in fp32, i.e. it is numerically the
// reference's Eigen GEMM (src/inference.cpp:86,127,143) up to summation order.
#include "common.cuh"
#include "kernels.cuh"

namespace umx
{

constexpr int BM = 64, BN = 64, BK = 16;

__device__ __forceinline__ float apply_act(float v, int act)
{
    if (act == kActTanh)
        return tanhf(v);
    if (act == kActRelu)
        return fmaxf(v, 0.f);
    return v;
}

__global__ void __launch_bounds__(256) gemm_simt_kernel(const GemmBatch batch)
{
    const GemmProblem &pb = batch.p[blockIdx.z];
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (m0 >= pb.M || n0 >= pb.N)
        return;
    __shared__ float As[BK][BM + 1];
    __shared__ float Bs[BK][BN + 1];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4; // 16 x 16 threads, 4x4 outputs each
    float acc[4][4] = {};

    for (int k0 = 0; k0 < pb.K; k0 += BK)
    {
        // 64 rows x 16 k = 1024 elements, 4 per thread
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            const int e = tid + 256 * i;
            const int rr = e >> 4, kk = e & 15;
            float a = 0.f, b = 0.f;
            if (m0 + rr < pb.M && k0 + kk < pb.K)
            {
                const long o = (long)(m0 + rr) * pb.A.ld + k0 + kk;
                a = __half2float(pb.A.hi[o]) + (pb.A.lo ? __half2float(pb.A.lo[o]) : 0.f);
            }
            if (n0 + rr < pb.N && k0 + kk < pb.K)
            {
                const long o = (long)(n0 + rr) * pb.B.ld + k0 + kk;
                b = __half2float(pb.B.hi[o]) + (pb.B.lo ? __half2float(pb.B.lo[o]) : 0.f);
            }
            As[kk][rr] = a;
            Bs[kk][rr] = b;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk)
        {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                a[i] = As[kk][ty * 4 + i];
                b[i] = Bs[kk][tx * 4 + i];
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }

#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        const int m = m0 + ty * 4 + i;
        if (m >= pb.M)
            continue;
#pragma unroll
        for (int j = 0; j < 4; ++j)
        {
            const int n = n0 + tx * 4 + j;
            if (n >= pb.N)
                continue;
            const float al = pb.alpha ? pb.alpha[n] : 1.f;
            const float be = pb.beta ? pb.beta[n] : 0.f;
            const float rv = pb.rowvec ? pb.rowvec[m] * pb.gamma[n] : 0.f;
            const float v = apply_act(fmaf(acc[i][j], al, be + rv), pb.act);
            const long o = (long)(n / pb.n_group) * pb.group_stride + (long)m * pb.ldo + (n % pb.n_group);
            if (pb.out_f32)
                pb.out_f32[o] = v;
            if (pb.out_hi)
            {
                __half h, l;
                split_half(v, h, l);
                pb.out_hi[o] = h;
                pb.out_lo[o] = l;
            }
        }
    }
}

int launch_gemm_simt(const GemmBatch &batch, cudaStream_t st)
{
    int maxM = 0, maxN = 0;
    for (int i = 0; i < batch.count; ++i)
    {
        maxM = max(maxM, batch.p[i].M);
        maxN = max(maxN, batch.p[i].N);
    }
    dim3 grid(ceil_div(maxN, BN), ceil_div(maxM, BM), batch.count);
    gemm_simt_kernel<<<grid, 256, 0, st>>>(batch);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx
