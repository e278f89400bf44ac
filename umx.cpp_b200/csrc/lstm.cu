// LSTM recurrence, simple form (reference: src/lstm.cpp:101-179).
//
// One CTA per (target, direction, track); one thread per hidden unit.  The input projection
// W_ih x_t + b_ih + b_hh for every frame comes from the GEMM (`xproj`); this kernel does the serial
// part: gates = xproj[t] + W_hh h_{t-1}, i/f/o = sigmoid, g = tanh, c = f*c + i*g, h = o*tanh(c),
// with h and c loaded from / stored to the session so the state streams across segments
// (lstm.cpp:160-161; never reset, backward direction included: SURVEY B8).
// W_hh stays quantised (uint8, [k][unit][gate]) and is dequantised in flight with the same
// fmaf(q, scale, offset) as the loader, so the weights are bit-identical to the reference's.
// This is the bring-up / validation kernel (session option "lstm_impl" = 1); lstm_cluster.cu holds
// the SMEM-resident cluster kernel.
#include "common.cuh"
#include "kernels.cuh"

namespace umx
{

__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.0f / (1.0f + expf(-x)); } // lstm.cpp:36-39

__global__ void lstm_simple_kernel(const LstmLayerArgs a)
{
    extern __shared__ float h_s[]; // Hh floats
    const int Hh = a.Hh;
    const int u = threadIdx.x;
    const int target = a.target_begin + blockIdx.x / 2;
    const int dir = blockIdx.x & 1;
    const int trk = blockIdx.y;
    const uchar4 *wq = reinterpret_cast<const uchar4 *>(a.whh_q[target][dir]);
    const float scale = a.whh_scale[target][dir], offset = a.whh_offset[target][dir];

    const long st = (long)trk * a.state_track_stride + ((long)target * 2 + dir) * Hh + u;
    float h = a.h_state[st], c = a.c_state[st];
    h_s[u] = h;
    __syncthreads();

    const float *xp = a.xproj + (long)target * a.xproj_target_stride + (long)dir * 4 * Hh + 4 * u; // gate-interleaved
    const int ldx = 8 * Hh;
    // first layer of a short chunk: frames from the first all-zero one on share one computed row of xproj (LstmLayerArgs::chunks)
    const int t_z = a.chunks ? frames_reaching(a.chunks[trk].n_wave, a.F) : a.F;
    for (int step = 0; step < a.F; ++step)
    {
        const int t = dir == 0 ? step : a.F - 1 - step;
        const long row = (long)trk * a.F + t;
        const float *x = xp + ((long)trk * a.F + min(t, t_z)) * ldx;
        const float4 xg = *reinterpret_cast<const float4 *>(x);
        const float gi = xg.x, gf = xg.y, gg = xg.z, go = xg.w;
        float ai = 0.f, af = 0.f, ag = 0.f, ao = 0.f;
#pragma unroll 8
        for (int k = 0; k < Hh; ++k)
        {
            const uchar4 q = __ldg(&wq[(long)k * Hh + u]);
            const float hk = h_s[k];
            ai = fmaf(fmaf((float)q.x, scale, offset), hk, ai);
            af = fmaf(fmaf((float)q.y, scale, offset), hk, af);
            ag = fmaf(fmaf((float)q.z, scale, offset), hk, ag);
            ao = fmaf(fmaf((float)q.w, scale, offset), hk, ao);
        }
        const float i_t = sigmoidf_ref(gi + ai);
        const float f_t = sigmoidf_ref(gf + af);
        const float g_t = tanhf(gg + ag);
        const float o_t = sigmoidf_ref(go + ao);
        c = f_t * c + i_t * g_t;
        h = o_t * tanhf(c);
        __syncthreads(); // everyone finished reading h_{t-1}
        h_s[u] = h;
        const long o = (long)target * a.out_target_stride + row * a.ldo + a.col_offset + dir * Hh + u;
        __half hh, hl;
        split_half(h, hh, hl);
        a.out_hi[o] = hh;
        a.out_lo[o] = hl;
        if (a.out_f32)
            a.out_f32[(long)target * a.out_f32_target_stride + row * (2 * Hh) + dir * Hh + u] = h;
        __syncthreads();
    }
    a.h_state[st] = h;
    a.c_state[st] = c;
}

int launch_lstm_simple(const LstmLayerArgs &a, cudaStream_t st)
{
    UMX_CHECK(a.Hh >= 1 && a.Hh <= 1024, "lstm_simple: hidden units per direction must be in [1,1024]");
    dim3 grid(2 * a.target_count, a.n_tracks);
    lstm_simple_kernel<<<grid, a.Hh, sizeof(float) * a.Hh, st>>>(a);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx
