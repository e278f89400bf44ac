// LSTM recurrence on the tensor cores (reference: the W_hh h_{t-1} + gate half of src/lstm.cpp:112-174).
//
// One thread-block CLUSTER of 8 CTAs per (target, direction, group of <= 16 tracks).  CTA r owns hidden
// units [64r, 64r+64): its 256 gate rows of W_hh stay RESIDENT IN SHARED MEMORY for the whole segment as
// the raw uint8 quantised weights (128 KB, pre-swizzled image built at model load), so a step never
// touches HBM for weights.  Per step:
//   * h_{t-1} of all 512 units x tracks sits in each CTA's SMEM as the B operand, written as four
//     signed base-128 digits per value (h = sum_i d_i 2^(-6-7i), exact to 2^-28);
//   * one thread issues tcgen05.mma kind::i8: D[row][track,digit] = sum_k q[row][k] d[k]  (uint8 x int8 ->
//     int32, exact), for the two 128-row halves, plus the same against an all-ones tile, which yields
//     sum_k h_k for the offset term of W = scale*q + offset;
//   * 8 epilogue warps read TMEM (lane = gate row, ordered 4*unit + gate so one quad of lanes holds the
//     four gates of a unit), add the input projection, apply the activations, update c and h, and
//     write the new h as digits straight into the B operand of ALL 8 CTAs through distributed shared
//     memory, then arrive (release.cluster) on the destination's mbarrier;
//   * the MMA thread waits for the 256 arrivals (acquire.cluster), fences the async proxy and goes on.
// There is no cluster-wide barrier in the loop; two h buffers alternate by step parity.
// h / c are loaded from and stored to the session so the state streams across segments (lstm.cpp:160-161).
#include "common.cuh"
#include "kernels.cuh"
#include "tc_common.cuh"

namespace umx
{
using namespace tc;

namespace
{
constexpr int kC = 8;            // CTAs per cluster
constexpr int kHhC = 512;        // hidden units per direction this kernel is built for
constexpr int kUnits = 64;       // units per CTA
constexpr int kMaxTracks = 16;   // tracks per cluster (N = 4 digits * tracks <= 64)
constexpr int kATile = 128 * 128;                 // one (half, k-block) tile of W_hh: 128 rows x 128 bytes
constexpr int kABytes = 2 * 4 * kATile;           // 128 KB
constexpr int kOnesOff = kABytes;                 // all-ones tile, 16 KB
constexpr int kBOff = kOnesOff + kATile;          // h digits: [2 buffers][4 k-blocks][64 rows][128 B]
constexpr int kBTile = 64 * 128;
constexpr int kBBuf = 4 * kBTile;
constexpr int kBarOff = kBOff + 2 * kBBuf;
constexpr int kSmemBytes = kBarOff + 64 + 1024;
constexpr int kThreads = 288;    // 8 epilogue warps + 1 control warp
constexpr uint32_t kTmemCols = 256;

__device__ __forceinline__ uint32_t mapa(uint32_t local_addr, uint32_t cta)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void st_cluster_v2(uint32_t addr, uint32_t a, uint32_t b)
{
    asm volatile("st.shared::cluster.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void arrive_remote_release(uint32_t addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ bool try_wait_cluster(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
                 "selp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(bar), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ void wait_cluster(uint32_t bar, uint32_t parity)
{
    if (try_wait_cluster(bar, parity))
        return;
    const long long t0 = clock64();
    while (!try_wait_cluster(bar, parity))
        if (clock64() - t0 > 4000000000LL)
            __trap();
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32"
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
}

// sum_i D_i * 2^(-6-7i): the integer dot products of the four digit planes back to one float
__device__ __forceinline__ float combine_digits(uint32_t d0, uint32_t d1, uint32_t d2, uint32_t d3)
{
    const float a = fmaf((float)(int)d1, 1.0f / 8192.0f, (float)(int)d0 * (1.0f / 64.0f));
    const float b = fmaf((float)(int)d3, 1.0f / 134217728.0f, (float)(int)d2 * (1.0f / 1048576.0f));
    return a + b;
}

struct ClusterArgs
{
    LstmLayerArgs a;
    const uint8_t *img[kTargets][2]; // pre-swizzled W_hh image per (target, dir): [8 CTAs][128 KB]
};

// Writes this warp's 8 units of one track (h identical in the 4 lanes of a quad) as digit bytes into
// the B operand buffer `buf` of every CTA of the cluster: lane (q, g) sends digit g to CTA q.
__device__ __forceinline__ void publish_h(float h, int q, int g, int track, uint32_t b_local_base, uint32_t chunk, uint32_t within)
{
    int v = __float2int_rn(h * 134217728.0f); // 2^27
    const int d3 = ((v + 64) & 127) - 64;
    v = (v - d3) >> 7;
    const int d2 = ((v + 64) & 127) - 64;
    v = (v - d2) >> 7;
    const int d1 = ((v + 64) & 127) - 64;
    v = (v - d1) >> 7;
    const int mine = g == 0 ? v : (g == 1 ? d1 : (g == 2 ? d2 : d3));
    const uint32_t b = (uint32_t)mine & 0xffu;
    uint32_t o = __shfl_xor_sync(0xffffffffu, b, 4);
    const uint32_t w2 = (q & 1) ? (o | (b << 8)) : (b | (o << 8));
    o = __shfl_xor_sync(0xffffffffu, w2, 8);
    const uint32_t w4 = (q & 2) ? (o | (w2 << 16)) : (w2 | (o << 16));
    o = __shfl_xor_sync(0xffffffffu, w4, 16);
    const uint32_t lo = (q & 4) ? o : w4, hi = (q & 4) ? w4 : o;
    const uint32_t n = 4u * track + g; // B row: (track, digit)
    const uint32_t la = b_local_base + n * 128u + ((chunk ^ (n & 7u)) << 4) + within;
    st_cluster_v2(mapa(la, (uint32_t)q), lo, hi);
}

template <int TP>
__global__ void __cluster_dims__(kC, 1, 1) __launch_bounds__(kThreads, 1) lstm_cluster_kernel(const ClusterArgs P)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const LstmLayerArgs &a = P.a;
    const int r = blockIdx.x; // rank in cluster (cluster spans x)
    const int target = a.target_begin + blockIdx.y / 2;
    const int dir = blockIdx.y & 1;
    const int trk0 = blockIdx.z * kMaxTracks;
    const int n_valid = min(kMaxTracks, a.n_tracks - trk0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int N = 4 * TP;

    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kBarOff);
    const uint32_t hfull0 = smem_u32(bars), mma_done = smem_u32(bars + 2);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 3);
    const uint32_t sA = smem_u32(smem), sOnes = sA + kOnesOff, sB = sA + kBOff;

    // ---- one-time setup: weights into SMEM, ones tile, zeroed h buffers, barriers, TMEM ----
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(P.img[target][dir] + (size_t)r * kABytes);
        uint4 *dst = reinterpret_cast<uint4 *>(smem);
        for (int i = threadIdx.x; i < kABytes / 16; i += kThreads)
            dst[i] = __ldg(&src[i]);
        uint4 *ones = reinterpret_cast<uint4 *>(smem + kOnesOff);
        for (int i = threadIdx.x; i < kATile / 16; i += kThreads)
            ones[i] = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
        uint4 *bz = reinterpret_cast<uint4 *>(smem + kBOff);
        for (int i = threadIdx.x; i < 2 * kBBuf / 16; i += kThreads)
            bz[i] = make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0)
    {
        mbar_init(hfull0, 256);
        mbar_init(hfull0 + 8, 256);
        mbar_init(mma_done, 1);
        fence_barrier_init();
    }
    if (warp == 8)
        tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    fence_proxy_async(); // generic-proxy writes of the weight image before the first MMA reads them
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    cluster_sync_all(); // every CTA's barriers and zeroed buffers exist before anyone writes remotely

    const int F = a.F;
    if (warp < 8)
    {
        // ================= epilogue warps =================
        const int qd = warp & 3, hh = warp >> 2, q = lane >> 2, g = lane & 3;
        const int k = kUnits * r + 32 * hh + 8 * qd + q; // hidden unit of this quad
        const uint32_t kk0 = (uint32_t)((kUnits * r + 32 * hh + 8 * qd) & 127); // first unit of the warp inside its k-block
        const uint32_t kb = (uint32_t)(kUnits * r) >> 7;
        const uint32_t chunk = kk0 >> 4, within = kk0 & 15u;
        const float scale = a.whh_scale[target][dir], offset = a.whh_offset[target][dir];
        const float gm = (g == 2) ? 1.0f : 0.5f, ga = (g == 2) ? 0.0f : 0.5f; // sigmoid(x) = 0.5 tanh(x/2) + 0.5
        const uint32_t t_lane = tmem_base + ((uint32_t)(32 * qd) << 16);
        const long st_base = ((long)target * 2 + dir) * kHhC + k;
        const float *xp_base = a.xproj + (long)target * a.xproj_target_stride + (long)dir * 4 * kHhC + g * kHhC + k;
        const int ldx = 8 * kHhC;

        float c[TP], xp[TP];
#pragma unroll
        for (int j = 0; j < TP; ++j)
        {
            const bool ok = j < n_valid;
            const long so = (long)(trk0 + j) * a.state_track_stride + st_base;
            c[j] = ok ? a.c_state[so] : 0.f;
            const float h0 = ok ? a.h_state[so] : 0.f;
            publish_h(h0, q, g, j, sB + kb * kBTile, chunk, within); // buffer 0
            const int t0 = dir == 0 ? 0 : F - 1;
            xp[j] = ok ? __ldg(xp_base + ((long)(trk0 + j) * F + t0) * ldx) : 0.f;
        }
        fence_proxy_async();
        arrive_remote_release(mapa(hfull0, (uint32_t)q));

        float hcur[TP];
        for (int step = 0; step < F; ++step)
        {
            const int t = dir == 0 ? step : F - 1 - step;
            const bool more = step + 1 < F;
            float xn[TP];
            if (more)
            {
                const int tn = dir == 0 ? step + 1 : F - 2 - step;
#pragma unroll
                for (int j = 0; j < TP; ++j)
                    xn[j] = (j < n_valid) ? __ldg(xp_base + ((long)(trk0 + j) * F + tn) * ldx) : 0.f;
            }
            mbar_wait(mma_done, (uint32_t)(step & 1));
            tc_fence_after();
            const uint32_t bnext = sB + (uint32_t)((step + 1) & 1) * kBBuf + kb * kBTile;
#pragma unroll
            for (int jc = 0; jc < TP / 4; ++jc)
            {
                uint32_t d[16], s[16];
                tmem_ld16(t_lane + hh * 64 + 16 * jc, d);
                tmem_ld16(t_lane + 128 + 16 * jc, s);
                tc_wait_ld();
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                {
                    const int j = 4 * jc + jj;
                    const float dot = combine_digits(d[4 * jj], d[4 * jj + 1], d[4 * jj + 2], d[4 * jj + 3]);
                    const float hs = combine_digits(s[4 * jj], s[4 * jj + 1], s[4 * jj + 2], s[4 * jj + 3]);
                    const float pre = xp[j] + fmaf(scale, dot, offset * hs);
                    const float act = fmaf(gm, tanhf(gm * pre), ga);
                    const int qb = lane & ~3;
                    const float i_t = __shfl_sync(0xffffffffu, act, qb);
                    const float f_t = __shfl_sync(0xffffffffu, act, qb + 1);
                    const float g_t = __shfl_sync(0xffffffffu, act, qb + 2);
                    const float o_t = __shfl_sync(0xffffffffu, act, qb + 3);
                    c[j] = f_t * c[j] + i_t * g_t;       // lstm.cpp:152-154
                    const float h = o_t * tanhf(c[j]);    // lstm.cpp:155
                    hcur[j] = h;
                    if (more)
                        publish_h(h, q, g, j, bnext, chunk, within);
                    if (j < n_valid)
                    {
                        const long row = (long)(trk0 + j) * F + t;
                        const long o = (long)target * a.out_target_stride + row * a.ldo + a.col_offset + dir * kHhC + k;
                        __half hh16, hl16;
                        split_half(h, hh16, hl16);
                        if (g == 0)
                            a.out_hi[o] = hh16;
                        else if (g == 1)
                            a.out_lo[o] = hl16;
                        else if (g == 2 && a.out_f32)
                            a.out_f32[(long)target * a.out_f32_target_stride + row * (2 * kHhC) + dir * kHhC + k] = h;
                    }
                }
            }
            if (more)
            {
                tc_fence_before();   // our TMEM reads are done before the next MMAs may overwrite the accumulators
                fence_proxy_async(); // our SMEM (DSMEM) writes become visible to the async proxy (MMA operand reads)
                arrive_remote_release(mapa(hfull0 + 8u * (uint32_t)((step + 1) & 1), (uint32_t)q));
#pragma unroll
                for (int j = 0; j < TP; ++j)
                    xp[j] = xn[j];
            }
        }
#pragma unroll
        for (int j = 0; j < TP; ++j)
            if (j < n_valid)
            {
                const long so = (long)(trk0 + j) * a.state_track_stride + st_base;
                if (g == 0)
                    a.h_state[so] = hcur[j];
                else if (g == 1)
                    a.c_state[so] = c[j];
            }
    }
    else if (lane == 0)
    {
        // ================= MMA issuer =================
        constexpr uint32_t idesc = idesc_i8(128, N, /*a_signed=*/0, /*b_signed=*/1);
        const uint64_t dOnes = smem_desc_sw128(sOnes);
        for (int step = 0; step < F; ++step)
        {
            const uint32_t p = (uint32_t)(step & 1);
            wait_cluster(hfull0 + 8 * p, (uint32_t)((step >> 1) & 1));
            fence_proxy_async();
            tc_fence_after();
            const uint32_t bb = sB + p * kBBuf;
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                for (int kb = 0; kb < 4; ++kb)
                {
                    const uint64_t da = smem_desc_sw128(sA + (h2 * 4 + kb) * kATile);
                    const uint64_t db = smem_desc_sw128(bb + kb * kBTile);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) // 32 bytes (K = 32) per MMA
                        mma_i8(tmem_base + h2 * 64, da + 2 * ks, db + 2 * ks, idesc, (kb | ks) ? 1u : 0u);
                }
#pragma unroll
            for (int kb = 0; kb < 4; ++kb)
            {
                const uint64_t db = smem_desc_sw128(bb + kb * kBTile);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                    mma_i8(tmem_base + 128, dOnes + 2 * ks, db + 2 * ks, idesc, (kb | ks) ? 1u : 0u);
            }
            tc_commit(mma_done);
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all(); // no CTA may exit while a peer can still write into its shared memory
    if (warp == 8)
    {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// byte o of the image: CTA r, tile (half h, k-block kb), row m = 4*unit' + gate, 16-byte chunks XOR-swizzled
__global__ void whh_image_kernel(const uint8_t *__restrict__ q, uint8_t *__restrict__ img)
{
    const long o = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= (long)kC * kABytes)
        return;
    const int r = (int)(o / kABytes);
    const int rem = (int)(o % kABytes);
    const int tile = rem / kATile, within = rem % kATile;
    const int h = tile >> 2, kb = tile & 3;
    const int m = within >> 7, pos = within & 127;
    const int chunk = (pos >> 4) ^ (m & 7), b = pos & 15;
    const int k = 128 * kb + 16 * chunk + b;
    const int up = m >> 2, g = m & 3;
    const int row = g * kHhC + kUnits * r + 32 * h + up;
    img[o] = q[(long)row * kHhC + k];
}

template <int TP>
int launch_tp(const ClusterArgs &P, int groups, cudaStream_t st)
{
    static bool attr_set[64] = {};
    int dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&dev));
    if (!attr_set[dev & 63])
    {
        UMX_CUDA_OK(cudaFuncSetAttribute(lstm_cluster_kernel<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
        attr_set[dev & 63] = true;
    }
    dim3 grid(kC, 2 * P.a.target_count, groups);
    lstm_cluster_kernel<TP><<<grid, kThreads, kSmemBytes, st>>>(P);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}
} // namespace

size_t whh_image_bytes() { return (size_t)kC * kABytes; }

int launch_whh_image(const uint8_t *q, uint8_t *img, cudaStream_t st)
{
    whh_image_kernel<<<ceil_div((long)kC * kABytes, 256), 256, 0, st>>>(q, img);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

bool lstm_cluster_supported(int Hh) { return Hh == kHhC; }

int launch_lstm_cluster(const LstmLayerArgs &a, const uint8_t *const (*img)[2], cudaStream_t st)
{
    UMX_CHECK(a.Hh == kHhC, "lstm_cluster: built for 512 hidden units per direction (UMX-L)");
    ClusterArgs P;
    P.a = a;
    for (int t = 0; t < kTargets; ++t)
        for (int d = 0; d < 2; ++d)
            P.img[t][d] = img[t][d];
    const int groups = ceil_div(a.n_tracks, kMaxTracks);
    const int per = std::min(a.n_tracks, kMaxTracks);
    if (per <= 4)
        return launch_tp<4>(P, groups, st);
    if (per <= 8)
        return launch_tp<8>(P, groups, st);
    return launch_tp<16>(P, groups, st);
}

} // namespace umx
