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
//     write the new h as digits into this CTA's own tile of the h buffer; one bulk DSMEM copy per peer
//     (cp.async.bulk shared::cta -> shared::cluster) then ships the tile, its bytes counted on the peer's
//     mbarrier: 7 transactions per CTA per step, async proxy end to end;
//   * the MMA thread posts expect_tx for the 7 peer tiles, waits, and issues the next step's MMAs.
// There is no cluster-wide barrier in the loop; two h buffers alternate by step parity.
// h / c are loaded from and stored to the session so the state streams across segments (lstm.cpp:160-161).
#include "common.cuh"
#include "kernels.cuh"
#include "tc_common.cuh"

#include <cstdlib>

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
constexpr int kBOff = kOnesOff + kATile;          // h digits: [2 buffers][8 source CTAs][64 rows][64 B], 64-byte swizzle
constexpr int kBTile = 64 * 64;                   // one source CTA's 64 units x up to 64 (track, digit) rows
constexpr int kBBuf = kC * kBTile;
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

__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// tcgen05.mma kind::i8 with the descriptors given as (lo, hi) words and a compile-time accumulate flag
__device__ __forceinline__ void mma_i8_lohi(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                            bool accumulate)
{
    if (accumulate)
        asm volatile("{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.eq.u32 p, 0, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
                     "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc)
                     : "memory");
    else
        asm volatile("{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\tmov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.ne.u32 p, 0, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
                     "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc)
                     : "memory");
}

// debug (compile with -DUMX_LSTM_TIMING): clock64 stamps of one step of CTA (0,0,0); UMX_LSTM_DEBUG=1 prints them
__device__ long long g_lstm_dbg[16];
#ifdef UMX_LSTM_TIMING
#define UMX_DBG_STAMP(i) do { if (dbg) g_lstm_dbg[i] = clock64(); } while (0)
#else
#define UMX_DBG_STAMP(i) do { (void)dbg; } while (0)
#endif

struct ClusterArgs
{
    LstmLayerArgs a;
    const uint8_t *img[kTargets][2]; // pre-swizzled W_hh image per (target, dir): [8 CTAs][128 KB]
};

// gate activation with per-lane constants: sigmoid = 1 * rcp(1 + 2^(-log2e x)) + 0, tanh = -2 * rcp(1 + 2^(2 log2e x)) + 1
__device__ __forceinline__ float act_rcp(float x, float A, float B, float Cc)
{
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * B));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return fmaf(A, r, Cc);
}
__device__ __forceinline__ float tanh_fast(float x) { return act_rcp(x, -2.0f, 2.885390081777927f, 1.0f); }

// One bulk copy of this CTA's freshly written h tile into the same place of a peer CTA's shared memory;
// the bytes are counted on the peer's mbarrier (async proxy end to end: no fence on the consumer side).
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t remote_dst, uint32_t local_src, uint32_t bytes, uint32_t remote_bar)
{
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(remote_dst),
                 "r"(local_src), "r"(bytes), "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint32_t a, uint32_t b)
{
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
// K-major tile with 64-byte rows and 64-byte swizzle (16-byte chunk index ^= (row >> 1) & 3); 8-row groups 512 B apart
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)4 << 61);
}

// Lane (q, g) holds h of (unit q of this warp, track 4*jc + g).  The 8 lanes with the same g collect the
// 8 units' digit words, transpose them to per-digit byte rows and write them into this CTA's own tile of the
// h buffer (B row n = 4*track + digit, 8 consecutive k bytes); publish_tile() then ships the tile to the peers.
__device__ __forceinline__ void publish_quad(float h, int g, int jc, uint32_t tile, uint32_t chunk, uint32_t within)
{
    int v = __float2int_rn(h * 134217728.0f); // 2^27
    const int d3 = ((v + 64) & 127) - 64;
    v = (v - d3) >> 7;
    const int d2 = ((v + 64) & 127) - 64;
    v = (v - d2) >> 7;
    const int d1 = ((v + 64) & 127) - 64;
    v = (v - d1) >> 7; // d0
    const uint32_t w = ((uint32_t)v & 0xffu) | (((uint32_t)d1 & 0xffu) << 8) | (((uint32_t)d2 & 0xffu) << 16) | ((uint32_t)d3 << 24);
    uint32_t W[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
        W[u] = __shfl_sync(0xffffffffu, w, 4 * u + g);
    uint32_t lo[4], hi[4];
    {
        const uint32_t a0 = __byte_perm(W[0], W[1], 0x5140), a1 = __byte_perm(W[0], W[1], 0x7362);
        const uint32_t b0 = __byte_perm(W[2], W[3], 0x5140), b1 = __byte_perm(W[2], W[3], 0x7362);
        lo[0] = __byte_perm(a0, b0, 0x5410);
        lo[1] = __byte_perm(a0, b0, 0x7632);
        lo[2] = __byte_perm(a1, b1, 0x5410);
        lo[3] = __byte_perm(a1, b1, 0x7632);
        const uint32_t c0 = __byte_perm(W[4], W[5], 0x5140), c1 = __byte_perm(W[4], W[5], 0x7362);
        const uint32_t e0 = __byte_perm(W[6], W[7], 0x5140), e1 = __byte_perm(W[6], W[7], 0x7362);
        hi[0] = __byte_perm(c0, e0, 0x5410);
        hi[1] = __byte_perm(c0, e0, 0x7632);
        hi[2] = __byte_perm(c1, e1, 0x5410);
        hi[3] = __byte_perm(c1, e1, 0x7632);
    }
    const uint32_t n0 = 4u * (4u * jc + g); // row of digit 0 of this lane's track
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        const uint32_t n = n0 + i;
        st_shared_v2(tile + n * 64u + ((chunk ^ ((n >> 1) & 3u)) << 4) + within, lo[i], hi[i]);
    }
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
    constexpr int NQ = TP / 4; // track quads
    constexpr uint32_t kTileBytes = 64u * N;               // one CTA's contribution per step: N rows x 64 units
    constexpr uint32_t kStepBytes = (kC - 1) * kTileBytes; // bytes arriving from the 7 peers per step

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
        mbar_init(hfull0, 2);     // MMA thread (expect_tx of the peers' bytes) + this CTA's own epilogue
        mbar_init(hfull0 + 8, 2);
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
        const uint32_t kk0 = (uint32_t)(32 * hh + 8 * qd); // first unit of this warp inside the CTA's 64-byte rows
        const uint32_t chunk = kk0 >> 4, within = kk0 & 15u;
        const float scale = a.whh_scale[target][dir], offset = a.whh_offset[target][dir];
        // per-lane activation constants: gate 2 (g) is tanh, gates 0,1,3 (i,f,o) are sigmoids
        const float actA = (g == 2) ? -2.0f : 1.0f, actB = (g == 2) ? 2.885390081777927f : -1.4426950408889634f,
                    actC = (g == 2) ? 1.0f : 0.0f;
        const uint32_t t_lane = tmem_base + ((uint32_t)(32 * qd) << 16);
        // this CTA's own tile (source index r) of h buffer 0; lanes 0..7 of warp 0 ship it to CTA `lane`
        const uint32_t tile0 = sB + (uint32_t)r * kBTile;
        const bool shipper = (warp == 0 && lane < kC);
        const uint32_t peer_tile0 = shipper ? mapa(tile0, (uint32_t)lane) : 0u;
        const uint32_t peer_bar0 = shipper ? mapa(hfull0, (uint32_t)lane) : 0u;
        // after every epilogue thread has written its digits: make them visible to the async proxy, meet on a
        // named barrier, then 7 bulk copies (one per peer) + one local arrive
        auto publish_tile = [&](uint32_t pbuf) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (shipper)
            {
                if (lane == r)
                    mbar_arrive(hfull0 + 8u * pbuf);
                else
                    bulk_copy_to_peer(peer_tile0 + pbuf * kBBuf, tile0 + pbuf * kBBuf, kTileBytes, peer_bar0 + 8u * pbuf);
            }
        };
        const int ldx = 8 * kHhC;
        const float *xrow = a.xproj + (long)target * a.xproj_target_stride + (long)dir * 4 * kHhC + g * kHhC + k;
        const long trk_x = (long)F * ldx; // floats between tracks in xproj
        // unit-level ownership: lane (q, g) owns tracks 4*jc + g of unit k
        const long st_base = ((long)target * 2 + dir) * kHhC + k;
        const long out_base = (long)target * a.out_target_stride + a.col_offset + dir * kHhC + k;
        const long trk_o = (long)F * a.ldo;

        float c[NQ], hlast[NQ], xp[TP];
#pragma unroll
        for (int jc = 0; jc < NQ; ++jc)
        {
            const int trk = min(trk0 + 4 * jc + g, a.n_tracks - 1); // padded tracks mirror a real one, never stored
            const long so = (long)trk * a.state_track_stride + st_base;
            c[jc] = a.c_state[so];
            hlast[jc] = a.h_state[so];
            publish_quad(hlast[jc], g, jc, tile0, chunk, within); // buffer 0
        }
        publish_tile(0);
        {
            const int t0 = dir == 0 ? 0 : F - 1;
#pragma unroll
            for (int j = 0; j < TP; ++j)
                xp[j] = __ldg(xrow + (long)min(trk0 + j, a.n_tracks - 1) * trk_x + (long)t0 * ldx);
        }

        for (int step = 0; step < F; ++step)
        {
            const int t = dir == 0 ? step : F - 1 - step;
            const int tn = dir == 0 ? min(step + 1, F - 1) : max(F - 2 - step, 0);
            float xn[TP];
#pragma unroll
            for (int j = 0; j < TP; ++j) // next step's input projection, in flight during this step
                xn[j] = __ldg(xrow + (long)min(trk0 + j, a.n_tracks - 1) * trk_x + (long)tn * ldx);
            const uint32_t pn = (uint32_t)((step + 1) & 1);
            const uint32_t tile = tile0 + pn * kBBuf;

            const bool dbg = (step == 200 && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
            UMX_DBG_STAMP(0);
            mbar_wait(mma_done, (uint32_t)(step & 1));
            tc_fence_after();
            UMX_DBG_STAMP(1);
#pragma unroll
            for (int jc = 0; jc < NQ; ++jc)
            {
                uint32_t d[16], s[16];
                tmem_ld16(t_lane + hh * 64 + 16 * jc, d);
                tmem_ld16(t_lane + 128 + 16 * jc, s);
                tc_wait_ld();
                float act[4];
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                {
                    const float dot = combine_digits(d[4 * jj], d[4 * jj + 1], d[4 * jj + 2], d[4 * jj + 3]);
                    const float hs = combine_digits(s[4 * jj], s[4 * jj + 1], s[4 * jj + 2], s[4 * jj + 3]);
                    const float pre = xp[4 * jc + jj] + fmaf(scale, dot, offset * hs); // lstm.cpp:132-140
                    act[jj] = act_rcp(pre, actA, actB, actC);
                }
                // 4x4 transpose inside the quad: afterwards lane g holds the four gates of track 4*jc + g
                const bool b0 = g & 1, b1 = g & 2;
                const float keepA = b0 ? act[1] : act[0], keepB = b0 ? act[3] : act[2];
                const float recvA = __shfl_xor_sync(0xffffffffu, b0 ? act[0] : act[1], 1);
                const float recvB = __shfl_xor_sync(0xffffffffu, b0 ? act[2] : act[3], 1);
                // now: track (b0) gates {g: keepA, g^1: recvA}, track (2+b0) gates {g: keepB, g^1: recvB}
                const float v0 = b1 ? keepB : keepA, v1 = b1 ? recvB : recvA;             // gates g, g^1 of track g
                const float v2 = __shfl_xor_sync(0xffffffffu, b1 ? keepA : keepB, 2);     // gate g^2
                const float v3 = __shfl_xor_sync(0xffffffffu, b1 ? recvA : recvB, 2);     // gate g^3
                // value with gate label x sits at index g ^ x
                const float w0 = b0 ? v1 : v0, w1 = b0 ? v0 : v1, w2 = b0 ? v3 : v2, w3 = b0 ? v2 : v3; // labels (g&2)+{0,1}, (g&2)^2+{0,1}
                const float i_t = b1 ? w2 : w0, f_t = b1 ? w3 : w1, g_t = b1 ? w0 : w2, o_t = b1 ? w1 : w3;
                c[jc] = fmaf(f_t, c[jc], i_t * g_t);          // lstm.cpp:152-154
                const float h = o_t * tanh_fast(c[jc]);       // lstm.cpp:155
                hlast[jc] = h;
                publish_quad(h, g, jc, tile, chunk, within);
                const int trk = trk0 + 4 * jc + g;
                if (trk < a.n_tracks)
                {
                    const long o = out_base + (long)trk * trk_o + (long)t * a.ldo;
                    __half hh16, hl16;
                    split_half(h, hh16, hl16);
                    a.out_hi[o] = hh16;
                    a.out_lo[o] = hl16;
                    if (a.out_f32)
                        a.out_f32[(long)target * a.out_f32_target_stride + ((long)trk * F + t) * (2 * kHhC) + dir * kHhC + k] = h;
                }
            }
            UMX_DBG_STAMP(2);
            publish_tile(pn);
            UMX_DBG_STAMP(3);
#pragma unroll
            for (int j = 0; j < TP; ++j)
                xp[j] = xn[j];
        }
#pragma unroll
        for (int jc = 0; jc < NQ; ++jc)
        {
            const int trk = trk0 + 4 * jc + g;
            if (trk < a.n_tracks)
            {
                const long so = (long)trk * a.state_track_stride + st_base;
                a.h_state[so] = hlast[jc];
                a.c_state[so] = c[jc];
            }
        }
    }
    else
    {
        // ================= MMA issuer: the whole warp runs the loop converged, one elected lane issues =================
        constexpr uint32_t idesc = idesc_i8(128, N, /*a_signed=*/0, /*b_signed=*/1);
        // descriptor low words (address >> 4 | LBO << 16); high words are constants per swizzle mode
        const uint32_t a_lo = (uint32_t)(smem_desc_sw128(sA) & 0xffffffffu), a_hi = (uint32_t)(smem_desc_sw128(sA) >> 32);
        const uint32_t o_lo = (uint32_t)(smem_desc_sw128(sOnes) & 0xffffffffu);
        const uint32_t b_lo0 = (uint32_t)(smem_desc_sw64(sB) & 0xffffffffu), b_hi = (uint32_t)(smem_desc_sw64(sB) >> 32);
        for (int step = 0; step < F; ++step)
        {
            const uint32_t p = (uint32_t)(step & 1);
            const bool dbg = ((step == 200 || step == 201) && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
            const int dbo = step == 200 ? 4 : 8;
            UMX_DBG_STAMP(dbo + 0);
            if (lane == 0)
                mbar_arrive_expect_tx(hfull0 + 8 * p, kStepBytes); // the 7 peers' bulk copies complete this phase
            wait_cluster(hfull0 + 8 * p, (uint32_t)((step >> 1) & 1));
            tc_fence_after();
            UMX_DBG_STAMP(dbo + 1);
            const uint32_t b_lo = b_lo0 + p * (kBBuf >> 4);
            if (elect_one())
            {
#pragma unroll
                for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) // 32 bytes (K = 32) per MMA; B = the h tile of source CTA 2*kb + ks/2
                            mma_i8_lohi(tmem_base + h2 * 64, a_lo + (((h2 * 4 + kb) * kATile) >> 4) + 2 * ks, a_hi,
                                        b_lo + (((2 * kb + (ks >> 1)) * kBTile) >> 4) + 2 * (ks & 1), b_hi, idesc, (kb | ks) != 0);
#pragma unroll
                for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        mma_i8_lohi(tmem_base + 128, o_lo + 2 * ks, a_hi, b_lo + (((2 * kb + (ks >> 1)) * kBTile) >> 4) + 2 * (ks & 1),
                                    b_hi, idesc, (kb | ks) != 0);
                tc_commit(mma_done);
            }
            __syncwarp();
            UMX_DBG_STAMP(dbo + 2);
        }
        // the last step also published (into the other buffer): absorb those bytes so no transaction is
        // in flight towards this CTA when it exits
        const uint32_t p = (uint32_t)(F & 1);
        if (lane == 0)
            mbar_arrive_expect_tx(hfull0 + 8 * p, kStepBytes);
        wait_cluster(hfull0 + 8 * p, (uint32_t)((F >> 1) & 1));
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
#ifdef UMX_LSTM_TIMING
    if (getenv("UMX_LSTM_DEBUG"))
    {
        long long h[16];
        UMX_CUDA_OK(cudaStreamSynchronize(st));
        UMX_CUDA_OK(cudaMemcpyFromSymbol(h, g_lstm_dbg, sizeof h));
        fprintf(stderr, "lstm_cluster<%d> step 200: epi wait_begin 0, mma_done +%lld, stored +%lld, shipped +%lld | mma200: wait_begin %+lld, "
                        "hfull %+lld, issued %+lld | mma201: wait_begin %+lld, hfull %+lld, issued %+lld\n",
                TP, h[1] - h[0], h[2] - h[0], h[3] - h[0], h[4] - h[0], h[5] - h[0], h[6] - h[0], h[8] - h[0], h[9] - h[0], h[10] - h[0]);
    }
#endif
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

// how many 8-CTA clusters of the recurrence kernel the current device can run at once (0 on error)
int lstm_cluster_max_active()
{
    int n = 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(kC, 64, 1);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = kSmemBytes;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = kC;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (cudaFuncSetAttribute(lstm_cluster_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes) != cudaSuccess)
        return 0;
    if (cudaOccupancyMaxActiveClusters(&n, lstm_cluster_kernel<16>, &cfg) != cudaSuccess)
        return 0;
    return n;
}

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
