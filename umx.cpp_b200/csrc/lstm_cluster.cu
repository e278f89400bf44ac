// LSTM recurrence on the tensor cores (reference: the W_hh h_{t-1} + gate half of src/lstm.cpp:112-174).
//
// One thread-block CLUSTER of 8 CTAs per (target, direction, group of <= 32 tracks).  CTA r owns hidden
// units [64r, 64r+64): its 256 gate rows of W_hh stay RESIDENT IN SHARED MEMORY for the whole segment as
// the raw uint8 quantised weights (128 KB, pre-swizzled image built at model load), so a step never
// touches HBM for weights.  Per step and per HALF of the group's tracks (two halves ping-pong, so one
// half's MMAs run while the other half's gates are computed):
//   * h_{t-1} of all 512 units x tracks sits in each CTA's SMEM as the MMA B operand, written as three
//     signed base-128 digits per value (h = sum_i d_i 2^(-6-7i), exact to 2^-21), one 64-byte-swizzled
//     tile per source CTA;
//   * one elected thread issues tcgen05.mma kind::i8: D[row][3*track+digit] = sum_k q[row][k] d[k]
//     (uint8 x int8 -> int32, exact) for the two 128-row halves of the CTA's gate rows;
//   * 8 epilogue warps read TMEM (lane = gate row, ordered 4*unit + gate so one quad of lanes holds the four
//     gates of a unit), add the input projection and offset * sum_k h_k (W = scale*q + offset), apply the
//     activations, transpose inside the quad so each lane owns one (unit, track), update c and h, and
//     write the new h digits into this CTA's own tile; a named barrier later, 8 lanes ship the tile (and
//     the CTA's partial sum_k h_k) to the 7 peers with bulk DSMEM copies (cp.async.bulk shared::cta ->
//     shared::cluster) whose bytes are counted on the peer's mbarrier: async proxy end to end;
//   * the MMA warp posts expect_tx for the 7 peer tiles, waits, and issues the next step's MMAs.
// There is no cluster-wide barrier in the loop; two h buffers per half alternate by step parity.
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
constexpr int kMaxHalf = 16;     // tracks per half
constexpr int kMaxTracks = 2 * kMaxHalf; // tracks per cluster
constexpr int kATile = 128 * 128;                 // one (row half, k-block) tile of W_hh: 128 rows x 128 bytes
constexpr int kABytes = 2 * 4 * kATile;           // 128 KB
constexpr int kBTile = 48 * 64;                   // one source CTA's tile: up to 48 (track, digit) rows x 64 units
constexpr int kBBuf = kC * kBTile;                // one (half, parity) h buffer: 24 KB
constexpr int kBOff = kABytes;                    // [half 2][parity 2][8 source CTAs][48 rows][64 B], 64-byte swizzle
constexpr int kSOff = kBOff + 4 * kBBuf;          // partial sums of h: [half 2][parity 2][8 source CTAs][16 tracks] f32
constexpr int kSBuf = kC * 64;
constexpr int kBarOff = kSOff + 4 * kSBuf;
constexpr int kSmemBytes = kBarOff + 64;          // 231,488 B of the 232,448 B one CTA may have
// (the per-warp partial sums of h, 8 warps x 16 tracks, borrow the CTA's own tile of the buffer the current
//  step has just consumed: nobody reads or writes that tile again before the next publish)
static_assert(kSmemBytes <= 232448, "dynamic shared memory budget of one CTA exceeded");
constexpr int kThreads = 320;    // 8 epilogue warps + 1 MMA warp + 1 shipper warp
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
// One bulk copy from this CTA's shared memory into the same place of a peer CTA's; the bytes are counted
// on the peer's mbarrier (async proxy end to end: no fence on the consumer side).
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
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v)
{
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float ld_shared_f32(uint32_t addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
// K-major tile with 64-byte rows and 64-byte swizzle (16-byte chunk index ^= (row >> 1) & 3); 8-row groups 512 B apart
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)4 << 61);
}
// gate activation with per-lane constants: sigmoid = 1 * rcp(1 + 2^(-log2e x)) + 0, tanh = -2 * rcp(1 + 2^(2 log2e x)) + 1
__device__ __forceinline__ float act_rcp(float x, float A, float B, float Cc)
{
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * B));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return fmaf(A, r, Cc);
}
__device__ __forceinline__ float tanh_fast(float x) { return act_rcp(x, -2.0f, 2.885390081777927f, 1.0f); }

// sum_i D_i * 2^(-6-7i): the integer dot products of the three digit planes back to one float
__device__ __forceinline__ float combine_digits(uint32_t d0, uint32_t d1, uint32_t d2)
{
    return fmaf((float)(int)d2, 1.0f / 1048576.0f, fmaf((float)(int)d1, 1.0f / 8192.0f, (float)(int)d0 * (1.0f / 64.0f)));
}

// debug (compile with -DUMX_LSTM_TIMING): clock64 stamps of one step of CTA (0,0,0); UMX_LSTM_DEBUG=1 prints them
#ifdef UMX_LSTM_TIMING
__device__ long long g_lstm_dbg[32];
#define UMX_DBG_STAMP(i) do { if (dbg) g_lstm_dbg[i] = clock64(); } while (0)
#else
#define UMX_DBG_STAMP(i) do { } while (0)
#endif

struct ClusterArgs
{
    LstmLayerArgs a;
    const uint8_t *img[kTargets][2]; // pre-swizzled W_hh image per (target, dir): [8 CTAs][128 KB]
};

// Lane (q, g) holds h of (unit q of this warp, track 4*jc + g of the half).  The 8 lanes with the same g collect
// the 8 units' digit words, transpose them to per-digit byte rows and write them into this CTA's own tile of the
// h buffer: B row n = 3*track + digit, 8 consecutive k bytes.
__device__ __forceinline__ void publish_quad(float h, int g, int jc, uint32_t tile, uint32_t chunk, uint32_t within)
{
    int v = __float2int_rn(h * 1048576.0f); // 2^20
    const int d2 = ((v + 64) & 127) - 64;
    v = (v - d2) >> 7;
    const int d1 = ((v + 64) & 127) - 64;
    v = (v - d1) >> 7; // d0 in [-64, 64]
    const uint32_t w = ((uint32_t)v & 0xffu) | (((uint32_t)d1 & 0xffu) << 8) | (((uint32_t)d2 & 0xffu) << 16);
    uint32_t W[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
        W[u] = __shfl_sync(0xffffffffu, w, 4 * u + g);
    const uint32_t a0 = __byte_perm(W[0], W[1], 0x5140), a1 = __byte_perm(W[0], W[1], 0x7362);
    const uint32_t b0 = __byte_perm(W[2], W[3], 0x5140), b1 = __byte_perm(W[2], W[3], 0x7362);
    const uint32_t c0 = __byte_perm(W[4], W[5], 0x5140), c1 = __byte_perm(W[4], W[5], 0x7362);
    const uint32_t e0 = __byte_perm(W[6], W[7], 0x5140), e1 = __byte_perm(W[6], W[7], 0x7362);
    const uint32_t lo[3] = {__byte_perm(a0, b0, 0x5410), __byte_perm(a0, b0, 0x7632), __byte_perm(a1, b1, 0x5410)};
    const uint32_t hi[3] = {__byte_perm(c0, e0, 0x5410), __byte_perm(c0, e0, 0x7632), __byte_perm(c1, e1, 0x5410)};
    const uint32_t n0 = 3u * (4u * jc + g); // row of digit 0 of this lane's track
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const uint32_t n = n0 + i;
        st_shared_v2(tile + n * 64u + ((chunk ^ ((n >> 1) & 3u)) << 4) + within, lo[i], hi[i]);
    }
}

template <int TP> // tracks per half: 4, 8 or 16
__global__ void __cluster_dims__(kC, 1, 1) __launch_bounds__(kThreads, 1) lstm_cluster_kernel(const ClusterArgs P)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    const LstmLayerArgs &a = P.a;
    const int r = blockIdx.x; // rank in cluster (cluster spans x)
    const int target = a.target_begin + blockIdx.y / 2;
    const int dir = blockIdx.y & 1;
    const int trk0 = blockIdx.z * kMaxTracks;
    const int n_group = min(2 * TP, a.n_tracks - trk0); // tracks of this cluster
    const int n_halves = n_group > TP ? 2 : 1;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr int N = (3 * TP + 15) / 16 * 16; // MMA N: (track, digit) rows, padded to a multiple of 16
    constexpr int NQ = TP / 4;                 // track quads per half
    constexpr uint32_t kTileBytes = 64u * N;
    constexpr uint32_t kStepBytes = (kC - 1) * (kTileBytes + 64u); // from the 7 peers: digit tile + partial sums

    if ((smem_u32(smem) & 1023u) != 0)
        __trap(); // the swizzled operand tiles need the 1024-byte alignment requested above

    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kBarOff);
    // hfull[half][parity] (4), mma_done[half] (2), then the TMEM base address
    const uint32_t hfull0 = smem_u32(bars), mma_done0 = smem_u32(bars + 4);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 6);
    const uint32_t sA = smem_u32(smem), sB = sA + kBOff, sS = sA + kSOff;

    // ---- one-time setup: weights into SMEM, zeroed h buffers, barriers, TMEM ----
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(P.img[target][dir] + (size_t)r * kABytes);
        uint4 *dst = reinterpret_cast<uint4 *>(smem);
        for (int i = threadIdx.x; i < kABytes / 16; i += kThreads)
            dst[i] = __ldg(&src[i]);
        uint4 *bz = reinterpret_cast<uint4 *>(smem + kBOff);
        for (int i = threadIdx.x; i < (4 * kBBuf + 4 * kSBuf) / 16; i += kThreads)
            bz[i] = make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0)
    {
        for (int i = 0; i < 4; ++i)
            mbar_init(hfull0 + 8 * i, 2); // MMA warp (expect_tx of the peers' bytes) + this CTA's own shipper
        mbar_init(mma_done0, 1);
        mbar_init(mma_done0 + 8, 1);
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
        const int ldx = 8 * kHhC;
        const float *xrow = a.xproj + (long)target * a.xproj_target_stride + (long)dir * 4 * kHhC + g * kHhC + k;
        const long trk_x = (long)F * ldx; // floats between tracks in xproj
        const long st_base = ((long)target * 2 + dir) * kHhC + k;
        const long out_base = (long)target * a.out_target_stride + a.col_offset + dir * kHhC + k;
        const long trk_o = (long)F * a.ldo;

        // After an epilogue thread has written its digits (and per-warp partial sums of h) of a half: make them
        // visible to the async proxy and ARRIVE (non-blocking) on the half's named barrier; the shipper warp
        // waits there, finishes the sum over warps and ships tile + sums to the 7 peers.
        auto publish_tile = [&](int half) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            if (half == 0)
                asm volatile("bar.arrive 1, 288;" ::: "memory");
            else
                asm volatile("bar.arrive 2, 288;" ::: "memory");
        };
        // per-warp partial sum over its 8 units of this lane's h (track 4*jc + g), left in the scratch tile
        auto store_partial = [&](float h, int jc, uint32_t part) {
            float ssum = h;
            ssum += __shfl_xor_sync(0xffffffffu, ssum, 4);
            ssum += __shfl_xor_sync(0xffffffffu, ssum, 8);
            ssum += __shfl_xor_sync(0xffffffffu, ssum, 16);
            if (q == 0)
                st_shared_f32(part + (uint32_t)(warp * 64 + (4 * jc + g) * 4), ssum);
        };

        // input projection of one step for the TP tracks of a half.  The per-track addresses are a dependent
        // chain on purpose (the empty asm keeps the compiler from hoisting 2*TP loop-invariant 64-bit offsets
        // into registers, which spilled the hot loop).
        auto load_xp = [&](float(&dst)[TP], int tb, int tt) {
            const float *pt = xrow + (long)tt * ldx + (long)tb * trk_x;
            const int nv = a.n_tracks - tb;
#pragma unroll
            for (int j = 0; j < TP; ++j)
            {
                dst[j] = (j < nv) ? __ldg(pt) : 0.f;
                pt += trk_x;
                asm volatile("" : "+l"(pt));
            }
        };

        float c[2][NQ], hlast[2][NQ], xp[2][TP];
#pragma unroll
        for (int half = 0; half < 2; ++half)
        {
            if (half >= n_halves)
                break;
            const int tb = trk0 + half * TP;
            // the scratch for the partial sums: this CTA's own tile of the OTHER parity buffer (not in use yet)
            const uint32_t part = sB + (uint32_t)(half * 2 + 1) * kBBuf + (uint32_t)r * kBTile;
#pragma unroll
            for (int jc = 0; jc < NQ; ++jc)
            {
                const int trk = min(tb + 4 * jc + g, a.n_tracks - 1); // padded tracks mirror a real one, never stored
                const long so = (long)trk * a.state_track_stride + st_base;
                c[half][jc] = a.c_state[so];
                hlast[half][jc] = a.h_state[so];
                publish_quad(hlast[half][jc], g, jc, sB + (uint32_t)(half * 2) * kBBuf + (uint32_t)r * kBTile, chunk, within);
                store_partial(hlast[half][jc], jc, part);
            }
            load_xp(xp[half], tb, dir == 0 ? 0 : F - 1);
            publish_tile(half);
        }

        for (int step = 0; step < F; ++step)
        {
            const int t = dir == 0 ? step : F - 1 - step;
            const int tn = dir == 0 ? min(step + 1, F - 1) : max(F - 2 - step, 0);
            const uint32_t p = (uint32_t)(step & 1), pn = p ^ 1u;
#pragma unroll
            for (int half = 0; half < 2; ++half)
            {
                if (half >= n_halves)
                    break;
                const int tb = trk0 + half * TP;
                const uint32_t slot_in = (uint32_t)(half * 2) + p, slot_out = (uint32_t)(half * 2) + pn;
                const uint32_t tile_out = sB + slot_out * kBBuf + (uint32_t)r * kBTile;
                const uint32_t part = sB + slot_in * kBBuf + (uint32_t)r * kBTile; // consumed input tile = scratch

#ifdef UMX_LSTM_TIMING
                const bool dbg = (step == 200 && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
#endif
                UMX_DBG_STAMP(half * 4 + 0);
                mbar_wait(mma_done0 + 8u * half, (uint32_t)(step & 1));
                tc_fence_after();
                UMX_DBG_STAMP(half * 4 + 1);
                // sum_k h_k per track: the 8 CTAs' partial sums arrived with the digit tiles (the phase of
                // hfull[slot_in] completed before the MMAs we just waited for were issued)
                float s_tot = 0.f;
                if (lane < 16)
                {
#pragma unroll
                    for (int src = 0; src < kC; ++src)
                        s_tot += ld_shared_f32(sS + slot_in * kSBuf + (uint32_t)(src * 64 + lane * 4));
                    s_tot *= offset;
                }
                long o = out_base + (long)t * a.ldo + (long)(tb + g) * trk_o;
#pragma unroll
                for (int jc = 0; jc < NQ; ++jc)
                {
                    uint32_t d[16];
                    tmem_ld16(t_lane + half * 128 + hh * 64 + 12 * jc, d);
                    tc_wait_ld();
                    float act[4];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)
                    {
                        const float dot = combine_digits(d[3 * jj], d[3 * jj + 1], d[3 * jj + 2]);
                        const float os = __shfl_sync(0xffffffffu, s_tot, 4 * jc + jj);
                        const float pre = xp[half][4 * jc + jj] + fmaf(scale, dot, os); // lstm.cpp:132-140
                        act[jj] = act_rcp(pre, actA, actB, actC);
                    }
                    // 4x4 transpose inside the quad: afterwards lane g holds the four gates of track 4*jc + g
                    const bool b0 = g & 1, b1 = g & 2;
                    const float keepA = b0 ? act[1] : act[0], keepB = b0 ? act[3] : act[2];
                    const float recvA = __shfl_xor_sync(0xffffffffu, b0 ? act[0] : act[1], 1);
                    const float recvB = __shfl_xor_sync(0xffffffffu, b0 ? act[2] : act[3], 1);
                    const float v0 = b1 ? keepB : keepA, v1 = b1 ? recvB : recvA;         // gates g, g^1 of track g
                    const float v2 = __shfl_xor_sync(0xffffffffu, b1 ? keepA : keepB, 2); // gate g^2
                    const float v3 = __shfl_xor_sync(0xffffffffu, b1 ? recvA : recvB, 2); // gate g^3
                    // the value with gate label x sits at index g ^ x
                    const float w0 = b0 ? v1 : v0, w1 = b0 ? v0 : v1, w2 = b0 ? v3 : v2, w3 = b0 ? v2 : v3;
                    const float i_t = b1 ? w2 : w0, f_t = b1 ? w3 : w1, g_t = b1 ? w0 : w2, o_t = b1 ? w1 : w3;
                    c[half][jc] = fmaf(f_t, c[half][jc], i_t * g_t);     // lstm.cpp:152-154
                    const float h = o_t * tanh_fast(c[half][jc]);        // lstm.cpp:155
                    hlast[half][jc] = h;
                    publish_quad(h, g, jc, tile_out, chunk, within);
                    store_partial(h, jc, part);
                    const int trk = tb + 4 * jc + g;
                    if (trk < a.n_tracks)
                    {
                        __half hh16, hl16;
                        split_half(h, hh16, hl16);
                        a.out_hi[o] = hh16;
                        a.out_lo[o] = hl16;
                        if (a.out_f32)
                            a.out_f32[(long)target * a.out_f32_target_stride + ((long)trk * F + t) * (2 * kHhC) + dir * kHhC + k] = h;
                    }
                    o += 4 * trk_o;
                    asm volatile("" : "+l"(o));
                }
                // next step's input projection for this half: in flight while the other half is processed
                load_xp(xp[half], tb, tn);
                UMX_DBG_STAMP(half * 4 + 2);
                publish_tile(half);
                UMX_DBG_STAMP(half * 4 + 3);
            }
        }
#pragma unroll
        for (int half = 0; half < 2; ++half)
        {
            if (half >= n_halves)
                break;
#pragma unroll
            for (int jc = 0; jc < NQ; ++jc)
            {
                const int trk = trk0 + half * TP + 4 * jc + g;
                if (trk < a.n_tracks)
                {
                    const long so = (long)trk * a.state_track_stride + st_base;
                    a.h_state[so] = hlast[half][jc];
                    a.c_state[so] = c[half][jc];
                }
            }
        }
    }
    else if (warp == 9)
    {
        // ================= shipper warp =================
        // Per half-step: wait until the 8 epilogue warps have arrived on the half's named barrier, ship the
        // digit tile to the 7 peers (lanes 0..7, one peer each), then sum the per-warp partial sums of h,
        // ship those 64 bytes too, and arrive on this CTA's own barrier for the local copy of the data.
        const bool shipper = lane < kC;
        const uint32_t peer_b = shipper ? mapa(sB + (uint32_t)r * kBTile, (uint32_t)lane) : 0u;
        const uint32_t peer_s = shipper ? mapa(sS + (uint32_t)r * 64u, (uint32_t)lane) : 0u;
        const uint32_t peer_bar = shipper ? mapa(hfull0, (uint32_t)lane) : 0u;
        for (int step = -1; step < F; ++step) // step -1 publishes the initial state into parity 0
        {
            const uint32_t pn = (uint32_t)((step + 1) & 1);
            for (int half = 0; half < n_halves; ++half)
            {
                if (half == 0)
                    asm volatile("bar.sync 1, 288;" ::: "memory");
                else
                    asm volatile("bar.sync 2, 288;" ::: "memory");
                const uint32_t slot = (uint32_t)(half * 2) + pn;
                // partial sums sit in the own tile of the buffer the step consumed (the OTHER parity)
                const uint32_t part = sB + ((uint32_t)(half * 2) + (pn ^ 1u)) * kBBuf + (uint32_t)r * kBTile;
                const uint32_t s_own = sS + slot * kSBuf + (uint32_t)r * 64u;
                if (shipper && lane != r)
                    bulk_copy_to_peer(peer_b + slot * kBBuf, sB + slot * kBBuf + (uint32_t)r * kBTile, kTileBytes, peer_bar + 8u * slot);
                if (lane < 16)
                {
                    float tot = 0.f;
#pragma unroll
                    for (int w = 0; w < 8; ++w)
                        tot += ld_shared_f32(part + (uint32_t)(w * 64 + lane * 4));
                    st_shared_f32(s_own + lane * 4u, tot);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (shipper)
                {
                    if (lane == r)
                        mbar_arrive(hfull0 + 8u * slot);
                    else
                        bulk_copy_to_peer(peer_s + slot * kSBuf, s_own, 64u, peer_bar + 8u * slot);
                }
            }
        }
    }
    else
    {
        // ================= MMA issuer: the whole warp runs the loop converged, one elected lane issues =================
        constexpr uint32_t idesc = idesc_i8(128, N, /*a_signed=*/0, /*b_signed=*/1);
        // descriptor low words (address >> 4 | LBO << 16); high words are constants per swizzle mode
        const uint32_t a_lo = (uint32_t)(smem_desc_sw128(sA) & 0xffffffffu), a_hi = (uint32_t)(smem_desc_sw128(sA) >> 32);
        const uint32_t b_lo0 = (uint32_t)(smem_desc_sw64(sB) & 0xffffffffu), b_hi = (uint32_t)(smem_desc_sw64(sB) >> 32);
        for (int step = 0; step <= F; ++step) // the extra round only absorbs the last publish
        {
            const uint32_t p = (uint32_t)(step & 1);
            for (int half = 0; half < n_halves; ++half)
            {
                const uint32_t slot = (uint32_t)(half * 2) + p;
#ifdef UMX_LSTM_TIMING
                const bool dbg = ((step == 200 || step == 201) && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
                const int dbo = 8 + (step - 200) * 8 + half * 4;
#endif
                UMX_DBG_STAMP(dbo + 0);
                if (lane == 0)
                    mbar_arrive_expect_tx(hfull0 + 8 * slot, kStepBytes); // the 7 peers' bulk copies complete this phase
                wait_cluster(hfull0 + 8 * slot, (uint32_t)((step >> 1) & 1));
                if (step == F)
                    continue; // nothing left to compute: the bytes of the final publish have landed
                tc_fence_after();
                UMX_DBG_STAMP(dbo + 1);
                const uint32_t b_lo = b_lo0 + slot * (kBBuf >> 4);
                if (elect_one())
                {
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                        for (int kb = 0; kb < 4; ++kb)
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) // 32 bytes (K = 32) per MMA; B = the tile of source CTA 2*kb + ks/2
                                mma_i8_lohi(tmem_base + half * 128 + h2 * 64, a_lo + (((h2 * 4 + kb) * kATile) >> 4) + 2 * ks, a_hi,
                                            b_lo + (((2 * kb + (ks >> 1)) * kBTile) >> 4) + 2 * (ks & 1), b_hi, idesc, (kb | ks) != 0);
                    tc_commit(mma_done0 + 8u * half);
                }
                __syncwarp();
                UMX_DBG_STAMP(dbo + 2);
            }
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
#ifdef UMX_LSTM_TIMING
    if (getenv("UMX_LSTM_DEBUG"))
    {
        long long h[32];
        UMX_CUDA_OK(cudaStreamSynchronize(st));
        UMX_CUDA_OK(cudaMemcpyFromSymbol(h, g_lstm_dbg, sizeof h));
        fprintf(stderr, "lstm_cluster<%d> step 200 (cycles rel. to epilogue X wait begin)\n  epi X: done %+lld computed %+lld shipped %+lld | epi Y: wait %+lld done %+lld computed %+lld shipped %+lld\n"
                        "  mma200 X: wait %+lld hfull %+lld issued %+lld | Y: wait %+lld hfull %+lld issued %+lld\n  mma201 X: wait %+lld hfull %+lld issued %+lld | Y: wait %+lld hfull %+lld issued %+lld\n",
                TP, h[1] - h[0], h[2] - h[0], h[3] - h[0], h[4] - h[0], h[5] - h[0], h[6] - h[0], h[7] - h[0], h[8] - h[0], h[9] - h[0], h[10] - h[0],
                h[12] - h[0], h[13] - h[0], h[14] - h[0], h[16] - h[0], h[17] - h[0], h[18] - h[0], h[20] - h[0], h[21] - h[0], h[22] - h[0]);
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
    // a cluster takes up to 32 tracks as two ping-pong halves of TP tracks each
    const int groups = ceil_div(a.n_tracks, kMaxTracks);
    const int per = std::min(a.n_tracks, kMaxTracks);
    if (per <= 4)
        return launch_tp<4>(P, groups, st); // one half of 4
    if (per <= 8)
        return launch_tp<4>(P, groups, st); // two halves of 4
    if (per <= 16)
        return launch_tp<8>(P, groups, st); // two halves of 8
    return launch_tp<16>(P, groups, st);
}

} // namespace umx
