// LSTM recurrence on the tensor cores (reference: the W_hh h_{t-1} + gate half of src/lstm.cpp:112-174).
//
// One thread-block CLUSTER of 8 CTAs per (target, direction, group of <= 32 tracks).  CTA r owns hidden
// units [64r, 64r+64): its 256 gate rows of W_hh stay RESIDENT IN TENSOR MEMORY for the whole segment as
// the raw uint8 quantised weights (128 KB = 256 of the SM's 512 TMEM columns; the other 256 hold the
// accumulators), so a step touches neither HBM nor shared memory for weights: the MMAs take A from TMEM and
// only the small h operand from SMEM.  Per step and per HALF of the group's tracks (two halves ping-pong,
// so one half's MMAs run while the other half's gates are computed):
//   * h_{t-1} of all 512 units x tracks sits in each CTA's SMEM as the MMA B operand, written as three
//     signed base-128 digits per value (h = sum_i d_i 2^(-6-7i), exact to 2^-21), one 64-byte-swizzled
//     tile per source CTA;
//   * one elected thread issues tcgen05.mma kind::i8: D[row][16*digit+track] = sum_k q[row][k] d[k]
//     (uint8 x int8 -> int32, exact) for the two 128-row halves of the CTA's gate rows;
//   * offset * sum_k h_k per track (W = scale*q + offset: the term the integer MMA cannot carry) comes from the digit
//     bytes themselves: the shipper warp sums its CTA's own tile with dp4a and ships the 16 partial sums on a second
//     mbarrier, a helper warp adds the 8 CTAs' partials while the MMAs run;
//   * 8 epilogue warps read TMEM with the 16x256b shape.  TMEM lane (gate row) 8*gate + unit' inside each 32-lane
//     quadrant and B row 16*digit + track make that shape hand thread (unit', c) all four gates and all three digits of
//     its unit for four tracks: no exchange between lanes.  They add the input projection and the offset term, apply the
//     activations, update c and h, and write the three digit bytes of the new h straight into this CTA's own tile; a
//     named barrier later, 8 lanes of the shipper warp send the tile to the 7 peers with bulk DSMEM copies
//     (cp.async.bulk shared::cta -> shared::cluster) whose bytes are counted on the peer's mbarrier: async proxy end
//     to end;
//   * the MMA warp posts expect_tx for the 7 peer tiles, waits, and issues the next step's MMAs;
//   * the MMA warp and two helper warps rebuild h from the shipped digits and write the layer output (fp16 hi/lo planes
//     for the next GEMM) off the critical path.
// There is no cluster-wide barrier in the loop; two h buffers per half alternate by step parity.
// h / c are loaded from and stored to the session so the state streams across segments (lstm.cpp:160-161).
// The input projection arrives gate-interleaved (column 4*unit + gate per direction): one 16-byte load per (unit, track).
#include "common.cuh"
#include "kernels.cuh"
#include "tc_common.cuh"

#include <cstdlib>
#include <type_traits>

namespace umx
{
using namespace tc;

namespace
{
constexpr int kUnits = 64;       // units per CTA
constexpr int kMaxNG = 4;        // at most four ping-pong groups of tracks per cluster
constexpr int kBTile = 48 * 64;  // one source CTA's tile: up to 48 (track, digit) rows x 64 units
// Everything that depends on the hidden size per direction HH: 512 (UMX-L) = clusters of 8 CTAs, 256 (UMX-HQ) = clusters of 4.
template <int HH>
struct Cfg
{
    static_assert(HH == 512 || HH == 256, "hidden units per direction this kernel is built for");
    static constexpr int kC = HH / kUnits;              // CTAs per cluster
    static constexpr int kABytes = 256 * HH;            // this CTA's 256 gate rows x HH k of W_hh, uint8 (128 / 64 KB), held in TMEM
    static constexpr int kACols = HH / 4;               // TMEM columns of one 128-row half of it (four k bytes per 32-bit column)
    static constexpr int kBBuf = kC * kBTile;           // one (half, parity) h buffer
    static constexpr int kBOff = 0;                     // [half][parity 2][source CTAs][48 rows][64 B], 64-byte swizzle
    static constexpr int kSOff = kBOff + 2 * kMaxNG * kBBuf;  // offset * sum_k h_k: [group][16 tracks] f32
    static constexpr int kPOff = kSOff + kMaxNG * 64;         // its per-CTA partial sums: [group][parity 2][source CTAs][16 tracks] f32
    static constexpr int kPBuf = kC * 64;
    static constexpr int kBarOff = kPOff + 2 * kMaxNG * kPBuf;
    static constexpr int kSmemBytes = kBarOff + 256;
    static_assert(kSmemBytes <= 232448, "dynamic shared memory budget of one CTA exceeded");
};
// 8 epilogue warps (one per TMEM lane quadrant and row half), 1 MMA warp, 1 shipper warp and 2 helper warps
constexpr int kThreads = 32 * 12; // 8 epilogue warps, MMA warp, shipper warp, 2 helper warps
constexpr uint32_t kTmemCols = 512; // [0,256): accumulators (half*128 + row half*64), [256, 256 + HH/2): W_hh (row half * HH/4 + k/4)
constexpr uint32_t kTmemA = 256;

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
// tcgen05.ld 16x256b: the warp reads 16 TMEM lanes x (8 * X) columns; thread t gets, for each group j of 8 columns,
// r[4j], r[4j+1] = lane t/4, columns 8j + 2(t%4) + {0,1} and r[4j+2], r[4j+3] = lane t/4 + 8, same columns
__device__ __forceinline__ void tmem_ld_16x256b_x4(uint32_t taddr, uint32_t *r)
{
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32"
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t taddr, uint32_t *r)
{
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_16x256b_x1(uint32_t taddr, uint32_t *r)
{
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// tcgen05.mma kind::i8 with A in tensor memory and the B descriptor given as (lo, hi) words
__device__ __forceinline__ void mma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, bool accumulate)
{
    if (accumulate)
        asm volatile("{\n\t.reg .b64 db;\n\t.reg .pred p;\n\tmov.b64 db, {%2, %3};\n\tsetp.eq.u32 p, 0, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], db, %4, p;\n\t}" ::"r"(d_tmem),
                     "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc)
                     : "memory");
    else
        asm volatile("{\n\t.reg .b64 db;\n\t.reg .pred p;\n\tmov.b64 db, {%2, %3};\n\tsetp.ne.u32 p, 0, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], db, %4, p;\n\t}" ::"r"(d_tmem),
                     "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc)
                     : "memory");
}
// 16 consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint4 &a, const uint4 &b, const uint4 &c, const uint4 &d)
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(taddr), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w), "r"(c.x), "r"(c.y),
                 "r"(c.z), "r"(c.w), "r"(d.x), "r"(d.y), "r"(d.z), "r"(d.w)
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
// prmt.b32 with sign-replicating selector nibbles (8 | byte): one byte of x, sign-extended (__byte_perm masks that bit off)
__device__ __forceinline__ int sext_byte(uint32_t x, uint32_t sel)
{
    int r;
    asm("prmt.b32 %0, %1, 0, %2;" : "=r"(r) : "r"(x), "r"(sel));
    return r;
}
__device__ __forceinline__ void st_shared_u8(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v)
{
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// K-major tile with 64-byte rows and 64-byte swizzle (16-byte chunk index ^= (row >> 1) & 3); 8-row groups 512 B apart
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)4 << 61);
}
// sigmoid(x) = 1 / (1 + 2^(-log2e x)), tanh(x) = 1 - 2 / (1 + 2^(2 log2e x))
__device__ __forceinline__ float sigmoid_fast(float x)
{
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return r;
}
__device__ __forceinline__ float tanh_fast(float x)
{
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.885390081777927f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return fmaf(-2.0f, r, 1.0f);
}

// base + scale * sum_i D_i 2^(-6-7i): the integer dot products of the three digit planes back to one float.
// D0 * 128 + D1 stays inside int32 (|D_i| <= 512 * 255 * 64); s13 = scale * 2^-13, s20 = scale * 2^-20.
__device__ __forceinline__ float gate_pre(uint32_t d0, uint32_t d1, uint32_t d2, float s13, float s20, float base)
{
    const int v = (int)d0 * 128 + (int)d1;
    return fmaf((float)v, s13, fmaf((float)(int)d2, s20, base));
}

// debug (compile with -DUMX_LSTM_TIMING): clock64 stamps of one step of CTA (0,0,0); UMX_LSTM_DEBUG=1 prints them
#ifdef UMX_LSTM_TIMING
__device__ long long g_lstm_dbg[32];
__device__ long long g_lstm_loop[256]; // cycles of the whole step loop per CTA
#define UMX_DBG_STAMP(i) do { if (dbg) g_lstm_dbg[i] = clock64(); } while (0)
#else
#define UMX_DBG_STAMP(i) do { } while (0)
#endif

struct ClusterArgs
{
    LstmLayerArgs a;
    const uint8_t *img[kTargets][2]; // pre-swizzled W_hh image per (target, dir): [8 CTAs][128 KB]
};

// The three digit bytes of h of (unit kk of this CTA, track trk of the half) go straight into this CTA's own tile of
// the h buffer: B row n = TP * digit + track, byte kk of the (swizzled) 64-byte row.  The 8 lanes of a track write
// 8 consecutive bytes; the four tracks of a warp land in different 16-byte chunks.
template <int TP>
__device__ __forceinline__ void publish_digits(float h, uint32_t trk, uint32_t tile, uint32_t kk)
{
    int v = __float2int_rn(h * 1048576.0f); // 2^20
    const int d2 = ((v + 64) & 127) - 64;
    v = (v - d2) >> 7;
    const int d1 = ((v + 64) & 127) - 64;
    v = (v - d1) >> 7; // d0 in [-64, 64]
    const int dig[3] = {v, d1, d2};
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
        const uint32_t n = (uint32_t)(TP * i) + trk;
        st_shared_u8(tile + n * 64u + ((((kk >> 4) ^ (n >> 1)) & 3u) << 4) + (kk & 15u), (uint32_t)dig[i]);
    }
}

// tracks per ping-pong group (8 or 16), groups per cluster (2 or 4), hidden units per direction; launched with a cluster of
// Cfg<HH>::kC CTAs along x (launch attribute)
template <int TP, int NG, int HH>
__global__ void __launch_bounds__(kThreads, 1) lstm_cluster_kernel(const ClusterArgs P)
{
    using C = Cfg<HH>;
    constexpr int kC = C::kC, kHhC = HH, kABytes = C::kABytes, kBBuf = C::kBBuf, kBOff = C::kBOff, kSOff = C::kSOff, kPOff = C::kPOff,
                  kPBuf = C::kPBuf, kBarOff = C::kBarOff;
    extern __shared__ __align__(1024) uint8_t smem[];
    const LstmLayerArgs &a = P.a;
    const int r = blockIdx.x; // rank in cluster (cluster spans x)
    const int target = a.target_begin + blockIdx.y / 2;
    const int dir = blockIdx.y & 1;
    static_assert(NG == 2 || NG == 4, "ping-pong groups");
    static_assert(NG * ((3 * TP + 15) / 16 * 16) * 2 <= 256, "accumulator columns");
    const int trk0 = blockIdx.z * (NG * TP);
    const int n_group = min(NG * TP, a.n_tracks - trk0); // tracks of this cluster
    const int n_halves = (n_group + TP - 1) / TP;       // ping-pong groups in use ("halves": there were two of them at first)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    static_assert(TP == 8 || TP == 16, "tracks per half");
    constexpr int N = (3 * TP + 15) / 16 * 16; // MMA N: rows TP * digit + track of the h operand, padded to a multiple of 16
    constexpr int NT = TP / 4;                 // tracks per group and epilogue thread
    constexpr int GPD = TP / 8;                // 8-column groups per digit
    constexpr int NE = 8;                      // epilogue warps; warp NE issues the MMAs, warp NE + 1 ships, NE + 2.. help
    constexpr int kBarCount = 32 * (NE + 1);   // named barrier: the epilogue warps arrive, the shipper waits
    constexpr uint32_t kTileBytes = 64u * N;
    constexpr uint32_t kStepBytes = (kC - 1) * kTileBytes; // the 7 peers' digit tiles

    if ((smem_u32(smem) & 1023u) != 0)
        __trap(); // the swizzled operand tiles need the 1024-byte alignment requested above

    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kBarOff);
    // hfull[half][parity] (2 NG), mma_done[half] (NG), sfull[half][parity] (2 NG), then the TMEM base address
    const uint32_t hfull0 = smem_u32(bars), mma_done0 = smem_u32(bars + 2 * NG), sfull0 = smem_u32(bars + 3 * NG);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 5 * NG);
    constexpr uint32_t kDCols = 256 / NG; // accumulator columns of one group: row half h2 at + kDCols / 2
    const uint32_t sB = smem_u32(smem) + kBOff, sS = smem_u32(smem) + kSOff, sP = smem_u32(smem) + kPOff;

    // ---- one-time setup: zeroed h buffers, barriers, TMEM, weights into TMEM ----
    {
        uint4 *bz = reinterpret_cast<uint4 *>(smem + kBOff);
        for (int i = threadIdx.x; i < (kBarOff - kBOff) / 16; i += kThreads)
            bz[i] = make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0)
    {
        for (int i = 0; i < 2 * NG; ++i)
            mbar_init(hfull0 + 8 * i, 2); // MMA warp (expect_tx of the peers' bytes) + this CTA's own shipper
        for (int i = 0; i < NG; ++i)
            mbar_init(mma_done0 + 8 * i, 2); // tcgen05.commit of the MMAs + the sum warp once the offset terms are written
        for (int i = 0; i < 2 * NG; ++i)
            mbar_init(sfull0 + 8 * i, 2); // sum warp (expect_tx of the peers' partial sums) + this CTA's own shipper
        fence_barrier_init();
    }
    if (warp == NE)
        tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (warp < 8)
    {
        // warp (qd, hh) fills TMEM lanes [32 qd, 32 qd + 32) of row half hh: lane = gate row 8*gate + unit', one
        // 512-byte row of the image per lane, four k bytes per 32-bit column
        const int qd = warp & 3, hh = warp >> 2;
        const uint4 *src = reinterpret_cast<const uint4 *>(P.img[target][dir] + (size_t)r * kABytes + (size_t)(hh * 128 + 32 * qd + lane) * kHhC);
        const uint32_t dst = tmem_base + ((uint32_t)(32 * qd) << 16) + kTmemA + (uint32_t)hh * (uint32_t)C::kACols;
#pragma unroll 2
        for (int c16 = 0; c16 < C::kACols / 16; ++c16)
            tmem_st16(dst + 16u * c16, __ldg(&src[4 * c16]), __ldg(&src[4 * c16 + 1]), __ldg(&src[4 * c16 + 2]), __ldg(&src[4 * c16 + 3]));
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cluster_sync_all(); // every CTA's barriers and zeroed buffers exist before anyone writes remotely

    const int F = a.F;
    // The layer output, off the critical path and shared by the four helper warps: h of one step for NC pairs of tracks
    // of a half, rebuilt from the digit bytes of this CTA's own tile (the exact value the recurrence continues with)
    // and split into the fp16 (hi, lo) planes the next GEMM reads.  A track's 64 units are one 128-byte row per plane;
    // lane (kk = 4 (lane % 16), sub = lane / 16) owns 4 units of track 2 c + sub of the half, two tracks per store.
    auto write_h = [&](int hstep, int whalf, uint32_t parity, int c0, auto nc_tag) {
        constexpr int NC = decltype(nc_tag)::value;
        const uint32_t wkk = 4u * (lane & 15), sub = lane >> 4;
        const int t = dir == 0 ? hstep : F - 1 - hstep;
        const uint8_t *tile = smem + kBOff + ((whalf * 2) + parity) * kBBuf + r * kBTile;
        uint32_t dg[NC][3];
#pragma unroll
        for (int c = 0; c < NC; ++c)
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
                const uint32_t n = (uint32_t)(TP * i + 2 * (c0 + c)) + sub;
                dg[c][i] = *reinterpret_cast<const uint32_t *>(tile + n * 64u + ((((wkk >> 4) ^ (n >> 1)) & 3u) << 4) + (wkk & 15u));
            }
#pragma unroll
        for (int c = 0; c < NC; ++c)
        {
            float hv[4];
            __half hi[4], lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e)
            {
                // byte e of each digit word, sign-extended (PRMT with the sign-replicate selector)
                const uint32_t sel = (uint32_t)e | ((8u | e) << 4) | ((8u | e) << 8) | ((8u | e) << 12);
                const int v = (sext_byte(dg[c][0], sel) << 14) + (sext_byte(dg[c][1], sel) << 7) + sext_byte(dg[c][2], sel);
                hv[e] = (float)v * (1.0f / 1048576.0f);
                split_half(hv[e], hi[e], lo[e]);
            }
            const int trk = trk0 + whalf * TP + 2 * (c0 + c) + (int)sub;
            if (trk < a.n_tracks)
            {
                const long o = (long)target * a.out_target_stride + ((long)trk * F + t) * a.ldo + a.col_offset + dir * kHhC + kUnits * r + (int)wkk;
                __half2 ph[2] = {__halves2half2(hi[0], hi[1]), __halves2half2(hi[2], hi[3])};
                __half2 pl[2] = {__halves2half2(lo[0], lo[1]), __halves2half2(lo[2], lo[3])};
                *reinterpret_cast<uint2 *>(a.out_hi + o) = *reinterpret_cast<uint2 *>(ph);
                *reinterpret_cast<uint2 *>(a.out_lo + o) = *reinterpret_cast<uint2 *>(pl);
                if (a.out_f32)
                    *reinterpret_cast<float4 *>(a.out_f32 + (long)target * a.out_f32_target_stride + ((long)trk * F + t) * (2 * kHhC) + dir * kHhC +
                                                kUnits * r + (int)wkk) = make_float4(hv[0], hv[1], hv[2], hv[3]);
            }
        }
    };
    // pairs of tracks (chunks) per half: [0, kW0) MMA warp, then kW1 for helper warp 0 (which also adds up the partial
    // sums) and kW2 for helper warp 1
    constexpr int kW0 = TP / 8, kW1 = (TP / 2 - kW0) / 2, kW2 = TP / 2 - kW0 - kW1; // MMA warp, helper 0, helper 1
    using W0Tag = std::integral_constant<int, kW0>;
    using W1Tag = std::integral_constant<int, kW1>;
    using W2Tag = std::integral_constant<int, kW2>;
    if (warp < NE)
    {
        // ================= epilogue warps =================
        // TMEM lane (= gate row of the MMA) 8*gate + unit' inside each 32-lane quadrant: the 16x256b load shape then
        // hands thread (unit' = lane / 4, cp = lane % 4) all four gates of its unit for the tracks
        // 8 (e / 2) + 2 cp + (e % 2), e < NT, and all three digits: no exchange between lanes is needed.
        const int qd = warp & 3, hh = warp >> 2, up = lane >> 2, cp = lane & 3;
        const uint32_t kk = (uint32_t)(32 * hh + 8 * qd + up); // hidden unit of this thread inside the CTA
        const int k = kUnits * r + (int)kk;
        const float scale = a.whh_scale[target][dir];
        const float s13 = scale * (1.0f / 8192.0f), s20 = scale * (1.0f / 1048576.0f);
        const uint32_t t_lane = tmem_base + ((uint32_t)(32 * qd) << 16);
        const int ldx = 8 * kHhC;
        const float *xrow = a.xproj + (long)target * a.xproj_target_stride + (long)dir * 4 * kHhC + 4 * k; // 4 gates of unit k
        const long trk_x = (long)F * ldx; // floats between tracks in xproj
        const long st_base = ((long)target * 2 + dir) * kHhC + k;
        auto trk_of = [&](int e) { return 8 * (e >> 1) + 2 * cp + (e & 1); };

        // After an epilogue thread has written its digits of a half: make them visible to the async proxy and
        // ARRIVE (non-blocking) on the half's named barrier; the shipper warp waits there and ships the tile.
        auto publish_tile = [&](int half) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            // (immediate barrier ids: a register id makes ptxas reserve all 16 named barriers and costs a few per cent)
            if (half == 0)
                asm volatile("bar.arrive 1, %0;" ::"n"(kBarCount) : "memory");
            else if (half == 1)
                asm volatile("bar.arrive 2, %0;" ::"n"(kBarCount) : "memory");
            else if (half == 2)
                asm volatile("bar.arrive 3, %0;" ::"n"(kBarCount) : "memory");
            else
                asm volatile("bar.arrive 4, %0;" ::"n"(kBarCount) : "memory");
        };
        // input projection (4 gates) of one step for this thread's tracks of a half
        // first layer of a short chunk: every frame from the first all-zero one on has the same input projection and only
        // that one row was computed (LstmLayerArgs::chunks); zrow = that frame per track of this thread, F when all rows exist
        int zrow[NG][NT];
#pragma unroll
        for (int half = 0; half < NG; ++half)
#pragma unroll
            for (int e = 0; e < NT; ++e)
            {
                const int trk = trk0 + half * TP + trk_of(e);
                zrow[half][e] = (a.chunks && trk < a.n_tracks) ? frames_reaching(a.chunks[trk].n_wave, F) : F;
            }
        auto load_xp = [&](float4(&dst)[NT], int half, int tt) {
            const int tb = trk0 + half * TP;
#pragma unroll
            for (int e = 0; e < NT; ++e)
            {
                const int trk = tb + trk_of(e);
                const int te = min(tt, zrow[half][e]);
                dst[e] = trk < a.n_tracks ? __ldg(reinterpret_cast<const float4 *>(xrow + (long)te * ldx + (long)trk * trk_x)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };

        float c[NG][NT], hlast[NG][NT];
        float4 xp[NG][NT];
#pragma unroll
        for (int half = 0; half < NG; ++half)
        {
            if (half >= n_halves)
                break;
            const int tb = trk0 + half * TP;
#pragma unroll
            for (int e = 0; e < NT; ++e)
            {
                const int trk = min(tb + trk_of(e), a.n_tracks - 1); // padded tracks mirror a real one, never stored
                const long so = (long)trk * a.state_track_stride + st_base;
                c[half][e] = a.c_state[so];
                hlast[half][e] = a.h_state[so];
                publish_digits<TP>(hlast[half][e], (uint32_t)trk_of(e), sB + (uint32_t)(half * 2) * kBBuf + (uint32_t)r * kBTile, kk);
            }
            load_xp(xp[half], half, dir == 0 ? 0 : F - 1);
            publish_tile(half);
        }

#ifdef UMX_LSTM_TIMING
        const long long loop_t0 = clock64();
#endif
        for (int step = 0; step < F; ++step)
        {
            const int tn = dir == 0 ? min(step + 1, F - 1) : max(F - 2 - step, 0);
            const uint32_t pn = (uint32_t)(step & 1) ^ 1u;
#pragma unroll
            for (int half = 0; half < NG; ++half)
            {
                if (half >= n_halves)
                    break;
                const uint32_t tile_out = sB + ((uint32_t)(half * 2) + pn) * kBBuf + (uint32_t)r * kBTile;

#ifdef UMX_LSTM_TIMING
                const bool dbg = (half < 2 && step == 200 && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
#endif
                UMX_DBG_STAMP(half * 4 + 0);
                // MMAs of this half-step complete, and offset * sum_k h_k per track written by the sum warps
                mbar_wait(mma_done0 + 8u * half, (uint32_t)(step & 1));
                tc_fence_after();
                UMX_DBG_STAMP(half * 4 + 1);
                // D[s][.]: 16-lane half s of the quadrant = gates 2s (lanes 0..7) and 2s + 1 (lanes 8..15); column
                // group j of 8 = digit (8 j) / TP, tracks (8 j) % TP + 2 cp + {0,1}
                uint32_t D[2][16 * GPD];
#pragma unroll
                for (int sgh = 0; sgh < 2; ++sgh)
                {
                    const uint32_t ta = t_lane + ((uint32_t)(16 * sgh) << 16) + (uint32_t)half * kDCols + (uint32_t)hh * (kDCols / 2);
                    if (TP == 16)
                    {
                        tmem_ld_16x256b_x4(ta, &D[sgh][0]);
                        tmem_ld_16x256b_x2(ta + 32, &D[sgh][16]);
                    }
                    else
                    {
                        tmem_ld_16x256b_x2(ta, &D[sgh][0]);
                        tmem_ld_16x256b_x1(ta + 16, &D[sgh][8]);
                    }
                }
                float os[NT];
#pragma unroll
                for (int e2 = 0; e2 < NT / 2; ++e2)
                {
                    float2 v;
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(sS + (uint32_t)(half * 64 + (8 * e2 + 2 * cp) * 4)) : "memory");
                    os[2 * e2] = v.x;
                    os[2 * e2 + 1] = v.y;
                }
                tc_wait_ld();
                float hn[NT];
#pragma unroll
                for (int e = 0; e < NT; ++e)
                {
                    // register of (gate, digit dg) of track e: loaded group dg * GPD + e / 2, element e % 2
                    auto Dv = [&](int gate, int dg) { return D[gate >> 1][4 * (dg * GPD + (e >> 1)) + 2 * (gate & 1) + (e & 1)]; };
                    const float xg[4] = {xp[half][e].x, xp[half][e].y, xp[half][e].z, xp[half][e].w};
                    float pre[4];
#pragma unroll
                    for (int gate = 0; gate < 4; ++gate)
                        pre[gate] = gate_pre(Dv(gate, 0), Dv(gate, 1), Dv(gate, 2), s13, s20, xg[gate] + os[e]); // lstm.cpp:132-140
                    const float i_t = sigmoid_fast(pre[0]), f_t = sigmoid_fast(pre[1]), g_t = tanh_fast(pre[2]), o_t = sigmoid_fast(pre[3]);
                    c[half][e] = fmaf(f_t, c[half][e], i_t * g_t); // lstm.cpp:152-154
                    hn[e] = o_t * tanh_fast(c[half][e]);           // lstm.cpp:155
                    hlast[half][e] = hn[e];
                }
#pragma unroll
                for (int e = 0; e < NT; ++e)
                    publish_digits<TP>(hn[e], (uint32_t)trk_of(e), tile_out, kk);
                // next step's input projection for this half: in flight while the other half is processed
                load_xp(xp[half], half, tn);
                UMX_DBG_STAMP(half * 4 + 2);
                publish_tile(half);
                UMX_DBG_STAMP(half * 4 + 3);
            }
        }
#ifdef UMX_LSTM_TIMING
        if (threadIdx.x == 0)
            g_lstm_loop[(blockIdx.x + 8 * (blockIdx.y + 8 * blockIdx.z)) & 255] = clock64() - loop_t0;
#endif
#pragma unroll
        for (int half = 0; half < NG; ++half)
        {
            if (half >= n_halves)
                break;
#pragma unroll
            for (int e = 0; e < NT; ++e)
            {
                const int trk = trk0 + half * TP + trk_of(e);
                if (trk < a.n_tracks)
                {
                    const long so = (long)trk * a.state_track_stride + st_base;
                    a.h_state[so] = hlast[half][e];
                    a.c_state[so] = c[half][e];
                }
            }
        }
    }
    else if (warp == NE + 1)
    {
        // ================= shipper warp =================
        // Per half-step: wait until the epilogue warps have arrived on the half's named barrier, ship the digit
        // tile to the 7 peers (lanes 0..7, one peer each) and arrive on this CTA's own barrier for the local copy.
        // Then sum the digit bytes of the tile (dp4a) into this CTA's share of offset * sum_k h_k per track
        // (W = scale*q + offset: the term the integer MMA cannot carry; exactly the h the MMAs see) and ship those
        // 64 bytes on a second barrier: they are needed only after the MMAs.
        const bool shipper = lane < kC;
        const uint32_t peer_b = shipper ? mapa(sB + (uint32_t)r * kBTile, (uint32_t)lane) : 0u;
        const uint32_t peer_p = shipper ? mapa(sP + (uint32_t)r * 64u, (uint32_t)lane) : 0u;
        const uint32_t peer_bar = shipper ? mapa(hfull0, (uint32_t)lane) : 0u;
        const uint32_t peer_sbar = shipper ? mapa(sfull0, (uint32_t)lane) : 0u;
        const float offset = a.whh_offset[target][dir];
        const float w0 = offset * (1.0f / 64.0f), w1 = offset * (1.0f / 8192.0f), w2 = offset * (1.0f / 1048576.0f);
        constexpr int kRounds = 3 * TP / 8; // 512-byte rounds of 8 rows: lane l holds a quarter of row 8 j + l / 4
        for (int step = -1; step < F; ++step) // step -1 publishes the initial state into parity 0
        {
            const uint32_t pn = (uint32_t)((step + 1) & 1);
            for (int half = 0; half < n_halves; ++half)
            {
                if (half == 0)
                    asm volatile("bar.sync 1, %0;" ::"n"(kBarCount) : "memory");
                else if (half == 1)
                    asm volatile("bar.sync 2, %0;" ::"n"(kBarCount) : "memory");
                else if (half == 2)
                    asm volatile("bar.sync 3, %0;" ::"n"(kBarCount) : "memory");
                else
                    asm volatile("bar.sync 4, %0;" ::"n"(kBarCount) : "memory");
                const uint32_t slot = (uint32_t)(half * 2) + pn;
                if (shipper)
                {
                    if (lane == r)
                        mbar_arrive(hfull0 + 8u * slot);
                    else
                        bulk_copy_to_peer(peer_b + slot * kBBuf, sB + slot * kBBuf + (uint32_t)r * kBTile, kTileBytes, peer_bar + 8u * slot);
                }
                // rows TP * digit + track: round j covers digit (8 j) / TP, tracks (8 j) % TP + l / 4
                const uint4 *tl = reinterpret_cast<const uint4 *>(smem + kBOff + slot * kBBuf + r * kBTile) + lane;
                uint4 v[kRounds];
#pragma unroll
                for (int j = 0; j < kRounds; ++j)
                    v[j] = tl[32 * j];
                int acc[kRounds];
#pragma unroll
                for (int j = 0; j < kRounds; ++j)
                {
                    acc[j] = __dp4a((int)v[j].x, 0x01010101, __dp4a((int)v[j].y, 0x01010101, 0)) +
                             __dp4a((int)v[j].z, 0x01010101, __dp4a((int)v[j].w, 0x01010101, 0));
                    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
                    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
                }
                const uint32_t s_own = sP + slot * kPBuf + (uint32_t)r * 64u;
                if ((lane & 3) == 0)
                {
#pragma unroll
                    for (int q = 0; q < TP / 8; ++q) // this lane's tracks: 8 q + l / 4
                        st_shared_f32(s_own + (uint32_t)(8 * q + (lane >> 2)) * 4u,
                                      fmaf((float)acc[q], w0, fmaf((float)acc[TP / 8 + q], w1, (float)acc[2 * (TP / 8) + q] * w2)));
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (shipper)
                {
                    if (lane == r)
                        mbar_arrive(sfull0 + 8u * slot);
                    else
                        bulk_copy_to_peer(peer_p + slot * kPBuf, s_own, 64u, peer_sbar + 8u * slot);
                }
            }
        }
    }
    else if (warp == NE)
    {
        // ================= MMA issuer: the whole warp runs the loop converged, one elected lane issues =================
        constexpr uint32_t idesc = idesc_i8(128, N, /*a_signed=*/0, /*b_signed=*/1);
        // B descriptor low word (address >> 4 | LBO << 16); the high word is constant
        const uint32_t b_lo0 = (uint32_t)(smem_desc_sw64(sB) & 0xffffffffu), b_hi = (uint32_t)(smem_desc_sw64(sB) >> 32);
        for (int step = 0; step <= F; ++step) // the extra round only absorbs the last publish
        {
            const uint32_t p = (uint32_t)(step & 1);
            for (int half = 0; half < n_halves; ++half)
            {
                const uint32_t slot = (uint32_t)(half * 2) + p;
#ifdef UMX_LSTM_TIMING
                const bool dbg = (half < 2 && (step == 200 || step == 201) && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
                const int dbo = 8 + (step - 200) * 8 + half * 4;
#endif
                UMX_DBG_STAMP(dbo + 0);
                if (lane == 0)
                    mbar_arrive_expect_tx(hfull0 + 8 * slot, kStepBytes); // the 7 peers' bulk copies complete this phase
                wait_cluster(hfull0 + 8 * slot, (uint32_t)((step >> 1) & 1));
                if (step == F)
                {
                    write_h(step - 1, half, p, 0, W0Tag{});
                    continue; // nothing left to compute: the bytes of the final publish have landed
                }
                tc_fence_after();
                UMX_DBG_STAMP(dbo + 1);
                const uint32_t b_lo = b_lo0 + slot * (kBBuf >> 4);
                if (elect_one())
                {
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
                        for (int ks = 0; ks < HH / 32; ++ks) // 32 bytes (K = 32) per MMA; B = the tile of source CTA ks / 2
                            mma_i8_ts(tmem_base + half * kDCols + h2 * (kDCols / 2), tmem_base + kTmemA + h2 * C::kACols + 8 * ks,
                                      b_lo + (((ks >> 1) * kBTile) >> 4) + 2 * (ks & 1), b_hi, idesc, ks != 0);
                    tc_commit(mma_done0 + 8u * half);
                }
                __syncwarp();
                UMX_DBG_STAMP(dbo + 2);
                if (step >= 1)
                    write_h(step - 1, half, p, 0, W0Tag{}); // its share of the previous step's h (own tile of this buffer)
            }
        }
    }
    else
    {
        // ================= helper warps =================
        // Warp 0: once the 8 CTAs' partial sums of a half-step have landed, add them up into offset * sum_k h_k per
        // track for the epilogue.  Both: their share of the layer output.
        const int sw = warp - (NE + 2);
        for (int step = 0; step <= F; ++step) // the extra round only writes out the last h
        {
            const uint32_t p = (uint32_t)(step & 1);
            for (int half = 0; half < n_halves; ++half)
            {
                const uint32_t slot = (uint32_t)(half * 2) + p;
#ifdef UMX_LSTM_TIMING
                const bool dbg = (half < 2 && (step == 200 || step == 201) && sw == 0 && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0);
                const int dbo = 8 + (step - 200) * 8 + half * 4;
#endif
                if (sw == 0)
                {
                    if (lane == 0)
                        mbar_arrive_expect_tx(sfull0 + 8 * slot, (kC - 1) * 64u);
                    wait_cluster(sfull0 + 8 * slot, (uint32_t)((step >> 1) & 1)); // (the round step == F only absorbs the last shipment)
                }
                if (sw == 0 && step < F)
                {
#ifdef UMX_LSTM_TIMING
                    if (dbg)
                        g_lstm_dbg[28 + (step - 200) * 2 + half] = clock64();
#endif
                    // lane l: track l % 16, source CTAs 4 (l / 16) .. + 3 (a cluster of 4 has only the first four)
                    const float *ps = reinterpret_cast<const float *>(smem + kPOff + slot * kPBuf) + (lane & 15) + (kC == 8 ? (lane >> 4) * 64 : 0);
                    float tot = (ps[0] + ps[16]) + (ps[32] + ps[48]);
                    if (kC == 8)
                        tot += __shfl_xor_sync(0xffffffffu, tot, 16);
                    if (lane < 16)
                        st_shared_f32(sS + (uint32_t)(half * 64 + lane * 4), tot);
                    __syncwarp();
                    UMX_DBG_STAMP(dbo + 3);
                    if (lane == 0)
                        mbar_arrive(mma_done0 + 8u * half); // release: the epilogue's wait acquires the sums
                }
                if (step == 0)
                    continue;
                // h of the previous step sits in this CTA's own tile of the buffer whose barrier completes now
                wait_cluster(hfull0 + 8 * slot, (uint32_t)((step >> 1) & 1));
                if (sw == 0)
                    write_h(step - 1, half, p, kW0, W1Tag{});
                else
                    write_h(step - 1, half, p, kW0 + kW1, W2Tag{});
#ifdef UMX_LSTM_TIMING
                if (dbg)
                    g_lstm_dbg[24 + (step - 200) * 2 + half] = clock64();
#endif
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all(); // no CTA may exit while a peer can still write into its shared memory
    if (warp == NE)
    {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// byte o of the image: CTA r, row half h, TMEM lane m, k: plain rows of HH bytes (one per TMEM lane)
__global__ void whh_image_kernel(const uint8_t *__restrict__ q, uint8_t *__restrict__ img, int HH)
{
    const long o = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int a_bytes = 256 * HH;
    if (o >= (long)(HH / kUnits) * a_bytes)
        return;
    const int r = (int)(o / a_bytes);
    const int rem = (int)(o % a_bytes);
    const int h = rem / (128 * HH), m = (rem / HH) & 127, k = rem % HH;
    const int qd = m >> 5, g = (m >> 3) & 3, up = 8 * qd + (m & 7); // lane 8*gate + unit' inside each 32-lane quadrant
    const int row = g * HH + kUnits * r + 32 * h + up;
    img[o] = q[(long)row * HH + k];
}

template <int TP, int NG, int HH>
int launch_tp(const ClusterArgs &P, cudaStream_t st)
{
    using C = Cfg<HH>;
    const int groups = ceil_div(P.a.n_tracks, TP * NG); // clusters per (target, direction)
    static PerDeviceOnce once; // one per template instantiation
    if (int rc = once.run([&](int) {
            UMX_CUDA_OK(cudaFuncSetAttribute(lstm_cluster_kernel<TP, NG, HH>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
            return 0;
        }))
        return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(C::kC, 2 * P.a.target_count, groups);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C::kC;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    UMX_CUDA_OK(cudaLaunchKernelEx(&cfg, lstm_cluster_kernel<TP, NG, HH>, P));
    UMX_CUDA_OK(cudaGetLastError());
#ifdef UMX_LSTM_TIMING
    if (getenv("UMX_LSTM_DEBUG"))
    {
        long long h[32];
        UMX_CUDA_OK(cudaStreamSynchronize(st));
        UMX_CUDA_OK(cudaMemcpyFromSymbol(h, g_lstm_dbg, sizeof h));
        {
            long long lp[256];
            UMX_CUDA_OK(cudaMemcpyFromSymbol(lp, g_lstm_loop, sizeof lp));
            fprintf(stderr, "cycles/step per cluster (CTA 0 .. 7):");
            for (int cl = 0; cl < 2 * P.a.target_count * groups; ++cl)
                fprintf(stderr, " [%lld..%lld]", lp[8 * cl] / P.a.F, lp[8 * cl + 7] / P.a.F);
            fprintf(stderr, "\n");
        }
        fprintf(stderr, "lstm_cluster<%d,%d> step 200 (cycles rel. to epilogue X wait begin)\n  epi X: done %+lld computed %+lld shipped %+lld | epi Y: wait %+lld done %+lld computed %+lld shipped %+lld\n"
                        "  mma200 X: wait %+lld hfull %+lld issued %+lld sums %+lld | Y: wait %+lld hfull %+lld issued %+lld sums %+lld\n  mma201 X: wait %+lld hfull %+lld issued %+lld sums %+lld | Y: wait %+lld hfull %+lld issued %+lld sums %+lld\n  writer done: %+lld %+lld %+lld %+lld | sum warp woke: %+lld %+lld %+lld %+lld\n",
                TP, NG, h[1] - h[0], h[2] - h[0], h[3] - h[0], h[4] - h[0], h[5] - h[0], h[6] - h[0], h[7] - h[0], h[8] - h[0], h[9] - h[0], h[10] - h[0], h[11] - h[0],
                h[12] - h[0], h[13] - h[0], h[14] - h[0], h[15] - h[0], h[16] - h[0], h[17] - h[0], h[18] - h[0], h[19] - h[0], h[20] - h[0], h[21] - h[0], h[22] - h[0], h[23] - h[0], h[24] - h[0], h[25] - h[0], h[26] - h[0], h[27] - h[0], h[28] - h[0], h[29] - h[0], h[30] - h[0], h[31] - h[0]);
    }
#endif
    return 0;
}
} // namespace

size_t whh_image_bytes(int Hh) { return (size_t)(Hh / kUnits) * 256 * Hh; }

int launch_whh_image(const uint8_t *q, uint8_t *img, int Hh, cudaStream_t st)
{
    UMX_CHECK(lstm_cluster_supported(Hh), "whh_image: unsupported hidden size");
    whh_image_kernel<<<ceil_div((long)whh_image_bytes(Hh), 256), 256, 0, st>>>(q, img, Hh);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

// UMX-L (512 units per direction: clusters of 8 CTAs) and UMX-HQ (256: clusters of 4)
bool lstm_cluster_supported(int Hh) { return Hh == 512 || Hh == 256; }

// how many 8-CTA clusters of the recurrence kernel the current device can run at once (0 on error)
int lstm_cluster_max_active()
{
    int n = 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(8, 64, 1);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Cfg<512>::kSmemBytes;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 8;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (cudaFuncSetAttribute(lstm_cluster_kernel<16, 2, 512>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<512>::kSmemBytes) != cudaSuccess)
        return 0;
    if (cudaOccupancyMaxActiveClusters(&n, lstm_cluster_kernel<16, 2, 512>, &cfg) != cudaSuccess)
        return 0;
    return n;
}

int launch_lstm_cluster(const LstmLayerArgs &a, const uint8_t *const (*img)[2], cudaStream_t st)
{
    UMX_CHECK(lstm_cluster_supported(a.Hh), "lstm_cluster: built for 512 (UMX-L) or 256 (UMX-HQ) hidden units per direction");
    ClusterArgs P;
    P.a = a;
    for (int t = 0; t < kTargets; ++t)
        for (int d = 0; d < 2; ++d)
            P.img[t][d] = img[t][d];
    // a cluster takes up to 32 tracks as two ping-pong groups: one group's MMAs and exchange run while the other's gates
    // are computed.  (Four groups of 8 — lstm_cluster_kernel<16, 2> — shorten the chain of a group but double the MMAs,
    // bulk copies and barrier traffic per step: measured 116 ms against 92 ms per bench step.)
    if (a.Hh == 512)
        return a.n_tracks <= 16 ? launch_tp<8, 2, 512>(P, st) : launch_tp<16, 2, 512>(P, st);
    return a.n_tracks <= 16 ? launch_tp<8, 2, 256>(P, st) : launch_tp<16, 2, 256>(P, st);
}

} // namespace umx
