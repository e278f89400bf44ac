// Tensor-core GEMM for the dense layers of the network (reference: the Eigen GEMMs at
// src/inference.cpp:86,127,143 and the W_ih x_t half of src/lstm.cpp:132-136, for ALL frames at once).
//
//   C[m,n] = act(alpha[n] * sum_k A[m,k] B[n,k] + rowvec[m] * gamma[n] + beta[n])
//
// sm_100a design: persistent CTAs (one per SM), 128x256 output tiles walked column-fastest (the CTAs running together
// share a band of row tiles, so the activations stream from HBM once and the weights stay in L2), fp16 operands
// staged by TMA (cp.async.bulk.tensor, 64/128-byte swizzle) through a 4..6-stage mbarrier ring, tcgen05.mma
// (M128 N<=256 K16, one issuing thread) accumulating in fp32 in TMEM, two accumulator buffers (2 x 256 TMEM columns)
// so the epilogue of tile i overlaps the MMAs of tile i+1.  Warp roles: 0 = TMA producer, 1 = MMA issuer
// (+ TMEM alloc), 2..9 = epilogue, two warps per TMEM lane quadrant (tcgen05.ld -> SMEM transpose in 16-column
// chunks -> BN/bias/activation -> fp32 or fp16 hi/lo stores), with one branch-free row loop per epilogue kind.
// Precision: every operand is a (hi, lo) fp16 pair; `terms` selects hi*hi [+ lo*hi [+ hi*lo]] into the
// same accumulator, i.e. 1, 2 or 3 MMA passes per k-block ("GEMM precision").  The W_ih GEMM passes the
// centred quantised integers as a single exact plane and folds the dequantisation offset in as the rank-1 term.
#include "common.cuh"
#include "kernels.cuh"
#include "tc_common.cuh"

#include <cstring>
#include <mutex>

namespace umx
{
using namespace tc;

namespace
{
constexpr int BM = 128, BN = 256;
// One pipeline stage holds, for one k-block, every operand plane the chosen precision needs, so each plane is
// fetched from L2 once per k-block however many MMA passes use it:
//   TERMS 1: A_hi, B_hi              k-block 64 (128-byte rows, SWIZZLE_128B)   48 KB/stage, 4 stages
//   TERMS 2: A_hi, A_lo, B_hi        k-block 32 ( 64-byte rows, SWIZZLE_64B)    32 KB/stage, 6 stages
//   TERMS 3: A_hi, A_lo, B_hi, B_lo  k-block 32 ( 64-byte rows, SWIZZLE_64B)    48 KB/stage, 4 stages
// PAIR: two CTAs of a cluster (one TPC) work on one 256 x 256 tile with tcgen05.mma.cta_group::2: each CTA stages its own 128
// rows of A and HALF of the B tile (128 of the 256 weight rows); the tensor cores of both SMs read both halves.  Per CTA and
// k-block the B traffic — TMA writes and MMA operand reads — halves, which is what bounds the K = 1024 launches.
template <int TERMS, bool PAIR = false>
struct Cfg
{
    static constexpr int BK = TERMS == 1 ? 64 : 32;
    static constexpr int nA = TERMS >= 2 ? 2 : 1, nB = TERMS == 3 ? 2 : 1;
    static constexpr int kBRows = PAIR ? BN / 2 : BN; // B rows this CTA stages
    static constexpr int kAPlane = BM * BK * 2, kBPlane = kBRows * BK * 2;
    static constexpr int kStageBytes = nA * kAPlane + nB * kBPlane;
    static constexpr int kStages = PAIR ? 6 : (TERMS == 2 ? 6 : 4);
    static constexpr int kTileM = PAIR ? 2 * BM : BM; // rows of one scheduled tile
};
constexpr int kEpiWarps = 8;  // two per TMEM lane quadrant, each takes one half of the tile's columns
// Transpose tile of one epilogue warp: 32 rows x 16 columns of fp32, 64-byte rows, the four 16-byte chunks of row r stored at
// chunk position c ^ ((r >> 1) & 3).  Both sides are then free of bank conflicts: the writer (lane = row, one chunk per
// instruction) covers all 8 16-byte bank groups with every 8 lanes, and so does the reader (4 lanes per row, 2 rows per 8 lanes).
// (A padded pitch of 20 floats was conflict-free for the writes only: 105 M conflicts per W_ih launch in ncu.)
constexpr int kEpiPitch = 16;
__device__ __forceinline__ uint32_t epi_swz(int row, int chunk) { return (uint32_t)(row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4)); }
constexpr int kEpiBytesPerWarp = 32 * kEpiPitch * 4;
constexpr int kEpiChunks = BN / 2 / 16; // 16-column chunks per epilogue warp and tile
constexpr int kMaxStages = 6;
constexpr int kSched = 4; // depth of the tile-index ring between the producer and the MMA / epilogue warps
constexpr int kSmemBytes = 4 * 49152 + kEpiWarps * kEpiBytesPerWarp + 256 + 1024; // stages + epilogue + barriers + align slack
constexpr int kThreads = 32 * (2 + kEpiWarps);
constexpr uint32_t kTmemCols = 512;

struct TcProblem
{
    int M, N, K;
    int act;
    const float *alpha;
    const float *beta;
    const float *rowvec;
    const float *gamma;
    float *out_f32;
    __half *out_hi;
    __half *out_lo;
    int ldo;
    int n_group;
    long group_stride;
    int m_tiles, n_tiles, tile_begin; // tiles [tile_begin, tile_begin + m_tiles*n_tiles)
    int rows_per_track;               // with chunks: rows are (track, frame); see GemmProblem
    int extra_rows;
    const IstftTrack *chunks;
};

struct alignas(64) TcParams
{
    CUtensorMap a_hi[kMaxGemmProblems], a_lo[kMaxGemmProblems], b_hi[kMaxGemmProblems], b_lo[kMaxGemmProblems];
    TcProblem p[kMaxGemmProblems];
    int count, total_tiles, terms, tile_m; // tile_m: rows of one scheduled tile (128, or 256 for CTA pairs)
    // null: tiles blockIdx.x, + gridDim.x, ...; else int[2] {next tile, CTAs finished}, both zero at launch: tiles are drawn from
    // the first, and the last CTA to finish puts both back to zero for the next launch (no memset between launches: a memset
    // node may be served by a copy engine and then queues behind whatever host<->device copies that engine has pending)
    int *tile_counter;
};

__device__ __forceinline__ float act_apply(float v, int act)
{
    if (act == kActTanh)
        return tanhf(v);
    if (act == kActRelu)
        return fmaxf(v, 0.f);
    return v;
}

// 64-byte-row K-major tile, 64-byte swizzle: 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t smem_desc_sw64(uint32_t smem_addr)
{
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)4 << 61);
}

// The rows of one 16-column chunk after the transpose through shared memory: lane (sub, c4) owns columns
// [c4, c4 + 4) of the rows 8 it + sub.  ACT < 0: activation / outputs / rank-1 term decided at run time.
struct EpiRows
{
    uint32_t stage_buf;
    int sub, c4, row0, M, N, n, ldo;
    long obase;
    float *__restrict__ of32;
    __half *__restrict__ ohi;
    __half *__restrict__ olo;
};
template <int ACT, bool RV, bool F32, bool HALF>
__device__ __forceinline__ void epi_rows(const EpiRows &e, const float (&al)[4], const float (&be)[4], const float (&ga)[4],
                                         const float (&rvv)[4], int act_rt = 0, bool rv_rt = true)
{
    const bool full4 = e.n + 3 < e.N;
#pragma unroll
    for (int it = 0; it < 4; ++it)
    {
        const int row = it * 8 + e.sub;
        const int m = e.row0 + row;
        float v[4];
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3])
                     : "r"(e.stage_buf + epi_swz(row, e.c4 >> 2))
                     : "memory");
#pragma unroll
        for (int c = 0; c < 4; ++c)
        {
            const float b = (RV && (ACT >= 0 || rv_rt)) ? fmaf(rvv[it], ga[c], be[c]) : be[c];
            v[c] = fmaf(v[c], al[c], b);
            if (ACT == kActTanh)
                v[c] = tanhf(v[c]);
            else if (ACT == kActRelu)
                v[c] = fmaxf(v[c], 0.f);
            else if (ACT < 0)
                v[c] = act_apply(v[c], act_rt);
        }
        if (m >= e.M || e.n >= e.N)
            continue;
        const long o = e.obase + (long)m * e.ldo;
        if (F32 && (ACT >= 0 || e.of32))
        {
            if (full4)
                *reinterpret_cast<float4 *>(e.of32 + o) = make_float4(v[0], v[1], v[2], v[3]);
            else
                for (int c = 0; c < 4; ++c)
                    if (e.n + c < e.N)
                        e.of32[o + c] = v[c];
        }
        if (HALF && (ACT >= 0 || e.ohi))
        {
            __half h[4], l[4];
#pragma unroll
            for (int c = 0; c < 4; ++c)
                split_half(v[c], h[c], l[c]);
            if (full4)
            {
                *reinterpret_cast<uint2 *>(e.ohi + o) = make_uint2((uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16),
                                                                   (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16));
                *reinterpret_cast<uint2 *>(e.olo + o) = make_uint2((uint32_t)__half_as_ushort(l[0]) | ((uint32_t)__half_as_ushort(l[1]) << 16),
                                                                   (uint32_t)__half_as_ushort(l[2]) | ((uint32_t)__half_as_ushort(l[3]) << 16));
            }
            else
                for (int c = 0; c < 4; ++c)
                    if (e.n + c < e.N)
                    {
                        e.ohi[o + c] = h[c];
                        e.olo[o + c] = l[c];
                    }
        }
    }
}

// ---- CTA-pair (cta_group::2) forms ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_cluster(uint32_t local_addr, uint32_t cta)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
    return r;
}
__device__ __forceinline__ void cluster_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA tile load whose bytes are counted on a barrier that may live in the peer CTA of the pair (the leader's)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar_cluster_addr)
{
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar_cluster_addr)
                 : "memory");
}
__device__ __forceinline__ void mma_f16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                 : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far have retired
__device__ __forceinline__ void tc_commit_pair(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void st_cluster_u32(uint32_t cluster_addr, uint32_t v)
{
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}
// wait on a LOCAL barrier whose arrivals (and the data they publish) come from the peer CTA: cluster-scope acquire
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity)
{
    const long long t0 = clock64();
    for (;;)
    {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(bar), "r"(parity)
                     : "memory");
        if (ok)
            return;
        if (clock64() - t0 > 4000000000LL) // a protocol bug must abort the kernel, never hang the GPU
            __trap();
    }
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_dst, uint32_t cols) // one warp of EACH CTA of the pair
{
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t cols)
{
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

template <int TERMS, bool PAIR>
__global__ void __launch_bounds__(kThreads, 1) gemm_tc_kernel(const __grid_constant__ TcParams P)
{
    using C = Cfg<TERMS, PAIR>;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u; // CTA of the pair: 0 = leader (issues the MMAs)
    constexpr int BK = C::BK, kStages = C::kStages, kStageBytes = C::kStageBytes, kAPlane = C::kAPlane, kBPlane = C::kBPlane;
    static_assert(kStages * kStageBytes <= 4 * 49152 && kStages <= kMaxStages, "stage ring does not fit");
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment for the 128-byte swizzle atoms
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t *epi_base = smem + 4 * 49152;
    uint64_t *bars = reinterpret_cast<uint64_t *>(epi_base + kEpiWarps * kEpiBytesPerWarp);
    // bars: full[kMaxStages], empty[kMaxStages], tmem_full[2], tmem_empty[2], then the TMEM base address
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kMaxStages);
    const uint32_t tfull0 = smem_u32(bars + 2 * kMaxStages), tempty0 = smem_u32(bars + 2 * kMaxStages + 2);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 4);
    // tile schedule: the producer publishes the index of every tile this CTA works on (or -1: no more) through a small ring, so
    // the tiles may come from a global counter (CTAs that get their SM late, or share the GPU with other kernels, take fewer)
    const uint32_t sfull0 = smem_u32(bars + 2 * kMaxStages + 5), sempty0 = sfull0 + 8 * kSched;
    volatile int *sched_tile = reinterpret_cast<volatile int *>(bars + 2 * kMaxStages + 5 + 2 * kSched);
    // CTA pairs drawing from the global counter: the LEADER draws and hands every tile index to both rings; the peer's ten
    // readers (its producer, its MMA thread, its 8 epilogue warps) release a slot on the leader's psempty barrier
    const uint32_t psempty0 = smem_u32(bars + 2 * kMaxStages + 5 + 2 * kSched + 2);
    const bool pair_dyn = PAIR && P.tile_counter != nullptr;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0)
    {
        for (int i = 0; i < kStages; ++i)
        {
            mbar_init(full0 + 8 * i, 1);
            mbar_init(empty0 + 8 * i, 1);
        }
        for (int i = 0; i < 2; ++i)
        {
            mbar_init(tfull0 + 8 * i, 1);
            mbar_init(tempty0 + 8 * i, PAIR ? 2 * kEpiWarps : kEpiWarps); // pair: the leader's barrier collects both CTAs' epilogues
        }
        for (int i = 0; i < kSched; ++i)
        {
            mbar_init(sfull0 + 8 * i, 1);
            mbar_init(sempty0 + 8 * i, 1 + kEpiWarps); // the MMA thread and one lane per epilogue warp
            mbar_init(psempty0 + 8 * i, 2 + kEpiWarps);
        }
        fence_barrier_init();
    }
    if (warp == 1)
    {
        if (PAIR)
            tmem_alloc_pair(smem_u32(tmem_slot), kTmemCols);
        else
            tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (PAIR)
        cluster_sync(); // both CTAs' barriers exist before a TMA load or an arrive of the peer can reach them
    const uint32_t tmem_base = *tmem_slot;

    auto locate = [&](int tile, int &pi, int &m0, int &n0) {
        pi = 0;
        while (pi + 1 < P.count && tile >= P.p[pi + 1].tile_begin)
            ++pi;
        const int local = tile - P.p[pi].tile_begin;
        // n fastest: the CTAs running together cover a band of ~grid / n_tiles row tiles x all column tiles, so the
        // activations (the big operand) stream from HBM once while the weights (a few MB) stay in L2
        n0 = (local % P.p[pi].n_tiles) * BN;
        m0 = (local / P.p[pi].n_tiles) * C::kTileM;
    };
    // A row tile none of whose rows is a frame that reaches its track's chunk is not computed (all three warp roles skip
    // the same tiles, so their pipeline counters stay in step).
    auto needed = [&](int pi, int m0) -> bool {
        const IstftTrack *ch = P.p[pi].chunks;
        if (!ch)
            return true;
        const int F = P.p[pi].rows_per_track;
        const int t0 = m0 / F, f0 = m0 - t0 * F;
        if (f0 < frames_reaching(ch[t0].n_wave, F) + P.p[pi].extra_rows)
            return true;
        const int t1 = (min(m0 + C::kTileM, P.p[pi].M) - 1) / F;
        for (int t = t0 + 1; t <= t1; ++t) // later tracks start inside this tile at frame 0
            if (ch[t].n_wave > 0 || P.p[pi].extra_rows)
                return true;
        return false;
    };


    // a reader of the tile ring is done with slot `sslot`
    auto release_slot = [&](int sslot) {
        if (pair_dyn && rank != 0)
            mbar_arrive_cluster(mapa_cluster(psempty0 + 8 * sslot, 0));
        else
            mbar_arrive(sempty0 + 8 * sslot);
    };
    auto wait_slot = [&](int sslot, uint32_t sphase) {
        if (pair_dyn && rank != 0)
            mbar_wait_cluster(sfull0 + 8 * sslot, sphase);
        else
            mbar_wait(sfull0 + 8 * sslot, sphase);
    };

    if (warp == 0)
    {
        // ================= TMA producer =================
        if (lane == 0)
        {
            int stage = 0;
            uint32_t phase = 0;
            int sslot = 0;
            uint32_t sphase = 0;
            // static stride: every CTA (both CTAs of a pair alike) derives its tiles itself.  Global counter: the CTA — of a pair,
            // the leader — draws them
            int *const counter = (PAIR && rank != 0) ? nullptr : P.tile_counter;
            const int stride = PAIR ? (int)gridDim.x / 2 : (int)gridDim.x;
            int cand = counter ? atomicAdd(counter, 1) : (PAIR ? (int)blockIdx.x / 2 : (int)blockIdx.x);
            for (;;)
            {
                int tile = -1, pi = 0, m0 = 0, n0 = 0;
                if (pair_dyn && rank != 0)
                { // the peer's producer reads what the leader drew
                    mbar_wait_cluster(sfull0 + 8 * sslot, sphase);
                    tile = sched_tile[sslot];
                    release_slot(sslot);
                    if (tile >= 0)
                        locate(tile, pi, m0, n0);
                }
                else
                {
                    // next tile that has to be computed (row tiles beyond every chunk are skipped here, once, for all warp roles)
                    while (cand < P.total_tiles)
                    {
                        locate(cand, pi, m0, n0);
                        const int c = cand;
                        cand = counter ? atomicAdd(counter, 1) : cand + stride; // the next draw is in flight during this tile's loads
                        if (needed(pi, m0))
                        {
                            tile = c;
                            break;
                        }
                    }
                    mbar_wait(sempty0 + 8 * sslot, sphase ^ 1);
                    sched_tile[sslot] = tile;
                    if (pair_dyn)
                    { // ... and the peer's ring: its readers are done with the slot, the index goes over, its full barrier follows
                        mbar_wait_cluster(psempty0 + 8 * sslot, sphase ^ 1);
                        st_cluster_u32(mapa_cluster(smem_u32(const_cast<int *>(sched_tile) + sslot), 1), (uint32_t)tile);
                        mbar_arrive_cluster(mapa_cluster(sfull0 + 8 * sslot, 1));
                    }
                    mbar_arrive(sfull0 + 8 * sslot);
                }
                if (++sslot == kSched)
                {
                    sslot = 0;
                    sphase ^= 1;
                }
                if (tile < 0)
                    break;
                const int kblocks = (P.p[pi].K + BK - 1) / BK;
                for (int kb = 0; kb < kblocks; ++kb)
                {
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                    if (PAIR)
                    { // this CTA's 128 rows of A and its half of the B tile; every byte is counted on the LEADER's full barrier
                        const uint32_t bar = mapa_cluster(full0 + 8 * stage, 0);
                        if (rank == 0)
                            mbar_arrive_expect_tx(full0 + 8 * stage, 2 * kStageBytes);
                        const int ma = m0 + (int)rank * BM, nb = n0 + (int)rank * C::kBRows;
                        tma_load_2d_pair(sa, &P.a_hi[pi], kb * BK, ma, bar);
                        if (C::nA == 2)
                            tma_load_2d_pair(sa + kAPlane, &P.a_lo[pi], kb * BK, ma, bar);
                        tma_load_2d_pair(sa + C::nA * kAPlane, &P.b_hi[pi], kb * BK, nb, bar);
                        if (C::nB == 2)
                            tma_load_2d_pair(sa + C::nA * kAPlane + kBPlane, &P.b_lo[pi], kb * BK, nb, bar);
                    }
                    else
                    {
                        const uint32_t bar = full0 + 8 * stage;
                        mbar_arrive_expect_tx(bar, kStageBytes);
                        tma_load_2d(sa, &P.a_hi[pi], kb * BK, m0, bar);
                        if (C::nA == 2)
                            tma_load_2d(sa + kAPlane, &P.a_lo[pi], kb * BK, m0, bar);
                        tma_load_2d(sa + C::nA * kAPlane, &P.b_hi[pi], kb * BK, n0, bar);
                        if (C::nB == 2)
                            tma_load_2d(sa + C::nA * kAPlane + kBPlane, &P.b_lo[pi], kb * BK, n0, bar);
                    }
                    if (++stage == kStages)
                    {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    }
    else if (warp == 1)
    {
        // ================= MMA issuer =================
        if (lane == 0)
        {
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            int sslot = 0;
            uint32_t sphase = 0;
            for (;;)
            {
                wait_slot(sslot, sphase);
                const int tile = sched_tile[sslot];
                release_slot(sslot);
                if (++sslot == kSched)
                {
                    sslot = 0;
                    sphase ^= 1;
                }
                if (tile < 0)
                    break;
                if (PAIR && rank != 0)
                    continue; // the leader issues for both CTAs; this thread only keeps the tile ring moving
                int pi, m0, n0;
                locate(tile, pi, m0, n0);
                const int kiters = (P.p[pi].K + BK - 1) / BK;
                // a ragged last column tile only pays for the columns it has (N of the MMA in steps of 16); a pair always issues
                // the full width: its B rows are split between the two CTAs at N / 2
                const uint32_t idesc = PAIR ? idesc_f16_f32acc(2 * BM, BN) : idesc_f16_f32acc(BM, min(BN, (P.p[pi].N - n0 + 15) & ~15));
                mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1); // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int it = 0; it < kiters; ++it)
                {
                    mbar_wait(full0 + 8 * stage, phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * kStageBytes);
                    const uint32_t sb = sa + C::nA * kAPlane;
                    auto desc = [](uint32_t addr) { return TERMS == 1 ? smem_desc_sw128(addr) : smem_desc_sw64(addr); };
                    const uint64_t a_hi = desc(sa), b_hi = desc(sb);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) // +32 bytes per K=16 step inside the swizzle row
                    {
                        auto mma = [&](uint64_t da, uint64_t db, uint32_t accum) {
                            if (PAIR)
                                mma_f16_pair(d_tmem, da, db, idesc, accum);
                            else
                                mma_f16(d_tmem, da, db, idesc, accum);
                        };
                        mma(a_hi + 2 * k, b_hi + 2 * k, (it > 0 || k > 0) ? 1u : 0u); // hi * hi
                        if (C::nA == 2)
                            mma(desc(sa + kAPlane) + 2 * k, b_hi + 2 * k, 1u);          // lo * hi
                        if (C::nB == 2)
                            mma(a_hi + 2 * k, desc(sb + kBPlane) + 2 * k, 1u);          // hi * lo
                    }
                    // frees the smem slot (in both CTAs of a pair) when these MMAs retire
                    if (PAIR)
                        tc_commit_pair(empty0 + 8 * stage);
                    else
                        tc_commit(empty0 + 8 * stage);
                    if (++stage == kStages)
                    {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (PAIR)
                    tc_commit_pair(tfull0 + 8 * acc); // accumulator complete, in both CTAs
                else
                    tc_commit(tfull0 + 8 * acc);
                if (++acc == 2)
                {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    }
    else
    {
        // ================= epilogue =================
        // TMEM lane = output row.  Each 16-column chunk is transposed through a per-warp SMEM tile with
        // 128-bit accesses, so that 4 lanes cover one row's 16 columns and a store instruction writes 8 rows x 64 bytes.
        const int q = warp & 3;           // TMEM lane quadrant this warp may read
        const int part = (warp - 2) >> 2; // which half of the tile's columns
        // shared-window address of this warp's transpose tile (explicit ld/st.shared: a generic pointer costs the
        // long-latency generic path on every access)
        const uint32_t stage_buf = smem_u32(epi_base + (warp - 2) * kEpiBytesPerWarp);
        const int sub = lane >> 2, c4 = (lane & 3) * 4;
        int acc = 0;
        uint32_t acc_phase = 0;
        int sslot = 0;
        uint32_t sphase = 0;
        for (;;)
        {
            wait_slot(sslot, sphase);
            const int tile = sched_tile[sslot];
            __syncwarp();
            if (lane == 0)
                release_slot(sslot);
            if (++sslot == kSched)
            {
                sslot = 0;
                sphase ^= 1;
            }
            if (tile < 0)
                break;
            int pi, m0, n0;
            locate(tile, pi, m0, n0);
            // hoist the problem description into registers once per tile
            const int M = P.p[pi].M, N = P.p[pi].N, ldo = P.p[pi].ldo, act = P.p[pi].act, n_group = P.p[pi].n_group;
            const long gstride = P.p[pi].group_stride;
            const float *__restrict__ alpha = P.p[pi].alpha;
            const float *__restrict__ beta = P.p[pi].beta;
            const float *__restrict__ rowvec = P.p[pi].rowvec;
            const float *__restrict__ gamma = P.p[pi].gamma;
            float *__restrict__ of32 = P.p[pi].out_f32;
            __half *__restrict__ ohi = P.p[pi].out_hi;
            __half *__restrict__ olo = P.p[pi].out_lo;
            const int row0 = m0 + (int)rank * BM + 32 * q; // a pair's second CTA holds rows 128..255 of the tile
            // the rank-1 term's row factors of this thread's 8 rows: loaded once per tile, before waiting for the MMAs
            float rvv[4];
#pragma unroll
            for (int it = 0; it < 4; ++it)
                rvv[it] = (rowvec && row0 + it * 8 + sub < M) ? __ldg(rowvec + row0 + it * 8 + sub) : 0.f;
            // a tile lies inside one output group when the groups are whole tiles (or there is only one)
            const int grp_tile = n_group >= N ? 0 : (n_group % BN == 0 ? n0 / n_group : -1);
            mbar_wait(tfull0 + 8 * acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + acc * BN + part * (BN / 2) + ((uint32_t)(32 * q) << 16);
#pragma unroll 1
            for (int ch = 0; ch < kEpiChunks; ++ch)
            {
                // per-column constants of this chunk: in flight while the accumulators come out of TMEM
                const int nc0 = n0 + part * (BN / 2) + ch * 16;
                const int n = nc0 + c4;
                float al[4], be[4], ga[4];
#pragma unroll
                for (int e = 0; e < 4; ++e)
                {
                    const bool ok = n + e < N;
                    al[e] = (alpha && ok) ? __ldg(alpha + n + e) : 1.f;
                    be[e] = (beta && ok) ? __ldg(beta + n + e) : 0.f;
                    ga[e] = (rowvec && ok) ? __ldg(gamma + n + e) : 0.f;
                }
                uint32_t r[16];
                tmem_ld16(taddr + ch * 16, r);
                tc_wait_ld();
                if (ch == kEpiChunks - 1)
                { // everything this warp needs has left TMEM
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0)
                    {
                        if (PAIR && rank != 0)
                            mbar_arrive_cluster(mapa_cluster(tempty0 + 8 * acc, 0));
                        else
                            mbar_arrive(tempty0 + 8 * acc);
                    }
                }
                if (nc0 >= N)
                    continue;
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(stage_buf + epi_swz(lane, i)), "r"(r[4 * i]),
                                 "r"(r[4 * i + 1]), "r"(r[4 * i + 2]), "r"(r[4 * i + 3])
                                 : "memory");
                __syncwarp();
                const int grp = grp_tile >= 0 ? grp_tile : n / n_group;
                const long obase = (long)grp * gstride + (n - grp * n_group);
                const EpiRows er{stage_buf, sub, c4, row0, M, N, n, ldo, obase, of32, ohi, olo};
                // one specialised, branch-free row loop per epilogue the pipeline uses; anything else takes the generic one
                if (of32 && !ohi && act == kActNone && rowvec)
                    epi_rows<kActNone, true, true, false>(er, al, be, ga, rvv);  // W_ih: exact-integer GEMM + rank-1 term
                else if (of32 && !ohi && act == kActRelu && !rowvec)
                    epi_rows<kActRelu, false, true, false>(er, al, be, ga, rvv); // fc3
                else if (!of32 && ohi && act == kActTanh && !rowvec)
                    epi_rows<kActTanh, false, false, true>(er, al, be, ga, rvv); // fc1
                else if (!of32 && ohi && act == kActRelu && !rowvec)
                    epi_rows<kActRelu, false, false, true>(er, al, be, ga, rvv); // fc2
                else
                    epi_rows<-1, true, true, true>(er, al, be, ga, rvv, act, rowvec != nullptr);
                __syncwarp();
            }
            if (++acc == 2)
            {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (PAIR)
        cluster_sync(); // no CTA of a pair leaves (or frees tensor memory) while its peer may still signal or read it
    if (warp == 1)
    {
        tc_fence_after();
        if (PAIR)
            tmem_dealloc_pair(tmem_base, kTmemCols);
        else
            tmem_dealloc(tmem_base, kTmemCols);
    }
    if (P.tile_counter && threadIdx.x == 0)
    { // this CTA's last draw is long done (the producer left its loop): the last CTA out re-arms the queue
        if (atomicAdd(P.tile_counter + 1, 1) == (int)gridDim.x - 1)
        {
            P.tile_counter[0] = 0;
            P.tile_counter[1] = 0;
            __threadfence();
        }
    }
}

// ---- host side -----------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 2-D fp16 K-major operand: dims (K, rows), row pitch ld elements, box (bk, box_rows), swizzle = row bytes
int make_map(CUtensorMap *map, const __half *ptr, int K, int rows, int ld, int box_rows, int bk)
{
    EncodeTiledFn fn = get_encode_fn();
    UMX_CHECK(fn != nullptr, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    UMX_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld % 8) == 0, "GEMM operand must be 16-byte aligned with a 16-byte row pitch");
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    const cuuint32_t box[2] = {(cuuint32_t)bk, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half *>(ptr), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
    {
        set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
        return -1;
    }
    return 0;
}
} // namespace

int launch_gemm_tc(const GemmBatch &batch, int terms, cudaStream_t st, int *tile_counter, bool pair)
{
    UMX_CHECK(batch.count >= 1 && batch.count <= kMaxGemmProblems, "gemm_tc: bad problem count");
    UMX_CHECK(terms >= 1 && terms <= 3, "gemm_tc: terms must be 1..3");
    TcParams P;
    std::memset(&P, 0, sizeof P);
    P.count = batch.count;
    P.terms = terms;
    int tiles = 0;
    for (int i = 0; i < batch.count; ++i)
    {
        const GemmProblem &g = batch.p[i];
        UMX_CHECK(g.M > 0 && g.N > 0 && g.K > 0, "gemm_tc: empty problem");
        UMX_CHECK(g.ldo % 4 == 0 && g.n_group % 4 == 0 && g.group_stride % 4 == 0,
                  "gemm_tc: output pitch, column group and group stride must be multiples of 4 (vector stores)");
        UMX_CHECK((terms < 2 || g.A.lo != nullptr) && (terms < 3 || g.B.lo != nullptr), "gemm_tc: the chosen terms need lo planes");
        UMX_CHECK(g.rowvec == nullptr || g.gamma != nullptr, "gemm_tc: rowvec needs gamma");
        const int bk = terms == 1 ? 64 : 32;
        const int b_rows = pair ? BN / 2 : BN; // a CTA of a pair stages half of the B tile
        if (make_map(&P.a_hi[i], g.A.hi, g.K, g.M, g.A.ld, BM, bk) || make_map(&P.b_hi[i], g.B.hi, g.K, g.N, g.B.ld, b_rows, bk) ||
            (terms >= 2 && make_map(&P.a_lo[i], g.A.lo, g.K, g.M, g.A.ld, BM, bk)) ||
            (terms >= 3 && make_map(&P.b_lo[i], g.B.lo, g.K, g.N, g.B.ld, b_rows, bk)))
            return -1;
        TcProblem &p = P.p[i];
        p.M = g.M;
        p.N = g.N;
        p.K = g.K;
        p.act = g.act;
        p.alpha = g.alpha;
        p.beta = g.beta;
        p.rowvec = g.rowvec;
        p.gamma = g.gamma;
        p.out_f32 = g.out_f32;
        p.out_hi = g.out_hi;
        p.out_lo = g.out_lo;
        p.ldo = g.ldo;
        p.n_group = g.n_group;
        p.group_stride = g.group_stride;
        p.chunks = g.chunks;
        p.rows_per_track = g.rows_per_track;
        p.extra_rows = g.extra_rows;
        UMX_CHECK(!g.chunks || g.rows_per_track > 0, "gemm_tc: rows_per_track missing");
        p.m_tiles = ceil_div(g.M, pair ? 2 * BM : BM);
        p.n_tiles = ceil_div(g.N, BN);
        p.tile_begin = tiles;
        tiles += p.m_tiles * p.n_tiles;
    }
    P.total_tiles = tiles;
    P.tile_counter = tile_counter;
    P.tile_m = pair ? 2 * BM : BM;

    static PerDeviceOnce once;
    static int sm_count[64] = {};
    if (int rc = once.run([&](int dev) {
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
            UMX_CUDA_OK(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
            return 0;
        }))
        return rc;
    int dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&dev));
    if (pair)
    { // clusters of two CTAs (one TPC); one pair per tile at a time
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(2 * std::min(tiles, sm_count[dev] / 2));
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = kSmemBytes;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        if (terms == 1)
            UMX_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<1, true>, P));
        else if (terms == 2)
            UMX_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<2, true>, P));
        else
            UMX_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<3, true>, P));
        UMX_CUDA_OK(cudaGetLastError());
        return 0;
    }
    const int grid = std::min(tiles, sm_count[dev]);
    if (terms == 1)
        gemm_tc_kernel<1, false><<<grid, kThreads, kSmemBytes, st>>>(P);
    else if (terms == 2)
        gemm_tc_kernel<2, false><<<grid, kThreads, kSmemBytes, st>>>(P);
    else
        gemm_tc_kernel<3, false><<<grid, kThreads, kSmemBytes, st>>>(P);
    UMX_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace umx
