// Host orchestration of the demixing path: umx_inference (src/inference.cpp:12-207) and
// shift_inference / split_inference (umx.cpp:99-295) as stream-ordered kernel launches.
#include "session.h"

#include <algorithm>
#include <cstring>

namespace umx
{

namespace
{
template <typename T>
int salloc(Session *s, T **p, size_t count, bool zero = false)
{
    void *q = nullptr;
    UMX_CUDA_OK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    s->allocs.push_back(q);
    if (zero)
        UMX_CUDA_OK(cudaMemset(q, 0, std::max<size_t>(count, 1) * sizeof(T)));
    *p = static_cast<T *>(q);
    return 0;
}
} // namespace

// ---- measurement -----------------------------------------------------------------------------------
// Stage timers are CUDA event pairs on the stream the stage is launched to.  The events come from a pool that only grows, so a
// profiled call creates no events once the pool is warm; with profiling off nothing is recorded at all.
static cudaEvent_t take_timer_event(Session *s)
{
    if (s->timer_cursor == s->timer_pool.size())
    {
        cudaEvent_t e;
        cudaEventCreate(&e);
        s->timer_pool.push_back(e);
    }
    return s->timer_pool[s->timer_cursor++];
}
// Timing events are only ever recorded on compute streams.  For the two copy streams (st_in, st_out) a plain event is recorded
// there and the timestamp is taken on a probe stream that waits for it: a timing event in a stream served by a copy engine
// held back the compute work waiting on that stream's later events until ALL queued copies had run.
static void record_timer(Session *s, cudaEvent_t timer, cudaStream_t st)
{
    if (st != s->st_in && st != s->st_out)
    {
        cudaEventRecord(timer, st);
        return;
    }
    if (s->probe_cursor == s->probe_pool.size())
    {
        cudaEvent_t e;
        cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        s->probe_pool.push_back(e);
    }
    cudaEvent_t x = s->probe_pool[s->probe_cursor++];
    cudaStream_t probe = st == s->st_in ? s->st_probe[0] : s->st_probe[1];
    cudaEventRecord(x, st);
    cudaStreamWaitEvent(probe, x, 0);
    cudaEventRecord(timer, probe);
}
StageScope::StageScope(Session *s_, const char *stage, cudaStream_t st_) : s(s_), st(st_ ? st_ : s_->st)
{
    if (!s->profiling)
        return;
    StageTimer t;
    t.stage = stage;
    t.a = take_timer_event(s);
    t.b = take_timer_event(s);
    record_timer(s, t.a, st);
    s->timers.push_back(t);
    idx = (int)s->timers.size() - 1;
}
StageScope::~StageScope()
{
    if (idx >= 0)
        record_timer(s, s->timers[idx].b, st);
}

void session_begin_profile(Session *s)
{
    s->launches = 0;
    s->stage_ms.clear();
    s->timers.clear();
    s->timer_cursor = 0;
    s->probe_cursor = 0;
    s->profile_resolved = true;
    if (s->profiling)
    { // common time base of this call: the stages run on several streams
        s->profile_base = take_timer_event(s);
        cudaEventRecord(s->profile_base, s->st);
    }
}

// Nothing is waited for here: the timers are read when somebody asks (session_resolve_profile), so a profiled call costs its
// event records and nothing else.
int session_end_profile(Session *s)
{
    s->profile_resolved = !s->profiling;
    return 0;
}

// Per-stage time of the most recent call = length of the UNION of the stage's [start, end] intervals (stages of different
// segments overlap in the pipelined schedules, and the wavefront runs several launches of one stage side by side).
int session_resolve_profile(Session *s)
{
    if (s->profile_resolved)
        return 0;
    UMX_CUDA_OK(cudaSetDevice(s->device));
    UMX_CUDA_OK(cudaStreamSynchronize(s->st)); // every internal stream was joined into s->st at the end of the call
    UMX_CUDA_OK(cudaStreamSynchronize(s->st_probe[0]));
    UMX_CUDA_OK(cudaStreamSynchronize(s->st_probe[1]));
    std::map<std::string, std::vector<std::pair<float, float>>> spans;
    s->timeline.clear();
    for (auto &t : s->timers)
    {
        float a = 0.f, b = 0.f;
        UMX_CUDA_OK(cudaEventElapsedTime(&a, s->profile_base, t.a));
        UMX_CUDA_OK(cudaEventElapsedTime(&b, s->profile_base, t.b));
        spans[t.stage].push_back({a, b});
        s->timeline.push_back({t.stage, a, b});
    }
    for (auto &kv : spans)
    {
        auto &v = kv.second;
        std::sort(v.begin(), v.end());
        double total = 0.0;
        float lo = v[0].first, hi = v[0].second;
        for (size_t i = 1; i < v.size(); ++i)
        {
            if (v[i].first > hi)
            {
                total += hi - lo;
                lo = v[i].first;
                hi = v[i].second;
            }
            else
                hi = std::max(hi, v[i].second);
        }
        s->stage_ms[kv.first] = total + (hi - lo);
    }
    s->profile_resolved = true;
    return 0;
}

// ---- lifetime ----------------------------------------------------------------------------------------
int session_create(Model *m, int max_tracks, int n_buf, Session **out)
{
    UMX_CHECK(m != nullptr && out != nullptr, "session_create: null argument");
    UMX_CHECK(max_tracks >= 1 && max_tracks <= 1024, "session_create: max_tracks must be in [1,1024]");
    UMX_CHECK(n_buf >= kNfft && n_buf % 2 == 0, "session_create: n_buf must be even and >= 4096");
    UMX_CUDA_OK(cudaSetDevice(m->device));
    if (int rc = init_fft_tables(m->device))
        return rc;
    Session *s = new Session();
    s->m = m;
    s->device = m->device;
    s->T = max_tracks;
    s->n_buf = n_buf;
    s->F = n_buf / kHop + 1;
    s->H = m->hidden;
    s->Hh = m->hidden / 2;
    s->Mcap = (long)s->T * s->F;
    if (!lstm_cluster_supported(s->Hh))
    { // never silent: the validation kernel is ~30x slower than the cluster kernel
        s->lstm_impl = 1;
        static bool warned = false;
        if (!warned)
            fprintf(stderr, "umx_cuda: hidden_size %d has no tcgen05 cluster recurrence kernel (built for hidden_size 1024 = UMX-L and 512 = UMX-HQ); "
                            "this session runs the slow validation recurrence kernel (option lstm_impl = 1)\n", s->H);
        warned = true;
    }
    const size_t M = (size_t)s->Mcap, H = (size_t)s->H;
    auto fail = [&](int rc) {
        session_destroy(s);
        return rc;
    };
    // Streams.  Priorities (greatest first): the front stage (STFT, fc1: short, and everything else of a segment waits for it; the
    // buffer rings bound how far it runs ahead), the post stage (short kernels, and the finished output ranges should leave for
    // the host as early as possible), then the deepest LSTM layer — the oldest segment in flight — layer 1, layer 0.
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    auto prio = [&](int rank_from_top) { return std::min(prio_lo, prio_hi + rank_from_top); };
    bool ok = cudaStreamCreateWithFlags(&s->own_st, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&s->st_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&s->st_out, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithPriority(&s->st_post, cudaStreamNonBlocking, prio(1)) == cudaSuccess &&
              cudaStreamCreateWithPriority(&s->st_front, cudaStreamNonBlocking, prio(0)) == cudaSuccess &&
              cudaStreamCreateWithPriority(&s->st_comm, cudaStreamNonBlocking, prio_hi) == cudaSuccess &&
              cudaStreamCreateWithFlags(&s->st_probe[0], cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&s->st_probe[1], cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&s->desc_copied, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&s->spare.desc_copied, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&s->call_done[0], cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&s->call_done[1], cudaEventDisableTiming) == cudaSuccess;
    for (int l = 0; ok && l < 3; ++l)
        for (int g = 0; ok && g < kTargets; ++g)
            ok = cudaStreamCreateWithPriority(&s->st_lane[l][g], cudaStreamNonBlocking, prio(4 - l)) == cudaSuccess;
    if (!ok)
    {
        set_error("session_create: cudaStreamCreate failed");
        return fail(-1);
    }
    s->st = s->own_st;
    int rc = 0;
    for (int i = 0; i < Session::kSpecRing; ++i)
    {
        rc |= salloc(s, &s->spec_ring[i], M * 2 * kBins);
        rc |= salloc(s, &s->maxabs_ring[i], (size_t)s->T * kWienerChunks, true);
    }
    rc |= salloc(s, &s->xh, M * kFc1InPad, true); // pad columns stay zero
    rc |= salloc(s, &s->xl, M * kFc1InPad, true);
    for (int i = 0; i < Session::kActRing; ++i)
    {
        rc |= salloc(s, &s->act_ring_h[i], 4 * M * 2 * H);
        rc |= salloc(s, &s->act_ring_l[i], 4 * M * 2 * H);
    }
    for (int l = 0; l < 2; ++l)
        for (int i = 0; i < Session::kLbRing; ++i)
        {
            rc |= salloc(s, &s->lb_ring_h[l][i], 4 * M * H);
            rc |= salloc(s, &s->lb_ring_l[l][i], 4 * M * H);
        }
    for (int l = 0; l < 3; ++l)
    {
        rc |= salloc(s, &s->xproj_l[l], 4 * M * 4 * H);
        rc |= salloc(s, &s->rowsum_l[l], 4 * M);
    }
    rc |= salloc(s, &s->fc2_h, 4 * M * H);
    rc |= salloc(s, &s->fc2_l, 4 * M * H);
    for (int i = 0; i < Session::kMaskRing; ++i)
        rc |= salloc(s, &s->mask_ring[i], 4 * M * kMaskLd);
    rc |= salloc(s, &s->partials, (size_t)s->T * kWienerChunks * 20 * kBins);
    rc |= salloc(s, &s->stats, (size_t)s->T * kWienerChunks * 16 * kBins);
    rc |= salloc(s, &s->h_state, (size_t)s->T * 3 * 8 * s->Hh, true);
    rc |= salloc(s, &s->c_state, (size_t)s->T * 3 * 8 * s->Hh, true);
    rc |= salloc(s, &s->tile_counters, 2 * 16, true);
    if (rc)
        return fail(-1);
    s->spec = s->spec_ring[0];
    s->mask = s->mask_ring[0];
    s->maxabs = s->maxabs_ring[0];
    s->act_h = s->act_ring_h[0];
    s->act_l = s->act_ring_l[0];
    s->acc_pool.assign(s->T, nullptr);
    s->in_pool.assign(s->T, nullptr);
    s->out_pool.assign(s->T, nullptr);
    s->acc_cap.assign(s->T, 0);
    s->in_cap.assign(s->T, 0);
    s->out_cap.assign(s->T, 0);
    s->spare.in_pool.assign(s->T, nullptr);
    s->spare.in_cap.assign(s->T, 0);
    if (ensure_desc_capacity(s, 16))
        return fail(-1);
    *out = s;
    return 0;
}

// The post stage of a segment (filter + inverse STFT + overlap-add, finalize, copy-out) is cut into kPostParts runs of frames, so
// the first finished output ranges of a segment are on their way to the host while the rest of it is still computed: the tail
// of a host-buffer call after the last kernel is one part's copy, not a whole segment's.
static int post_part_frame(const Session *s, int q)
{
    if (s->wiener_niter > 1) // more than one EM iteration runs unfused, whole segments at a time: everything in the last part
        return q < kPostParts ? 0 : s->F;
    const int per = ((s->F + kPostParts - 1) / kPostParts + 31) / 32 * 32; // whole CTAs of the fused kernel
    return std::min(s->F, q * per);
}

int ensure_iter_buffers(Session *s)
{
    UMX_CUDA_OK(cudaSetDevice(s->device));
    for (auto &p : s->stems_iter)
        if (!p && salloc(s, &p, (size_t)4 * s->Mcap * 2 * kBins))
            return -1;
    return 0;
}

int ensure_desc_capacity(Session *s, int segs)
{
    if (segs <= s->seg_cap)
        return 0;
    UMX_CUDA_OK(cudaDeviceSynchronize()); // nothing may still be reading the old tables
    s->seg_cap = 0;
    const size_t n = (size_t)segs * s->T;
    s->h_desc_bytes = (sizeof(StftTrack) + sizeof(IstftTrack) + sizeof(FinalizeTrack) * kPostParts) * n;
    auto renew = [&](StftTrack *&a, IstftTrack *&b, FinalizeTrack *&c, void *&h) -> int {
        cudaFree(a);
        cudaFree(b);
        cudaFree(c);
        if (h)
            cudaFreeHost(h);
        a = nullptr;
        b = nullptr;
        c = nullptr;
        h = nullptr;
        UMX_CUDA_OK(cudaMalloc(&a, sizeof(StftTrack) * n));
        UMX_CUDA_OK(cudaMalloc(&b, sizeof(IstftTrack) * n));
        UMX_CUDA_OK(cudaMalloc(&c, sizeof(FinalizeTrack) * n * kPostParts));
        UMX_CUDA_OK(cudaMallocHost(&h, s->h_desc_bytes));
        return 0;
    };
    if (renew(s->d_stft, s->d_istft, s->d_final, s->h_desc) ||
        renew(s->spare.d_stft, s->spare.d_istft, s->spare.d_final, s->spare.h_desc))
        return -1;
    s->seg_cap = segs;
    return 0;
}

void session_destroy(Session *s)
{
    if (!s)
        return;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    session_begin_profile(s);
    tp_destroy(s);
    for (void *p : s->allocs)
        cudaFree(p);
    for (auto *p : s->acc_pool)
        cudaFree(p);
    for (auto *p : s->in_pool)
        cudaFree(p);
    for (auto *p : s->out_pool)
        cudaFree(p);
    for (auto *p : s->spare.in_pool)
        cudaFree(p);
    cudaFree(s->d_stft);
    cudaFree(s->d_istft);
    cudaFree(s->d_final);
    cudaFree(s->spare.d_stft);
    cudaFree(s->spare.d_istft);
    cudaFree(s->spare.d_final);
    if (s->h_desc)
        cudaFreeHost(s->h_desc);
    if (s->spare.h_desc)
        cudaFreeHost(s->spare.h_desc);
    for (cudaEvent_t e : {s->spare.desc_copied, s->call_done[0], s->call_done[1]})
        if (e)
            cudaEventDestroy(e);
    for (auto e : s->ev_pool)
        cudaEventDestroy(e);
    for (auto e : s->timer_pool)
        cudaEventDestroy(e);
    if (s->desc_copied)
        cudaEventDestroy(s->desc_copied);
    for (auto e : s->probe_pool)
        cudaEventDestroy(e);
    for (cudaStream_t st : {s->st_in, s->st_out, s->st_post, s->st_front, s->st_comm, s->st_probe[0], s->st_probe[1], s->own_st})
        if (st)
            cudaStreamDestroy(st);
    for (auto &row : s->st_lane)
        for (cudaStream_t st : row)
            if (st)
                cudaStreamDestroy(st);
    delete s;
}

int session_reset(Session *s)
{
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const size_t n = sizeof(float) * (size_t)s->T * 3 * 8 * s->Hh;
    UMX_CUDA_OK(cudaMemsetAsync(s->h_state, 0, n, s->st));
    UMX_CUDA_OK(cudaMemsetAsync(s->c_state, 0, n, s->st));
    return 0;
}

// ---- the network (src/inference.cpp:58-186) ------------------------------------------------------------
static int run_gemm(Session *s, const GemmBatch &b, cudaStream_t st, int terms = 0, int *tile_counter = nullptr)
{
    s->launches++;
    if (s->gemm_impl == 1)
        return launch_gemm_simt(b, st);
    // CTA pairs (256 x 256 tiles, B split across the pair) when the GEMM has the GPU to itself — the in-order schedule, where
    // they are measured faster even for fc3 (in-order GEMM time per bench step 105.4 ms with pairs, 111.4 ms without).  In the
    // segment wavefront (GEMMs of several lanes and the recurrence clusters of 8 share the SMs; the launch passes a tile queue)
    // single-CTA tiles pack better: 226.5 ms per bench step against 229 ms with pairs, which need both SMs of a TPC at once.
    // "gemm_impl" 2 forces pairs everywhere, 3 single-CTA tiles everywhere.
    const bool pair = s->gemm_impl == 2 || (s->gemm_impl == 0 && tile_counter == nullptr);
    return launch_gemm_tc(b, terms ? terms : s->gemm_terms, st, tile_counter, pair);
}

static cudaEvent_t take_event(Session *s, size_t &cursor)
{
    if (cursor == s->ev_pool.size())
    {
        cudaEvent_t e;
        cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        s->ev_pool.push_back(e);
    }
    return s->ev_pool[cursor++];
}

// the buffers the in-order schedules and the kernel-level entry points use: slot 0 of every ring, and the current spec / mask
static NetView current_view(Session *s)
{
    NetView v{};
    v.spec = s->spec;
    v.maxabs = s->maxabs;
    v.mask = s->mask;
    v.act_h = s->act_ring_h[0];
    v.act_l = s->act_ring_l[0];
    for (int l = 0; l < 2; ++l)
    {
        v.lb_h[l] = s->lb_ring_h[l][0];
        v.lb_l[l] = s->lb_ring_l[l][0];
    }
    v.chunks = s->cur_chunks;
    return v;
}

// fc1 + BN1 + tanh for this rank's targets in one GEMM (inference.cpp:75-99)
static int net_fc1(Session *s, const NetView &v, int n_tracks, cudaStream_t st, int *tile_counter)
{
    const int H = s->H;
    const long M = (long)n_tracks * s->F, Mc = s->Mcap;
    Model *m = s->m;
    StageScope sc(s, "gemm", st);
    GemmBatch b{};
    b.count = 1;
    GemmProblem &p = b.p[0];
    const size_t t0 = (size_t)s->tp_t0; // target-parallel ranks only compute their own targets' columns
    p.A = {s->xh, s->xl, kFc1InPad, 0};
    p.B = {m->fc1_hi + t0 * H * kFc1InPad, m->fc1_lo + t0 * H * kFc1InPad, kFc1InPad, 0};
    p.M = (int)M;
    p.N = s->tp_nt * H;
    p.K = kFc1InPad;
    p.act = kActTanh;
    p.chunks = v.chunks; // all-zero frames beyond the chunk: only the first one is computed
    p.rows_per_track = s->F;
    p.extra_rows = 1;
    p.alpha = m->fc1_alpha + t0 * H;
    p.beta = m->fc1_beta + t0 * H;
    p.out_hi = v.act_h + t0 * Mc * 2 * H;
    p.out_lo = v.act_l + t0 * Mc * 2 * H;
    p.ldo = 2 * H;
    p.n_group = H;
    p.group_stride = Mc * 2 * H;
    return run_gemm(s, b, st, 0, tile_counter);
}

// W_ih x + b_ih + b_hh of layer l for all frames (the GEMM half of src/lstm.cpp:132-136), targets [t0, t0 + nt)
static int net_layer_gemm(Session *s, const NetView &v, int l, int n_tracks, int t0, int nt, cudaStream_t st, int *tile_counter)
{
    const int H = s->H;
    const long M = (long)n_tracks * s->F, Mc = s->Mcap;
    Model *m = s->m;
    StageScope sc(s, "gemm", st);
    GemmBatch b{};
    b.count = nt;
    for (int i = 0; i < nt; ++i)
    {
        const int t = t0 + i;
        GemmProblem &p = b.p[i];
        if (l == 0)
            p.A = {v.act_h + (size_t)t * Mc * 2 * H, v.act_l + (size_t)t * Mc * 2 * H, 2 * H, 0};
        else
            p.A = {v.lb_h[l - 1] + (size_t)t * Mc * H, v.lb_l[l - 1] + (size_t)t * Mc * H, H, 0};
        // W_ih as exact integers (one fp16 plane): x.W = scale*(x.q) + offset*rowsum(x) + bias
        s->launches++;
        if (int rc = launch_rowsum(p.A.hi, p.A.lo, M, H, p.A.ld, s->rowsum_l[l] + (size_t)t * Mc, st))
            return rc;
        p.B = {m->t[t].wih_q[l], nullptr, H, 0};
        p.M = (int)M;
        p.N = 4 * H;
        p.K = H;
        p.act = kActNone;
        if (l == 0)
        { // the first all-zero frame's input projection stands for all later ones (LstmLayerArgs::chunks)
            p.chunks = v.chunks;
            p.rows_per_track = s->F;
            p.extra_rows = 1;
        }
        p.alpha = m->t[t].wih_alpha[l];
        p.gamma = m->t[t].wih_gamma[l];
        p.rowvec = s->rowsum_l[l] + (size_t)t * Mc;
        p.beta = m->t[t].wih_beta[l];
        p.out_f32 = s->xproj_l[l] + (size_t)t * Mc * 4 * H;
        p.ldo = 4 * H;
        p.n_group = 4 * H;
        p.group_stride = 0;
    }
    return run_gemm(s, b, st, std::min(s->gemm_terms, 2), tile_counter); // no weight lo plane: at most hi*q + lo*q
}

// the recurrence of layer l (src/lstm.cpp:112-174) for targets [t0, t0 + nt); h/c stream on in the session
static int net_layer_rec(Session *s, const NetView &v, int l, int n_tracks, int t0, int nt, bool want_f32, cudaStream_t st)
{
    const int H = s->H, Hh = s->Hh;
    const long Mc = s->Mcap;
    Model *m = s->m;
    StageScope sc(s, "lstm", st);
    LstmLayerArgs a{};
    for (int t = 0; t < kTargets; ++t)
        for (int d = 0; d < 2; ++d)
        {
            a.whh_q[t][d] = m->t[t].whh_q[l][d];
            a.whh_scale[t][d] = m->t[t].whh_scale[l][d];
            a.whh_offset[t][d] = m->t[t].whh_offset[l][d];
        }
    a.xproj = s->xproj_l[l];
    a.xproj_target_stride = Mc * 4 * H;
    a.h_state = s->h_state + (size_t)l * 8 * Hh;
    a.c_state = s->c_state + (size_t)l * 8 * Hh;
    a.state_track_stride = 3 * 8 * Hh;
    if (l < 2)
    {
        a.out_hi = v.lb_h[l];
        a.out_lo = v.lb_l[l];
        a.out_target_stride = Mc * H;
        a.ldo = H;
        a.col_offset = 0;
    }
    else
    {
        a.out_hi = v.act_h;
        a.out_lo = v.act_l;
        a.out_target_stride = Mc * 2 * H;
        a.ldo = 2 * H;
        a.col_offset = H;
    }
    a.out_f32 = (want_f32 && l == 2) ? s->lstm_f32 : nullptr;
    a.out_f32_target_stride = Mc * H;
    a.F = s->F;
    a.Hh = Hh;
    a.n_tracks = n_tracks;
    a.target_begin = t0;
    a.target_count = nt;
    a.chunks = l == 0 ? v.chunks : nullptr;
    if (s->lstm_reset)
    { // upstream: every segment starts from zero state (the reference streams h/c across segments, src/lstm.cpp:160-161)
        const long off = (long)t0 * 2 * Hh;
        if (launch_zero_strided(a.h_state, a.state_track_stride, off, nt * 2 * Hh, n_tracks, st) ||
            launch_zero_strided(a.c_state, a.state_track_stride, off, nt * 2 * Hh, n_tracks, st))
            return -1;
    }
    s->launches++;
    if (s->lstm_impl == 0)
    {
        UMX_CHECK(lstm_cluster_supported(Hh), "the cluster recurrence kernel does not support this hidden size: set option lstm_impl = 1");
        const uint8_t *img[kTargets][2];
        for (int t = 0; t < kTargets; ++t)
            for (int d = 0; d < 2; ++d)
                img[t][d] = m->t[t].whh_img[l][d];
        return launch_lstm_cluster(a, img, st);
    }
    return launch_lstm_simple(a, st);
}

// skip concat + fc2 + BN2 + ReLU (inference.cpp:118-140), fc3 + BN3 + output affine + ReLU -> mask (inference.cpp:143-166)
static int net_head(Session *s, const NetView &v, int n_tracks, int t0, int nt, cudaStream_t st, int *tile_counter)
{
    const int H = s->H;
    const long M = (long)n_tracks * s->F, Mc = s->Mcap;
    Model *m = s->m;
    StageScope sc(s, "gemm", st);
    {
        GemmBatch b{};
        b.count = nt;
        for (int t = t0; t < t0 + nt; ++t)
        {
            GemmProblem &p = b.p[t - t0];
            p.A = {v.act_h + (size_t)t * Mc * 2 * H, v.act_l + (size_t)t * Mc * 2 * H, 2 * H, 0};
            p.B = {m->t[t].fc2_hi, m->t[t].fc2_lo, 2 * H, 0};
            p.M = (int)M;
            p.N = H;
            p.K = 2 * H;
            p.act = kActRelu;
            p.alpha = m->t[t].fc2_alpha;
            p.beta = m->t[t].fc2_beta;
            p.chunks = v.chunks;
            p.rows_per_track = s->F;
            p.out_hi = s->fc2_h + (size_t)t * Mc * H;
            p.out_lo = s->fc2_l + (size_t)t * Mc * H;
            p.ldo = H;
            p.n_group = H;
            p.group_stride = 0;
        }
        if (int rc = run_gemm(s, b, st, 0, tile_counter))
            return rc;
    }
    GemmBatch b{};
    b.count = nt;
    for (int t = t0; t < t0 + nt; ++t)
    {
        GemmProblem &p = b.p[t - t0];
        p.A = {s->fc2_h + (size_t)t * Mc * H, s->fc2_l + (size_t)t * Mc * H, H, 0};
        p.B = {m->t[t].fc3_hi, m->t[t].fc3_lo, H, 0};
        p.M = (int)M;
        p.N = kFc3Out;
        p.K = H;
        p.act = kActRelu;
        p.alpha = m->t[t].fc3_alpha;
        p.beta = m->t[t].fc3_beta;
        p.chunks = v.chunks;
        p.rows_per_track = s->F;
        p.out_f32 = v.mask + (size_t)t * Mc * kMaskLd;
        p.ldo = kMaskLd;
        p.n_group = kMaskLd;
        p.group_stride = 0;
    }
    return run_gemm(s, b, st, 0, tile_counter);
}

int session_lstm(Session *s, int n_tracks, int target_begin, int target_count, bool want_f32)
{
    const NetView v = current_view(s);
    for (int l = 0; l < 3; ++l)
    {
        if (int rc = net_layer_gemm(s, v, l, n_tracks, target_begin, target_count, s->st, nullptr))
            return rc;
        if (int rc = net_layer_rec(s, v, l, n_tracks, target_begin, target_count, want_f32, s->st))
            return rc;
    }
    return 0;
}

int session_network(Session *s, int n_tracks)
{
    const NetView v = current_view(s);
    if (int rc = net_fc1(s, v, n_tracks, s->st, nullptr))
        return rc;
    if (int rc = session_lstm(s, n_tracks, s->tp_t0, s->tp_nt, false))
        return rc;
    return net_head(s, v, n_tracks, s->tp_t0, s->tp_nt, s->st, nullptr);
}

static int run_stft(Session *s, int seg, int n_tracks, const NetView &v, cudaStream_t st)
{
    StageScope sc(s, "stft", st);
    // max|X| per track starts at zero: a kernel, not a memset node (a copy engine may serve memsets, and there they queue behind
    // the host<->device copies of the whole call)
    if (int rc = launch_zero_f32(v.maxabs, s->T * kWienerChunks, st))
        return rc;
    s->launches++;
    return launch_stft(s->d_stft + (size_t)seg * s->T, n_tracks, s->n_buf, v.spec, s->xh, s->xl, v.maxabs, st, s->skip_zero_frames,
                       s->stft_reflect, s->wiener_win_len, kWienerChunks);
}

int session_front(Session *s, int seg, int n_tracks)
{
    if (int rc = run_stft(s, seg, n_tracks, current_view(s), s->st))
        return rc;
    s->cur_chunks = s->skip_zero_frames ? s->d_istft + (size_t)seg * s->T : nullptr;
    const int rc_net = session_network(s, n_tracks);
    s->cur_chunks = nullptr;
    if (rc_net)
        return rc_net;
    if (s->tp_world > 1)
    { // the one exchange step of the target-parallel mode: every rank needs all four sources' masks
        StageScope sc(s, "allgather");
        if (int rc = tp_allgather_masks(s, s->mask, s->st))
            return rc;
    }
    return 0;
}

// Wiener-EM statistics (reduce + finalize), then the filter pass fused into the inverse STFT + overlap-add
static WienerArgs post_args(Session *s, int seg, int n_tracks, const NetView &v)
{
    WienerArgs w{};
    w.spec = v.spec;
    w.mask = v.mask;
    w.mask_target_stride = s->Mcap * kMaskLd;
    w.mask_ld = kMaskLd;
    w.mask_is_mag = 0;
    w.maxabs = v.maxabs;
    w.partials = s->partials;
    w.stats = s->stats;
    w.stat_windows = kWienerChunks;
    w.psd_abs2 = s->wiener_psd_abs2;
    w.reg_once = s->wiener_reg_once;
    w.win_len = s->wiener_win_len;
    w.no_em = s->wiener_niter == 0;
    w.out = nullptr; // statistics only
    w.out_stem_stride = 0;
    w.mix_out = nullptr;
    w.F = s->F;
    w.n_tracks = n_tracks;
    w.stem_begin = s->tp_t0;
    w.stem_count = s->tp_nt;
    w.chunks = s->d_istft + (size_t)seg * s->T;
    return w;
}
static int run_post_stats(Session *s, int seg, int n_tracks, const NetView &v, cudaStream_t st)
{
    if (s->wiener_niter > 1)
        return 0; // run_post_iterated does everything, in the last part
    StageScope sc(s, "wiener", st);
    s->launches += 2;
    return launch_wiener(post_args(s, seg, n_tracks, v), st);
}
// More than one EM iteration (option "wiener_niter" > 1; the reference's loop at src/wiener.cpp:175 with WIENER_ITERATIONS > 1):
// every iteration needs all sources' estimates of the one before, so the filtered spectrograms are materialised
// (Session::stems_iter) and the last iteration feeds the stand-alone inverse STFT kernel.
static int run_post_iterated(Session *s, int seg, int n_tracks, const NetView &v, cudaStream_t st)
{
    UMX_CHECK(s->stems_iter[0] && s->stems_iter[1], "wiener_niter > 1: iteration buffers are not allocated");
    const long stride = s->Mcap * 2 * kBins;
    for (int it = 0; it < s->wiener_niter; ++it)
    {
        StageScope sc(s, "wiener", st);
        WienerArgs w = post_args(s, seg, n_tracks, v);
        if (it > 0)
        {
            w.mask = nullptr;
            w.y_in = s->stems_iter[(it - 1) & 1];
            w.y_in_stem_stride = stride;
        }
        w.out = s->stems_iter[it & 1];
        w.out_stem_stride = stride;
        const bool last = it + 1 == s->wiener_niter;
        w.out_normalized = !last;
        if (!last)
        { // every iteration but the last needs all four sources' outputs
            w.stem_begin = 0;
            w.stem_count = kTargets;
        }
        s->launches += 3;
        if (int rc = launch_wiener(w, st))
            return rc;
    }
    StageScope sc(s, "istft", st);
    s->launches++;
    return launch_istft(s->stems_iter[(s->wiener_niter - 1) & 1], stride, s->tp_nt, s->d_istft + (size_t)seg * s->T, n_tracks, s->n_buf, st,
                        s->tp_t0);
}
static int run_post_part(Session *s, int seg, int n_tracks, const NetView &v, cudaStream_t st, int part)
{
    if (s->wiener_niter > 1)
        return part + 1 == kPostParts ? run_post_iterated(s, seg, n_tracks, v, st) : 0;
    const int fa = post_part_frame(s, part), fb = part + 1 == kPostParts ? s->F : post_part_frame(s, part + 1);
    if (fa >= fb)
        return 0;
    StageScope sc(s, "istft", st);
    s->launches++;
    return launch_wiener_istft(post_args(s, seg, n_tracks, v), s->d_istft + (size_t)seg * s->T, s->n_buf, st, fa, fb);
}

int session_back(Session *s, int seg, int n_tracks)
{
    const NetView v = current_view(s);
    if (int rc = run_post_stats(s, seg, n_tracks, v, s->st))
        return rc;
    for (int q = 0; q < kPostParts; ++q)
        if (int rc = run_post_part(s, seg, n_tracks, v, s->st, q))
            return rc;
    return 0;
}

int session_run_segment(Session *s, int seg, int n_tracks)
{
    if (int rc = session_front(s, seg, n_tracks))
        return rc;
    return session_back(s, seg, n_tracks);
}

// ---- umx_inference on host buffers (src/inference.cpp:12-207) ------------------------------------------
static int grow(float2 **p, long *cap, long need)
{
    if (need <= *cap)
        return 0;
    if (*p)
        cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    UMX_CUDA_OK(cudaMalloc(p, sizeof(float2) * (size_t)need));
    *cap = need;
    return 0;
}

int session_infer_segment(Session *s, const float *audio, int n, float *out)
{
    UMX_CHECK(n >= 0 && n <= s->n_buf, "umx_cuda_infer_segment: n must be in [0, n_buf]");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    if (int rc = session_wait(s, 0)) // calls enqueued with umx_cuda_demix_async use the same buffers
        return rc;
    session_begin_profile(s);
    if (grow(&s->in_pool[0], &s->in_cap[0], std::max(n, 1)) || grow(&s->out_pool[0], &s->out_cap[0], 4L * std::max(n, 1)))
        return -1;
    {
        StageScope sc(s, "h2d");
        UMX_CUDA_OK(cudaMemcpyAsync(s->in_pool[0], audio, sizeof(float2) * (size_t)n, cudaMemcpyHostToDevice, s->st));
    }
    // descriptors go through the session's pinned staging buffer: no synchronisation before the kernels are queued
    UMX_CUDA_OK(cudaEventSynchronize(s->desc_copied)); // (a previous call's upload has long left the staging buffer)
    StftTrack *ha = static_cast<StftTrack *>(s->h_desc);
    IstftTrack *hb = reinterpret_cast<IstftTrack *>(ha + 1);
    *ha = StftTrack{s->in_pool[0], n, 0, n, 0};
    *hb = IstftTrack{s->out_pool[0], n, 0, n, 0, nullptr, 0, 0, 0, 0};
    UMX_CUDA_OK(cudaMemcpyAsync(s->d_stft, ha, sizeof *ha, cudaMemcpyHostToDevice, s->st));
    UMX_CUDA_OK(cudaMemcpyAsync(s->d_istft, hb, sizeof *hb, cudaMemcpyHostToDevice, s->st));
    UMX_CUDA_OK(cudaEventRecord(s->desc_copied, s->st));
    if (int rc = session_run_segment(s, 0, 1))
        return rc;
    {
        StageScope sc(s, "d2h");
        UMX_CUDA_OK(cudaMemcpyAsync(out, s->out_pool[0], sizeof(float2) * 4 * (size_t)n, cudaMemcpyDeviceToHost, s->st));
    }
    UMX_CUDA_OK(cudaStreamSynchronize(s->st));
    return session_end_profile(s);
}

// ---- shift_inference + split_inference (umx.cpp:99-295) ------------------------------------------------
static int stride_of(int n_buf) { return (int)((1 - 0.25f) * n_buf); } // umx.cpp:180

// the finished range [i_begin, i_end) of every stem of this rank, one strided copy per track, on the output copy stream
static int ship_ranges(Session *s, const FinalizeTrack *fin_host, int n_tracks, const int64_t *n_samples, float *const *d_out,
                       float *const *host_out)
{
    StageScope sc(s, "d2h", s->st_out);
    for (int i = 0; i < n_tracks; ++i)
    {
        const FinalizeTrack &f = fin_host[i];
        if (f.ship_end <= f.ship_begin)
            continue;
        const size_t o = 2 * ((size_t)s->tp_t0 * n_samples[i] + f.ship_begin), pitch = sizeof(float2) * (size_t)n_samples[i];
        UMX_CUDA_OK(cudaMemcpy2DAsync(host_out[i] + o, pitch, d_out[i] + o, pitch, sizeof(float2) * (size_t)(f.ship_end - f.ship_begin),
                                      (size_t)s->tp_nt, cudaMemcpyDeviceToHost, s->st_out));
    }
    return 0;
}

// the samples segment sg reads (original indices below sg*stride - shift + n_buf) on the input copy stream
static int ship_inputs(Session *s, int sg, int stride, int shift_offset, int n_tracks, const int64_t *n_samples,
                       const float *const *d_audio, const float *const *host_in, std::vector<long> &copied_in)
{
    StageScope sc(s, "h2d", s->st_in);
    for (int i = 0; i < n_tracks; ++i)
    {
        const long need = std::max(0L, std::min((long)n_samples[i], (long)sg * stride - shift_offset + s->n_buf));
        if (need > copied_in[i])
        {
            UMX_CUDA_OK(cudaMemcpyAsync(const_cast<float *>(d_audio[i]) + 2 * copied_in[i], host_in[i] + 2 * copied_in[i],
                                        sizeof(float2) * (size_t)(need - copied_in[i]), cudaMemcpyHostToDevice, s->st_in));
            copied_in[i] = need;
        }
    }
    return 0;
}

// ---- the segment wavefront (option "pipeline" = 3) ----------------------------------------------------------
// The recurrences are what a track's latency is made of: 3 layers x 2584 dependent steps per segment, and the h/c carry chains
// the segments (src/lstm.cpp:112-174, state kept at :160-161).  But the carry is PER LAYER: layer l of segment s+1 needs only
// layer l of segment s and layer l-1 of segment s+1.  So the three layers run as three pipeline stages on different segments at
// the same time, and — the four targets being independent until the Wiener filter (src/inference.cpp:70-186) — every
// (layer, target group) has its own stream:
//     front   STFT(s), fc1(s)                                             st_front
//     lane l  W_ih GEMM(s), recurrence(s) of layer l  [+ fc2, fc3 on l=2] st_lane[l][group]
//     post    (mask all-gather,) Wiener(s), iSTFT(s), finalize(s)          st_post (lowest priority: fills idle SMs)
// Buffers that cross stages are rings indexed by segment (Session::*_ring); a producer of slot s % depth first waits for the
// event of the last consumer of segment s - depth.  Every kernel computes exactly what it computes in the in-order schedule, so
// the result is bit-identical (tests/test_gpu_parity.py).
static int demix_wavefront(Session *s, const float *const *d_audio, const int64_t *n_samples, int n_tracks, int shift_offset,
                           float *const *d_out, const float *const *host_in, float *const *host_out, int max_segs, int stride,
                           const std::vector<FinalizeTrack> &hf, const std::vector<long> &max_range, const std::vector<char> &ship_any)
{
    size_t &ev = s->ev_cursor;
    cudaStream_t main_st = s->st;
    // targets per recurrence launch: one (2 clusters) packs the 15 cluster slots of the GPU best when a launch is short (few
    // tracks); with many tracks per launch the grouping makes no measurable difference and fewer launches are issued
    const int lt_opt = s->lane_targets > 0 ? s->lane_targets : (n_tracks <= 16 ? 1 : 4);
    const int lt = std::max(1, std::min(lt_opt, s->tp_nt));
    const int G = (s->tp_nt + lt - 1) / lt; // target groups = lanes per layer
    auto grp_t0 = [&](int g) { return s->tp_t0 + g * lt; };
    auto grp_nt = [&](int g) { return std::min(lt, s->tp_t0 + s->tp_nt - grp_t0(g)); };
    auto wait = [&](cudaStream_t st, cudaEvent_t e) -> int {
        UMX_CUDA_OK(cudaStreamWaitEvent(st, e, 0));
        return 0;
    };
    auto record = [&](cudaEvent_t &e, cudaStream_t st) -> int {
        e = take_event(s, ev);
        UMX_CUDA_OK(cudaEventRecord(e, st));
        return 0;
    };
    // events per segment
    std::vector<cudaEvent_t> ev_fc1(max_segs), ev_wiener(max_segs), back_done(max_segs), ev_gather(max_segs);
    std::vector<cudaEvent_t> ev_wih[3][kTargets], ev_rec[3][kTargets], ev_net[kTargets];
    for (int g = 0; g < G; ++g)
    {
        ev_net[g].resize(max_segs);
        for (int l = 0; l < 3; ++l)
        {
            ev_wih[l][g].resize(max_segs);
            ev_rec[l][g].resize(max_segs);
        }
    }
    // fork: every internal stream starts after what the caller's stream holds so far (state reset, accumulators, descriptors)
    cudaEvent_t fork;
    if (record(fork, main_st))
        return -1;
    if (wait(s->st_front, fork) || wait(s->st_post, fork) || wait(s->st_comm, fork))
        return -1;
    for (int l = 0; l < 3; ++l)
        for (int g = 0; g < G; ++g)
            if (wait(s->st_lane[l][g], fork))
                return -1;
    std::vector<long> copied_in(n_tracks, 0);
    // the previous call's finished ranges may still be on their way out of the output buffers this call's post stage writes
    if (host_out && s->calls_enqueued > 0 && wait(s->st_post, s->call_done[s->calls_enqueued & 1]))
        return -1;

    for (int sg = 0; sg < max_segs; ++sg)
    {
        NetView v{};
        v.spec = s->spec_ring[sg % Session::kSpecRing];
        v.maxabs = s->maxabs_ring[sg % Session::kSpecRing];
        v.mask = s->mask_ring[sg % Session::kMaskRing];
        v.act_h = s->act_ring_h[sg % Session::kActRing];
        v.act_l = s->act_ring_l[sg % Session::kActRing];
        for (int l = 0; l < 2; ++l)
        {
            v.lb_h[l] = s->lb_ring_h[l][sg % Session::kLbRing];
            v.lb_l[l] = s->lb_ring_l[l][sg % Session::kLbRing];
        }
        v.chunks = s->skip_zero_frames ? s->d_istft + (size_t)sg * s->T : nullptr;

        // ---- front: STFT + fc1
        if (host_in)
        {
            if (int rc = ship_inputs(s, sg, stride, shift_offset, n_tracks, n_samples, d_audio, host_in, copied_in))
                return rc;
            cudaEvent_t e;
            if (record(e, s->st_in) || wait(s->st_front, e))
                return -1;
        }
        if (sg >= Session::kSpecRing && wait(s->st_front, ev_wiener[sg - Session::kSpecRing])) // spec / maxabs slot is free again
            return -1;
        if (int rc = run_stft(s, sg, n_tracks, v, s->st_front))
            return rc;
        if (sg >= Session::kActRing) // fc2 of segment sg - 3 has read the activation slot fc1 is about to overwrite
            for (int g = 0; g < G; ++g)
                if (wait(s->st_front, ev_net[g][sg - Session::kActRing]))
                    return -1;
        if (int rc = net_fc1(s, v, n_tracks, s->st_front, s->tile_counters + 2 * 12))
            return rc;
        if (record(ev_fc1[sg], s->st_front))
            return -1;

        // ---- lanes: layer l of this segment for every target group
        for (int l = 0; l < 3; ++l)
            for (int g = 0; g < G; ++g)
            {
                cudaStream_t st = s->st_lane[l][g];
                int *counter = s->tile_counters + 2 * (4 * l + g);
                if (wait(st, l == 0 ? ev_fc1[sg] : ev_rec[l - 1][g][sg]))
                    return -1;
                if (int rc = net_layer_gemm(s, v, l, n_tracks, grp_t0(g), grp_nt(g), st, counter))
                    return rc;
                if (record(ev_wih[l][g][sg], st))
                    return -1;
                // the layer output slot: the next layer's GEMM of segment sg - 2 has read it
                if (l < 2 && sg >= Session::kLbRing && wait(st, ev_wih[l + 1][g][sg - Session::kLbRing]))
                    return -1;
                if (int rc = net_layer_rec(s, v, l, n_tracks, grp_t0(g), grp_nt(g), false, st))
                    return rc;
                if (record(ev_rec[l][g][sg], st))
                    return -1;
                if (l == 2)
                {
                    if (sg >= Session::kMaskRing && wait(st, ev_wiener[sg - Session::kMaskRing])) // mask slot is free again
                        return -1;
                    if (int rc = net_head(s, v, n_tracks, grp_t0(g), grp_nt(g), st, counter))
                        return rc;
                    if (record(ev_net[g][sg], st))
                        return -1;
                }
            }

        // ---- post: (all-gather,) Wiener, iSTFT + overlap-add, finalize, copy-out
        cudaStream_t join_st = s->tp_world > 1 ? s->st_comm : s->st_post;
        for (int g = 0; g < G; ++g)
            if (wait(join_st, ev_net[g][sg]))
                return -1;
        if (s->tp_world > 1)
        { // the one exchange step of the target-parallel mode, on its own stream: it overlaps the next segment's network
            if (sg >= Session::kMaskRing && wait(s->st_comm, ev_wiener[sg - Session::kMaskRing]))
                return -1;
            {
                StageScope sc(s, "allgather", s->st_comm);
                if (int rc = tp_allgather_masks(s, v.mask, s->st_comm))
                    return rc;
            }
            if (record(ev_gather[sg], s->st_comm) || wait(s->st_post, ev_gather[sg]))
                return -1;
        }
        if (int rc = run_post_stats(s, sg, n_tracks, v, s->st_post))
            return rc;
        for (int q = 0; q < kPostParts; ++q)
        {
            if (int rc = run_post_part(s, sg, n_tracks, v, s->st_post, q))
                return rc;
            if (q + 1 == kPostParts && record(ev_wiener[sg], s->st_post)) // this segment's spec / mask / maxabs are consumed
                return -1;
            const size_t d = ((size_t)sg * kPostParts + q) * s->T;
            if (max_range[(size_t)sg * kPostParts + q] > 0)
            {
                StageScope sc(s, "finalize", s->st_post);
                s->launches++;
                if (int rc = launch_finalize(s->d_final + d, n_tracks, max_range[(size_t)sg * kPostParts + q], shift_offset, s->n_buf,
                                             stride, s->tp_nt, s->st_post, s->tp_t0))
                    return rc;
            }
            if (q + 1 == kPostParts && record(back_done[sg], s->st_post))
                return -1;
            if (host_out && ship_any[(size_t)sg * kPostParts + q])
            {
                cudaEvent_t e;
                if (record(e, s->st_post) || wait(s->st_out, e))
                    return -1;
                if (int rc = ship_ranges(s, hf.data() + d, n_tracks, n_samples, d_out, host_out))
                    return rc;
            }
        }
    }
    // join: the post stage of the last segment is the last thing to finish; the lanes and the front stream finished before it
    if (wait(main_st, back_done[max_segs - 1]))
        return -1;
    return 0; // (the caller, session_demix_host, marks the end of the output copies)
}

int session_demix_device(Session *s, const float *const *d_audio, const int64_t *n_samples, int n_tracks,
                         int shift_offset, float *const *d_out, const float *const *host_in, float *const *host_out)
{
    UMX_CHECK(n_tracks >= 1 && n_tracks <= s->T, "umx_cuda_demix: n_tracks must be in [1, max_tracks]");
    // umx.cpp:120-122 places N samples at column `offset` of an (N + 22050 - offset)-column buffer: only
    // offsets <= 11025 stay inside it (the reference's own rand() % 22050 draw is 4033)
    UMX_CHECK(shift_offset >= 0 && shift_offset <= kMaxShift / 2, "umx_cuda_demix: shift_offset must be in [0, 11025]");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const int stride = stride_of(s->n_buf);
    std::vector<long> L(n_tracks);
    int max_segs = 0;
    for (int i = 0; i < n_tracks; ++i)
    {
        UMX_CHECK(n_samples[i] >= 0 && n_samples[i] < (1LL << 40), "umx_cuda_demix: bad track length");
        L[i] = (long)n_samples[i] + kMaxShift - shift_offset; // umx.cpp:120-121
        max_segs = std::max(max_segs, (int)((L[i] + stride - 1) / stride)); // offsets 0, stride, ... < L
    }
    if (ensure_desc_capacity(s, max_segs))
        return -1;
    if (int rc = session_reset(s)) // a fresh create_lstm_data per track (umx.cpp:167-171)
        return rc;
    // descriptors for every (segment, track); tracks that have ended run an empty chunk
    std::vector<StftTrack> hs((size_t)max_segs * s->T);
    std::vector<IstftTrack> hi((size_t)max_segs * s->T);
    std::vector<FinalizeTrack> hf((size_t)max_segs * kPostParts * s->T);
    // output range that is final once segment sg has been added: shifted positions below (sg+1)*stride
    auto final_end = [&](int i, int sg) {
        const long e = (sg == max_segs - 1) ? (long)n_samples[i] : (long)(sg + 1) * stride - shift_offset;
        return std::max(0L, std::min((long)n_samples[i], e));
    };
    std::vector<long> max_range((size_t)max_segs * kPostParts, 0);
    std::vector<char> ship_any((size_t)max_segs * kPostParts, 0); // some track has a finished range in this part
    for (int i = 0; i < n_tracks; ++i)
    {
        if (grow(&s->acc_pool[i], &s->acc_cap[i], 4 * L[i]))
            return -1;
        UMX_CUDA_OK(cudaMemsetAsync(s->acc_pool[i], 0, sizeof(float2) * 4 * (size_t)L[i], s->st));
        for (int sg = 0; sg < max_segs; ++sg)
        {
            const long off = (long)sg * stride;
            const int clen = off < L[i] ? (int)std::min<long>(s->n_buf, L[i] - off) : 0; // umx.cpp:217-218
            hs[(size_t)sg * s->T + i] = {reinterpret_cast<const float2 *>(d_audio[i]), (long)n_samples[i],
                                         off - shift_offset, clen, 0};
            // chunk samples no other segment covers: after the overlap with the previous segment, before the next segment starts
            const int ovl = s->n_buf - stride;
            const int ex_lo = sg == 0 ? 0 : std::min(clen, ovl);
            const int ex_hi = off + stride < L[i] ? std::min(clen, stride) : clen;
            const long x_lo = off + ex_lo - shift_offset, x_hi = off + ex_hi - shift_offset; // the same range as output indices
            // part q of the segment's post stage emits the chunk samples below 1024 * (its last frame + 1) - 2048
            long pb[kPostParts], pe[kPostParts];
            long b = sg == 0 ? 0 : final_end(i, sg - 1);
            bool direct = s->direct_final && ex_hi > ex_lo && s->wiener_niter <= 1; // (the stand-alone inverse STFT only accumulates)
            for (int q = 0; q < kPostParts; ++q)
            {
                long e = final_end(i, sg);
                if (q + 1 < kPostParts)
                    e = std::min(e, std::max(b, off + (long)kHop * post_part_frame(s, q + 1) - kPad - shift_offset));
                pb[q] = b;
                pe[q] = e;
                if (b < x_lo && e > x_hi) // finalize_kernel takes one range per part: it cannot skip a hole in the middle
                    direct = false;
                b = e;
            }
            hi[(size_t)sg * s->T + i] = {s->acc_pool[i], L[i], off, clen, 1, direct ? reinterpret_cast<float2 *>(d_out[i]) : nullptr,
                                         (long)n_samples[i], off - shift_offset, direct ? ex_lo : 0, direct ? ex_hi : 0};
            for (int q = 0; q < kPostParts; ++q)
            {
                const size_t d = ((size_t)sg * kPostParts + q) * s->T + i;
                // finalize_kernel only has to visit what the inverse STFT did not write itself
                long fb = pb[q], fe = pe[q];
                if (direct)
                {
                    if (fb >= x_lo)
                        fb = std::max(fb, std::min(fe, x_hi));
                    else
                        fe = std::min(fe, x_lo);
                    fe = std::max(fe, fb);
                }
                hf[d] = {s->acc_pool[i], reinterpret_cast<float2 *>(d_out[i]), L[i], (long)n_samples[i], fb, fe, pb[q], pe[q]};
                max_range[(size_t)sg * kPostParts + q] = std::max(max_range[(size_t)sg * kPostParts + q], fe - fb);
                ship_any[(size_t)sg * kPostParts + q] = ship_any[(size_t)sg * kPostParts + q] || pe[q] > pb[q];
            }
        }
    }
    s->ev_cursor = 0;
    { // upload through the pinned staging buffer (no host synchronisation with the stream; at most one call of run-ahead)
        UMX_CUDA_OK(cudaEventSynchronize(s->desc_copied));
        uint8_t *hp = static_cast<uint8_t *>(s->h_desc);
        const size_t b0 = sizeof(StftTrack) * hs.size(), b1 = sizeof(IstftTrack) * hi.size(), b2 = sizeof(FinalizeTrack) * hf.size();
        UMX_CHECK(b0 + b1 + b2 <= s->h_desc_bytes, "umx_cuda_demix: descriptor staging too small");
        std::memcpy(hp, hs.data(), b0);
        std::memcpy(hp + b0, hi.data(), b1);
        std::memcpy(hp + b0 + b1, hf.data(), b2);
        // The tables travel at the HEAD of the input copy stream, not on the compute stream: a copy engine serves one stream's
        // queued copies to the end before it turns to another stream's, so three small uploads issued elsewhere would sit behind
        // every input copy of the call (measured: the first kernel then starts when the last input byte has arrived).
        // Host-buffer calls alternate between two sets of tables and input buffers (session_demix_host): the set in hand was
        // last used two calls ago, so nothing is waited for and the uploads run under the previous call's kernels.
        if (!host_in)
        {
            cudaEvent_t prev = take_event(s, s->ev_cursor);
            UMX_CUDA_OK(cudaEventRecord(prev, s->st)); // the previous call's kernels (stream order) are done with the old tables
            UMX_CUDA_OK(cudaStreamWaitEvent(s->st_in, prev, 0));
        }
        UMX_CUDA_OK(cudaMemcpyAsync(s->d_stft, hp, b0, cudaMemcpyHostToDevice, s->st_in));
        UMX_CUDA_OK(cudaMemcpyAsync(s->d_istft, hp + b0, b1, cudaMemcpyHostToDevice, s->st_in));
        UMX_CUDA_OK(cudaMemcpyAsync(s->d_final, hp + b0 + b1, b2, cudaMemcpyHostToDevice, s->st_in));
        UMX_CUDA_OK(cudaEventRecord(s->desc_copied, s->st_in));
        UMX_CUDA_OK(cudaStreamWaitEvent(s->st, s->desc_copied, 0));
    }

    size_t &ev = s->ev_cursor;
    std::vector<long> copied_in(n_tracks, 0);
    if (s->pipeline == 3 && max_segs > 1)
        return demix_wavefront(s, d_audio, n_samples, n_tracks, shift_offset, d_out, host_in, host_out, max_segs, stride, hf, max_range, ship_any);
    // in order on the caller's stream: h/c stream from segment to segment
    if (host_out && s->calls_enqueued > 0) // the previous call's output copies have left the output buffers
        UMX_CUDA_OK(cudaStreamWaitEvent(s->st, s->call_done[s->calls_enqueued & 1], 0));
    for (int sg = 0; sg < max_segs; ++sg)
    {
        if (host_in)
        {
            if (int rc = ship_inputs(s, sg, stride, shift_offset, n_tracks, n_samples, d_audio, host_in, copied_in))
                return rc;
            cudaEvent_t e = take_event(s, ev);
            UMX_CUDA_OK(cudaEventRecord(e, s->st_in));
            UMX_CUDA_OK(cudaStreamWaitEvent(s->st, e, 0));
        }
        if (int rc = session_front(s, sg, n_tracks))
            return rc;
        const NetView v = current_view(s);
        if (int rc = run_post_stats(s, sg, n_tracks, v, s->st))
            return rc;
        for (int q = 0; q < kPostParts; ++q)
        {
            if (int rc = run_post_part(s, sg, n_tracks, v, s->st, q))
                return rc;
            const size_t d = ((size_t)sg * kPostParts + q) * s->T;
            if (max_range[(size_t)sg * kPostParts + q] > 0)
            {
                StageScope sc(s, "finalize");
                s->launches++;
                if (int rc = launch_finalize(s->d_final + d, n_tracks, max_range[(size_t)sg * kPostParts + q], shift_offset, s->n_buf,
                                             stride, s->tp_nt, s->st, s->tp_t0))
                    return rc;
            }
            if (host_out && ship_any[(size_t)sg * kPostParts + q])
            { // ship the finished range of every stem while later parts and segments compute
                cudaEvent_t e = take_event(s, ev);
                UMX_CUDA_OK(cudaEventRecord(e, s->st));
                UMX_CUDA_OK(cudaStreamWaitEvent(s->st_out, e, 0));
                if (int rc = ship_ranges(s, hf.data() + d, n_tracks, n_samples, d_out, host_out))
                    return rc;
            }
        }
    }
    return 0;
}

int session_wait(Session *s, long ticket)
{
    if (ticket <= 0 || ticket > s->calls_enqueued)
        ticket = s->calls_enqueued;
    if (ticket <= s->calls_completed)
        return 0;
    UMX_CUDA_OK(cudaSetDevice(s->device));
    for (long t = std::max(s->calls_completed + 1, ticket - 1); t <= ticket; ++t) // at most two calls are in flight
        UMX_CUDA_OK(cudaEventSynchronize(s->call_done[t & 1]));
    s->calls_completed = ticket;
    return 0;
}

int session_demix_host(Session *s, const float *const *audio, const int64_t *n_samples, int n_tracks, int shift_offset,
                       float *const *out, long *ticket)
{
    UMX_CHECK(n_tracks >= 1 && n_tracks <= s->T, "umx_cuda_demix: n_tracks must be in [1, max_tracks]");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const long me = s->calls_enqueued + 1;
    if (me > 2)
        if (int rc = session_wait(s, me - 2)) // the tables and input buffers this call takes over belong to the call before the last
        return rc;
    std::swap(s->d_stft, s->spare.d_stft);
    std::swap(s->d_istft, s->spare.d_istft);
    std::swap(s->d_final, s->spare.d_final);
    std::swap(s->h_desc, s->spare.h_desc);
    std::swap(s->desc_copied, s->spare.desc_copied);
    s->in_pool.swap(s->spare.in_pool);
    s->in_cap.swap(s->spare.in_cap);
    // stage timers belong to one call at a time: a call enqueued behind a running one is not profiled
    const bool was_profiling = s->profiling;
    if (ticket && s->calls_completed < s->calls_enqueued)
        s->profiling = false;
    session_begin_profile(s);
    std::vector<const float *> din(n_tracks);
    std::vector<float *> dout(n_tracks);
    int rc = 0;
    for (int i = 0; i < n_tracks && !rc; ++i)
    {
        const long n = std::max<long>((long)n_samples[i], 1);
        rc = grow(&s->in_pool[i], &s->in_cap[i], n) || grow(&s->out_pool[i], &s->out_cap[i], 4 * n);
        din[i] = reinterpret_cast<const float *>(s->in_pool[i]);
        dout[i] = reinterpret_cast<float *>(s->out_pool[i]);
    }
    // inputs stream in and finished output ranges stream out on their own copy streams, segment by segment
    if (!rc)
        rc = session_demix_device(s, din.data(), n_samples, n_tracks, shift_offset, dout.data(), audio, out);
    if (!rc)
        rc = session_end_profile(s);
    s->profiling = was_profiling;
    // the end of the call: its kernels (joined into s->st) and every copy queued on the output stream.  Recorded even after a
    // failure, so that whatever was enqueued is waited for before its buffers are touched again.
    cudaEvent_t joined = take_event(s, s->ev_cursor);
    if (cudaEventRecord(joined, s->st) != cudaSuccess || cudaStreamWaitEvent(s->st_out, joined, 0) != cudaSuccess ||
        cudaEventRecord(s->call_done[me & 1], s->st_out) != cudaSuccess)
    {
        cudaDeviceSynchronize();
        set_error("umx_cuda_demix: could not mark the end of the call");
        return -1;
    }
    s->calls_enqueued = me;
    if (rc || !ticket)
    {
        const int wrc = session_wait(s, me);
        return rc ? rc : wrc;
    }
    *ticket = me;
    return 0;
}

} // namespace umx
