// Per-track streaming state and work buffers for up to T tracks demixed in lockstep
// (replaces lstm_data src/lstm.hpp:10-16, stft_buffers src/dsp.hpp:21-59 and the temporaries of
// umx_inference src/inference.cpp:12-207).
#pragma once
#include "kernels.cuh"
#include "model.h"
#include <map>
#include <string>
#include <vector>

namespace umx
{

constexpr int kPostParts = 4; // runs of frames a segment's post stage is cut into (session.cu)

struct StageTimer
{
    std::string stage;
    cudaEvent_t a, b;
};

// buffers one segment's pass through the network reads and writes (a slot of every ring)
struct NetView
{
    float2 *spec;
    float *maxabs, *mask;
    __half *act_h, *act_l;
    __half *lb_h[2], *lb_l[2];
    const IstftTrack *chunks; // chunk descriptors of the segment (frames beyond a chunk are skipped), or null
};

struct Session
{
    Model *m = nullptr;
    int device = 0;
    int T = 1;      // max tracks in lockstep
    int n_buf = 0;  // samples per segment window
    int F = 0;      // frames per window
    int H = 0, Hh = 0;
    long Mcap = 0;  // T * F rows
    cudaStream_t st = nullptr;      // stream every launch goes to
    cudaStream_t own_st = nullptr;  // the session's own stream (st may be replaced by a caller's)

    // activations (HBM layout documented in DESIGN.md).  Buffers that cross the stages of the segment wavefront (option
    // "pipeline" = 3) are rings indexed by segment; the plain pointers below are ring slot 0, which is what the in-order
    // schedules and the kernel-level entry points use.
    static constexpr int kSpecRing = 4, kActRing = 3, kLbRing = 2, kMaskRing = 2;
    float2 *spec_ring[kSpecRing] = {};       // [Mcap][2][2049]                    STFT -> Wiener
    float *maxabs_ring[kSpecRing] = {};      // [T]
    __half *act_ring_h[kActRing] = {}, *act_ring_l[kActRing] = {}; // [4][Mcap][2H]  (fc1 out | lstm out)   fc1 -> fc2
    __half *lb_ring_h[2][kLbRing] = {}, *lb_ring_l[2][kLbRing] = {}; // [layer 0/1][slot]: [4][Mcap][H] LSTM layer outputs
    float *mask_ring[kMaskRing] = {};        // [4][Mcap][kMaskLd]                 fc3 -> Wiener
    float *xproj_l[3] = {};                  // one input projection [4][Mcap][4H] per layer: the three layers of different
    float *rowsum_l[3] = {};                 //   segments run side by side in the wavefront;  [4][Mcap] row sums per layer
    float2 *spec = nullptr;                  // current slots (the in-order schedules swap spec / mask / maxabs)
    __half *xh = nullptr, *xl = nullptr;     // [Mcap][2976]
    __half *act_h = nullptr, *act_l = nullptr; // = act_ring[0]
    __half *fc2_h = nullptr, *fc2_l = nullptr; // [4][Mcap][H]
    float *mask = nullptr;
    float *maxabs = nullptr;                 // [T]
    float *partials = nullptr, *stats = nullptr;
    float *h_state = nullptr, *c_state = nullptr; // [T][3][4][2][Hh]
    float *lstm_f32 = nullptr;               // optional fp32 LSTM output (kernel-level API), [4][Mcap][H]
    int *tile_counters = nullptr;            // [16][2] GEMM tile queues, one per stream of the wavefront

    StftTrack *d_stft = nullptr;   // descriptor arrays [seg_cap][T]
    const IstftTrack *cur_chunks = nullptr; // chunk descriptors of the segment the network is running for (frames beyond a chunk are skipped)
    int skip_zero_frames = 1;               // option "skip_zero_frames": 0 computes every frame of every window like the reference does
    IstftTrack *d_istft = nullptr;
    FinalizeTrack *d_final = nullptr;
    int seg_cap = 0;
    cudaStream_t st_in = nullptr, st_out = nullptr; // host<->device copy streams (overlap with compute)
    cudaStream_t st_post = nullptr;                 // wavefront: Wiener statistics, filter + iSTFT, finalize of a segment
    // option "pipeline": 0 = every kernel in order on one stream, 3 (default) = segment wavefront: layer l of segment s+1 runs
    // beside layer l+1 of segment s (1 and 2, earlier two-stream schedules, are accepted and mean 3)
    int pipeline = 3;
    int lane_targets = 0;                           // option "lane_targets": targets per recurrence launch in the wavefront (1, 2, 4; 0 = by batch size)
    cudaStream_t st_front = nullptr, st_comm = nullptr;
    cudaStream_t st_lane[3][kTargets] = {};         // wavefront: one stream per (LSTM layer, target group)
    void *h_desc = nullptr;                         // pinned staging for the descriptor tables
    size_t h_desc_bytes = 0;
    cudaEvent_t desc_copied = nullptr;              // the previous call's descriptor upload has left the staging buffer
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_cursor = 0;
    // Calls on host buffers alternate between two sets of descriptor tables, staging and input buffers, so that a call can be
    // enqueued (umx_cuda_demix_async) while the previous one is still running: its inputs and tables go up meanwhile, and the
    // previous call's last output ranges travel to the host under its first kernels.  At most two calls are in flight.
    struct CallSet
    {
        StftTrack *d_stft = nullptr;
        IstftTrack *d_istft = nullptr;
        FinalizeTrack *d_final = nullptr;
        void *h_desc = nullptr;
        cudaEvent_t desc_copied = nullptr;
        std::vector<float2 *> in_pool;
        std::vector<long> in_cap;
    } spare; // the set the previous host-buffer call used
    cudaEvent_t call_done[2] = {};             // [ticket & 1]: everything of that call, copies included, has finished
    long calls_enqueued = 0, calls_completed = 0; // tickets count from 1
    // grow-only pools for whole-track demixing
    std::vector<float2 *> acc_pool, in_pool, out_pool;
    std::vector<long> acc_cap, in_cap, out_cap;

    std::vector<void *> allocs;
    // target-parallel mode (BASELINE config 5): this rank runs targets [tp_t0, tp_t0 + tp_nt) and the masks are
    // all-gathered over NCCL before the Wiener stage; tp_world == 1 means off
    int tp_rank = 0, tp_world = 1, tp_t0 = 0, tp_nt = kTargets;
    void *tp_comm = nullptr; // ncclComm_t
    // options
    int gemm_impl = 0, gemm_terms = 3, lstm_impl = 0;
    // upstream-faithful variants (all 0 = what the reference does; SURVEY Appendix B lists the reference's deviations)
    int stft_reflect = 0;      // "stft_pad" 1: true reflect padding (torch.stft) instead of the symmetric edge copy (src/dsp.cpp:109-128)
    int wiener_psd_abs2 = 0;   // "wiener_psd" 1: v = mean_ch |y|^2 instead of ((re+im)_L^2 + (re+im)_R^2)/2 (src/wiener.cpp:187-202)
    int wiener_reg_once = 0;   // "wiener_reg" 1: sqrt(eps) I added once instead of once per source (src/wiener.cpp:301-325)
    int wiener_win_len = 0;    // "wiener_win_len" N: EM statistics per window of N frames (upstream 300) instead of per segment
    int wiener_niter = 1;      // "wiener_niter" 0: no EM iteration (y = mask X); 1: one iteration (src/wiener.hpp:14); 2..8: more (unfused path)
    float2 *stems_iter[2] = {}; // [4][Mcap][2][2049] each, allocated when wiener_niter > 1: estimates between EM iterations
    int direct_final = 1;      // "direct_final" 1 (default): samples no other segment covers are written finalised by the inverse STFT itself
    int lstm_reset = 0;        // "lstm_reset" 1: h/c zeroed before every segment (upstream) instead of streamed (src/lstm.cpp:160-161)
    // measurement
    bool profiling = false;
    std::vector<StageTimer> timers;
    std::vector<cudaEvent_t> timer_pool; // timing events are created once and reused
    size_t timer_cursor = 0;
    std::vector<cudaEvent_t> probe_pool; // plain events that carry copy-stream positions over to the probe streams
    size_t probe_cursor = 0;
    cudaStream_t st_probe[2] = {};       // timestamps of the copy streams' progress are taken here (session.cu: record_timer)
    cudaEvent_t profile_base = nullptr;  // time base of the most recent profiled call
    bool profile_resolved = true;        // stage_ms holds the most recent call's times
    std::map<std::string, double> stage_ms;
    struct Span
    {
        std::string stage;
        float begin_ms, end_ms;
    };
    std::vector<Span> timeline; // every stage interval of the most recent profiled call, relative to its start
    long launches = 0;
};

int session_create(Model *m, int max_tracks, int n_buf, Session **out);
void session_destroy(Session *s);
int session_reset(Session *s);
// one lockstep pass over descriptor row `seg` for n_tracks tracks (stft .. istft)
int session_run_segment(Session *s, int seg, int n_tracks);
int session_front(Session *s, int seg, int n_tracks); // STFT + networks (+ mask all-gather)
int session_back(Session *s, int seg, int n_tracks);  // Wiener-EM + iSTFT / overlap-add
int session_network(Session *s, int n_tracks);
int session_lstm(Session *s, int n_tracks, int target_begin, int target_count, bool want_f32);
// host_in / host_out (optional): stream the inputs in and the finished output ranges out, segment by segment,
// on the copy streams while the compute stream keeps going
int session_demix_device(Session *s, const float *const *d_audio, const int64_t *n_samples, int n_tracks,
                         int shift_offset, float *const *d_out, const float *const *host_in = nullptr,
                         float *const *host_out = nullptr);
// ticket (optional): the call only enqueues the work and returns its ticket for session_wait; null: returns when the outputs are there
int session_demix_host(Session *s, const float *const *audio, const int64_t *n_samples, int n_tracks,
                       int shift_offset, float *const *out, long *ticket = nullptr);
int session_wait(Session *s, long ticket); // 0: every call enqueued so far
int session_infer_segment(Session *s, const float *audio, int n, float *out);
void session_begin_profile(Session *s);
int session_end_profile(Session *s);
int session_resolve_profile(Session *s); // waits for the most recent call and fills stage_ms
int ensure_desc_capacity(Session *s, int segs);
int ensure_iter_buffers(Session *s); // the two stem-spectrogram buffers EM iterations beyond the first need
// tp.cu: NCCL is loaded with dlopen on first use, so the library has no link-time dependency on it
int tp_unique_id(void *id128);
int tp_init(Session *s, const void *id128, int rank, int world);
int tp_allgather_masks(Session *s, float *mask, cudaStream_t st);
void tp_destroy(Session *s);

struct StageScope
{
    Session *s;
    int idx = -1;
    cudaStream_t st;
    StageScope(Session *s_, const char *stage, cudaStream_t st_ = nullptr); // null: the session's current stream
    ~StageScope();
};

} // namespace umx
