/* umx_cuda.h — C ABI of the B200-native (sm_100a) UMX-L demixing hot path.
 *
 * Drop-in boundary for sevagh/umx.cpp's CPU path.  Every entry point cites the reference interface it
 * replaces (paths relative to the reference tree).  Plain pointers and sizes only; no C++ or torch
 * types.  All calls return 0 on success and a negative status on failure (never throw, never exit);
 * umx_cuda_last_error() then describes the failure.  There is NO CPU fallback: without a CUDA device
 * every compute call fails with a clear error.
 *
 * Memory orders at the boundary are the reference's own (Eigen column-major), so a reference-side
 * binding passes Eigen's .data() pointers straight through:
 *   waveform  Eigen::MatrixXf(2, N)              -> float[N][2]          interleaved L,R
 *   spec      Eigen::Tensor3dXcf(2, F, 2049)     -> float[2049][F][2][2] (re,im)
 *   mags      Eigen::Tensor3dXf(2, F, 2049)      -> float[2049][F][2]
 *   matrices  Eigen::MatrixXf(r, c)              -> float[c][r]
 * Host pointers unless a parameter is named d_* (device pointer).
 *
 * Threading: like the reference's lstm_data / stft_buffers, a session belongs to one host thread at a time; different sessions
 * (and models) may be used from different threads.  umx_cuda_last_error() is per thread.
 */
#ifndef UMX_CUDA_H
#define UMX_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct umx_cuda_model umx_cuda_model;     /* weights resident on one GPU       (struct umx_model, src/model.hpp:12-56) */
typedef struct umx_cuda_session umx_cuda_session; /* per-track streaming state + work buffers (lstm_data src/lstm.hpp:10-16, stft_buffers src/dsp.hpp:21-59) */

/* reference constants (src/dsp.hpp:16-19, src/inference.hpp:13-16, umx.cpp:112-113,156-157,180) */
#define UMX_SAMPLE_RATE 44100
#define UMX_NFFT 4096
#define UMX_HOP 1024
#define UMX_BINS 2049
#define UMX_SEGMENT_SAMPLES 2646000
#define UMX_STRIDE_SAMPLES 1984500
#define UMX_MAX_SHIFT 22050
#define UMX_DEFAULT_SHIFT_OFFSET 4033 /* rand() % 22050 of an unseeded glibc process, umx.cpp:115 */

/* ---- errors / device ------------------------------------------------------------------------- */
const char *umx_cuda_last_error(void);
int umx_cuda_device_count(void);
/* frames of an STFT over a buffer of n_buf samples: n_buf/1024 + 1 (src/dsp.hpp:46) */
int umx_cuda_nb_frames(int n_buf);

/* ---- model: replaces load_umx_model (src/model.hpp:58, src/model.cpp:42-665) ----------------- */
/* ggml stream (magic 0x756d7867), gzip-compressed or raw, parsed on the host, dequantised on the GPU
 * with one fmaf per element (bit-exact with src/model.cpp:614,660). */
int umx_cuda_model_create(const void *ggml_bytes, size_t n_bytes, int device, umx_cuda_model **out);
int umx_cuda_model_create_from_file(const char *path, int device, umx_cuda_model **out);
void umx_cuda_model_destroy(umx_cuda_model *m);
int umx_cuda_model_hidden_size(const umx_cuda_model *m); /* umx_model::hidden_size */
int umx_cuda_model_device(const umx_cuda_model *m);
/* Dequantised fp32 tensor as the reference holds it (same member names as struct umx_model, e.g.
 * "fc1_w", "bn1_rv", "lstm_hh_w", "input_mean"; layer/dir only for lstm_*).  Column-major (rows, cols).
 * out may be NULL to query the shape.  Returns the element count, or <0. */
long umx_cuda_model_tensor(const umx_cuda_model *m, const char *name, int target, int layer, int dir,
                           float *out, long capacity, int *rows, int *cols);

/* ---- session: streaming state for up to max_tracks tracks processed in lockstep ------------- */
/* n_buf = samples per segment buffer (UMX_SEGMENT_SAMPLES in the CLI, umx.cpp:161-171).
 * Creating a session zeroes the LSTM h/c carry (create_lstm_data, src/lstm.cpp:41-84). */
int umx_cuda_session_create(umx_cuda_model *m, int max_tracks, int n_buf, umx_cuda_session **out);
void umx_cuda_session_destroy(umx_cuda_session *s);
int umx_cuda_session_reset(umx_cuda_session *s); /* umx_lstm_set_zero, src/lstm.cpp:86-99 */
/* h / c carry of (track, target, layer, dir): lstm_hidden floats each (lstm_data::h/c, src/lstm.hpp:14-15) */
int umx_cuda_session_get_state(umx_cuda_session *s, int track, int target, int layer, int dir, float *h, float *c);

/* ---- umx_inference (src/inference.hpp:20-23, src/inference.cpp:12-207) ----------------------- */
/* One chunk of n <= n_buf samples, processed as a full n_buf-sample window (always nb_frames(n_buf)
 * frames, zero-extended: SURVEY F4), LSTM state carried in the session (track slot 0).
 * audio: float[n][2]; out: 4 stems x float[n][2] (bass, drums, other, vocals). */
int umx_cuda_infer_segment(umx_cuda_session *s, const float *audio, int n, float *out);

/* ---- whole tracks: shift_inference + split_inference (umx.cpp:99-295) ------------------------ */
/* n_tracks <= max_tracks independent tracks, each shifted by shift_offset in [0, 11025] (umx.cpp:112-122;
 * a larger offset overruns the reference's own (N + 22050 - offset)-column buffer), cut into
 * n_buf-sample segments at 0.75*n_buf stride, demixed with streaming LSTM state, recombined with the
 * triangular weights and sum_weight normalisation (umx.cpp:197-273), trimmed (umx.cpp:136-147).
 * Session state is reset first (a fresh process per track in the reference).
 * audio[i]: float[n_samples[i]][2]; out[i]: 4 x float[n_samples[i]][2]. Host memory (pinned is faster). */
int umx_cuda_demix(umx_cuda_session *s, const float *const *audio, const int64_t *n_samples, int n_tracks,
                   int shift_offset, float *const *out);
/* The same call, enqueued only: it returns once the work is queued and hands back a ticket; out[] holds the stems after
 * umx_cuda_session_wait(s, ticket) (ticket 0: every call enqueued so far), and audio[] / out[] must stay valid (and should be
 * page-locked) until then.  Up to two calls are in flight per session — a third call first waits for the oldest — so the
 * inputs of call k+1 go up, and the last output ranges of call k come down, while the other call's kernels run: a host feeding
 * a queue of batches (one shift_inference after another, umx.cpp:99-147) is not idle between them.  Give consecutive calls
 * different out[] buffers.  Every other entry point of the session first waits for the calls in flight. */
int umx_cuda_demix_async(umx_cuda_session *s, const float *const *audio, const int64_t *n_samples, int n_tracks,
                         int shift_offset, float *const *out, int64_t *ticket);
int umx_cuda_session_wait(umx_cuda_session *s, int64_t ticket);
/* Same with inputs/outputs already resident in device memory (no host<->device copies). */
int umx_cuda_demix_device(umx_cuda_session *s, const float *const *d_audio, const int64_t *n_samples,
                          int n_tracks, int shift_offset, float *const *d_out);
/* Page-locked host buffers for umx_cuda_demix / umx_cuda_infer_segment (hosts that do not link the CUDA runtime themselves):
 * with them the copies run at PCIe speed and overlap the kernels; pageable buffers work too, slower. */
int umx_cuda_host_alloc(size_t bytes, void **out);
void umx_cuda_host_free(void *p);
/* Segment table of split_inference for a track of n samples shifted by shift_offset (umx.cpp:214-218):
 * writes up to cap (offset, chunk_length) pairs, returns the segment count. */
int umx_cuda_segment_plan(int64_t n_samples, int shift_offset, int n_buf, int64_t *offsets, int *lengths, int cap);

/* ---- kernel-level entry points (parity tests, microbenchmarks, the Eigen-typed shim) --------- */
/* stft (src/dsp.hpp:114, src/dsp.cpp:141-176): audio float[n][2] zero-extended to n_buf;
 * spec: complex float[2049][F][2], F = nb_frames(n_buf). */
int umx_cuda_stft(const float *audio, int n, int n_buf, float *spec);
/* The same with true reflect padding (reflect_pad = 1: upstream's torch.stft; 0 = umx_cuda_stft). */
int umx_cuda_stft_opts(const float *audio, int n, int n_buf, float *spec, int reflect_pad);
/* istft (src/dsp.hpp:115, src/dsp.cpp:178-207): spec complex float[2049][F][2] -> wave float[n_wave][2] */
int umx_cuda_istft(const float *spec, int n_buf, int n_wave, float *wave);
/* umx_lstm_forward (src/lstm.hpp:22-24, src/lstm.cpp:101-179) for `target`, state of track slot 0:
 * input column-major (F, hidden) -> out column-major (F, hidden); F <= nb_frames(n_buf). */
int umx_cuda_lstm_forward(umx_cuda_session *s, int target, const float *input, int F, float *out);
/* wiener_filter (src/wiener.hpp:19-21, src/wiener.cpp:92-425): mix complex float[2049][F][2] (IN/OUT:
 * left divided by max_abs like the reference), mags 4 x float[2049][F][2] -> out 4 x complex [2049][F][2] */
int umx_cuda_wiener(float *mix, const float *mags, int F, float *out);
/* wiener_filter with the upstream-faithful variants of the session options (all zero and niter = 1: umx_cuda_wiener):
 * psd_abs2 / reg_once / win_len / niter as "wiener_psd" / "wiener_reg" / "wiener_win_len" / "wiener_niter" below. */
int umx_cuda_wiener_opts(float *mix, const float *mags, int F, float *out, int psd_abs2, int reg_once, int win_len, int niter);
/* The network only (src/inference.cpp:29-186): spec complex [2049][F][2] -> target magnitudes
 * 4 x float[2049][F][2]; LSTM state of track slot 0 advances. F == nb_frames(n_buf) of the session. */
int umx_cuda_target_mags(umx_cuda_session *s, const float *spec, float *mags);

/* Wiener-EM that allocates nothing.  NO COUNTERPART IN THE REFERENCE: it exists for callers that may not
 * call cudaMalloc when they need the filter (a real-time audio thread, which cannot take cudaMalloc's
 * device-wide synchronisation and unbounded latency; a plugin inside a process whose GPU memory belongs
 * entirely to another allocator).  Same arithmetic as umx_cuda_wiener, but inputs and outputs are DEVICE
 * buffers the caller already owns, in the library's internal order, and the kernels' scratch is a static
 * device array inside the library (about 6.5 MB, held by every GPU that loads the library):
 *   d_mix   complex float[n_tracks*F][2][2049] (left untouched), d_mags 4 x float[n_tracks*F][2][2049],
 *   d_out   4 x complex float[n_tracks*F][2][2049], stream a cudaStream_t (NULL = default stream).
 * n_tracks > UMX_WIENER_NOALLOC_MAX_TRACKS is rejected.  The first call on each device resolves the
 * array's address: make it before any stream capture.  All calls share the array (its max|X| slots are
 * re-zeroed per call, the rest is overwritten before use), so calls that overlap on the GPU race:
 * serialise them.  The buffers and the stream must belong to the device that is current at the call (checked);
 * F <= 1,048,560.  The resolved address is cached per device for the life of the process: after a
 * cudaDeviceReset the library must be reloaded before this entry point is used again. */
#define UMX_WIENER_NOALLOC_MAX_TRACKS 2
int umx_cuda_wiener_noalloc(const float *d_mix, const float *d_mags, int F, int n_tracks, float *d_out, void *stream);

/* ---- target-parallel mode over NCCL (no counterpart in the single-process reference; BASELINE config 5) ------ */
/* The 4 stem networks are independent until wiener_filter (src/inference.cpp:70-186): rank j of a 1/2/4-GPU group
 * runs targets [j*4/world, (j+1)*4/world); one ncclAllGather of the fc3 masks per segment precedes the Wiener
 * stage, which every rank then runs for its own stems.  After umx_cuda_demix* each rank's output buffers hold
 * ITS stems only.  One process per GPU; rank 0 creates the 128-byte id and hands it to the others. */
int umx_cuda_nccl_unique_id(void *id128);
int umx_cuda_session_init_target_parallel(umx_cuda_session *s, const void *id128, int rank, int world);

/* ---- measurement ------------------------------------------------------------------------------ */
/* Per-stage device time of the most recent demix / infer_segment call, from CUDA events recorded on
 * the session's own stream.  Stage names: "stft", "gemm", "lstm", "wiener", "istft", "finalize",
 * "h2d", "d2h", "total".  Returns milliseconds, or <0 if unknown.  launches: kernels launched. */
int umx_cuda_session_set_profiling(umx_cuda_session *s, int enabled);
double umx_cuda_session_stage_ms(umx_cuda_session *s, const char *stage);
long umx_cuda_session_launches(umx_cuda_session *s);
/* Every stage interval of the most recent profiled call (the stages of different segments overlap): up to cap entries of a
 * 16-byte NUL-padded stage name, begin and end in ms since the call started.  Returns the number of intervals. */
int umx_cuda_session_timeline(umx_cuda_session *s, char *stages, float *begin_ms, float *end_ms, int cap);
/* Launch everything on the caller's cudaStream_t (NULL restores the session's own stream), so a caller
 * can bracket the work with its own CUDA events. */
int umx_cuda_session_set_stream(umx_cuda_session *s, void *stream);
/* How many 8-CTA clusters of the recurrence kernel the current device can hold at once (diagnostic). */
int umx_cuda_lstm_max_active_clusters(void);
/* Tuning knobs (documented in DESIGN.md): "gemm_terms" 1..3 fp16 split terms per tensor-core GEMM,
 * "gemm_impl" 0 = tcgen05 (CTA pairs, cta_group::2 on 256 x 256 tiles, in the in-order schedule; single-CTA 128 x 256 tiles in
 * the segment wavefront), 1 = SIMT fp32 validation kernel, 2 = CTA pairs always, 3 = single-CTA tiles always (0, 2 and 3 give
 * identical results), "lstm_impl" 0 = cluster kernel, 1 = simple,
 * "pipeline" 0 = every kernel in order on one stream, 3 (default) = segment wavefront: LSTM layer l of segment s+1 runs beside
 * layer l+1 of segment s (the h/c carry of src/lstm.cpp:160-161 is per layer), one stream per (layer, target group) with
 * "lane_targets" = 1, 2 or 4 targets per group (0, the default: 1 up to 16 tracks, else 4); results are identical,
 * "skip_zero_frames" 1 (default) = frames of a window beyond a short chunk are computed once (first layer) or not at all
 * (after the network), 0 = every frame of every 60 s window is computed as the reference does; results are identical.
 * "direct_final" 1 (default) = output samples that only one segment covers are written finalised, (w y)/w, by the inverse
 * STFT kernel itself instead of going through the track accumulator and the finalize kernel; results are identical.
 * Upstream-faithful variants, all 0 by default = what the reference does (its deviations from Open-Unmix):
 * "stft_pad" 1 = true reflect padding (torch.stft) instead of the symmetric edge copy of src/dsp.cpp:109-128,
 * "wiener_psd" 1 = PSD |y|^2 instead of (re+im)^2 (src/wiener.cpp:187-202), "wiener_reg" 1 = regulariser added once instead
 * of once per source (src/wiener.cpp:301-325), "wiener_win_len" N = EM statistics and max|X| per window of N frames
 * (upstream 300; at most 19 windows per segment) instead of per segment, "wiener_niter" 0 = no EM iteration, 2..8 = more of them, unfused (default 1 =
 * src/wiener.hpp:14), "lstm_reset" 1 = LSTM state zeroed before every segment instead of streamed (src/lstm.cpp:160-161). */
int umx_cuda_session_set_option(umx_cuda_session *s, const char *name, int value);

#ifdef __cplusplus
}
#endif
#endif /* UMX_CUDA_H */
