// extern "C" boundary (include/umx_cuda.h).  Every function catches nothing and throws nothing:
// failures come back as negative codes with a message in umx_cuda_last_error().
#include "../../include/umx_cuda.h"
#include "session.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace umx
{
// The segment wavefront (session.cu) runs on up to 18 CUDA streams.  The driver maps streams onto CUDA_DEVICE_MAX_CONNECTIONS
// hardware work queues (default 8); streams that share a queue serialise behind each other's event waits.  Every queue costs
// about 50 ms of context creation and teardown (measured: 2.1 s / 2.6 s / 3.7 s wall for a 3-minute CLI run at 8 / 16 / 32), so
// the library asks for 16 at its first call that touches the device, unless the host has chosen a value (so a host's own setenv
// before that call wins: the one-track CLI sets 8; a long-running host that batches tracks may export 32).  It takes effect if
// the CUDA context of the device does not exist yet.
namespace
{
void connections_default()
{
    static std::once_flag once;
    std::call_once(once, [] { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "16", 0); });
}
} // namespace

static thread_local std::string g_err;
void set_error(const std::string &msg) { g_err = msg; }
} // namespace umx

using namespace umx;

struct umx_cuda_model
{
    Model *m;
};
struct umx_cuda_session
{
    Session *s;
};

namespace
{
struct DevBuf
{ // scoped device allocation for the kernel-level entry points
    void *p = nullptr;
    ~DevBuf()
    {
        if (p)
            cudaFree(p);
    }
    int alloc(size_t bytes)
    {
        UMX_CUDA_OK(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
        return 0;
    }
    template <typename T>
    T *as() { return static_cast<T *>(p); }
};

int require_device()
{
    connections_default();
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1)
    {
        set_error("no CUDA device available: this library has no CPU fallback");
        return -1;
    }
    return 0;
}
} // namespace

extern "C"
{

const char *umx_cuda_last_error(void) { return g_err.c_str(); }

int umx_cuda_device_count(void)
{
    connections_default();
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

int umx_cuda_nb_frames(int n_buf) { return n_buf / kHop + 1; }

// ---- model -------------------------------------------------------------------------------------------
int umx_cuda_model_create(const void *ggml_bytes, size_t n_bytes, int device, umx_cuda_model **out)
{
    if (require_device())
        return -1;
    UMX_CHECK(out != nullptr, "umx_cuda_model_create: out is null");
    Model *m = nullptr;
    if (int rc = model_create(ggml_bytes, n_bytes, device, &m))
        return rc;
    *out = new umx_cuda_model{m};
    return 0;
}

int umx_cuda_model_create_from_file(const char *path, int device, umx_cuda_model **out)
{
    UMX_CHECK(path != nullptr, "umx_cuda_model_create_from_file: path is null");
    FILE *f = std::fopen(path, "rb");
    if (!f)
    {
        set_error(std::string("load_umx_model: failed to open ") + path); // src/model.cpp:59-64
        return -3;
    }
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> buf((size_t)std::max(sz, 0L));
    const size_t got = std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    UMX_CHECK(got == buf.size(), "load_umx_model: short read");
    return umx_cuda_model_create(buf.data(), buf.size(), device, out);
}

void umx_cuda_model_destroy(umx_cuda_model *m)
{
    if (!m)
        return;
    model_destroy(m->m);
    delete m;
}

int umx_cuda_model_hidden_size(const umx_cuda_model *m) { return m ? m->m->hidden : -2; }
int umx_cuda_model_device(const umx_cuda_model *m) { return m ? m->m->device : -2; }

long umx_cuda_model_tensor(const umx_cuda_model *mm, const char *name, int target, int layer, int dir, float *out,
                           long capacity, int *rows, int *cols)
{
    UMX_CHECK(mm && name, "umx_cuda_model_tensor: null argument");
    UMX_CHECK(target >= 0 && target < kTargets && layer >= 0 && layer < 3 && dir >= 0 && dir < 2,
              "umx_cuda_model_tensor: index out of range");
    Model *m = mm->m;
    const std::string g = ref_name_to_ggml(name, layer, dir);
    auto it = m->t[target].f32.find(g);
    UMX_CHECK(!g.empty() && it != m->t[target].f32.end(), "umx_cuda_model_tensor: unknown tensor name");
    const DevTensor &dt = it->second;
    // the reference duplicates the 1487/2049-long affine vectors for L and R (src/model.cpp:240-290) and
    // holds 1-D tensors other than LSTM biases as row vectors (transposeInPlace)
    const bool dup = g == "input_mean" || g == "input_scale" || g == "output_mean" || g == "output_scale";
    const bool rowvec = dt.cols == 1 && g.rfind("lstm.", 0) != 0;
    const long count = dup ? 2 * dt.n : dt.n;
    if (rows)
        *rows = rowvec ? 1 : dt.rows;
    if (cols)
        *cols = rowvec ? (int)count : dt.cols;
    if (out)
    {
        UMX_CHECK(capacity >= count, "umx_cuda_model_tensor: output buffer too small");
        UMX_CUDA_OK(cudaSetDevice(m->device));
        UMX_CUDA_OK(cudaMemcpy(out, dt.d, sizeof(float) * dt.n, cudaMemcpyDeviceToHost));
        if (dup)
            std::memcpy(out + dt.n, out, sizeof(float) * dt.n);
    }
    return count;
}

// ---- session -----------------------------------------------------------------------------------------
int umx_cuda_session_create(umx_cuda_model *m, int max_tracks, int n_buf, umx_cuda_session **out)
{
    UMX_CHECK(m && out, "umx_cuda_session_create: null argument");
    Session *s = nullptr;
    if (int rc = session_create(m->m, max_tracks, n_buf, &s))
        return rc;
    *out = new umx_cuda_session{s};
    return 0;
}

void umx_cuda_session_destroy(umx_cuda_session *s)
{
    if (!s)
        return;
    session_destroy(s->s);
    delete s;
}

int umx_cuda_session_reset(umx_cuda_session *s)
{
    UMX_CHECK(s, "umx_cuda_session_reset: null session");
    if (int rc = session_reset(s->s))
        return rc;
    UMX_CUDA_OK(cudaStreamSynchronize(s->s->st));
    return 0;
}

int umx_cuda_session_get_state(umx_cuda_session *ss, int track, int target, int layer, int dir, float *h, float *c)
{
    UMX_CHECK(ss && h && c, "umx_cuda_session_get_state: null argument");
    Session *s = ss->s;
    UMX_CHECK(track >= 0 && track < s->T && target >= 0 && target < 4 && layer >= 0 && layer < 3 && dir >= 0 && dir < 2,
              "umx_cuda_session_get_state: index out of range");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const size_t o = ((((size_t)track * 3 + layer) * 4 + target) * 2 + dir) * s->Hh;
    UMX_CUDA_OK(cudaStreamSynchronize(s->st));
    UMX_CUDA_OK(cudaMemcpy(h, s->h_state + o, sizeof(float) * s->Hh, cudaMemcpyDeviceToHost));
    UMX_CUDA_OK(cudaMemcpy(c, s->c_state + o, sizeof(float) * s->Hh, cudaMemcpyDeviceToHost));
    return 0;
}

int umx_cuda_infer_segment(umx_cuda_session *s, const float *audio, int n, float *out)
{
    UMX_CHECK(s && audio && out, "umx_cuda_infer_segment: null argument");
    return session_infer_segment(s->s, audio, n, out);
}

int umx_cuda_demix(umx_cuda_session *s, const float *const *audio, const int64_t *n_samples, int n_tracks,
                   int shift_offset, float *const *out)
{
    UMX_CHECK(s && audio && n_samples && out, "umx_cuda_demix: null argument");
    return session_demix_host(s->s, audio, n_samples, n_tracks, shift_offset, out);
}

int umx_cuda_demix_async(umx_cuda_session *s, const float *const *audio, const int64_t *n_samples, int n_tracks,
                         int shift_offset, float *const *out, int64_t *ticket)
{
    UMX_CHECK(s && audio && n_samples && out && ticket, "umx_cuda_demix_async: null argument");
    long t = 0;
    if (int rc = session_demix_host(s->s, audio, n_samples, n_tracks, shift_offset, out, &t))
        return rc;
    *ticket = t;
    return 0;
}

int umx_cuda_session_wait(umx_cuda_session *s, int64_t ticket)
{
    UMX_CHECK(s, "umx_cuda_session_wait: null argument");
    return session_wait(s->s, (long)ticket);
}

int umx_cuda_demix_device(umx_cuda_session *s, const float *const *d_audio, const int64_t *n_samples, int n_tracks,
                          int shift_offset, float *const *d_out)
{
    UMX_CHECK(s && d_audio && n_samples && d_out, "umx_cuda_demix_device: null argument");
    if (int rc = session_wait(s->s, 0))
        return rc;
    session_begin_profile(s->s);
    if (int rc = session_demix_device(s->s, d_audio, n_samples, n_tracks, shift_offset, d_out))
        return rc;
    UMX_CUDA_OK(cudaStreamSynchronize(s->s->st));
    return session_end_profile(s->s);
}

int umx_cuda_segment_plan(int64_t n_samples, int shift_offset, int n_buf, int64_t *offsets, int *lengths, int cap)
{
    UMX_CHECK(n_samples >= 0 && shift_offset >= 0 && shift_offset <= kMaxShift / 2 && n_buf >= kNfft,
              "umx_cuda_segment_plan: bad argument");
    const long L = (long)n_samples + kMaxShift - shift_offset;
    const int stride = (int)((1 - 0.25f) * n_buf);
    int count = 0;
    for (long off = 0; off < L; off += stride, ++count) // umx.cpp:214-218
        if (count < cap)
        {
            if (offsets)
                offsets[count] = off;
            if (lengths)
                lengths[count] = (int)std::min<long>(n_buf, L - off);
        }
    return count;
}

// ---- measurement ---------------------------------------------------------------------------------------
int umx_cuda_session_set_profiling(umx_cuda_session *s, int enabled)
{
    UMX_CHECK(s, "umx_cuda_session_set_profiling: null session");
    s->s->profiling = enabled != 0;
    return 0;
}

double umx_cuda_session_stage_ms(umx_cuda_session *s, const char *stage)
{
    if (!s || !stage)
        return -1.0;
    if (session_resolve_profile(s->s)) // waits for the most recent call if its timers have not been read yet
        return -1.0;
    if (std::string(stage) == "total")
    {
        double tot = 0;
        for (auto &kv : s->s->stage_ms)
            tot += kv.second;
        return tot;
    }
    auto it = s->s->stage_ms.find(stage);
    return it == s->s->stage_ms.end() ? -1.0 : it->second;
}

int umx_cuda_host_alloc(size_t bytes, void **out)
{
    UMX_CHECK(out != nullptr, "umx_cuda_host_alloc: null argument");
    if (require_device())
        return -1;
    UMX_CUDA_OK(cudaHostAlloc(out, std::max<size_t>(bytes, 1), cudaHostAllocDefault));
    return 0;
}

void umx_cuda_host_free(void *p)
{
    if (p)
        cudaFreeHost(p);
}

int umx_cuda_session_timeline(umx_cuda_session *s, char *stages, float *begin_ms, float *end_ms, int cap)
{
    UMX_CHECK(s, "umx_cuda_session_timeline: null session");
    if (int rc = session_resolve_profile(s->s))
        return rc;
    const auto &tl = s->s->timeline;
    for (int i = 0; i < cap && i < (int)tl.size(); ++i)
    {
        if (stages)
        {
            std::memset(stages + 16 * i, 0, 16);
            std::strncpy(stages + 16 * i, tl[i].stage.c_str(), 15);
        }
        if (begin_ms)
            begin_ms[i] = tl[i].begin_ms;
        if (end_ms)
            end_ms[i] = tl[i].end_ms;
    }
    return (int)tl.size();
}

int umx_cuda_session_set_stream(umx_cuda_session *s, void *stream)
{
    UMX_CHECK(s, "umx_cuda_session_set_stream: null session");
    UMX_CUDA_OK(cudaSetDevice(s->s->device));
    if (int rc = session_resolve_profile(s->s))
        return rc;
    UMX_CUDA_OK(cudaStreamSynchronize(s->s->st));
    s->s->st = stream ? static_cast<cudaStream_t>(stream) : s->s->own_st;
    return 0;
}

int umx_cuda_lstm_max_active_clusters(void) { return lstm_cluster_max_active(); }

int umx_cuda_nccl_unique_id(void *id128)
{
    UMX_CHECK(id128, "umx_cuda_nccl_unique_id: null argument");
    return tp_unique_id(id128);
}

int umx_cuda_session_init_target_parallel(umx_cuda_session *s, const void *id128, int rank, int world)
{
    UMX_CHECK(s && (id128 || world == 1), "umx_cuda_session_init_target_parallel: null argument");
    return tp_init(s->s, id128, rank, world);
}

long umx_cuda_session_launches(umx_cuda_session *s) { return s ? s->s->launches : -1; }

int umx_cuda_session_set_option(umx_cuda_session *s, const char *name, int value)
{
    UMX_CHECK(s && name, "umx_cuda_session_set_option: null argument");
    if (int rc = session_wait(s->s, 0)) // options change between calls, not under one
        return rc;
    const std::string n(name);
    if (n == "gemm_terms")
    {
        UMX_CHECK(value >= 1 && value <= 3, "gemm_terms must be 1..3");
        s->s->gemm_terms = value;
    }
    else if (n == "gemm_impl")
    {
        UMX_CHECK(value >= 0 && value <= 3, "gemm_impl must be 0 (tcgen05: CTA pairs in order, single-CTA tiles in the wavefront), 1 (SIMT validation), 2 (CTA pairs always) or 3 (single-CTA tiles always)");
        s->s->gemm_impl = value;
    }
    else if (n == "pipeline")
    {
        UMX_CHECK(value >= 0 && value <= 3, "pipeline must be 0 (in order) or 3 (segment wavefront)");
        s->s->pipeline = value == 0 ? 0 : 3; // 1 and 2 were two-stream schedules the wavefront replaced
    }
    else if (n == "lane_targets")
    {
        UMX_CHECK(value == 0 || value == 1 || value == 2 || value == 4, "lane_targets must be 0 (chosen by batch size), 1, 2 or 4");
        s->s->lane_targets = value;
    }
    else if (n == "skip_zero_frames")
    {
        UMX_CHECK(value == 0 || value == 1, "skip_zero_frames must be 0 or 1");
        s->s->skip_zero_frames = value;
    }
    else if (n == "lstm_impl")
    {
        UMX_CHECK(value == 0 || value == 1, "lstm_impl must be 0 (cluster) or 1 (simple)");
        UMX_CHECK(value == 1 || lstm_cluster_supported(s->s->Hh), "lstm_impl = 0: the cluster recurrence kernel does not support this model's hidden size");
        s->s->lstm_impl = value;
    }
    else if (n == "direct_final")
    {
        UMX_CHECK(value == 0 || value == 1, "direct_final must be 0 or 1");
        s->s->direct_final = value;
    }
    else if (n == "stft_pad")
    {
        UMX_CHECK(value == 0 || value == 1, "stft_pad must be 0 (the reference's symmetric edge copy) or 1 (reflect)");
        s->s->stft_reflect = value;
    }
    else if (n == "wiener_psd")
    {
        UMX_CHECK(value == 0 || value == 1, "wiener_psd must be 0 (the reference's (re+im)^2) or 1 (|y|^2)");
        s->s->wiener_psd_abs2 = value;
    }
    else if (n == "wiener_reg")
    {
        UMX_CHECK(value == 0 || value == 1, "wiener_reg must be 0 (regulariser once per source, as the reference) or 1 (once)");
        s->s->wiener_reg_once = value;
    }
    else if (n == "wiener_win_len")
    {
        UMX_CHECK(value == 0 || (value > 0 && (s->s->F + value - 1) / value <= kWienerChunks),
                  "wiener_win_len must be 0 (whole segment) or give at most 19 windows per segment");
        s->s->wiener_win_len = value;
    }
    else if (n == "wiener_niter")
    {
        UMX_CHECK(value >= 0 && value <= 8, "wiener_niter must be in [0, 8] (1 = the reference)");
        if (value > 1)
            if (int rc = ensure_iter_buffers(s->s))
                return rc;
        s->s->wiener_niter = value;
    }
    else if (n == "lstm_reset")
    {
        UMX_CHECK(value == 0 || value == 1, "lstm_reset must be 0 (state streams across segments, as the reference) or 1");
        s->s->lstm_reset = value;
    }
    else
    {
        set_error("umx_cuda_session_set_option: unknown option " + n);
        return -2;
    }
    return 0;
}

} // extern "C"

// ---- kernel-level entry points (host buffers in the reference's memory order) -------------------------
extern "C"
{

int umx_cuda_stft(const float *audio, int n, int n_buf, float *spec) { return umx_cuda_stft_opts(audio, n, n_buf, spec, 0); }

int umx_cuda_stft_opts(const float *audio, int n, int n_buf, float *spec, int reflect_pad)
{
    if (require_device())
        return -1;
    UMX_CHECK(audio && spec, "umx_cuda_stft: null argument");
    UMX_CHECK(n_buf >= kNfft && n >= 0 && n <= n_buf, "umx_cuda_stft: need 0 <= n <= n_buf and n_buf >= 4096");
    int dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&dev));
    if (int rc = init_fft_tables(dev))
        return rc;
    const int F = n_buf / kHop + 1;
    const size_t spec_elems = (size_t)F * 2 * kBins;
    DevBuf d_audio, d_spec, d_ref, d_desc;
    if (d_audio.alloc(sizeof(float2) * (size_t)n) || d_spec.alloc(sizeof(float2) * spec_elems) ||
        d_ref.alloc(sizeof(float2) * spec_elems) || d_desc.alloc(sizeof(StftTrack)))
        return -1;
    UMX_CUDA_OK(cudaMemcpy(d_audio.p, audio, sizeof(float2) * (size_t)n, cudaMemcpyHostToDevice));
    StftTrack tk{d_audio.as<float2>(), n, 0, n, 0};
    UMX_CUDA_OK(cudaMemcpy(d_desc.p, &tk, sizeof tk, cudaMemcpyHostToDevice));
    if (int rc = launch_stft(d_desc.as<StftTrack>(), 1, n_buf, d_spec.as<float2>(), nullptr, nullptr, nullptr, 0, 0, reflect_pad != 0))
        return rc;
    if (int rc = launch_to_ref_layout_c64(d_spec.as<float2>(), d_ref.as<float2>(), F, 0))
        return rc;
    UMX_CUDA_OK(cudaMemcpy(spec, d_ref.p, sizeof(float2) * spec_elems, cudaMemcpyDeviceToHost));
    return 0;
}

int umx_cuda_istft(const float *spec, int n_buf, int n_wave, float *wave)
{
    if (require_device())
        return -1;
    UMX_CHECK(spec && wave, "umx_cuda_istft: null argument");
    UMX_CHECK(n_buf >= kNfft && n_wave >= 0 && n_wave <= n_buf, "umx_cuda_istft: need 0 <= n_wave <= n_buf, n_buf >= 4096");
    int dev = 0;
    UMX_CUDA_OK(cudaGetDevice(&dev));
    if (int rc = init_fft_tables(dev))
        return rc;
    const int F = n_buf / kHop + 1;
    const size_t spec_elems = (size_t)F * 2 * kBins;
    DevBuf d_ref, d_spec, d_wave, d_desc;
    if (d_ref.alloc(sizeof(float2) * spec_elems) || d_spec.alloc(sizeof(float2) * spec_elems) ||
        d_wave.alloc(sizeof(float2) * (size_t)n_wave) || d_desc.alloc(sizeof(IstftTrack)))
        return -1;
    UMX_CUDA_OK(cudaMemcpy(d_ref.p, spec, sizeof(float2) * spec_elems, cudaMemcpyHostToDevice));
    if (int rc = launch_from_ref_layout_c64(d_ref.as<float2>(), d_spec.as<float2>(), F, 0))
        return rc;
    IstftTrack tk{d_wave.as<float2>(), n_wave, 0, n_wave, 0};
    UMX_CUDA_OK(cudaMemcpy(d_desc.p, &tk, sizeof tk, cudaMemcpyHostToDevice));
    if (int rc = launch_istft(d_spec.as<float2>(), 0, 1, d_desc.as<IstftTrack>(), 1, n_buf, 0))
        return rc;
    UMX_CUDA_OK(cudaMemcpy(wave, d_wave.p, sizeof(float2) * (size_t)n_wave, cudaMemcpyDeviceToHost));
    return 0;
}

int umx_cuda_wiener(float *mix, const float *mags, int F, float *out) { return umx_cuda_wiener_opts(mix, mags, F, out, 0, 0, 0, 1); }

int umx_cuda_wiener_opts(float *mix, const float *mags, int F, float *out, int psd_abs2, int reg_once, int win_len, int niter)
{
    if (require_device())
        return -1;
    UMX_CHECK(mix && mags && out && F >= 1, "umx_cuda_wiener: bad argument");
    UMX_CHECK(niter >= 0 && niter <= 8, "umx_cuda_wiener_opts: niter must be in [0, 8]");
    const int n_win = win_len > 0 ? (F + win_len - 1) / win_len : 1;
    UMX_CHECK(win_len >= 0 && n_win <= kWienerChunks, "umx_cuda_wiener_opts: at most 19 windows");
    const size_t e = (size_t)F * 2 * kBins;
    DevBuf d_ref, d_mix, d_mixo, d_mref, d_mag, d_max, d_part, d_stats, d_out, d_desc;
    if (d_ref.alloc(sizeof(float2) * e * 4) || d_mix.alloc(sizeof(float2) * e) || d_mixo.alloc(sizeof(float2) * e) ||
        d_mref.alloc(sizeof(float) * e * 4) || d_mag.alloc(sizeof(float) * e * 4) || d_max.alloc(sizeof(float) * kWienerChunks) ||
        d_part.alloc(sizeof(float) * kWienerChunks * 20 * kBins) || d_stats.alloc(sizeof(float) * kWienerChunks * 16 * kBins) ||
        d_out.alloc(sizeof(float2) * e * 4))
        return -1;
    UMX_CUDA_OK(cudaMemcpy(d_ref.p, mix, sizeof(float2) * e, cudaMemcpyHostToDevice));
    if (int rc = launch_from_ref_layout_c64(d_ref.as<float2>(), d_mix.as<float2>(), F, 0))
        return rc;
    UMX_CUDA_OK(cudaMemcpy(d_mref.p, mags, sizeof(float) * e * 4, cudaMemcpyHostToDevice));
    for (int j = 0; j < 4; ++j)
        if (int rc = launch_from_ref_layout_f32(d_mref.as<float>() + j * e, d_mag.as<float>() + j * e, F, 0))
            return rc;
    { // max |X| per window; the reference order of mix is [bin][frame][channel]
        std::vector<float> mx(kWienerChunks, 0.f);
        for (int b = 0; b < kBins; ++b)
            for (int f = 0; f < F; ++f)
                for (int c = 0; c < 2; ++c)
                {
                    const float *z = mix + 2 * (((size_t)b * F + f) * 2 + c);
                    float &m = mx[win_len > 0 ? f / win_len : 0];
                    m = std::max(m, std::sqrt(z[0] * z[0] + z[1] * z[1]));
                }
        UMX_CUDA_OK(cudaMemcpy(d_max.p, mx.data(), sizeof(float) * kWienerChunks, cudaMemcpyHostToDevice));
    }
    WienerArgs w{};
    w.spec = d_mix.as<float2>();
    w.mask = d_mag.as<float>();
    w.mask_target_stride = (long)e;
    w.mask_ld = 2 * kBins;
    w.mask_is_mag = 1;
    w.maxabs = d_max.as<float>();
    w.partials = d_part.as<float>();
    w.stats = d_stats.as<float>();
    w.stat_windows = kWienerChunks;
    w.psd_abs2 = psd_abs2 != 0;
    w.reg_once = reg_once != 0;
    w.win_len = win_len;
    w.no_em = niter == 0;
    w.out = d_out.as<float2>();
    w.out_stem_stride = (long)e;
    w.mix_out = d_mixo.as<float2>();
    w.F = F;
    w.n_tracks = 1;
    w.stem_begin = 0;
    w.stem_count = 4;
    // more than one EM iteration: every iteration but the last leaves its outputs divided by max_abs for the next one
    DevBuf d_out2;
    if (niter > 1 && d_out2.alloc(sizeof(float2) * e * 4))
        return -1;
    float2 *bufs[2] = {d_out.as<float2>(), niter > 1 ? d_out2.as<float2>() : nullptr};
    const float2 *final_out = bufs[0];
    for (int it = 0; it < std::max(niter, 1); ++it)
    {
        if (it > 0)
        {
            w.mask = nullptr;
            w.y_in = bufs[(it - 1) & 1];
            w.y_in_stem_stride = (long)e;
            w.mix_out = nullptr; // the mixture was left divided by max_abs by the first iteration
        }
        w.out = bufs[it & 1];
        w.out_normalized = it + 1 < niter;
        final_out = w.out;
        if (int rc = launch_wiener(w, 0))
            return rc;
    }
    for (int j = 0; j < 4; ++j)
    {
        if (int rc = launch_to_ref_layout_c64(final_out + j * e, d_ref.as<float2>() + j * e, F, 0))
            return rc;
    }
    UMX_CUDA_OK(cudaMemcpy(out, d_ref.p, sizeof(float2) * e * 4, cudaMemcpyDeviceToHost));
    if (int rc = launch_to_ref_layout_c64(d_mixo.as<float2>(), d_mix.as<float2>(), F, 0))
        return rc;
    UMX_CUDA_OK(cudaMemcpy(mix, d_mix.p, sizeof(float2) * e, cudaMemcpyDeviceToHost));
    return 0;
}

int umx_cuda_lstm_forward(umx_cuda_session *ss, int target, const float *input, int F, float *out)
{
    UMX_CHECK(ss && input && out, "umx_cuda_lstm_forward: null argument");
    Session *s = ss->s;
    UMX_CHECK(target >= 0 && target < 4, "umx_cuda_lstm_forward: target out of range");
    UMX_CHECK(F == s->F, "umx_cuda_lstm_forward: F must equal nb_frames(n_buf) of the session");
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const int H = s->H;
    if (!s->lstm_f32)
    {
        void *q = nullptr;
        UMX_CUDA_OK(cudaMalloc(&q, sizeof(float) * 4 * (size_t)s->Mcap * H));
        s->allocs.push_back(q);
        s->lstm_f32 = static_cast<float *>(q);
    }
    DevBuf d_in, d_out;
    if (d_in.alloc(sizeof(float) * (size_t)F * H) || d_out.alloc(sizeof(float) * (size_t)F * H))
        return -1;
    UMX_CUDA_OK(cudaMemcpyAsync(d_in.p, input, sizeof(float) * (size_t)F * H, cudaMemcpyHostToDevice, s->st));
    // the LSTM input lives in the left half of the target's activation buffer (pitch 2H)
    if (int rc = launch_colmajor_to_planes(d_in.as<float>(), F, H, s->act_h + (size_t)target * s->Mcap * 2 * H,
                                           s->act_l + (size_t)target * s->Mcap * 2 * H, 2 * H, s->st))
        return rc;
    session_begin_profile(s);
    if (int rc = session_lstm(s, 1, target, 1, true))
        return rc;
    if (int rc = launch_rowmajor_to_colmajor(s->lstm_f32 + (size_t)target * s->Mcap * H, F, H, H, d_out.as<float>(), s->st))
        return rc;
    UMX_CUDA_OK(cudaMemcpyAsync(out, d_out.p, sizeof(float) * (size_t)F * H, cudaMemcpyDeviceToHost, s->st));
    UMX_CUDA_OK(cudaStreamSynchronize(s->st));
    return session_end_profile(s);
}

int umx_cuda_target_mags(umx_cuda_session *ss, const float *spec, float *mags)
{
    UMX_CHECK(ss && spec && mags, "umx_cuda_target_mags: null argument");
    Session *s = ss->s;
    UMX_CUDA_OK(cudaSetDevice(s->device));
    const int F = s->F;
    const size_t e = (size_t)F * 2 * kBins;
    DevBuf d_ref, d_mag, d_mref;
    if (d_ref.alloc(sizeof(float2) * e) || d_mag.alloc(sizeof(float) * e) || d_mref.alloc(sizeof(float) * e))
        return -1;
    UMX_CUDA_OK(cudaMemcpyAsync(d_ref.p, spec, sizeof(float2) * e, cudaMemcpyHostToDevice, s->st));
    if (int rc = launch_from_ref_layout_c64(d_ref.as<float2>(), s->spec, F, s->st))
        return rc;
    // network input planes from |spec| (inference.cpp:29,58-68)
    if (int rc = launch_spec_to_x(s->spec, F, s->xh, s->xl, s->st))
        return rc;
    session_begin_profile(s);
    if (int rc = session_network(s, 1))
        return rc;
    for (int j = 0; j < 4; ++j)
    {
        if (int rc = launch_mask_to_mag(s->mask + (size_t)j * s->Mcap * kMaskLd, kMaskLd, s->spec, d_mag.as<float>(), F, s->st))
            return rc;
        if (int rc = launch_to_ref_layout_f32(d_mag.as<float>(), d_mref.as<float>(), F, s->st))
            return rc;
        UMX_CUDA_OK(cudaMemcpyAsync(mags + j * e, d_mref.p, sizeof(float) * e, cudaMemcpyDeviceToHost, s->st));
        UMX_CUDA_OK(cudaStreamSynchronize(s->st));
    }
    return session_end_profile(s);
}

} // extern "C"
