// Drop-in replacement for the reference's library translation units (src/dsp.cpp, src/inference.cpp,
// src/lstm.cpp, src/model.cpp, src/wiener.cpp): the SAME umxcpp:: signatures, declared by the reference's own
// headers, implemented on top of the C ABI (include/umx_cuda.h).  Compile it against the reference's
// headers (+ vendored Eigen) and link libumx_cuda.so instead of umx.cpp.lib; the reference's driver
// (umx.cpp) then runs unchanged on the GPU path (oracle/Makefile target `dropin` does exactly that).
//
// State: the reference keeps the streaming LSTM state in the caller's std::array<lstm_data, 4>; here it lives
// on the GPU inside a session that is found again through a cookie left in the caller's lstm_data.
#include "dsp.hpp"
#include "inference.hpp"
#include "lstm.hpp"
#include "model.hpp"
#include "wiener.hpp"

#include "umx_cuda.h"

#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>

namespace
{
std::mutex g_mu;
std::map<const umxcpp::umx_model *, umx_cuda_model *> g_models;
struct SessionSlot
{
    umx_cuda_session *s = nullptr;
    int n_buf = 0;
    const umxcpp::umx_model *model = nullptr;
};
std::vector<SessionSlot> g_sessions;
constexpr float kCookie = 1.0e30f; // cannot occur in a zero-initialised lstm_data (src/lstm.cpp:41-99)

[[noreturn]] void fail(const char *what)
{
    throw std::runtime_error(std::string(what) + ": " + umx_cuda_last_error());
}

umx_cuda_session *session_for(umxcpp::umx_model &model, std::array<umxcpp::lstm_data, 4> &st, int n_buf)
{
    std::lock_guard<std::mutex> lk(g_mu);
    Eigen::MatrixXf &tag = st[0].output_per_direction[0][0]; // scratch in the reference, never read back by callers
    if (tag.size() >= 2 && tag(0, 0) == kCookie)
    {
        const int id = (int)tag(1, 0);
        if (id >= 0 && id < (int)g_sessions.size() && g_sessions[id].n_buf == n_buf && g_sessions[id].model == &model)
            return g_sessions[id].s;
    }
    // a lstm_data without cookie is a fresh create_lstm_data(): new (or recycled) session with zero state
    auto it = g_models.find(&model);
    if (it == g_models.end())
        throw std::runtime_error("umx_inference: model was not loaded with load_umx_model");
    int id = -1;
    for (size_t i = 0; i < g_sessions.size(); ++i)
        if (g_sessions[i].model == &model && g_sessions[i].n_buf == n_buf)
            id = (int)i; // one track at a time per model in the reference's driver: reuse the buffers
    if (id < 0)
    {
        SessionSlot slot;
        slot.n_buf = n_buf;
        slot.model = &model;
        if (umx_cuda_session_create(it->second, 1, n_buf, &slot.s) != 0)
            fail("umx_cuda_session_create");
        g_sessions.push_back(slot);
        id = (int)g_sessions.size() - 1;
    }
    else if (umx_cuda_session_reset(g_sessions[id].s) != 0)
        fail("umx_cuda_session_reset");
    if (tag.size() >= 2)
    {
        tag(0, 0) = kCookie;
        tag(1, 0) = (float)id;
    }
    return g_sessions[id].s;
}
} // namespace

namespace umxcpp
{

// ---- src/model.hpp:58 --------------------------------------------------------------------------------------
bool load_umx_model(const std::string &model_file, struct umx_model *model)
{
    umx_cuda_model *m = nullptr;
    model->load_progress = 0.1f;
    if (umx_cuda_model_create_from_file(model_file.c_str(), 0, &m) != 0)
    {
        std::cerr << "load_umx_model: " << umx_cuda_last_error() << std::endl;
        return false;
    }
    model->hidden_size = umx_cuda_model_hidden_size(m);
    // keep the Eigen members valid for callers that inspect them (shapes as src/model.cpp:140-186)
    auto fetch = [&](const char *name, int t, int l, int d, Eigen::MatrixXf &dst) {
        int r = 0, c = 0;
        const long n = umx_cuda_model_tensor(m, name, t, l, d, nullptr, 0, &r, &c);
        dst.resize(r, c);
        umx_cuda_model_tensor(m, name, t, l, d, dst.data(), n, nullptr, nullptr);
    };
    for (int t = 0; t < 4; ++t)
    {
        fetch("input_mean", t, 0, 0, model->input_mean[t]);
        fetch("input_scale", t, 0, 0, model->input_scale[t]);
        fetch("output_mean", t, 0, 0, model->output_mean[t]);
        fetch("output_scale", t, 0, 0, model->output_scale[t]);
        fetch("bn1_w", t, 0, 0, model->bn1_w[t]);
        fetch("bn1_b", t, 0, 0, model->bn1_b[t]);
        fetch("bn1_rm", t, 0, 0, model->bn1_rm[t]);
        fetch("bn1_rv", t, 0, 0, model->bn1_rv[t]);
        fetch("bn2_w", t, 0, 0, model->bn2_w[t]);
        fetch("bn2_b", t, 0, 0, model->bn2_b[t]);
        fetch("bn2_rm", t, 0, 0, model->bn2_rm[t]);
        fetch("bn2_rv", t, 0, 0, model->bn2_rv[t]);
        fetch("bn3_w", t, 0, 0, model->bn3_w[t]);
        fetch("bn3_b", t, 0, 0, model->bn3_b[t]);
        fetch("bn3_rm", t, 0, 0, model->bn3_rm[t]);
        fetch("bn3_rv", t, 0, 0, model->bn3_rv[t]);
        model->load_progress += 0.2f;
        // the large matrices (fc*, lstm_*) stay on the GPU; umx_cuda_model_tensor fetches them on demand
    }
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_models[model] = m;
    }
    model->load_progress = 1.0f;
    model->is_initialized = true;
    return true;
}

// ---- src/lstm.hpp:18-24 --------------------------------------------------------------------------------------
struct lstm_data create_lstm_data(int hidden_size, int seq_len)
{
    // the recurrence state lives on the GPU; the caller's struct only carries the session cookie
    struct lstm_data ret;
    (void)seq_len;
    for (int l = 0; l < 3; ++l)
    {
        ret.output[l] = Eigen::MatrixXf::Zero(1, 1);
        for (int d = 0; d < 2; ++d)
        {
            ret.output_per_direction[l][d] = Eigen::MatrixXf::Zero(2, 1);
            ret.h[l][d] = Eigen::MatrixXf::Zero(hidden_size, 1);
            ret.c[l][d] = Eigen::MatrixXf::Zero(hidden_size, 1);
        }
    }
    return ret;
}

void umx_lstm_set_zero(struct lstm_data *data)
{
    for (int l = 0; l < 3; ++l)
    {
        data->output[l].setZero();
        for (int d = 0; d < 2; ++d)
        {
            data->output_per_direction[l][d].setZero(); // also drops the cookie: the next call starts from zero state
            data->h[l][d].setZero();
            data->c[l][d].setZero();
        }
    }
}

// ---- src/inference.hpp:20-23 ---------------------------------------------------------------------------------
std::vector<Eigen::MatrixXf> umx_inference(struct umx_model &model, const Eigen::MatrixXf audio,
                                           struct stft_buffers reusable_stft_buf,
                                           std::array<struct lstm_data, 4> &streaming_lstm_data)
{
    const int n_buf = (int)reusable_stft_buf.padded_waveform_mono_in.size() - FFT_WINDOW_SIZE; // stft_buffers(n_samples)
    umx_cuda_session *s = session_for(model, streaming_lstm_data, n_buf);
    const int n = (int)audio.cols();
    std::vector<float> out((size_t)8 * n);
    if (umx_cuda_infer_segment(s, audio.data(), n, out.data()) != 0)
        fail("umx_inference");
    std::vector<Eigen::MatrixXf> ret;
    for (int j = 0; j < 4; ++j)
        ret.push_back(Eigen::Map<Eigen::MatrixXf>(out.data() + (size_t)j * 2 * n, 2, n));
    return ret;
}

// ---- src/dsp.hpp:106-115 -------------------------------------------------------------------------------------
void stft(struct stft_buffers &b)
{
    const int n_buf = (int)b.padded_waveform_mono_in.size() - FFT_WINDOW_SIZE;
    if (umx_cuda_stft(b.waveform.data(), (int)b.waveform.cols(), n_buf, reinterpret_cast<float *>(b.spec.data())) != 0)
        fail("stft");
}

void istft(struct stft_buffers &b)
{
    const int n_buf = (int)b.padded_waveform_mono_in.size() - FFT_WINDOW_SIZE;
    if (umx_cuda_istft(reinterpret_cast<const float *>(b.spec.data()), n_buf, (int)b.waveform.cols(), b.waveform.data()) != 0)
        fail("istft");
}

Eigen::Tensor3dXcf polar_to_complex(const Eigen::Tensor3dXf &magnitude, const Eigen::Tensor3dXf &phase)
{
    Eigen::Tensor3dXcf out(magnitude.dimension(0), magnitude.dimension(1), magnitude.dimension(2));
    for (long i = 0; i < magnitude.size(); ++i)
        out.data()[i] = std::polar(magnitude.data()[i], phase.data()[i]);
    return out;
}

std::array<Eigen::Tensor3dXcf, 4> wiener_filter(Eigen::Tensor3dXcf &mix, const std::vector<Eigen::Tensor3dXf> &mags)
{
    const int F = (int)mix.dimension(1);
    const size_t e = (size_t)mix.size();
    std::vector<float> m(4 * e), y(8 * e);
    for (int j = 0; j < 4; ++j)
        std::memcpy(m.data() + j * e, mags[j].data(), sizeof(float) * e);
    if (umx_cuda_wiener(reinterpret_cast<float *>(mix.data()), m.data(), F, y.data()) != 0)
        fail("wiener_filter");
    std::array<Eigen::Tensor3dXcf, 4> ret;
    for (int j = 0; j < 4; ++j)
    {
        ret[j] = Eigen::Tensor3dXcf(mix.dimension(0), mix.dimension(1), mix.dimension(2));
        std::memcpy(reinterpret_cast<void *>(ret[j].data()), y.data() + 2 * j * e, sizeof(std::complex<float>) * e);
    }
    return ret;
}

// ---- WAV io (src/dsp.cpp:18-101), RIFF float32 / PCM16 only --------------------------------------------------
Eigen::MatrixXf load_audio(std::string filename)
{
    FILE *f = std::fopen(filename.c_str(), "rb");
    if (!f)
    {
        std::cerr << "[ERROR] cannot open " << filename << std::endl;
        exit(1);
    }
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> buf((size_t)sz);
    if (std::fread(buf.data(), 1, buf.size(), f) != buf.size())
        exit(1);
    std::fclose(f);
    auto r32 = [&](size_t o) { uint32_t v; std::memcpy(&v, &buf[o], 4); return v; };
    auto r16 = [&](size_t o) { uint16_t v; std::memcpy(&v, &buf[o], 2); return v; };
    int tag = 0, ch = 0, rate = 0, bits = 0;
    size_t data = 0, data_len = 0;
    for (size_t pos = 12; pos + 8 <= buf.size();)
    {
        const uint32_t len = r32(pos + 4);
        if (!std::memcmp(&buf[pos], "fmt ", 4))
        {
            tag = r16(pos + 8);
            ch = r16(pos + 10);
            rate = (int)r32(pos + 12);
            bits = r16(pos + 22);
        }
        else if (!std::memcmp(&buf[pos], "data", 4))
        {
            data = pos + 8;
            data_len = std::min<size_t>(len, buf.size() - data);
            break;
        }
        pos += 8 + len + (len & 1);
    }
    if (rate != SUPPORTED_SAMPLE_RATE)
    {
        std::cerr << "[ERROR] umx.cpp only supports the following sample rate (Hz): " << SUPPORTED_SAMPLE_RATE << std::endl;
        exit(1);
    }
    if ((ch != 1 && ch != 2) || !data || !((tag == 3 && bits == 32) || (tag == 1 && bits == 16)))
    {
        std::cerr << "[ERROR] umx.cpp only supports mono and stereo audio" << std::endl;
        exit(1);
    }
    const size_t N = data_len / (bits / 8) / ch;
    Eigen::MatrixXf ret(2, N);
    for (size_t i = 0; i < N; ++i)
        for (int c = 0; c < 2; ++c)
        {
            const size_t idx = ch * i + (ch == 2 ? c : 0);
            float v;
            if (tag == 3)
                std::memcpy(&v, &buf[data + 4 * idx], 4);
            else
                v = (float)(int16_t)r16(data + 2 * idx) / 32767.f;
            ret(c, i) = v;
        }
    return ret;
}

void write_audio_file(const Eigen::MatrixXf &waveform, std::string filename)
{
    FILE *f = std::fopen(filename.c_str(), "wb");
    if (!f)
        return;
    const uint32_t bytes = (uint32_t)(waveform.cols() * 2 * sizeof(float));
    uint8_t h[44];
    auto w32 = [&](int o, uint32_t v) { std::memcpy(h + o, &v, 4); };
    auto w16 = [&](int o, uint16_t v) { std::memcpy(h + o, &v, 2); };
    std::memcpy(h, "RIFF", 4);
    w32(4, 36 + bytes);
    std::memcpy(h + 8, "WAVEfmt ", 8);
    w32(16, 16);
    w16(20, 3);
    w16(22, 2);
    w32(24, SUPPORTED_SAMPLE_RATE);
    w32(28, SUPPORTED_SAMPLE_RATE * 8);
    w16(32, 8);
    w16(34, 32);
    std::memcpy(h + 36, "data", 4);
    w32(40, bytes);
    std::fwrite(h, 1, sizeof h, f);
    std::fwrite(waveform.data(), 1, bytes, f); // (2, N) column-major == interleaved
    std::fclose(f);
    std::cout << "Encoder Status: 0" << std::endl;
}

} // namespace umxcpp
