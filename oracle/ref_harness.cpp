// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// Thin extern "C" wrappers around the UNMODIFIED reference sources
// (/root/reference/src/*.cpp), compiled where they lie by oracle/Makefile into
// oracle/_ref/libumxref.so.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load that library.
//
// All buffers are raw float32 in the reference's own memory order (Eigen
// column-major), so the Python side can view them with numpy without copies:
//   waveform  MatrixXf(2, N)          -> data[c + 2*n]           (interleaved LRLR)
//   spec      Tensor3dXcf(2, F, B)    -> data[c + 2*(f + F*b)]   (complex64)
//   mags      Tensor3dXf(2, F, B)     -> data[c + 2*(f + F*b)]
//   matrices  MatrixXf(r, c)          -> data[i + r*j]
#include "dsp.hpp"
#include "inference.hpp"
#include "lstm.hpp"
#include "model.hpp"
#include "wiener.hpp"

#include <array>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

using namespace umxcpp;

namespace
{
struct quiet_scope
{
    bool on;
    explicit quiet_scope(bool q) : on(q)
    {
        if (on)
            std::cout.setstate(std::ios_base::failbit);
    }
    ~quiet_scope()
    {
        if (on)
            std::cout.clear();
    }
};
bool g_quiet = true;

struct lstm_set
{
    std::array<lstm_data, 4> d;
};
} // namespace

// the reference's own CLI entry (umx.cpp:26), renamed by -Dmain=umx_ref_main
int umx_ref_main(int argc, const char **argv);

extern "C"
{

void ref_set_quiet(int q) { g_quiet = q != 0; }

// ---- model (src/model.cpp:42) -------------------------------------------
void *ref_model_load(const char *path)
{
    quiet_scope qs(g_quiet);
    auto *m = new umx_model{};
    if (!load_umx_model(path, m))
    {
        delete m;
        return nullptr;
    }
    return m;
}
void ref_model_free(void *h) { delete static_cast<umx_model *>(h); }
int ref_model_hidden(void *h) { return static_cast<umx_model *>(h)->hidden_size; }

static const Eigen::MatrixXf *find_tensor(umx_model *m, const char *name, int t,
                                          int l, int d)
{
    std::string n(name);
    if (n == "input_mean") return &m->input_mean[t];
    if (n == "input_scale") return &m->input_scale[t];
    if (n == "output_mean") return &m->output_mean[t];
    if (n == "output_scale") return &m->output_scale[t];
    if (n == "fc1_w") return &m->fc1_w[t];
    if (n == "fc2_w") return &m->fc2_w[t];
    if (n == "fc3_w") return &m->fc3_w[t];
    if (n == "bn1_w") return &m->bn1_w[t];
    if (n == "bn1_b") return &m->bn1_b[t];
    if (n == "bn1_rm") return &m->bn1_rm[t];
    if (n == "bn1_rv") return &m->bn1_rv[t];
    if (n == "bn2_w") return &m->bn2_w[t];
    if (n == "bn2_b") return &m->bn2_b[t];
    if (n == "bn2_rm") return &m->bn2_rm[t];
    if (n == "bn2_rv") return &m->bn2_rv[t];
    if (n == "bn3_w") return &m->bn3_w[t];
    if (n == "bn3_b") return &m->bn3_b[t];
    if (n == "bn3_rm") return &m->bn3_rm[t];
    if (n == "bn3_rv") return &m->bn3_rv[t];
    if (n == "lstm_ih_w") return &m->lstm_ih_w[t][l][d];
    if (n == "lstm_ih_b") return &m->lstm_ih_b[t][l][d];
    if (n == "lstm_hh_w") return &m->lstm_hh_w[t][l][d];
    if (n == "lstm_hh_b") return &m->lstm_hh_b[t][l][d];
    return nullptr;
}

// returns element count (0 if unknown); fills rows/cols; copies when out != NULL
long ref_model_tensor(void *h, const char *name, int target, int layer, int dir,
                      float *out, int *rows, int *cols)
{
    const Eigen::MatrixXf *p =
        find_tensor(static_cast<umx_model *>(h), name, target, layer, dir);
    if (!p)
        return 0;
    if (rows) *rows = (int)p->rows();
    if (cols) *cols = (int)p->cols();
    if (out)
        std::memcpy(out, p->data(), sizeof(float) * p->size());
    return (long)p->size();
}

// ---- stft / istft (src/dsp.cpp:141,178) ----------------------------------
int ref_nb_frames(int n_buf) { return n_buf / FFT_HOP_SIZE + 1; }

// audio: (2, n) ; buffers sized for n_buf >= n (umx.cpp:161 + inference.cpp:23)
void ref_stft(const float *audio, int n, int n_buf, float *spec_out)
{
    quiet_scope qs(g_quiet);
    stft_buffers buf(n_buf);
    buf.waveform = Eigen::Map<const Eigen::MatrixXf>(audio, 2, n);
    stft(buf);
    std::memcpy(spec_out, buf.spec.data(),
                sizeof(std::complex<float>) * buf.spec.size());
}

// spec: (2, F(n_buf), 2049) ; result (2, n_wave) as umx_inference leaves it
void ref_istft(const float *spec, int n_buf, int n_wave, float *wave_out)
{
    quiet_scope qs(g_quiet);
    stft_buffers buf(n_buf);
    // keep the waveform at its full n_buf columns: a shorter one is overrun by istft (see ref_umx_inference)
    std::memcpy((void *)buf.spec.data(), spec,
                sizeof(std::complex<float>) * buf.spec.size());
    istft(buf);
    for (int i = 0; i < n_wave; ++i)
    {
        wave_out[2 * i] = buf.waveform(0, i);
        wave_out[2 * i + 1] = buf.waveform(1, i);
    }
}

// ---- lstm (src/lstm.cpp:41,101) -------------------------------------------
void *ref_lstm_create(int lstm_hidden, int seq_len)
{
    auto *s = new lstm_set{{create_lstm_data(lstm_hidden, seq_len),
                            create_lstm_data(lstm_hidden, seq_len),
                            create_lstm_data(lstm_hidden, seq_len),
                            create_lstm_data(lstm_hidden, seq_len)}};
    return s;
}
void ref_lstm_free(void *s) { delete static_cast<lstm_set *>(s); }

// input (F, hidden) col-major ; out (F, hidden) col-major
void ref_lstm_forward(void *model, void *state, int target, const float *input,
                      int F, float *out)
{
    quiet_scope qs(g_quiet);
    auto *m = static_cast<umx_model *>(model);
    auto *s = static_cast<lstm_set *>(state);
    Eigen::MatrixXf in = Eigen::Map<const Eigen::MatrixXf>(input, F, m->hidden_size);
    Eigen::MatrixXf o =
        umx_lstm_forward(m, target, in, &s->d[target], m->hidden_size / 2);
    std::memcpy(out, o.data(), sizeof(float) * o.size());
}

// h/c carry of (target, layer, dir): two vectors of lstm_hidden floats
void ref_lstm_get_state(void *state, int target, int layer, int dir, float *h,
                        float *c)
{
    auto *s = static_cast<lstm_set *>(state);
    const auto &H = s->d[target].h[layer][dir];
    const auto &C = s->d[target].c[layer][dir];
    std::memcpy(h, H.data(), sizeof(float) * H.size());
    std::memcpy(c, C.data(), sizeof(float) * C.size());
}

// ---- wiener (src/wiener.cpp:92) -------------------------------------------
// mix (2,F,B) complex IN/OUT (the reference leaves it divided by max_abs);
// mags: 4 x (2,F,B) float ; out: 4 x (2,F,B) complex
void ref_wiener(float *mix, const float *mags, int F, float *out)
{
    quiet_scope qs(g_quiet);
    const int B = FFT_WINDOW_SIZE / 2 + 1;
    Eigen::Tensor3dXcf mixt(2, F, B);
    std::memcpy((void *)mixt.data(), mix, sizeof(std::complex<float>) * mixt.size());
    std::vector<Eigen::Tensor3dXf> mg;
    for (int j = 0; j < 4; ++j)
    {
        Eigen::Tensor3dXf t(2, F, B);
        std::memcpy(t.data(), mags + (size_t)j * t.size(), sizeof(float) * t.size());
        mg.push_back(t);
    }
    auto y = wiener_filter(mixt, mg);
    std::memcpy(mix, mixt.data(), sizeof(std::complex<float>) * mixt.size());
    for (int j = 0; j < 4; ++j)
        std::memcpy(out + (size_t)j * 2 * y[j].size(), y[j].data(),
                    sizeof(std::complex<float>) * y[j].size());
}

// ---- umx_inference (src/inference.cpp:12) ---------------------------------
// audio (2,n); buffers for n_buf; state from ref_lstm_create(hidden/2, frames(n_buf))
// out: 4 x (2,n)
void ref_umx_inference(void *model, void *state, const float *audio, int n,
                       int n_buf, float *out)
{
    quiet_scope qs(g_quiet);
    auto *m = static_cast<umx_model *>(model);
    auto *s = static_cast<lstm_set *>(state);
    stft_buffers buf(n_buf);
    // A chunk shorter than the buffer makes the reference's istft assign an n_buf-long row into an n-long one
    // (src/dsp.cpp:203-205): with -DNDEBUG that runs past the end of the waveform matrix and corrupts the heap
    // of whoever hosts this library.  The zero-extended chunk gives bit-identical results for the first n
    // samples (the buffers are zero past the chunk anyway) without the overflow.
    Eigen::MatrixXf a = Eigen::MatrixXf::Zero(2, n_buf);
    a.leftCols(n) = Eigen::Map<const Eigen::MatrixXf>(audio, 2, n);
    auto w = umx_inference(*m, a, buf, s->d);
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < n; ++i)
        {
            out[(size_t)j * 2 * n + 2 * i] = w[j](0, i);
            out[(size_t)j * 2 * n + 2 * i + 1] = w[j](1, i);
        }
}

// ---- the reference CLI itself (umx.cpp:26) --------------------------------
int ref_main(int argc, const char **argv)
{
    quiet_scope qs(g_quiet);
    return umx_ref_main(argc, argv);
}

} // extern "C"
