// umx.cpp.main <model file(.gz)> <wav file> <out dir> — the reference's CLI contract (umx.cpp:26-97) on the GPU path.
// C++17 host code over the C ABI (include/umx_cuda.h); no Eigen, no libnyquist: WAV in/out is handled here
// (PCM16/24/32 or float32 in, float32 stereo 44.1 kHz out as write_audio_file does, src/dsp.cpp:75-101).
// Writes target_0..3.wav = bass, drums, other, vocals.  Optional 4th argument: shift offset (default 4033, the
// first rand() % 22050 of an unseeded glibc process, umx.cpp:112-115).
#include "../include/umx_cuda.h"

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

namespace
{
uint32_t rd32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }

// load_audio (src/dsp.cpp:18-73): (2, N) interleaved float, mono duplicated, 44.1 kHz enforced
bool load_audio(const std::string &path, std::vector<float> &interleaved, size_t &n)
{
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f)
    {
        std::cerr << "[ERROR] cannot open " << path << std::endl;
        return false;
    }
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> buf((size_t)sz);
    const size_t got = std::fread(buf.data(), 1, buf.size(), f);
    std::fclose(f);
    if (got != buf.size() || sz < 12 || std::memcmp(buf.data(), "RIFF", 4) || std::memcmp(buf.data() + 8, "WAVE", 4))
    {
        std::cerr << "[ERROR] not a RIFF/WAVE file: " << path << std::endl;
        return false;
    }
    int tag = 0, ch = 0, rate = 0, bits = 0;
    const uint8_t *pcm = nullptr;
    size_t pcm_bytes = 0;
    for (size_t pos = 12; pos + 8 <= buf.size();)
    {
        const uint32_t len = rd32(&buf[pos + 4]);
        const uint8_t *body = &buf[pos + 8];
        if (!std::memcmp(&buf[pos], "fmt ", 4))
        {
            tag = rd16(body);
            ch = rd16(body + 2);
            rate = (int)rd32(body + 4);
            bits = rd16(body + 14);
            if (tag == 0xFFFE && len >= 26)
                tag = rd16(body + 24);
        }
        else if (!std::memcmp(&buf[pos], "data", 4))
        {
            pcm = body;
            pcm_bytes = std::min<size_t>(len, buf.size() - (pos + 8));
            break;
        }
        pos += 8 + len + (len & 1);
    }
    if (!pcm || !ch)
    {
        std::cerr << "[ERROR] missing fmt/data chunk" << std::endl;
        return false;
    }
    if (rate != UMX_SAMPLE_RATE)
    {
        std::cerr << "[ERROR] umx.cpp only supports the following sample rate (Hz): " << UMX_SAMPLE_RATE << std::endl;
        return false;
    }
    if (ch != 1 && ch != 2)
    {
        std::cerr << "[ERROR] umx.cpp only supports mono and stereo audio" << std::endl;
        return false;
    }
    const size_t total = pcm_bytes / (bits / 8);
    n = total / ch;
    std::cout << "Input Samples: " << total << "\nLength in seconds: " << (double)n / rate << "\nNumber of channels: " << ch << std::endl;
    auto sample = [&](size_t i) -> float {
        if (tag == 3 && bits == 32)
        {
            float v;
            std::memcpy(&v, pcm + 4 * i, 4);
            return v;
        }
        if (tag == 1 && bits == 16)
            return (float)(int16_t)rd16(pcm + 2 * i) / 32767.f;
        if (tag == 1 && bits == 24)
        {
            int32_t v = pcm[3 * i] | (pcm[3 * i + 1] << 8) | (pcm[3 * i + 2] << 16);
            if (v & 0x800000)
                v |= ~0xFFFFFF;
            return (float)v / 8388608.f;
        }
        if (tag == 1 && bits == 32)
            return (float)(int32_t)rd32(pcm + 4 * i) / 2147483648.f;
        return 0.f;
    };
    if (!((tag == 3 && bits == 32) || (tag == 1 && (bits == 16 || bits == 24 || bits == 32))))
    {
        std::cerr << "[ERROR] unsupported WAV sample format" << std::endl;
        return false;
    }
    interleaved.resize(2 * n);
    for (size_t i = 0; i < n; ++i)
    {
        interleaved[2 * i] = sample(ch * i);
        interleaved[2 * i + 1] = sample(ch * i + (ch - 1));
    }
    return true;
}

// write_audio_file (src/dsp.cpp:75-101): float32 stereo 44.1 kHz
bool write_wav(const std::string &path, const float *interleaved, size_t n)
{
    // RIFF sizes are 32-bit: longer than ~3.1 hours of float stereo does not fit a plain WAV
    if (n > (size_t)(UINT32_MAX - 36) / (2 * sizeof(float)))
    {
        std::cerr << "[ERROR] " << path << ": " << n << " samples do not fit a RIFF/WAV file (4 GiB limit)" << std::endl;
        return false;
    }
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f)
        return false;
    const uint32_t data_bytes = (uint32_t)(n * 2 * sizeof(float));
    uint8_t h[44];
    auto w32 = [&](int o, uint32_t v) { std::memcpy(h + o, &v, 4); };
    auto w16 = [&](int o, uint16_t v) { std::memcpy(h + o, &v, 2); };
    std::memcpy(h, "RIFF", 4);
    w32(4, 36 + data_bytes);
    std::memcpy(h + 8, "WAVEfmt ", 8);
    w32(16, 16);
    w16(20, 3);
    w16(22, 2);
    w32(24, UMX_SAMPLE_RATE);
    w32(28, UMX_SAMPLE_RATE * 8);
    w16(32, 8);
    w16(34, 32);
    std::memcpy(h + 36, "data", 4);
    w32(40, data_bytes);
    const bool ok = std::fwrite(h, 1, sizeof h, f) == sizeof h && std::fwrite(interleaved, 1, data_bytes, f) == data_bytes;
    return (std::fclose(f) == 0) && ok;
}
} // namespace

int main(int argc, const char **argv)
{
    if (argc != 4 && argc != 5)
    {
        std::cerr << "Usage: " << argv[0] << " <model dir> <wav file> <out dir> [shift offset]" << std::endl;
        return 1;
    }
    // one track per process: the driver's default number of hardware work queues is enough, and every extra queue costs ~50 ms
    // of start-up and teardown (the library's own default suits long-running hosts; see csrc/capi.cu).  Set before any CUDA call.
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "8", 0);
    std::cout << "umx.cpp Main driver program (B200 / CUDA path)" << std::endl;
    const std::string model_file = argv[1], wav_file = argv[2], out_dir = argv[3];
    const int shift = argc == 5 ? std::atoi(argv[4]) : UMX_DEFAULT_SHIFT_OFFSET;

    // the WAV is decoded on a second thread while the model loads (gunzip + parse + operand preparation on the GPU)
    std::vector<float> audio;
    size_t n = 0;
    bool audio_ok = false;
    std::thread audio_thread([&] { audio_ok = load_audio(wav_file, audio, n); });

    const auto t0 = std::chrono::steady_clock::now();
    umx_cuda_model *model = nullptr;
    const int rc = umx_cuda_model_create_from_file(model_file.c_str(), 0, &model);
    std::cout << "umx_model_load returned " << (rc == 0 ? "true" : "false") << std::endl;
    if (rc != 0)
    {
        audio_thread.join();
        std::cerr << "Error loading model: " << umx_cuda_last_error() << std::endl;
        return 1;
    }
    const auto t1 = std::chrono::steady_clock::now();
    umx_cuda_session *sess = nullptr;
    if (umx_cuda_session_create(model, 1, UMX_SEGMENT_SAMPLES, &sess) != 0)
    {
        audio_thread.join();
        std::cerr << "Error: " << umx_cuda_last_error() << std::endl;
        return 1;
    }
    const auto t1b = std::chrono::steady_clock::now();
    audio_thread.join();
    if (!audio_ok)
        return 1;
    // page-locked staging for both directions: the copies then run at PCIe speed and overlap the kernels
    float *pin_in = nullptr, *pin_out = nullptr;
    if (umx_cuda_host_alloc(sizeof(float) * 2 * std::max<size_t>(n, 1), reinterpret_cast<void **>(&pin_in)) != 0 ||
        umx_cuda_host_alloc(sizeof(float) * 8 * std::max<size_t>(n, 1), reinterpret_cast<void **>(&pin_out)) != 0)
    {
        std::cerr << "Error: " << umx_cuda_last_error() << std::endl;
        return 1;
    }
    std::memcpy(pin_in, audio.data(), sizeof(float) * 2 * n);
    const float *in_ptr = pin_in;
    float *out_ptr = pin_out;
    const int64_t n64 = (int64_t)n;
    const auto t2 = std::chrono::steady_clock::now();
    if (umx_cuda_demix(sess, &in_ptr, &n64, 1, shift, &out_ptr) != 0)
    {
        std::cerr << "Error: " << umx_cuda_last_error() << std::endl;
        return 1;
    }
    const auto t3 = std::chrono::steady_clock::now();
    auto secs = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
    std::cout << "Loaded model in " << secs(t0, t1) << " s; demixed " << (double)n / UMX_SAMPLE_RATE << " s of audio in " << secs(t2, t3)
              << " s" << std::endl;
    if (std::getenv("UMX_CLI_TIMES"))
        std::cout << "phases: model " << secs(t0, t1) << " s, session " << secs(t1, t1b) << " s, wait for audio + pinned staging "
                  << secs(t1b, t2) << " s, demix " << secs(t2, t3) << " s" << std::endl;

    // the four stems are written by four threads (the reference: `#pragma omp parallel for` over targets, umx.cpp:75)
    std::filesystem::create_directories(out_dir);
    bool ok[4] = {false, false, false, false};
    std::vector<std::thread> writers;
    for (int target = 0; target < 4; ++target)
    {
        const std::string p = (std::filesystem::path(out_dir) / ("target_" + std::to_string(target) + ".wav")).string();
        std::cout << "Writing wav file " << p << " to " << out_dir << std::endl;
        writers.emplace_back([=, &ok] { ok[target] = write_wav(p, pin_out + (size_t)target * 2 * n, n); });
    }
    for (auto &w : writers)
        w.join();
    if (std::getenv("UMX_CLI_TIMES"))
        std::cout << "phases: write " << secs(t3, std::chrono::steady_clock::now()) << " s" << std::endl;
    for (int target = 0; target < 4; ++target)
        if (!ok[target])
        {
            std::cerr << "Error writing target_" << target << ".wav" << std::endl;
            return 1;
        }
    umx_cuda_host_free(pin_in);
    umx_cuda_host_free(pin_out);
    umx_cuda_session_destroy(sess);
    umx_cuda_model_destroy(model);
    return 0;
}
