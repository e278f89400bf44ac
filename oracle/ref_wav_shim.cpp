// TEST INFRASTRUCTURE ONLY.
//
// The reference's dsp.cpp (load_audio / write_audio_file, src/dsp.cpp:18-101) calls two
// libnyquist entry points.  libnyquist needs its own cmake build, which the oracle recipe
// must not run, so this file supplies just those two symbols for RIFF/WAVE files:
//   nqr::NyquistIO::Load(AudioData*, const std::string&)   (vendor/libnyquist/include/libnyquist/Decoders.h:61)
//   nqr::encode_wav_to_disk(EncoderParams, const AudioData*, const std::string&)   (Encoders.h:36)
// PCM16 / PCM24 / PCM32 / float32 input, float32 output (the only format dsp.cpp:97 asks for).
#include <libnyquist/Common.h>
#include <libnyquist/Decoders.h>
#include <libnyquist/Encoders.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace nqr
{

NyquistIO::NyquistIO() {}
NyquistIO::~NyquistIO() {}

static uint32_t rd32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }

void NyquistIO::Load(AudioData *data, const std::string &path)
{
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f)
        throw std::runtime_error("wav shim: cannot open " + path);
    std::fseek(f, 0, SEEK_END);
    long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> buf((size_t)sz);
    if (std::fread(buf.data(), 1, buf.size(), f) != buf.size())
    {
        std::fclose(f);
        throw std::runtime_error("wav shim: short read");
    }
    std::fclose(f);
    if (sz < 12 || std::memcmp(buf.data(), "RIFF", 4) || std::memcmp(buf.data() + 8, "WAVE", 4))
        throw std::runtime_error("wav shim: not a RIFF/WAVE file");

    int fmt_tag = 0, channels = 0, rate = 0, bits = 0;
    const uint8_t *pcm = nullptr;
    size_t pcm_bytes = 0;
    size_t pos = 12;
    while (pos + 8 <= buf.size())
    {
        uint32_t len = rd32(&buf[pos + 4]);
        const uint8_t *body = &buf[pos + 8];
        if (!std::memcmp(&buf[pos], "fmt ", 4))
        {
            fmt_tag = rd16(body);
            channels = rd16(body + 2);
            rate = (int)rd32(body + 4);
            bits = rd16(body + 14);
            if (fmt_tag == 0xFFFE && len >= 26) // WAVE_FORMAT_EXTENSIBLE: sub-format GUID's first word
                fmt_tag = rd16(body + 24);
        }
        else if (!std::memcmp(&buf[pos], "data", 4))
        {
            pcm = body;
            pcm_bytes = std::min<size_t>(len, buf.size() - (pos + 8));
            break;
        }
        pos += 8 + len + (len & 1);
    }
    if (!pcm || !channels)
        throw std::runtime_error("wav shim: missing fmt/data chunk");

    const size_t n = pcm_bytes / (bits / 8);
    data->channelCount = channels;
    data->sampleRate = rate;
    data->frameSize = (size_t)channels * bits;
    data->lengthSeconds = (double)n / channels / rate;
    data->samples.resize(n);
    if (fmt_tag == 3 && bits == 32)
    {
        data->sourceFormat = PCM_FLT;
        std::memcpy(data->samples.data(), pcm, n * 4);
    }
    else if (fmt_tag == 1 && bits == 16)
    {
        data->sourceFormat = PCM_16;
        for (size_t i = 0; i < n; ++i)
            data->samples[i] = int16_to_float32((int16_t)rd16(pcm + 2 * i));
    }
    else if (fmt_tag == 1 && bits == 24)
    {
        data->sourceFormat = PCM_24;
        for (size_t i = 0; i < n; ++i)
            data->samples[i] = int24_to_float32(Pack(pcm[3 * i], pcm[3 * i + 1], pcm[3 * i + 2]));
    }
    else if (fmt_tag == 1 && bits == 32)
    {
        data->sourceFormat = PCM_32;
        for (size_t i = 0; i < n; ++i)
            data->samples[i] = int32_to_float32((int32_t)rd32(pcm + 4 * i));
    }
    else
        throw std::runtime_error("wav shim: unsupported sample format");
}

int encode_wav_to_disk(const EncoderParams p, const AudioData *d, const std::string &path)
{
    if (p.targetFormat != PCM_FLT)
        return EncoderError::UnsupportedBitdepth;
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f)
        return EncoderError::FileIOError;
    const uint32_t data_bytes = (uint32_t)(d->samples.size() * sizeof(float));
    const uint16_t ch = (uint16_t)d->channelCount;
    const uint32_t rate = (uint32_t)d->sampleRate;
    uint8_t h[58];
    std::memset(h, 0, sizeof h);
    auto w32 = [&](int o, uint32_t v) { std::memcpy(h + o, &v, 4); };
    auto w16 = [&](int o, uint16_t v) { std::memcpy(h + o, &v, 2); };
    std::memcpy(h, "RIFF", 4);
    w32(4, 50 + data_bytes);
    std::memcpy(h + 8, "WAVEfmt ", 8);
    w32(16, 18);
    w16(20, 3); // IEEE float
    w16(22, ch);
    w32(24, rate);
    w32(28, rate * ch * 4);
    w16(32, (uint16_t)(ch * 4));
    w16(34, 32);
    w16(36, 0);
    std::memcpy(h + 38, "fact", 4);
    w32(42, 4);
    w32(46, (uint32_t)(d->samples.size() / ch));
    std::memcpy(h + 50, "data", 4);
    w32(54, data_bytes);
    std::fwrite(h, 1, sizeof h, f);
    std::fwrite(d->samples.data(), 1, data_bytes, f);
    std::fclose(f);
    return EncoderError::NoError;
}

} // namespace nqr
