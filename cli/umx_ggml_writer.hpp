// C++ writer for the ggml model file the reference loader parses (src/model.cpp:93-553), i.e. the output format of
// scripts/convert-umx-pth-to-ggml.py:96-160 without Python:
//     int32 magic 0x756d7867 ; int32 hidden_size
//     per target (bass, drums, other, vocals) and tensor, in state_dict order:
//         float32 scale, float32 offset, int32 n_dims, int32 name_len, int32 dims[n_dims] (REVERSED PyTorch shape), name bytes,
//         payload: uint16 for names containing bn2 / bn3 / fc2 / fc3, uint8 otherwise (convert...py:146-150), PyTorch C order
// Quantisation as convert...py:13-26: scale = (max - min) / (qmax - 1), offset = min, q = round((x - offset) / scale), in float32.
// Header-only, no dependencies beyond zlib (optional gzip output): a host that holds fp32 weights can write the file itself.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <zlib.h>

namespace umxggml
{

struct Writer
{
    std::vector<uint8_t> bytes;

    explicit Writer(int hidden_size)
    {
        put_i32(0x756d7867);
        put_i32(hidden_size);
    }

    static bool is_wide(const std::string &name)
    {
        for (const char *k : {"bn2", "bn3", "fc2", "fc3"})
            if (name.find(k) != std::string::npos)
                return true;
        return false;
    }

    // One tensor: `shape` is the PyTorch shape (1 or 2 dims), `data` in PyTorch C order.
    void add(const std::string &name, const std::vector<int> &shape, const float *data)
    {
        size_t n = 1;
        for (int d : shape)
            n *= (size_t)d;
        const bool wide = is_wide(name);
        const float qmax = wide ? 65535.f : 255.f;
        float lo = data[0], hi = data[0];
        for (size_t i = 1; i < n; ++i)
        {
            lo = std::min(lo, data[i]);
            hi = std::max(hi, data[i]);
        }
        const float scale = (hi - lo) / (qmax - 1.f), offset = lo;
        put_f32(scale);
        put_f32(offset);
        put_i32((int32_t)shape.size());
        put_i32((int32_t)name.size());
        for (size_t i = 0; i < shape.size(); ++i)
            put_i32(shape[shape.size() - 1 - i]);
        bytes.insert(bytes.end(), name.begin(), name.end());
        for (size_t i = 0; i < n; ++i)
        {
            // numpy: np.round((a - offset) / scale) in float32, round-half-to-even
            const float q = std::nearbyint((data[i] - offset) / scale);
            if (wide)
            {
                const uint16_t v = (uint16_t)q;
                bytes.push_back((uint8_t)(v & 0xff));
                bytes.push_back((uint8_t)(v >> 8));
            }
            else
                bytes.push_back((uint8_t)q);
        }
    }

    bool save(const std::string &path, bool gzip) const
    {
        if (gzip)
        {
            gzFile f = gzopen(path.c_str(), "wb1");
            if (!f)
                return false;
            size_t pos = 0;
            while (pos < bytes.size())
            {
                const unsigned chunk = (unsigned)std::min<size_t>(bytes.size() - pos, 1u << 24);
                if (gzwrite(f, bytes.data() + pos, chunk) != (int)chunk)
                {
                    gzclose(f);
                    return false;
                }
                pos += chunk;
            }
            return gzclose(f) == Z_OK;
        }
        FILE *f = std::fopen(path.c_str(), "wb");
        if (!f)
            return false;
        const bool ok = std::fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
        return (std::fclose(f) == 0) && ok;
    }

  private:
    void put_i32(int32_t v)
    {
        const uint8_t *p = reinterpret_cast<const uint8_t *>(&v);
        bytes.insert(bytes.end(), p, p + 4);
    }
    void put_f32(float v)
    {
        const uint8_t *p = reinterpret_cast<const uint8_t *>(&v);
        bytes.insert(bytes.end(), p, p + 4);
    }
};

// (name, PyTorch shape) of the 43 tensors of one target in state_dict order, for a given hidden size
inline std::vector<std::pair<std::string, std::vector<int>>> tensor_list(int H)
{
    std::vector<std::pair<std::string, std::vector<int>>> t = {
        {"input_mean", {1487}}, {"input_scale", {1487}}, {"output_scale", {2049}}, {"output_mean", {2049}}, {"fc1.weight", {H, 2974}}};
    auto bn = [&](const std::string &p, int n) {
        for (const char *f : {"weight", "bias", "running_mean", "running_var"})
            t.push_back({p + "." + f, {n}});
    };
    bn("bn1", H);
    for (int l = 0; l < 3; ++l)
        for (const char *sfx : {"", "_reverse"})
        {
            const std::string s = "_l" + std::to_string(l) + sfx;
            t.push_back({"lstm.weight_ih" + s, {2 * H, H}});
            t.push_back({"lstm.weight_hh" + s, {2 * H, H / 2}});
            t.push_back({"lstm.bias_ih" + s, {2 * H}});
            t.push_back({"lstm.bias_hh" + s, {2 * H}});
        }
    t.push_back({"fc2.weight", {H, 2 * H}});
    bn("bn2", H);
    t.push_back({"fc3.weight", {4098, H}});
    bn("bn3", 4098);
    return t;
}

} // namespace umxggml
