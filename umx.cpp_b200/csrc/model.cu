// ggml model loading (reference: src/model.cpp:42-665; format written by
// scripts/convert-umx-pth-to-ggml.py:96-160).  The host gunzips (in memory: no ./temp.decompressed
// round trip) and parses headers; the GPU dequantises and builds the GEMM / recurrence operands.
#include "model.h"
#include "kernels.cuh"

#include <cmath>
#include <cstring>
#include <zlib.h>

namespace umx
{

namespace
{

constexpr uint32_t kMagic = 0x756d7867; // "umxg", src/model.cpp:101

bool gunzip(const uint8_t *src, size_t n, std::vector<uint8_t> &out)
{
    z_stream zs;
    std::memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, 16 + MAX_WBITS) != Z_OK)
        return false;
    zs.next_in = const_cast<Bytef *>(src);
    zs.avail_in = (uInt)n;
    out.resize(std::max<size_t>(n * 3, 1 << 20));
    size_t produced = 0;
    int rc = Z_OK;
    while (rc != Z_STREAM_END)
    {
        if (produced == out.size())
            out.resize(out.size() * 2);
        zs.next_out = out.data() + produced;
        const size_t room = std::min<size_t>(out.size() - produced, 1u << 30);
        zs.avail_out = (uInt)room;
        rc = inflate(&zs, Z_NO_FLUSH);
        produced += room - zs.avail_out;
        if (rc != Z_OK && rc != Z_STREAM_END)
        {
            inflateEnd(&zs);
            return false;
        }
        if (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0)
            break; // truncated stream
    }
    inflateEnd(&zs);
    out.resize(produced);
    return rc == Z_STREAM_END;
}

// The 43 tensor names of one target, matched EXACTLY (the reference's loader compares whole names too and answers anything
// else with "failed to load", src/model.cpp:240-546).  kind: what the tensor is; layer/dir only for lstm.* names.
struct TensorId
{
    enum Kind { kUnknown, kInAffine, kOutAffine, kFc1, kFc2, kFc3, kBn12, kBn3, kWih, kWhh, kLstmBias } kind = kUnknown;
    int layer = 0, dir = 0;
};

TensorId identify_tensor(const std::string &name)
{
    TensorId id;
    static const char *bn_fields[] = {"weight", "bias", "running_mean", "running_var"};
    if (name == "input_mean" || name == "input_scale") id.kind = TensorId::kInAffine;
    else if (name == "output_mean" || name == "output_scale") id.kind = TensorId::kOutAffine;
    else if (name == "fc1.weight") id.kind = TensorId::kFc1;
    else if (name == "fc2.weight") id.kind = TensorId::kFc2;
    else if (name == "fc3.weight") id.kind = TensorId::kFc3;
    else if (name.size() > 4 && name.compare(0, 2, "bn") == 0 && name[3] == '.' && name[2] >= '1' && name[2] <= '3')
    {
        for (const char *f : bn_fields)
            if (name.compare(4, std::string::npos, f) == 0)
                id.kind = name[2] == '3' ? TensorId::kBn3 : TensorId::kBn12;
    }
    else if (name.compare(0, 5, "lstm.") == 0)
    { // lstm.{weight_ih,weight_hh,bias_ih,bias_hh}_l{0,1,2}[_reverse]
        static const struct { const char *stem; TensorId::Kind kind; } stems[] = {
            {"lstm.weight_ih_l", TensorId::kWih}, {"lstm.weight_hh_l", TensorId::kWhh},
            {"lstm.bias_ih_l", TensorId::kLstmBias}, {"lstm.bias_hh_l", TensorId::kLstmBias}};
        for (const auto &st : stems)
        {
            const size_t len = std::strlen(st.stem);
            if (name.compare(0, len, st.stem) != 0 || name.size() <= len)
                continue;
            const char digit = name[len];
            const std::string tail = name.substr(len + 1);
            if (digit < '0' || digit > '2' || (!tail.empty() && tail != "_reverse"))
                continue;
            id.kind = st.kind;
            id.layer = digit - '0';
            id.dir = tail.empty() ? 0 : 1;
        }
    }
    return id;
}

struct Expect
{
    int d0, d1; // ggml dims (reversed PyTorch shape); d1 == 1 for vectors
};

// shapes the reference allocates (src/model.cpp:140-186); 0 = unknown tensor name
Expect expected_shape(const TensorId &id, int H)
{
    switch (id.kind)
    {
    case TensorId::kInAffine: return {kBinsCropped, 1};
    case TensorId::kOutAffine: return {kBins, 1};
    case TensorId::kFc1: return {kFc1In, H};
    case TensorId::kBn12: return {H, 1};
    case TensorId::kBn3: return {kFc3Out, 1};
    case TensorId::kFc2: return {2 * H, H};
    case TensorId::kFc3: return {H, kFc3Out};
    case TensorId::kWih: return {H, 2 * H};
    case TensorId::kWhh: return {H / 2, 2 * H};
    case TensorId::kLstmBias: return {2 * H, 1};
    default: return {0, 0};
    }
}

template <typename T>
int dev_alloc(Model *m, T **p, size_t count)
{
    void *q = nullptr;
    UMX_CUDA_OK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    m->allocs.push_back(q);
    *p = static_cast<T *>(q);
    return 0;
}

int upload(Model *m, float **p, const std::vector<float> &v)
{
    if (dev_alloc(m, p, v.size()))
        return -1;
    UMX_CUDA_OK(cudaMemcpy(*p, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
    return 0;
}

} // namespace

std::string ref_name_to_ggml(const std::string &name, int layer, int dir)
{
    static const std::map<std::string, std::string> plain = {
        {"input_mean", "input_mean"},   {"input_scale", "input_scale"}, {"output_mean", "output_mean"},
        {"output_scale", "output_scale"}, {"fc1_w", "fc1.weight"},      {"fc2_w", "fc2.weight"},
        {"fc3_w", "fc3.weight"},        {"bn1_w", "bn1.weight"},        {"bn1_b", "bn1.bias"},
        {"bn1_rm", "bn1.running_mean"}, {"bn1_rv", "bn1.running_var"},  {"bn2_w", "bn2.weight"},
        {"bn2_b", "bn2.bias"},          {"bn2_rm", "bn2.running_mean"}, {"bn2_rv", "bn2.running_var"},
        {"bn3_w", "bn3.weight"},        {"bn3_b", "bn3.bias"},          {"bn3_rm", "bn3.running_mean"},
        {"bn3_rv", "bn3.running_var"}};
    auto it = plain.find(name);
    if (it != plain.end())
        return it->second;
    static const std::map<std::string, std::string> lstm = {{"lstm_ih_w", "weight_ih"}, {"lstm_hh_w", "weight_hh"},
                                                            {"lstm_ih_b", "bias_ih"},   {"lstm_hh_b", "bias_hh"}};
    auto jt = lstm.find(name);
    if (jt == lstm.end())
        return "";
    return "lstm." + jt->second + "_l" + std::to_string(layer) + (dir ? "_reverse" : "");
}

void model_destroy(Model *m)
{
    if (!m)
        return;
    cudaSetDevice(m->device);
    for (void *p : m->allocs)
        cudaFree(p);
    delete m;
}

static int build_operands(Model *m, std::map<std::string, uint8_t *> (&whh_raw)[kTargets]);

int model_create(const void *bytes, size_t n, int device, Model **out)
{
    UMX_CHECK(bytes != nullptr && n >= 8 && out != nullptr, "model_create: bad arguments");
    int ndev = 0;
    UMX_CUDA_OK(cudaGetDeviceCount(&ndev));
    UMX_CHECK(device >= 0 && device < ndev, "model_create: no such CUDA device");
    UMX_CUDA_OK(cudaSetDevice(device));

    const uint8_t *raw = static_cast<const uint8_t *>(bytes);
    std::vector<uint8_t> inflated;
    if (raw[0] == 0x1f && raw[1] == 0x8b)
    {
        UMX_CHECK(gunzip(raw, n, inflated), "model_create: gzip stream is corrupt");
        raw = inflated.data();
        n = inflated.size();
    }
    uint32_t magic, hidden;
    std::memcpy(&magic, raw, 4);
    std::memcpy(&hidden, raw + 4, 4);
    UMX_CHECK(magic == kMagic, "load_umx_model: invalid model data (bad magic)");
    UMX_CHECK(hidden >= 2 && hidden <= 4096 && hidden % 2 == 0, "load_umx_model: unsupported hidden size");

    Model *m = new Model();
    m->device = device;
    m->hidden = (int)hidden;
    const int H = m->hidden;
    std::map<std::string, uint8_t *> whh_raw[kTargets];
    uint8_t *staging = nullptr;
    size_t staging_cap = 0;
    auto fail = [&](int rc) {
        if (staging)
            cudaFree(staging);
        model_destroy(m);
        return rc;
    };

    size_t pos = 8;
    int target = 0;
    while (pos + 16 <= n) // EOF ends the stream (src/model.cpp:229-232)
    {
        float scale, offset;
        int32_t n_dims, name_len;
        std::memcpy(&scale, raw + pos, 4);
        std::memcpy(&offset, raw + pos + 4, 4);
        std::memcpy(&n_dims, raw + pos + 8, 4);
        std::memcpy(&name_len, raw + pos + 12, 4);
        pos += 16;
        if (n_dims < 1 || n_dims > 2 || name_len <= 0 || name_len > 256 || pos + 4 * n_dims + name_len > n)
        {
            set_error("load_umx_model: corrupt tensor header");
            return fail(-3);
        }
        int32_t ne[2] = {1, 1};
        for (int i = 0; i < n_dims; ++i)
            std::memcpy(&ne[i], raw + pos + 4 * i, 4);
        pos += 4 * n_dims;
        const std::string name(reinterpret_cast<const char *>(raw + pos), name_len);
        pos += name_len;
        if (target >= kTargets)
        {
            set_error("load_umx_model: more than 4 targets in model file");
            return fail(-3);
        }
        const TensorId id = identify_tensor(name);
        const Expect ex = expected_shape(id, H);
        if (ex.d0 == 0)
        {
            set_error("load_umx_model: failed to load " + name); // src/model.cpp:541-546
            return fail(-3);
        }
        if (m->t[target].f32.count(name))
        {
            set_error("load_umx_model: tensor '" + name + "' appears twice in target " + std::to_string(target));
            return fail(-3);
        }
        if (ne[0] != ex.d0 || ne[1] != ex.d1)
        {
            set_error("load_single_matrix: tensor '" + name + "' has wrong size in model file"); // :582-591
            return fail(-3);
        }
        // u16 payloads: fc2 / fc3 / bn2.* / bn3.* (src/model.cpp:390-539 call load_single_matrix_uint16 for exactly these)
        const bool wide = id.kind == TensorId::kFc2 || id.kind == TensorId::kFc3 || id.kind == TensorId::kBn3 ||
                          (id.kind == TensorId::kBn12 && name[2] == '2');
        const long count = (long)ne[0] * ne[1];
        const size_t nbytes = (size_t)count * (wide ? 2 : 1);
        if (pos + nbytes > n)
        {
            set_error("load_umx_model: truncated payload for " + name);
            return fail(-3);
        }
        if (nbytes > staging_cap)
        {
            if (staging)
                cudaFree(staging);
            staging_cap = std::max<size_t>(nbytes, 16u << 20);
            if (cudaMalloc(&staging, staging_cap) != cudaSuccess)
            {
                staging = nullptr;
                set_error("model_create: cudaMalloc(staging) failed");
                return fail(-1);
            }
        }
        DevTensor dt;
        dt.rows = ne[0];
        dt.cols = ne[1];
        dt.n = count;
        if (dev_alloc(m, &dt.d, (size_t)count))
            return fail(-1);
        if (cudaMemcpy(staging, raw + pos, nbytes, cudaMemcpyHostToDevice) != cudaSuccess)
        {
            set_error("model_create: payload upload failed");
            return fail(-1);
        }
        const int rc = wide ? launch_dequant_u16(reinterpret_cast<const uint16_t *>(staging), scale, offset, dt.d, count, 0)
                            : launch_dequant_u8(staging, scale, offset, dt.d, count, 0);
        if (rc)
            return fail(rc);
        if (ne[1] == 1)
        { // vectors: host copy with the same single fused multiply-add
            dt.host.resize(count);
            for (long i = 0; i < count; ++i)
            {
                uint16_t q = wide ? (uint16_t)(raw[pos + 2 * i] | (raw[pos + 2 * i + 1] << 8)) : raw[pos + i];
                dt.host[i] = std::fmaf((float)q, scale, offset);
            }
        }
        if (id.kind == TensorId::kWih)
        { // keep the quantised input weights: the GEMM consumes the exact integers
            uint8_t *q = nullptr;
            if (dev_alloc(m, &q, nbytes))
                return fail(-1);
            if (cudaMemcpy(q, staging, nbytes, cudaMemcpyDeviceToDevice) != cudaSuccess)
                return fail(-1);
            whh_raw[target][name] = q;
            const int layer = id.layer, dir = id.dir; // validated by identify_tensor: layer in 0..2
            m->t[target].wih_scale[layer][dir] = scale;
            m->t[target].wih_offset[layer][dir] = offset;
        }
        if (id.kind == TensorId::kWhh)
        { // keep the quantised recurrent weights
            uint8_t *q = nullptr;
            if (dev_alloc(m, &q, nbytes))
                return fail(-1);
            if (cudaMemcpy(q, staging, nbytes, cudaMemcpyDeviceToDevice) != cudaSuccess)
                return fail(-1);
            whh_raw[target][name] = q;
            const int layer = id.layer, dir = id.dir; // validated by identify_tensor: layer in 0..2
            m->t[target].whh_scale[layer][dir] = scale;
            m->t[target].whh_offset[layer][dir] = offset;
        }
        if (cudaDeviceSynchronize() != cudaSuccess) // staging is reused by the next tensor
        {
            set_error("model_create: dequantisation kernel failed");
            return fail(-1);
        }
        m->t[target].f32[name] = std::move(dt);
        m->payload_bytes += nbytes;
        m->n_tensors++;
        pos += nbytes;
        if (name == "bn3.running_var")
            target++; // src/model.cpp:530-539
    }
    if (staging)
        cudaFree(staging);
    staging = nullptr;
    if (target != kTargets)
    {
        set_error("load_umx_model: model file holds " + std::to_string(target) + " complete targets, expected 4");
        return fail(-3);
    }
    for (int t = 0; t < kTargets; ++t)
        if (m->t[t].f32.size() != 43)
        {
            set_error("load_umx_model: target " + std::to_string(t) + " is missing tensors");
            return fail(-3);
        }
    const int rc = build_operands(m, whh_raw);
    if (rc)
        return fail(rc);
    *out = m;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// GEMM / recurrence operands derived from the dequantised tensors
// ------------------------------------------------------------------------------------------------
static int build_operands(Model *m, std::map<std::string, uint8_t *> (&whh_raw)[kTargets])
{
    const int H = m->hidden, Hh = H / 2;
    auto bn_fold = [&](TargetWeights &tw, const char *p, std::vector<float> &alpha, std::vector<float> &shift) {
        // ((x - rm) / sqrt(rv + 1e-5)) * w + b  ==  x*alpha + shift      (src/inference.cpp:91-99)
        const auto &w = tw.f32[std::string(p) + ".weight"].host;
        const auto &b = tw.f32[std::string(p) + ".bias"].host;
        const auto &rm = tw.f32[std::string(p) + ".running_mean"].host;
        const auto &rv = tw.f32[std::string(p) + ".running_var"].host;
        alpha.resize(w.size());
        shift.resize(w.size());
        for (size_t i = 0; i < w.size(); ++i)
        {
            alpha[i] = w[i] / std::sqrt(rv[i] + kBnEps);
            shift[i] = b[i] - rm[i] * alpha[i];
        }
    };

    // ---- fc1, all targets in one operand; x*input_scale + input_mean folded in (inference.cpp:78-99)
    if (dev_alloc(m, &m->fc1_hi, (size_t)4 * H * kFc1InPad) || dev_alloc(m, &m->fc1_lo, (size_t)4 * H * kFc1InPad))
        return -1;
    std::vector<float> a1(4 * H), b1(4 * H);
    float *d_tmp = nullptr;
    UMX_CUDA_OK(cudaMalloc(&d_tmp, sizeof(float) * H));
    std::vector<float> bias(H);
    for (int t = 0; t < kTargets; ++t)
    {
        TargetWeights &tw = m->t[t];
        const float *w = tw.f32["fc1.weight"].d;
        int rc = launch_make_planes(w, H, kFc1In, tw.f32["input_scale"].d, kBinsCropped,
                                    m->fc1_hi + (size_t)t * H * kFc1InPad, m->fc1_lo + (size_t)t * H * kFc1InPad, H,
                                    kFc1InPad, 0);
        if (!rc)
            rc = launch_row_dot(w, H, kFc1In, tw.f32["input_mean"].d, kBinsCropped, d_tmp, 0);
        if (rc)
        {
            cudaFree(d_tmp);
            return rc;
        }
        if (cudaMemcpy(bias.data(), d_tmp, sizeof(float) * H, cudaMemcpyDeviceToHost) != cudaSuccess)
        {
            cudaFree(d_tmp);
            set_error("model_create: fc1 bias download failed");
            return -1;
        }
        std::vector<float> al, sh;
        bn_fold(tw, "bn1", al, sh);
        for (int o = 0; o < H; ++o)
        {
            a1[t * H + o] = al[o];
            b1[t * H + o] = bias[o] * al[o] + sh[o];
        }
    }
    cudaFree(d_tmp);
    if (upload(m, &m->fc1_alpha, a1) || upload(m, &m->fc1_beta, b1))
        return -1;

    for (int t = 0; t < kTargets; ++t)
    {
        TargetWeights &tw = m->t[t];
        // ---- LSTM: W_ih of both directions stacked (rows dir*2H + 4*unit + gate), bias b_ih + b_hh
        for (int l = 0; l < 3; ++l)
        {
            if (dev_alloc(m, &tw.wih_q[l], (size_t)4 * H * H))
                return -1;
            std::vector<float> beta(4 * H), alpha(4 * H), gamma(4 * H);
            for (int d = 0; d < 2; ++d)
            {
                const std::string sfx = "_l" + std::to_string(l) + (d ? "_reverse" : "");
                int rc = launch_q_to_half(whh_raw[t]["lstm.weight_ih" + sfx], Hh, H, tw.wih_q[l] + (size_t)d * 2 * H * H, 0);
                if (rc)
                    return rc;
                for (int g = 0; g < 2 * H; ++g)
                {
                    alpha[d * 2 * H + g] = tw.wih_scale[l][d];
                    gamma[d * 2 * H + g] = std::fmaf(127.0f, tw.wih_scale[l][d], tw.wih_offset[l][d]); // W = s*(q-127) + (o + 127 s)
                }
                const auto &bi = tw.f32["lstm.bias_ih" + sfx].host;
                const auto &bh = tw.f32["lstm.bias_hh" + sfx].host;
                for (int g = 0; g < 4; ++g)
                    for (int u = 0; u < Hh; ++u)
                        beta[d * 2 * H + 4 * u + g] = bi[g * Hh + u] + bh[g * Hh + u]; // lstm.cpp:132-140, gate-interleaved rows
                uint8_t *packed = nullptr;
                if (dev_alloc(m, &packed, (size_t)4 * Hh * Hh))
                    return -1;
                rc = launch_pack_whh(whh_raw[t]["lstm.weight_hh" + sfx], Hh, packed, 0);
                if (rc)
                    return rc;
                tw.whh_q[l][d] = packed;
                if (lstm_cluster_supported(Hh))
                {
                    uint8_t *img = nullptr;
                    if (dev_alloc(m, &img, whh_image_bytes(Hh)))
                        return -1;
                    rc = launch_whh_image(whh_raw[t]["lstm.weight_hh" + sfx], img, Hh, 0);
                    if (rc)
                        return rc;
                    tw.whh_img[l][d] = img;
                }
            }
            if (upload(m, &tw.wih_beta[l], beta) || upload(m, &tw.wih_alpha[l], alpha) || upload(m, &tw.wih_gamma[l], gamma))
                return -1;
        }
        // ---- fc2 + BN2 (inference.cpp:127-140)
        if (dev_alloc(m, &tw.fc2_hi, (size_t)H * 2 * H) || dev_alloc(m, &tw.fc2_lo, (size_t)H * 2 * H))
            return -1;
        int rc = launch_make_planes(tw.f32["fc2.weight"].d, H, 2 * H, nullptr, 1, tw.fc2_hi, tw.fc2_lo, H, 2 * H, 0);
        if (rc)
            return rc;
        std::vector<float> al, sh;
        bn_fold(tw, "bn2", al, sh);
        if (upload(m, &tw.fc2_alpha, al) || upload(m, &tw.fc2_beta, sh))
            return -1;
        // ---- fc3 + BN3 + x*output_scale + output_mean (inference.cpp:143-166); vectors repeat for L,R
        if (dev_alloc(m, &tw.fc3_hi, (size_t)kMaskLd * H) || dev_alloc(m, &tw.fc3_lo, (size_t)kMaskLd * H))
            return -1;
        rc = launch_make_planes(tw.f32["fc3.weight"].d, kFc3Out, H, nullptr, 1, tw.fc3_hi, tw.fc3_lo, kMaskLd, H, 0);
        if (rc)
            return rc;
        bn_fold(tw, "bn3", al, sh);
        const auto &os = tw.f32["output_scale"].host;
        const auto &om = tw.f32["output_mean"].host;
        std::vector<float> a3(kMaskLd, 0.f), b3(kMaskLd, 0.f);
        for (int n = 0; n < kFc3Out; ++n)
        {
            a3[n] = al[n] * os[n % kBins];
            b3[n] = sh[n] * os[n % kBins] + om[n % kBins];
        }
        if (upload(m, &tw.fc3_alpha, a3) || upload(m, &tw.fc3_beta, b3))
            return -1;
    }
    UMX_CUDA_OK(cudaDeviceSynchronize());
    return 0;
}

} // namespace umx
