// umx.make-ggml <out file(.gz)> [hidden_size=1024] [seed=1234] [--dump-f32 <file>]
// Writes a ggml model file with synthetic weights (the distributions of SURVEY Appendix D) through cli/umx_ggml_writer.hpp:
// fixtures for the loader, the CLI and the benchmarks without Python and without the (unavailable) zenodo checkpoints.
// --dump-f32 also writes the fp32 values before quantisation (all tensors back to back) so a test can bound the round trip.
#include "umx_ggml_writer.hpp"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <random>

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        std::cerr << "Usage: " << argv[0] << " <out file(.gz)> [hidden_size=1024] [seed=1234] [--dump-f32 <file>]" << std::endl;
        return 1;
    }
    const std::string out = argv[1];
    int hidden = 1024;
    unsigned seed = 1234;
    std::string dump;
    int pos = 0;
    for (int i = 2; i < argc; ++i)
    {
        if (!std::strcmp(argv[i], "--dump-f32") && i + 1 < argc)
            dump = argv[++i];
        else if (pos == 0)
            hidden = std::atoi(argv[i]), ++pos;
        else
            seed = (unsigned)std::atoi(argv[i]), ++pos;
    }
    if (hidden < 2 || hidden % 2)
    {
        std::cerr << "hidden_size must be even and >= 2" << std::endl;
        return 1;
    }
    std::mt19937 rng(seed);
    auto fill = [&](std::vector<float> &v, float lo, float hi) {
        std::uniform_real_distribution<float> d(lo, hi);
        for (auto &x : v)
            x = d(rng);
    };
    umxggml::Writer w(hidden);
    FILE *df = dump.empty() ? nullptr : std::fopen(dump.c_str(), "wb");
    for (int target = 0; target < 4; ++target)
        for (const auto &[name, shape] : umxggml::tensor_list(hidden))
        {
            size_t n = 1;
            for (int d : shape)
                n *= (size_t)d;
            std::vector<float> v(n);
            auto has = [&](const char *k) { return name.find(k) != std::string::npos; };
            if (name == "input_mean") fill(v, -0.6f, -0.1f);
            else if (name == "input_scale") fill(v, 0.5f, 2.0f);
            else if (name == "output_scale") fill(v, 0.5f, 1.5f);
            else if (name == "output_mean") fill(v, 0.2f, 0.8f);
            else if (name == "fc1.weight") { const float b = 1.f / std::sqrt(2974.f); fill(v, -b, b); }
            else if (has("lstm.")) { const float b = 1.f / std::sqrt((float)(hidden / 2)); fill(v, -b, b); }
            else if (name == "fc2.weight") { const float b = 1.f / std::sqrt(2.f * hidden); fill(v, -b, b); }
            else if (name == "fc3.weight") { const float b = 1.f / std::sqrt((float)hidden); fill(v, -b, b); }
            else if (has("running_var") || has(".weight")) fill(v, 0.5f, 1.5f); // BatchNorm scale / variance
            else fill(v, -0.1f, 0.1f);                                          // BatchNorm bias / running mean
            w.add(name, shape, v.data());
            if (df)
                std::fwrite(v.data(), sizeof(float), n, df);
        }
    if (df)
        std::fclose(df);
    const bool gz = out.size() > 3 && out.compare(out.size() - 3, 3, ".gz") == 0;
    if (!w.save(out, gz))
    {
        std::cerr << "failed to write " << out << std::endl;
        return 1;
    }
    std::cout << "wrote " << out << ": 172 tensors, " << w.bytes.size() << " bytes" << (gz ? " (gzip)" : "") << ", hidden_size " << hidden << std::endl;
    return 0;
}
