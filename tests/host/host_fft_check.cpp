// CPU check of the exact FFT code the CUDA kernels run (umx.cpp_b200/csrc/fft4096.cuh):
// the 256 "threads" of a CTA are emulated pass by pass, and the result is compared with a
// double-precision DFT.  Also checks the symmetric-padding index map against a literal restatement
// of pad_signal (src/dsp.cpp:109-128).  Built and run by tests/test_host_logic.py.
#include "../../umx.cpp_b200/csrc/fft4096.cuh"
#include <array>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

using namespace umx;

template <bool INV>
static double run(unsigned seed, bool small_table = false)
{
    std::vector<float2> tw(4096), x(4096), s(kFftSmemFloat2), out(4096);
    for (int j = 0; j < 4096; ++j)
    {
        double a = 2.0 * M_PI * j / 4096.0;
        tw[j] = make_float2((float)cos(a), (float)-sin(a));
    }
    srand(seed);
    for (auto &v : x)
        v = make_float2(rand() / (float)RAND_MAX - 0.5f, rand() / (float)RAND_MAX - 0.5f);
    std::vector<std::array<float2, 16>> regs(256);
    auto R = [&](int t) -> float2(&)[16] { return *reinterpret_cast<float2(*)[16]>(regs[t].data()); };
    std::vector<float2> t256(256);
    for (int m = 0; m < 256; ++m)
        t256[m] = tw[16 * m];
    for (int t = 0; t < 256; ++t)
    {
        for (int a = 0; a < 16; ++a)
            R(t)[a] = x[256 * a + t];
        if (small_table)
            fft_pass1_t256<INV>(t, R(t), s.data(), t256.data());
        else
            fft_pass1<INV>(t, R(t), s.data(), tw.data());
    }
    for (int t = 0; t < 256; ++t)
        fft_pass2_load(t, R(t), s.data());
    for (int t = 0; t < 256; ++t)
    {
        if (small_table)
            fft_pass2_store_t256<INV>(t, R(t), s.data(), t256.data(), tw[(t >> 4) * (t & 15)]);
        else
            fft_pass2_store<INV>(t, R(t), s.data(), tw.data());
    }
    for (int t = 0; t < 256; ++t)
    {
        fft_pass3<INV>(t, R(t), s.data());
        for (int r = 0; r < 16; ++r)
            out[t + 256 * r] = R(t)[r];
    }
    // reference DFT in double (O(N^2) on a subset of bins)
    double maxerr = 0, maxref = 0;
    for (int k = 0; k < 4096; k += 37)
    {
        std::complex<double> acc = 0;
        for (int n = 0; n < 4096; ++n)
        {
            double a = (INV ? 2.0 : -2.0) * M_PI * (double)((long)n * k % 4096) / 4096.0;
            acc += std::complex<double>(x[n].x, x[n].y) * std::complex<double>(cos(a), sin(a));
        }
        maxerr = std::max(maxerr, std::abs(acc - std::complex<double>(out[k].x, out[k].y)));
        maxref = std::max(maxref, std::abs(acc));
    }
    return maxerr / maxref;
}

static int check_pad(int n_buf)
{
    const int size = n_buf + 4096, pad = 2048;
    std::vector<int> p(size, -1);
    for (int i = 0; i < n_buf; ++i)
        p[pad + i] = i;
    // literal restatement of pad_signal: copy, reverse, copy back
    std::vector<int> ps(p.begin() + pad, p.begin() + 2 * pad), pe(p.end() - 2 * pad, p.end() - pad);
    std::reverse(ps.begin(), ps.end());
    std::reverse(pe.begin(), pe.end());
    std::copy(ps.begin(), ps.end(), p.begin());
    std::copy(pe.begin(), pe.end(), p.end() - pad);
    for (int P = 0; P < size; ++P)
        if (padded_to_sample(P, n_buf) != p[P])
        {
            printf("pad mismatch n_buf=%d P=%d got %d want %d\n", n_buf, P, padded_to_sample(P, n_buf), p[P]);
            return 1;
        }
    return 0;
}

int main()
{
    double ef = run<false>(1), ei = run<true>(2);
    printf("fft4096 forward rel err %.3e inverse rel err %.3e\n", ef, ei);
    int bad = (ef > 2e-6) || (ei > 2e-6);
    double sf = run<false>(3, true), si = run<true>(4, true); // 256-entry twiddle table variants
    printf("fft4096 (256-entry twiddle table) forward rel err %.3e inverse rel err %.3e\n", sf, si);
    bad |= (sf > 2e-6) || (si > 2e-6);
    bad |= check_pad(4096) | check_pad(20000) | check_pad(2646000);
    printf(bad ? "FAIL\n" : "OK\n");
    return bad;
}
