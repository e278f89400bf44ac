// Shared constants and helpers for the sm_100a UMX-L demixing path.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>

namespace umx
{

// reference constants: src/dsp.hpp:16-19, src/inference.hpp:13-16, src/wiener.hpp:12-17
constexpr int kSampleRate = 44100;
constexpr int kNfft = 4096;
constexpr int kHop = 1024;
constexpr int kBins = kNfft / 2 + 1;      // 2049
constexpr int kPad = kNfft / 2;           // 2048
// Frames of an F-frame window that can reach one of the first n_wave samples: sample i sits at padded position i + kPad and
// frame f covers [kHop f, kHop f + kNfft).  Every later frame of a zero-extended chunk has an all-zero spectrum (so it adds
// nothing to the Wiener statistics whatever the network says there) and is never emitted: the post-network stages skip it.
__host__ __device__ inline int frames_reaching(int n_wave, int F)
{
    const int f = (n_wave > 0 ? n_wave + kPad + kHop - 1 : 0) / kHop;
    return f < F ? f : F;
}
constexpr int kBinsCropped = 1487;        // src/inference.cpp:41-55 (16 kHz bandwidth)
constexpr int kFc1In = 2 * kBinsCropped;  // 2974
constexpr int kFc1InPad = 2976;           // row pitch of the fc1 A operand (16-byte aligned rows)
constexpr int kFc3Out = 2 * kBins;        // 4098
constexpr int kMaskLd = 4104;             // row pitch of the fc3 output / mask matrix
constexpr int kTargets = 4;
constexpr int kSegmentSamples = 2646000;  // (int)(60.0f * 44100)                 umx.cpp:156-157
constexpr int kStrideSamples = 1984500;   // (int)((1 - 0.25f) * 2646000)          umx.cpp:180
constexpr int kMaxShift = 22050;          // (int)(0.5 * 44100)                    umx.cpp:112-113
constexpr float kBnEps = 1e-5f;           // src/inference.cpp:95
constexpr float kWienerEps = 1e-10f;      // src/wiener.hpp:12
constexpr float kWienerScale = 10.0f;     // src/wiener.hpp:13

// last error string, returned by umx_cuda_last_error()
void set_error(const std::string &msg);

#define UMX_CUDA_OK(expr)                                                                     \
    do                                                                                        \
    {                                                                                         \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
        {                                                                                     \
            ::umx::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +      \
                             " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")");         \
            return -1;                                                                        \
        }                                                                                     \
    } while (0)

#define UMX_CHECK(cond, msg)                                                                  \
    do                                                                                        \
    {                                                                                         \
        if (!(cond))                                                                          \
        {                                                                                     \
            ::umx::set_error(std::string(msg) + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
            return -2;                                                                        \
        }                                                                                     \
    } while (0)

// One-time set-up per device (cudaFuncSetAttribute, device attributes) that may be reached from several host threads at once:
// sessions may be driven from different threads (umx_cuda.h).  `done` is only read and written under the mutex.
struct PerDeviceOnce
{
    std::mutex mu;
    bool done[64] = {};
    template <typename Fn>
    int run(Fn &&fn) // fn(dev) returns 0 on success; it runs once per device, later calls only take the lock
    {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
        {
            set_error("cudaGetDevice failed or device index >= 64");
            return -1;
        }
        std::lock_guard<std::mutex> lk(mu);
        if (done[dev])
            return 0;
        if (int rc = fn(dev))
            return rc;
        done[dev] = true;
        return 0;
    }
};

inline int ceil_div(long a, long b) { return (int)((a + b - 1) / b); }

// fp32 -> (hi, lo) fp16 pair with hi + lo == x to ~2^-22 relative (both round-to-nearest-even)
__device__ __forceinline__ void split_half(float x, __half &hi, __half &lo)
{
    hi = __float2half_rn(x);
    lo = __float2half_rn(x - __half2float(hi));
}

} // namespace umx
