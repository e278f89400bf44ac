// Device-resident UMX model (replaces struct umx_model + load_umx_model, src/model.hpp:12-58).
#pragma once
#include "common.cuh"
#include <map>
#include <string>
#include <vector>

namespace umx
{

struct DevTensor
{
    float *d = nullptr;        // dequantised fp32, PyTorch C-order == Eigen column-major of (rows, cols)
    int rows = 0, cols = 0;    // the shape the reference's Eigen matrix has (ggml dims[0], dims[1])
    long n = 0;
    std::vector<float> host;   // host copy of 1-D tensors (used to fold BN / affine constants)
};

struct TargetWeights
{
    std::map<std::string, DevTensor> f32; // by ggml tensor name
    // recurrent weights stay quantised: [k][unit][gate] uint8 + (scale, offset)
    uint8_t *whh_q[3][2] = {};
    float whh_scale[3][2] = {}, whh_offset[3][2] = {};
    float wih_scale[3][2] = {}, wih_offset[3][2] = {};
    uint8_t *whh_img[3][2] = {}; // pre-swizzled SMEM image for the cluster kernel (hidden 1024 only)
    // GEMM operands (fp16 hi/lo planes, K-major) and fused-epilogue constants
    // W_ih of both directions stacked as the EXACT quantised integers in fp16 ([2 dirs * 2H][H]); the GEMM epilogue
    // applies W = scale*q + offset: x.W = scale*(x.q) + offset*rowsum(x) + (b_ih + b_hh)
    __half *wih_q[3] = {};
    float *wih_alpha[3] = {}, *wih_gamma[3] = {}, *wih_beta[3] = {}; // scale, offset, b_ih + b_hh per output row
    __half *fc2_hi = nullptr, *fc2_lo = nullptr; // [H][2H]
    float *fc2_alpha = nullptr, *fc2_beta = nullptr;
    __half *fc3_hi = nullptr, *fc3_lo = nullptr; // [4104][H] (rows >= 4098 zero)
    float *fc3_alpha = nullptr, *fc3_beta = nullptr;
};

struct Model
{
    int device = 0;
    int hidden = 0; // umx_model::hidden_size
    TargetWeights t[kTargets];
    // fc1 for all four targets as one operand: rows target*H + o, input affine folded in
    __half *fc1_hi = nullptr, *fc1_lo = nullptr; // [4H][kFc1InPad]
    float *fc1_alpha = nullptr, *fc1_beta = nullptr; // [4H]
    std::vector<void *> allocs; // everything cudaMalloc'ed, for destroy
    size_t payload_bytes = 0;
    int n_tensors = 0;
};

int model_create(const void *bytes, size_t n, int device, Model **out);
void model_destroy(Model *m);
// reference member name ("fc1_w", "lstm_hh_w", ...) -> ggml tensor name
std::string ref_name_to_ggml(const std::string &name, int layer, int dir);

} // namespace umx
