// Internal kernel interfaces of the demixing path (device pointers, explicit stream).
#pragma once
#include "common.cuh"

namespace umx
{

// ---- stft.cu -------------------------------------------------------------------------------------
struct StftTrack
{
    const float2 *audio; // interleaved stereo source track (device)
    long n_src;          // samples in the source track
    long src_shift;      // source index of chunk sample 0 (segment offset - shift offset; may be < 0)
    int n_valid;         // chunk length; samples >= n_valid of the n_buf window are zero
    int pad_;
};

struct IstftTrack
{
    float2 *out;          // chunk output, or the track accumulator when accumulate != 0
    long out_stem_stride; // float2 elements between stems
    long out_offset;      // segment offset inside the accumulator
    int n_wave;           // samples to emit (chunk length)
    int accumulate;       // 1: out[off+k] += w[k]*y[k] (umx.cpp:234-249), 0: plain store
    // Chunk samples [ex_lo, ex_hi) are covered by no other segment of the track: out/sum_weight there is (w y)/w with nothing
    // else added (umx.cpp:234-273), so wiener_istft_kernel writes it straight to the track's output (sample k of the chunk is
    // output sample final_off + k when that lies in [0, final_n)) and neither the accumulator nor finalize_kernel sees it.
    float2 *final_out;    // [n_stems][final_n], or null
    long final_n;
    long final_off;
    int ex_lo, ex_hi;
};

struct FinalizeTrack
{
    const float2 *acc; // [n_stems][L] weighted overlap-add accumulator of the shifted track
    float2 *out;       // [n_stems][n]
    long L, n;
    long i_begin, i_end; // output sample range to finalise
    // (host side only) the range that is complete — finalised here or written by the inverse STFT itself — once this entry's
    // kernels have run: what leaves for the host
    long ship_begin, ship_end;
};

constexpr int kIstftFramesPerCta = 32; // + 3 halo frames recomputed per CTA

int init_fft_tables(int device);
int launch_zero_f32(float *p, int n, cudaStream_t st); // p[0..n) = 0 by a kernel (stays on the compute engines)
int launch_zero_strided(float *p, long stride, long off, int count, int n_tracks, cudaStream_t st); // p[trk*stride + off + i] = 0
// skip_zero: frames beyond the first all-zero one of a short chunk are not computed (nothing downstream reads them)
// maxabs: [track][stat_windows] max |X| per Wiener window (window = frame / win_len; win_len = 0: one window, slot 0)
// reflect_pad: 0 = the reference's symmetric edge copy (src/dsp.cpp:109-128), 1 = true reflect padding (upstream torch.stft)
int launch_stft(const StftTrack *d_tracks, int n_tracks, int n_buf, float2 *spec, __half *xh, __half *xl,
                float *maxabs, cudaStream_t st, int skip_zero = 0, int reflect_pad = 0, int win_len = 0, int stat_windows = 1);
int launch_istft(const float2 *stems, long stem_stride, int n_stems, const IstftTrack *d_tracks, int n_tracks,
                 int n_buf, cudaStream_t st, int stem_begin = 0);
int launch_finalize(const FinalizeTrack *d_tracks, int n_tracks, long max_range, int shift, int n_buf, int stride, int n_stems,
                    cudaStream_t st, int stem_begin = 0);
int launch_to_ref_layout_c64(const float2 *in, float2 *out, int F, cudaStream_t st);
int launch_from_ref_layout_c64(const float2 *in, float2 *out, int F, cudaStream_t st);
int launch_to_ref_layout_f32(const float *in, float *out, int F, cudaStream_t st);
int launch_from_ref_layout_f32(const float *in, float *out, int F, cudaStream_t st);

// ---- gemm ----------------------------------------------------------------------------------------
// C[m, n] = act(alpha[n] * sum_k A[m,k] * B[n,k] + rowvec[m] * gamma[n] + beta[n]);  A, B are fp16 (hi, lo) plane pairs, K-major.
enum GemmAct
{
    kActNone = 0,
    kActTanh = 1,
    kActRelu = 2
};

struct GemmOperand
{
    const __half *hi;
    const __half *lo;
    int ld; // elements between rows
    int pad_;
};

struct GemmProblem
{
    GemmOperand A; // (M, K)
    GemmOperand B; // (N, K)
    int M, N, K;
    int act;
    const float *alpha; // per output column (may be null = 1)
    const float *beta;  // per output column (may be null = 0)
    // optional rank-1 term rowvec[m] * gamma[n] (exact-integer weight operands: W = scale*q + offset gives
    // x.W = scale * (x.q) + offset * sum_k x_k; alpha carries scale, gamma the offset, rowvec the row sums)
    const float *rowvec;
    const float *gamma;
    // output: column n goes to group g = n / n_group at  base + g*group_stride + m*ldo + n % n_group
    float *out_f32;  // fp32 output (or null)
    __half *out_hi;  // fp16 plane pair output (or null)
    __half *out_lo;
    int ldo;
    int n_group;
    long group_stride;
    // optional: rows are (track, frame) with rows_per_track frames per track; only the frames reaching the track's chunk are
    // needed (frames_reaching), row tiles without such a row are skipped by the tensor-core kernel
    const IstftTrack *chunks;
    int rows_per_track;
    int extra_rows; // 1: also the first all-zero frame (its row stands for every later one, see LstmLayerArgs::chunks)
};

constexpr int kMaxGemmProblems = 4;
struct GemmBatch
{
    GemmProblem p[kMaxGemmProblems];
    int count;
};

int launch_gemm_simt(const GemmBatch &batch, cudaStream_t st);
// terms: 1 = A_hi*B_hi ; 2 = + A_lo*B_hi ; 3 = + A_hi*B_lo
// tile_counter (optional, device int[2], zero before the first launch; every launch leaves it zero): tiles are drawn from it at
// run time instead of the static blockIdx stride, so a launch that shares the GPU with other kernels finishes as soon as the
// CTAs that did get an SM have worked through all tiles.  Launches that may overlap need counters of their own.
// pair: CTA pairs (tcgen05.mma.cta_group::2) on 256 x 256 tiles, static tile stride
int launch_gemm_tc(const GemmBatch &batch, int terms, cudaStream_t st, int *tile_counter = nullptr, bool pair = false);

// ---- lstm ----------------------------------------------------------------------------------------
struct LstmLayerArgs
{
    // per target: W_hh as [k][unit][gate] uint8 + its (scale, offset)
    const uint8_t *whh_q[kTargets][2];
    float whh_scale[kTargets][2];
    float whh_offset[kTargets][2];
    const float *xproj;    // [target][M][2*4*Hh] : W_ih x + b_ih + b_hh, column dir*4Hh + 4*unit + gate (gates i,f,g,o)
    long xproj_target_stride;
    float *h_state;        // [track][target][dir][Hh] for this layer
    float *c_state;
    long state_track_stride; // floats between tracks
    __half *out_hi;        // [target][M][ldo] + col_offset + dir*Hh + unit
    __half *out_lo;
    long out_target_stride;
    int ldo;
    int col_offset;
    float *out_f32;        // optional [target][M][2*Hh]
    long out_f32_target_stride;
    int F;                 // frames per track
    int Hh;                // hidden units per direction
    int n_tracks;
    int target_begin, target_count;
    // optional (first layer): frames >= z = frames_reaching(chunk) have an all-zero spectrum, hence identical fc1 outputs and
    // identical input projections; only row z of xproj is computed and every later step reads it
    const IstftTrack *chunks;
};
int launch_lstm_simple(const LstmLayerArgs &a, cudaStream_t st);
// lstm_cluster.cu: W_hh resident in tensor memory across a cluster of Hh / 64 CTAs, tcgen05 kind::i8 (Hh == 512 or 256)
bool lstm_cluster_supported(int Hh);
int lstm_cluster_max_active();
size_t whh_image_bytes(int Hh);
int launch_whh_image(const uint8_t *q, uint8_t *img, int Hh, cudaStream_t st);
int launch_lstm_cluster(const LstmLayerArgs &a, const uint8_t *const (*img)[2], cudaStream_t st);

// ---- wiener --------------------------------------------------------------------------------------
struct WienerArgs
{
    const float2 *spec;   // [M][2][2049]
    const float *mask;    // [4][M][ld] (ld = kMaskLd), or target magnitudes when mask_is_mag
    long mask_target_stride;
    int mask_ld;
    int mask_is_mag;      // 1: inputs are |Y_j| (kernel-level API), 0: inputs are the masks
    const float *maxabs;  // [track][stat_windows] max |X| per window (slot 0 when win_len == 0)
    float *partials;      // [track][chunks][20][2049]
    float *stats;         // [track][stat_windows][16][2049]
    int stat_windows;     // window slots per track in maxabs / stats: >= number of windows (1 is enough when win_len == 0)
    float2 *out;          // [4][M][2][2049]; null: statistics only (the filter pass is fused into the inverse STFT)
    long out_stem_stride;
    float2 *mix_out;      // optional: mix / max_abs (the reference mutates its argument)
    int F;
    int n_tracks;
    int stem_begin, stem_count; // stems whose output is written (all 4 unless target-parallel)
    const IstftTrack *chunks; // optional, per track: only the frames that reach its chunk are reduced and filtered
    // upstream-faithful variants (all 0 = the reference's behaviour, src/wiener.cpp):
    int psd_abs2;  // 1: PSD v = mean_ch |y|^2 (upstream filtering.py) instead of ((re+im)_L^2 + (re+im)_R^2)/2 (wiener.cpp:187-202)
    int reg_once;  // 1: sqrt(eps) I added to Cxx once (upstream) instead of once per source (wiener.cpp:301-325)
    int win_len;   // > 0: statistics and max|X| per window of win_len frames (upstream wiener_win_len = 300), <= kWienerChunks
                   //      windows; 0: one window = the whole segment (the reference)
    int no_em;     // 1: no EM iteration (upstream niter = 0): y_j = mask_j X
    // second and later EM iterations (upstream niter > 1; the reference's loop at src/wiener.cpp:175 with WIENER_ITERATIONS > 1):
    // the source estimates are the previous iteration's outputs, still divided by max_abs, instead of mask / magnitude times mix
    const float2 *y_in;   // [4][M][2][2049] or null
    long y_in_stem_stride;
    int out_normalized;   // 1: leave the outputs divided by max_abs (they feed another iteration)
};
constexpr int kWienerChunks = 19;
int launch_wiener(const WienerArgs &a, cudaStream_t st);
// stft.cu: filter pass + inverse STFT + overlap-add in one kernel; `w` as for launch_wiener after its statistics are in w.stats
// frames [frame_begin, frame_end) only (frame_end < 0: all): the launch emits the samples of those frames' hops, so a segment's
// post stage can be cut into parts whose finished output ranges leave for the host while the next part computes
int launch_wiener_istft(const WienerArgs &w, const IstftTrack *d_tracks, int n_buf, cudaStream_t st, int frame_begin = 0,
                        int frame_end = -1);

// ---- model prep ----------------------------------------------------------------------------------
int launch_dequant_u8(const uint8_t *q, float scale, float offset, float *out, long n, cudaStream_t st);
int launch_dequant_u16(const uint16_t *q, float scale, float offset, float *out, long n, cudaStream_t st);
// planes[n][k] = split(W[n][k] * (kscale ? kscale[k % kscale_mod] : 1)), zero padded to (n_pad, ld)
int launch_make_planes(const float *w, int N, int K, const float *kscale, int kscale_mod, __half *hi, __half *lo,
                       int n_pad, int ld, cudaStream_t st);
// out[n] = sum_k v[k % vmod] * W[n][k]
int launch_row_dot(const float *w, int N, int K, const float *v, int vmod, float *out, cudaStream_t st);
// whh_q [4*Hh][Hh] (PyTorch rows = gates i,f,g,o) -> [k][unit][gate]
int launch_pack_whh(const uint8_t *q, int Hh, uint8_t *out, cudaStream_t st);
// fp32 (F, H) column-major <-> fp16 plane pair row-major with pitch ld
int launch_colmajor_to_planes(const float *in, int F, int H, __half *hi, __half *lo, int ld, cudaStream_t st);
int launch_rowmajor_to_colmajor(const float *in, int F, int H, int ld, float *out, cudaStream_t st);
// mags[c][f][b] = mask[f][c*2049+b] * |spec[f][c][b]|  (src/inference.cpp:173-185), internal layouts
// rowsum[r] = sum_k (hi + lo)[r][k] over K columns of `rows` rows with pitch ld (one warp per row)
int launch_rowsum(const __half *hi, const __half *lo, long rows, int K, int ld, float *out, cudaStream_t st);
// fp16 plane of the centred quantised integers q - 127 (exact for uint8)
int launch_q_to_half(const uint8_t *q, int Hh, int K, __half *out, cudaStream_t st);
int launch_spec_to_x(const float2 *spec, long M, __half *xh, __half *xl, cudaStream_t st);
int launch_mask_to_mag(const float *mask, int ld, const float2 *spec, float *mag, long M, cudaStream_t st);

} // namespace umx
