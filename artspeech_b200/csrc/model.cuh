// Internal layout of asb_model and the launchers the forward pass is assembled from.
#pragma once

#include "asb_common.cuh"

namespace asb {

constexpr int kH = 128;        // hidden size the kernels are specialised for (models.py:104)
constexpr int kFc = 256;       // predictor width (models.py:12-16)
constexpr int kHeadPad = 128;  // x/y heads stacked (2 * n_samples = 100 rows) and padded to 128

// Every pointer lives in one device arena owned by the model.
struct PackedWeights {
  // embedding fused with the layer-0 input projection: tab0[dir][v][G*H] = E[v] . W_ih_l0^T + b_ih
  // (+ b_hh for the gates where it can be folded: GRU r,z; all LSTM gates)
  float* tab0;
  // recurrent weights [layer][dir]
  uint32_t* whh_frag[2][2];  // bf16, mma.m16n8k16 A-fragment order (fast mode)
  float* whh_t[2][2];        // fp32 [K = H][G*H] transposed (fp32 mode)
  float* bhh_n[2][2];        // GRU: b_hh of the n gate (cannot be folded); LSTM: zeros
  // Dense weights are stored row-major [N][K] like nn.Linear.weight; in fast mode they are
  // pre-rounded to tf32 (cvt.rna), in fp32 mode they are the raw fp32 values (split on the fly).
  // layer-1 input projection, both directions stacked: [2*G*H][2H], bias with foldable b_hh added
  float* wih1; float* bih1;
  // linear.0 [H][2H]
  float* lin_w; float* lin_b;
  // predictors, stacked over articulators
  float* ln0_w; float* ln0_b;     // [N][H]
  float* fc1_w; float* fc1_b;     // [N][256][H], [N][256]
  float* ln1_w; float* ln1_b;     // [N][256]
  float* fc2_w; float* fc2_b;     // [N][256][256], [N][256]
  float* ln2_w; float* ln2_b;     // [N][256]
  float* head_w; float* head_b;   // [N][128][256] (x rows 0..D-1, y rows D..2D-1, rest zero), [N][128]
  // pre-swizzled tf32 weight image streamed by the tcgen05 decoder (decoder_umma.cu); fast mode only
  float* umma_image;
  // the fp16-operand variant of that image (decoder_umma16.cu): half the bytes, kind::f16 at twice the tf32 rate
  float* umma16_image;
  // pre-swizzled tf32 image of wih1 for the tcgen05 layer-1 projection (proj_umma.cu); fast mode only
  float* proj_image;
  // fp16-operand image of W_ih_l1 (proj_umma16.cu); both precisions (fp32: hi + lo pieces)
  float* proj16_image;
  // bf16 W_hh images (K-major, 128-byte swizzle) kept in shared memory by the tcgen05 recurrent
  // kernel (gru_umma.cu); fast mode only
  void* whh_umma[2][2];
};

}  // namespace asb

#include <array>
#include <vector>

struct asb_model {
  int V, N, E, H, D, G, precision;
  void* arena;
  size_t arena_bytes;
  asb::PackedWeights w;
  // optional per-stage timing (asb_model_profile_*): one event set per profiled forward
  int decoder_impl = 0;  // 0 = mma.sync (decoder.cu), 1 = tcgen05/TMEM tf32 (decoder_umma.cu, proj_umma.cu), 2 = tcgen05 fp16 operands (decoder_umma16.cu)
  int rnn_impl = 0;      // 0 = register-resident mma.sync / fp32 (gru.cu), 1 = tcgen05, W_hh in tensor memory (gru_umma.cu; fast-mode default)
  // out-of-range token ids (nn.Embedding raises IndexError, models.py:135): the kernels that read tokens
  // store {1, offending id} here and clamp; the word lives in mapped pinned host memory so that the host
  // reads it without a device synchronisation (asb_model_token_status)
  int* tok_flag_host = nullptr;
  int* tok_flag_dev = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;   // asb_eval_step: fork / join of the target-side branch
  bool profile = false;
  std::vector<std::array<cudaEvent_t, ASB_N_STAGES + 1>> pending;
  std::vector<std::array<cudaEvent_t, ASB_N_STAGES + 1>> free_sets;
  double stage_ms[ASB_N_STAGES] = {0, 0, 0, 0, 0};
  int stage_calls[ASB_N_STAGES] = {0, 0, 0, 0, 0};
};

namespace asb {

// ---- recurrent layer (gru.cu) -------------------------------------------------------
// gi: (B*T, 2, G*H) input pre-activations; out: (B, T, 2H), pre-zeroed; lengths: device (B).
int launch_recurrent_layer(const asb_model* m, int layer, const float* gi, const int* lengths, float* out, int B,
                           int T, cudaStream_t st);
// gi[(b*T+t), dir, :] = tab0[dir][x[b, t]]   (embedding + layer-0 projection as a gather)
int launch_gather_gi0(const asb_model* m, const long long* x, int T_in, float* gi, int B, int T, cudaStream_t st);
int pack_recurrent_weights(asb_model* m, const asb_weights* w, cudaStream_t st);

// ---- dense layers (decoder.cu) ------------------------------------------------------
// C[M][N] = A[M][K] . W[N][K]^T + bias[N]   (tf32 single pass or 3xTF32)
int launch_linear(const asb_model* m, const float* A, const float* W, const float* bias, float* C, int M, int N,
                  int K, cudaStream_t st);
// rnn_out (M, 2H) -> contours (M, N_art, 2, D): linear.0 + ReLU, then every predictor, sigmoid
int launch_decoder(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st);
int pack_dense_weights(asb_model* m, const asb_weights* w, cudaStream_t st);

// ---- tcgen05 decoder (decoder_umma.cu) ----------------------------------------------
size_t umma_image_floats(int n_art);
bool umma_decoder_supported(const asb_model* m);
int pack_umma_image(asb_model* m, const asb_weights* w, cudaStream_t st);
int launch_decoder_umma(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st);

// ---- tcgen05 decoder, fp16 operands (decoder_umma16.cu) -------------------------------
size_t umma16_image_floats(int n_art, bool split);
bool umma16_decoder_supported(const asb_model* m);
int pack_umma16_image(asb_model* m, const asb_weights* w, cudaStream_t st);
int launch_decoder_umma16(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st);

// ---- tcgen05 layer-1 input projection (proj_umma.cu) --------------------------------
size_t proj_image_floats(const asb_model* m);
bool proj_umma_supported(const asb_model* m);
int pack_proj_image(asb_model* m, cudaStream_t st);   // after pack_dense_weights (reads w.wih1)
int launch_proj_umma(const asb_model* m, const float* A, float* C, int M, cudaStream_t st);
// ---- the same with fp16 operands (split operands in fp32 precision), proj_umma16.cu ------
size_t proj16_image_floats(const asb_model* m);
bool proj_umma16_supported(const asb_model* m);
int pack_proj16_image(asb_model* m, const asb_weights* src, cudaStream_t st);
int launch_proj_umma16(const asb_model* m, const float* A, float* C, int M, cudaStream_t st);

// ---- tcgen05 recurrent layer (gru_umma.cu) -------------------------------------------
size_t whh_umma_bytes(const asb_model* m);
bool rnn_umma_supported(const asb_model* m);
int pack_whh_umma(asb_model* m, const asb_weights* w, cudaStream_t st);
// tokens != null (layer 0 only): gi is not read, the kernel looks tab0 up itself
int launch_rnn_umma(const asb_model* m, int layer, const float* gi, const long long* tokens, int T_in,
                    const int* lengths, float* out, int B, int T, cudaStream_t st);
bool rnn_umma_reads_tokens(const asb_model* m, int B, int T);

}  // namespace asb
