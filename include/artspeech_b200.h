/*
 * artspeech_b200 - C ABI of the B200-native phoneme->articulation hot path.
 *
 * The reference (vribeiro1/artspeech) is pure Python and has no FFI layer; the
 * boundary it exposes for this path is its Python API.  Each entry point below
 * names the reference function it replaces (paths relative to the reference
 * checkout).  The Python package `artspeech_b200` binds these with ctypes and
 * re-exports the reference's names and signatures.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *  - the caller owns all buffers; the library allocates only inside asb_model_create
 *    (re-packed weight arena) and frees it in asb_model_destroy;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *    all work is enqueued on it, nothing synchronises unless stated;
 *  - return value: 0 = ok, otherwise an asb_status (>0) ; asb_last_error_string()
 *    gives a thread-local description.  CUDA launch errors are reported as
 *    ASB_ERR_CUDA with the cudaGetErrorString text.
 */
#ifndef ARTSPEECH_B200_H
#define ARTSPEECH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASB_ABI_VERSION 1
#if defined(__GNUC__)
#define ASB_API __attribute__((visibility("default")))
#else
#define ASB_API
#endif

typedef enum asb_status {
  ASB_OK = 0,
  ASB_ERR_INVALID = 1,   /* bad argument (shape, null pointer, alignment)      -> Python ValueError   */
  ASB_ERR_UNSORTED = 2,  /* lengths not sorted descending / out of range        -> Python RuntimeError
                            (what pack_padded_sequence raises, models.py:136)                          */
  ASB_ERR_CUDA = 3,      /* CUDA runtime error                                  -> Python RuntimeError */
  ASB_ERR_WORKSPACE = 4, /* workspace too small                                 -> Python ValueError   */
  ASB_ERR_UNSUPPORTED = 5
} asb_status;

ASB_API int asb_abi_version(void);
ASB_API const char* asb_last_error_string(void);
/* Hash of the sources, headers and compiler flags this library was built from (artspeech_b200/build.py
 * source_hash()); the Python binding refuses a library whose hash differs from the sources beside it. */
ASB_API const char* asb_build_hash(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
ASB_API int64_t asb_launch_count(void);

/* ------------------------------------------------------------------ geometry */

/* Point layouts of a contour of n points */
#define ASB_LAYOUT_PLANAR 0      /* (2, n): x row then y row - the model's output layout (models.py:31)   */
#define ASB_LAYOUT_INTERLEAVED 1 /* (n, 2): what MeanP2CPDistance.forward receives (metrics.py:35-36)     */

/* MeanP2CPDistance.forward with reduction="none", phoneme_to_articulation/metrics.py:32-46;
 * with ASB_LAYOUT_PLANAR it is metrics.py:38-52 `p2cp_distance` without the transposes.
 * u: items contours of n points, v: items contours of m points, out: (items) fp32.
 * out[k] = ( sum_i min_j |u_i-v_j| / n + sum_j min_i |u_i-v_j| / m ) / 2, distances evaluated
 * in fp32 with the direct (dx^2+dy^2) form, sqrt after the minimum. */
ASB_API int asb_p2cp(const float* u, const float* v, float* out, int64_t items, int n, int m,
             int layout, void* stream);

/* EuclideanDistance(reduction="none")(...).mean(-1), metrics.py:54-68 over
 * phoneme_to_articulation/metrics.py:10-24.  Planar (2,n) contours -> out (items). */
ASB_API int asb_mean_euclidean(const float* u, const float* v, float* out, int64_t items, int n,
                       void* stream);

/* EuclideanDistance.forward with reduction "none", phoneme_to_articulation/metrics.py:5-24: planar (2,n) contours
 * -> out (items, n) = per-point distance; and its gradient wrt u (the training criterion,
 * train_phoneme_to_articulation.py:86): grad_u (items, 2, n) = grad_out * (u - v) / distance (0 where the points
 * coincide). */
ASB_API int asb_point_distance(const float* u, const float* v, float* out, int64_t items, int n, void* stream);
ASB_API int asb_point_distance_backward(const float* u, const float* v, const float* grad_out, float* grad_u,
                                int64_t items, int n, void* stream);

/* pearsons_correlation, root metrics.py:9-35 (time axis = dim 1; the x targets are centred with the x OUTPUT
 * mean, :22): outputs / targets (B, T, n_art, 2, n_points) fp32 -> x_corr, y_corr (B, n_art, n_points). */
ASB_API int asb_pearson(const float* outputs, const float* targets, int B, int T, int n_art, int n_points,
                float* x_corr, float* y_corr, void* stream);

/* Closest pair of two planar contours per item: u (items, 2, n), v (items, 2, m) fp32 ->
 * out_d (items) = min_{i,j} |u_i - v_j|, out_ij (items, 2) int32 = the pair attaining it (smallest
 * i*m + j among ties; may be NULL).  The reduction inside CriticalLoss.forward,
 * phoneme_to_articulation/principal_components/losses.py:86-96 (cdist -> reshape -> min). */
ASB_API int asb_min_distance(const float* u, const float* v, int64_t items, int n, int m, float* out_d,
                     int32_t* out_ij, void* stream);

/* calculate_vocal_tract_variables, tract_variables.py:73-125 (LA, TTCD, TBCD, VEL via
 * _calculate_TV :23-35) for `sets` contour sets of shape (n_art, 2, 50), planar.
 * art_index_host[6]: rows of lower-lip, pharynx, soft-palate-midline, tongue, upper-incisor,
 * upper-lip inside a set (host array).
 * val (sets,4) fp32 distance; idx (sets,4,2) int32 = (i into arr1, j into arr2) with the
 * reference's tie-break (smallest j, then smallest i); poc (sets,4,2,2) fp32 = arr1[i], arr2[j].
 * TV order: LA, TTCD, TBCD, VEL. */
ASB_API int asb_tract_variables(const float* sets_ptr, int64_t sets, int n_art,
                        const int32_t* art_index_host, float* val, int32_t* idx, float* poc,
                        void* stream);

/* Same, with the upper-incisor contour taken from outside the sets: evaluation.py:93-109 injects a
 * fixed reference contour that the model does not predict; passing it separately avoids building
 * the concatenated (sets, n_art+1, 2, 50) tensor.  upper_incisor: (2,50) planar contour(s);
 * set k reads upper_incisor + k * upper_incisor_stride floats (stride 0 = one shared contour).
 * art_index_host[4] (upper-incisor) is ignored. */
ASB_API int asb_tract_variables_ref(const float* sets_ptr, int64_t sets, int n_art,
                            const int32_t* art_index_host, const float* upper_incisor,
                            int64_t upper_incisor_stride, float* val, int32_t* idx, float* poc,
                            void* stream);

/* Same with a cap on the grid (`max_ctas` CTAs of 8 warps; 0 = one full wave) and an optional
 * upper incisor (NULL = row art_index_host[4] of each set).  A call that runs BESIDE other kernels
 * (the target's tract variables next to the forward pass, evaluation.py:57 vs __init__.py:261) leaves
 * SMs free this way: a full wave of these CTAs takes every register of every SM. */
ASB_API int asb_tract_variables_ex(const float* sets_ptr, int64_t sets, int n_art, const int32_t* art_index_host,
                           const float* upper_incisor, int64_t upper_incisor_stride, int max_ctas,
                           float* val, int32_t* idx, float* poc, void* stream);

/* intersect_semipolar_grid -> area_function -> evenly_spaced_fx, area_function.py:175-223,
 * :124-142 (mid_point :113-121), :145-159, for `frames` independent frames.
 * walls_int / walls_ext: (frames, n_wall, 2) fp32 interleaved (x, y) wall polylines, 2 <= n_wall <= 64;
 * grid_ends: (n_lines, 2, 2) fp64 DEVICE array, [line][internal|external end][x|y] = first and last
 * sample of every line of build_semipolar_grid (:31-110), 1 <= n_lines <= 128.
 * xfx (frames, 2, n_samples) fp32 = [linspace(0, total mid-line length, n_samples); alpha * r^beta
 * interpolated there] (NaN when fewer than two grid lines touch a wall); K (frames) int32 = number
 * of grid lines that touch a wall.  Optional (NULL to skip): pts (frames, n_lines, 2, 2) fp64 =
 * the K selected [internal|external] intersection points in grid order, zero padded (what
 * intersect_semipolar_grid returns); lines (frames, n_lines) int32 = their grid-line indices, -1 padded.
 * Decisions (which lines touch, which intersection is kept) are those of fp64 arithmetic. */
ASB_API int asb_area_function(const float* walls_int, const float* walls_ext, int64_t frames, int n_wall,
                      const double* grid_ends, int n_lines, double alpha, double beta, int n_samples,
                      float* xfx, int32_t* K, double* pts, int32_t* lines, void* stream);

/* area_function alone, area_function.py:124-142: already paired wall points int_pts / ext_pts
 * (frames, K, 2) fp64 -> dists (frames, K) = cumulative distance between successive mid-points
 * (mid_point :113-121), fx (frames, K) = alpha * (|int - ext| / 2)^beta, both fp64. */
ASB_API int asb_area_from_pairs(const double* int_pts, const double* ext_pts, int64_t frames, int K,
                        double alpha, double beta, double* dists, double* fx, void* stream);

/* evenly_spaced_fx alone, area_function.py:145-159: polylines (x, fx) of K points each (x
 * non-decreasing), fp64 -> out (frames, 2, n_samples) fp32 = [np.linspace(x[0], x[K-1], n_samples);
 * the piecewise-linear fx there]. */
ASB_API int asb_evenly_spaced_fx(const double* x, const double* fx, int64_t frames, int K, int n_samples,
                         float* out, void* stream);

/* Length-masked aggregation of one batch's per-frame metrics, as run_test does per sentence
 * (evaluation.py:68-84,142-156): p2cp (B,T,n_art), tv_pred/tv_target (B,T,4) fp32, lengths_dev (B)
 * DEVICE int32.  scratch: B*(n_art+12) doubles.  acc (n_art+14 doubles) is OVERWRITTEN with
 * [0:n) sum over utterances of per-utterance mean P2CP; [n] utterances; [n+1:n+5) sum of predicted
 * TVs over valid frames; [n+5:n+9) target TVs; [n+9:n+13) |pred-target|; [n+13] valid frames.
 * Fixed summation order (bit-reproducible). */
ASB_API int asb_eval_accumulate(const float* p2cp, const float* tv_pred, const float* tv_target,
                        const int32_t* lengths_dev, int B, int T, int n_art, double* scratch,
                        double* acc, void* stream);

/* Same, and the batch's vector is also ADDED into `total` (n_art+14 doubles, may be NULL; the caller zeroes it
 * before the first batch): run_test aggregates over all batches ONCE at the end (evaluation.py:142-159), so a
 * multi-GPU evaluation keeps this running sum on the device and all-reduces it once per evaluation. */
ASB_API int asb_eval_accumulate_total(const float* p2cp, const float* tv_pred, const float* tv_target,
                              const int32_t* lengths_dev, int B, int T, int n_art, double* scratch,
                              double* acc, double* total, void* stream);

/* asb_eval_accumulate WITHOUT caller scratch and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  The per-utterance partials live in a file-scope device array
 * of ASB_EVAL_STATIC_MAX_DOUBLES doubles (256 KiB, held by every GPU that loads the library until exit); the
 * call returns ASB_ERR_WORKSPACE when B * (n_art + 12) exceeds it (e.g. B <= 1489 at n_art = 10).  Same inputs,
 * same `acc` (n_art+14 doubles, overwritten) and the same bits as asb_eval_accumulate.  It calls no cudaMalloc /
 * cudaFree / allocating routine; the array's device address is looked up once per device (cudaGetSymbolAddress)
 * on the first call there.  All calls on one GPU SHARE the array: calls that overlap on the GPU (different
 * streams without an ordering between them) race; stream-ordered calls are safe, nothing needs clearing. */
#define ASB_EVAL_STATIC_MAX_DOUBLES 32768
ASB_API int asb_eval_accumulate_static(const float* p2cp, const float* tv_pred, const float* tv_target,
                               const int32_t* lengths_dev, int B, int T, int n_art, double* acc, void* stream);

/* The other per-sentence numbers of run_test (evaluation.py:68-84): outputs/targets
 * (B,T,n_art,2,n_points) fp32 planar, lengths_dev (B) DEVICE int32.  Per sentence and articulator:
 * the mean over valid frames of euclidean_distance (metrics.py:54-68) and, over the valid frames as
 * the time axis, pearsons_correlation (metrics.py:9-35, with its quirk: x targets centred by the x
 * OUTPUT mean, :22) averaged over the points.  scratch: B*n_art*3 doubles.  acc (3*n_art doubles) is
 * OVERWRITTEN with the sums over the B sentences of [0:n) MED, [n:2n) x_corr, [2n:3n) y_corr (divide
 * by the sentence count for run_test's means, evaluation.py:150-156).  fp64, fixed summation order. */
ASB_API int asb_eval_sentence_metrics(const float* outputs, const float* targets, const int32_t* lengths_dev,
                              int B, int T, int n_art, int n_points, double* scratch, double* acc,
                              void* stream);

/* ------------------------------------------------------------------ model */

typedef struct asb_model asb_model;

#define ASB_CELL_GRU 3  /* nn.GRU, gate rows r, z, n (models.py:111) - both precisions                       */
#define ASB_CELL_LSTM 4 /* nn.LSTM, gate rows i, f, g, o (RNNType.LSTM, phoneme_to_articulation/__init__.py:47-49)
                           - both precisions (fast: tcgen05 kernel; fp32: lstm_fp32_kernel) */
#define ASB_PRECISION_FAST 0 /* fp16 GEMM operands on tcgen05, fp32 accumulate/state (2e-3 gate) */
#define ASB_PRECISION_FP32 1 /* split fp16 operands (hi + lo, three MMAs per product) on tcgen05; FFMA / 3xTF32 kernels
                                for the LSTM, n_samples != 50 and set_decoder / set_recurrent (1e-5 gate) */

/* Device pointers to fp32 parameters in the reference's state_dict layout (SURVEY.md 8b;
 * models.py:109-120).  Arrays of per-articulator pointers are HOST arrays of device pointers. */
typedef struct asb_weights {
  const float* embedding;          /* (V, E) */
  const float* rnn_w_ih[2][2];     /* [layer][dir] (G*H, E or 2H) */
  const float* rnn_w_hh[2][2];     /* (G*H, H) */
  const float* rnn_b_ih[2][2];     /* (G*H) */
  const float* rnn_b_hh[2][2];     /* (G*H) */
  const float* linear_w;           /* (H, 2H)  linear.0 */
  const float* linear_b;           /* (H) */
  const float* const* ln0_w; const float* const* ln0_b;   /* [N] (H)      predictors.i.linear.0 */
  const float* const* fc1_w; const float* const* fc1_b;   /* [N] (256,H)  predictors.i.linear.1 */
  const float* const* ln1_w; const float* const* ln1_b;   /* [N] (256)    predictors.i.linear.3 */
  const float* const* fc2_w; const float* const* fc2_b;   /* [N] (256,256) predictors.i.linear.4 */
  const float* const* ln2_w; const float* const* ln2_b;   /* [N] (256)    predictors.i.linear.6 */
  const float* const* xh_w;  const float* const* xh_b;    /* [N] (D,256)  predictors.i.x_coords */
  const float* const* yh_w;  const float* const* yh_b;    /* [N] (D,256)  predictors.i.y_coords */
} asb_weights;

/* Re-packs the weights into the kernels' layouts (synchronises `stream`).  Replaces
 * ArtSpeech.__init__ + load_state_dict + .to(device), models.py:99-120. */
ASB_API int asb_model_create(const asb_weights* w, int vocab, int n_art, int embed_dim, int hidden,
                     int n_samples, int cell, int precision, void* stream, asb_model** out);
ASB_API void asb_model_destroy(asb_model* m);

/* Bytes of scratch asb_model_forward needs for a (B, T) batch. */
ASB_API size_t asb_model_workspace_bytes(const asb_model* m, int B, int T);

/* ArtSpeech.forward(x, lengths), models.py:126-145.
 * x (B, T_in) int64 tokens (row stride T_in); lengths_host (B) sorted descending, each in
 * [1, T_in]; out (B, T, n_art, 2, n_samples) fp32 contiguous with T = lengths_host[0].
 * Rows past a sequence's length receive the constant padded frame, as the reference does. */
ASB_API int asb_model_forward(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host, int B,
                      float* out, void* workspace, size_t workspace_bytes, void* stream);

/* Same, for callers that already hold the lengths on the device (no host->device copy is enqueued,
 * which makes the call capturable in a CUDA graph).  lengths_host is still used for validation and
 * for T; lengths_dev (B) int32 must hold the same values. */
ASB_API int asb_model_forward_resident(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host,
                               const int32_t* lengths_dev, int B, float* out, void* workspace,
                               size_t workspace_bytes, void* stream);

/* The recurrent stack alone: embedding -> 2-layer bidirectional GRU/LSTM with packed-sequence
 * semantics (models.py:135-138; principal_components/models/rnn.py:94-104): same arguments as
 * asb_model_forward_resident, but `hidden` (B, T, 2*hidden) fp32 receives the padded recurrent output
 * (zeros at t >= length) and the decoder is not run.  lengths_dev may be NULL. */
ASB_API int asb_model_forward_hidden(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host,
                             const int32_t* lengths_dev, int B, float* hidden, void* workspace,
                             size_t workspace_bytes, void* stream);

/* Tail of PrincipalComponentsArtSpeech (phoneme_to_articulation/principal_components/models/rnn.py:
 * 11-33, 70-79, 106-108): Linear(2H,H)+ReLU -> LayerNorm(H) -> Linear(H,256)+ReLU -> LayerNorm(256) ->
 * Linear(256,128)+ReLU -> LayerNorm(128) -> Linear(128,n_components) -> tanh, fused in one kernel
 * (64-frame tiles, activations in shared memory).  All pointers are device fp32, row-major like
 * nn.Linear.weight. */
typedef struct asb_pc_head asb_pc_head;
typedef struct asb_pc_weights {
  const float* lin_w; const float* lin_b;   /* (H, 2H), (H)        linear.0 */
  const float* ln0_w; const float* ln0_b;   /* (H)                 predictor.linear.0 */
  const float* fc1_w; const float* fc1_b;   /* (256, H), (256)     predictor.linear.1 */
  const float* ln1_w; const float* ln1_b;   /* (256)               predictor.linear.3 */
  const float* fc2_w; const float* fc2_b;   /* (128, 256), (128)   predictor.linear.4 */
  const float* ln2_w; const float* ln2_b;   /* (128)               predictor.linear.6 */
  const float* fc3_w; const float* fc3_b;   /* (n_components, 128), (n_components)  predictor.linear.7 */
} asb_pc_weights;
ASB_API int asb_pc_head_create(const asb_pc_weights* w, int hidden, int n_components, int precision, void* stream,
                       asb_pc_head** out);
ASB_API void asb_pc_head_destroy(asb_pc_head* h);
/* features (rows, 2H) -> out (rows, n_components) */
ASB_API int asb_pc_head_forward(asb_pc_head* h, const float* features, int64_t rows, float* out, void* stream);

/* ArticulatorPredictor.forward on its own (phoneme_to_articulation/encoder_decoder/models.py:7-33): the
 * class is also used outside ArtSpeech - the transformer model builds its heads from it
 * (phoneme_to_articulation/transformer/models.py:6,337-339, in_features = embed_dim).
 * LayerNorm(C) -> Linear(C,256) -> ReLU -> LayerNorm(256) -> Linear(256,256) -> ReLU -> LayerNorm(256) ->
 * x_coords / y_coords stacked on dim 2; NO sigmoid.  in_features: multiple of 32 in [32, 256];
 * n_samples <= 64.  All weight pointers are device fp32, row-major like nn.Linear.weight. */
typedef struct asb_predictor asb_predictor;
typedef struct asb_predictor_weights {
  const float* ln0_w; const float* ln0_b;   /* (C)             linear.0 */
  const float* fc1_w; const float* fc1_b;   /* (256, C), (256) linear.1 */
  const float* ln1_w; const float* ln1_b;   /* (256)           linear.3 */
  const float* fc2_w; const float* fc2_b;   /* (256,256),(256) linear.4 */
  const float* ln2_w; const float* ln2_b;   /* (256)           linear.6 */
  const float* xh_w; const float* xh_b;     /* (D, 256), (D)   x_coords */
  const float* yh_w; const float* yh_b;     /* (D, 256), (D)   y_coords */
} asb_predictor_weights;
ASB_API int asb_predictor_create(const asb_predictor_weights* w, int in_features, int n_samples, int precision,
                         void* stream, asb_predictor** out);
ASB_API void asb_predictor_destroy(asb_predictor* p);
/* features (rows, in_features) -> out (rows, 2, n_samples) */
ASB_API int asb_predictor_forward(asb_predictor* p, const float* features, int64_t rows, float* out, void* stream);

/* Decoder alone on caller-supplied feature rows: features (rows, 2*hidden) fp32 ->
 * out (rows, n_art, 2, n_samples) = sigmoid(stack(predictor_i(ReLU(linear.0(features))))),
 * models.py:140-145.  SimpleArtSpeech (models.py:53-96) runs through this entry: its features are
 * embedding rows zero-padded to 2*hidden and its Linear(embed_dim, hidden) zero-padded likewise. */
ASB_API int asb_model_decode(asb_model* m, const float* features, int64_t rows, float* out, void* stream);

/* out[r, :] = table[tokens[r], :] for r < rows; table (vocab, row_floats) fp32, row_floats % 4 == 0,
 * table/out 16-byte aligned.  nn.Embedding (models.py:63,135) and, for SimpleArtSpeech, the whole
 * forward pass as a lookup in the per-token contour table (models.py:84-96).  A token outside
 * [0, vocab) yields a row of NaNs. */
ASB_API int asb_gather_rows(const float* table, const int64_t* tokens, int64_t rows, int vocab, int row_floats,
                    float* out, void* stream);

/* Same with a status word: token_flag (device-accessible int32[2], may be NULL) receives {1, id} when a token
 * id lies outside [0, vocab) - what nn.Embedding reports as IndexError (models.py:63,135). */
ASB_API int asb_gather_rows_flagged(const float* table, const int64_t* tokens, int64_t rows, int vocab,
                            int row_floats, float* out, int32_t* token_flag, void* stream);

/* Token-id range status of a model: nn.Embedding raises IndexError for ids outside [0, vocab)
 * (models.py:135).  The kernels that read tokens clamp such an id (memory safety) and store {1, id} in a
 * status word in mapped pinned host memory.  asb_model_token_status reads it WITHOUT synchronising (it
 * reflects the forwards that have completed): *bad_out = 1 if an id was out of range, *token_out (may be
 * NULL) one offending id; reset != 0 clears it.  asb_model_token_flag_ptr is the word's device address
 * (for asb_gather_rows_flagged). */
ASB_API int asb_model_token_status(asb_model* m, int32_t* bad_out, int64_t* token_out, int reset);
ASB_API int32_t* asb_model_token_flag_ptr(asb_model* m);

/* One evaluation step of run_test as one call (encoder_decoder/evaluation.py:53-128 with the frame driver
 * phoneme_to_articulation/__init__.py:201-297 batched): forward -> P2CP of every (frame, articulator) ->
 * tract variables of prediction and target -> length-masked accumulation.  All buffers are the caller's:
 *   outputs (B,T,N,2,50) fp32; p2cp (B,T,N) fp32; tv_* (B,T,4) fp32 with idx (B,T,4,2) int32 and poc (B,T,4,2,2)
 *   fp32 as asb_tract_variables; scratch B*(N+12) doubles; acc N+14 doubles (overwritten); total N+14 doubles or
 *   NULL (added to, asb_eval_accumulate_total); workspace >= asb_model_workspace_bytes(m, B, T) bytes,
 * with T = lengths_host[0].  tokens (B, T_in) int64; targets (B, T, N, 2, 50) fp32 CONTIGUOUS; lengths as for
 * asb_model_forward_resident (lengths_dev required); art_index_host / upper_incisor / upper_incisor_stride as for
 * asb_tract_variables_ex.  side_stream (may be NULL or == stream): the target's tract variables are forked onto
 * it and joined before the accumulation, so that they run beside the forward pass.  Nothing synchronises. */
typedef struct asb_eval_buffers {
  float* outputs;
  float* p2cp;
  float* tv_pred; int32_t* tv_pred_idx; float* tv_pred_poc;
  float* tv_target; int32_t* tv_target_idx; float* tv_target_poc;
  double* scratch; double* acc; double* total;
  void* workspace; size_t workspace_bytes;
} asb_eval_buffers;
ASB_API int asb_eval_step(asb_model* m, const int64_t* tokens, int T_in, const int32_t* lengths_host,
                  const int32_t* lengths_dev, int B, const float* targets, const int32_t* art_index_host,
                  const float* upper_incisor, int64_t upper_incisor_stride, const asb_eval_buffers* buffers,
                  void* stream, void* side_stream);

/* ------------------------------------------------------------------ training (row N2)
 * train_phoneme_to_articulation.py:84-96 runs loss.backward() through the model.  The fused inference kernels
 * save no activations; training runs an unfused fp32 path built from the primitives below (forward with saved
 * activations + backward), composed by artspeech_b200/training.py.  All pointers are device fp32 unless noted. */

/* C[i, j] (+)= sum_l A[i*a_i + l*a_l] * B[l*b_l + j*b_j] (+ bias[j]), i < M, j < N, l < K; C row stride ldc.
 * act: 0 none, 1 ReLU, 2 sigmoid (applied after bias / accumulation).  The three forms of nn.Linear:
 * y = x W^T + b (a_l = 1, b_l = 1, b_j = K), dx = dy W (b_l = K', b_j = 1), dW = dy^T x (a_i = 1, a_l = N'). */
ASB_API int asb_gemm(const float* A, const float* B, float* C, int64_t M, int N, int K, int64_t a_i, int64_t a_l,
             int64_t b_l, int64_t b_j, int64_t ldc, const float* bias, int act, int accumulate, void* stream);
/* out[j] (+)= sum_i X[i*ldx + j]: bias gradients */
ASB_API int asb_colsum(const float* X, int64_t M, int N, int64_t ldx, float* out, int accumulate, void* stream);
/* nn.LayerNorm(C) forward over rows (biased variance, eps 1e-5; models.py:11,14,17): saves xhat (M,C), rstd (M) */
ASB_API int asb_layernorm_fwd(const float* x, int64_t M, int C, const float* gamma, const float* beta, float* xhat,
                      float* rstd, float* y, void* stream);
/* its backward: dx (+)= rstd (g - mean g - xhat mean(g xhat)), g = dy gamma, zeroed where relu_src <= 0 (relu_src
 * may be NULL); dgamma / dbeta (may be NULL) (+)= column sums.  dx must not alias dy. */
ASB_API int asb_layernorm_bwd(const float* dy, const float* xhat, const float* rstd, const float* gamma,
                      const float* relu_src, int64_t M, int C, float* dx, int accumulate_dx, float* dgamma,
                      float* dbeta, int accumulate_params, void* stream);
ASB_API int asb_sigmoid_bwd(const float* dy, const float* y, float* dz, int64_t n, void* stream);   /* dz = dy y (1-y) */
ASB_API int asb_relu_mask(float* dx, const float* f, int64_t n, void* stream);                     /* dx = 0 where f <= 0 */
/* dst[tokens[r], :] += src[r, :]: gradient of nn.Embedding (models.py:110,135) */
ASB_API int asb_scatter_add_rows(const float* src, const int64_t* tokens, int64_t rows, int width, int vocab,
                         float* dst, void* stream);
/* inverted dropout (nn.GRU(dropout=p) between the layers, models.py:111; nn.Dropout, models.py:65): out = x * keep /
 * (1 - p) with keep drawn from a counter-based generator keyed by (seed, element index); calling it again on the
 * gradient with the same seed applies the same mask. */
ASB_API int asb_dropout(const float* x, float* out, int64_t n, float p, uint64_t seed, void* stream);
/* One bidirectional GRU layer with packed-sequence semantics (models.py:111,136-138), training forward:
 * gi (B*T, 2, 3H) = W_ih x + b_ih per direction; whh (3H, H), bhh (3H) as nn.GRU stores them; lengths_dev sorted
 * descending.  out (B, T, 2H) must be zeroed by the caller (padded steps stay zero); gates (B*T, 2, 4, H) receives
 * r, z, n and hn = W_hn h + b_hn of every valid step. */
ASB_API int asb_gru_train_forward(const float* whh_fwd, const float* whh_bwd, const float* bhh_fwd, const float* bhh_bwd,
                          const float* gi, const int32_t* lengths_dev, int B, int T, int hidden, float* out,
                          float* gates, void* stream);
/* BPTT through that layer: dout (B, T, 2H) gradient wrt its output -> dgi (B*T, 2, 3H) gradient wrt gi (= wrt b_ih
 * rows), dgh (B*T, 2, 3H) gradient wrt W_hh h + b_hh, hprev (B*T, 2, H) the h each step started from; all three
 * must be zeroed by the caller (padded steps stay zero).  Weight gradients follow as GEMMs over all steps:
 * dW_ih = dgi^T x, dW_hh = dgh^T hprev, db_ih = colsum dgi, db_hh = colsum dgh, dx = dgi W_ih. */
ASB_API int asb_gru_train_backward(const float* whh_fwd, const float* whh_bwd, const float* dout, const float* out,
                           const float* gates, const int32_t* lengths_dev, int B, int T, int hidden, float* dgi,
                           float* dgh, float* hprev, void* stream);

/* The same pair for nn.LSTM (RNNType.LSTM, phoneme_to_articulation/__init__.py:47-49): whh (4H, H), bhh (4H), gate
 * rows i, f, g, o; gi (B*T, 2, 4H); gates (B*T, 2, 5, H) receives i, f, g, o and the new cell state.  Every gate
 * pre-activation is gi + W_hh h + b_hh, so one gradient array serves both: dgi (B*T, 2, 4H), zeroed by the caller. */
ASB_API int asb_lstm_train_forward(const float* whh_fwd, const float* whh_bwd, const float* bhh_fwd, const float* bhh_bwd,
                           const float* gi, const int32_t* lengths_dev, int B, int T, int hidden, float* out,
                           float* gates, void* stream);
ASB_API int asb_lstm_train_backward(const float* whh_fwd, const float* whh_bwd, const float* dout, const float* out,
                            const float* gates, const int32_t* lengths_dev, int B, int T, int hidden, float* dgi,
                            float* hprev, void* stream);

/* Decoder implementation: 2 = tcgen05/TMEM kernel with fp16 operands (default with n_samples = 50; in fp32
 * precision the same kernel with split hi + lo operands), 1 = the tcgen05 kernel with tf32 operands (round 1; fast
 * precision only), 0 = mma.sync kernel (single-pass tf32, 3xTF32 in fp32 precision; any n_samples).  All compute
 * models.py:140-145.  The layer-1 input projection follows the same switch (proj_umma16.cu / proj_umma.cu /
 * mma.sync linear256). */
ASB_API int asb_model_set_decoder(asb_model* m, int impl);

/* Recurrent-layer implementation: 1 = W_hh resident in TENSOR MEMORY, tcgen05.mma kind::f16 (TS form) with the
 * gates fused behind tcgen05.ld (default; fast precision: fp16 operands, GRU and LSTM, 0.55 us per time step at
 * batch 256 on B200; fp32 precision: split hi + lo operands, GRU only, 1.07 us per step), 0 = fast precision:
 * W_hh resident in registers, mma.sync bf16 (GRU only, 0.84-0.91 us per step); fp32 precision: the FFMA kernels
 * (GRU 2.9 us per step; the LSTM's only implementation there).  All compute models.py:111,136-138. */
ASB_API int asb_model_set_recurrent(asb_model* m, int impl);

/* Per-stage device timing of asb_model_forward (CUDA events on `stream`): enable, run forwards,
 * then read the accumulated milliseconds and launch counts (synchronises the recorded events).
 * Stages: 0 gather (embedding + layer-0 projection), 1 GRU layer 0, 2 layer-1 projection GEMM,
 * 3 GRU layer 1, 4 fused decoder. */
#define ASB_N_STAGES 5
ASB_API int asb_model_profile_enable(asb_model* m, int enable);
ASB_API int asb_model_profile_read(asb_model* m, float* stage_ms_out, int32_t* stage_calls_out, int reset);

#ifdef __cplusplus
}
#endif
#endif /* ARTSPEECH_B200_H */
