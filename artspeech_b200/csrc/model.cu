// asb_model: weight re-packing and the forward pass of ArtSpeech
// (phoneme_to_articulation/encoder_decoder/models.py:99-145) assembled from the kernels in
// gru.cu and decoder.cu:
//
//   tokens --gather(tab0 = E.W_ih_l0^T + b)--> gi0 --GRU layer 0 (fwd | bwd CTAs)--> out0
//   out0 --[M,256]x[256,768] tf32 GEMM--> gi1 --GRU layer 1--> out1 (zeros at padded steps)
//   out1 --fused decoder (linear.0 + 10 predictors + sigmoid)--> contours (B, T, N, 2, D)
#include <new>
#include <vector>

#include "model.cuh"

namespace {

using namespace asb;

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct ArenaPlan {
  size_t total = 0;
  size_t take(size_t bytes) {
    const size_t off = total;
    total += align_up(bytes, 256);
    return off;
  }
};

struct Workspace {
  int* lengths;
  float* gi;
  float* out0;
  float* out1;
  size_t bytes;
};

Workspace plan_workspace(const asb_model* m, void* base, int B, int T) {
  ArenaPlan p;
  const size_t frames = (size_t)B * T;
  const size_t o_len = p.take(sizeof(int) * (size_t)B);
  const size_t o_gi = p.take(sizeof(float) * frames * 2 * m->G * kH);
  const size_t o_out0 = p.take(sizeof(float) * frames * 2 * kH);
  const size_t o_out1 = p.take(sizeof(float) * frames * 2 * kH);
  char* b = static_cast<char*>(base);
  Workspace w;
  w.lengths = reinterpret_cast<int*>(b + o_len);
  w.gi = reinterpret_cast<float*>(b + o_gi);
  w.out0 = reinterpret_cast<float*>(b + o_out0);
  w.out1 = reinterpret_cast<float*>(b + o_out1);
  w.bytes = p.total;
  return w;
}

}  // namespace

extern "C" int asb_model_create(const asb_weights* src, int vocab, int n_art, int embed_dim, int hidden,
                                int n_samples, int cell, int precision, void* stream, asb_model** out) {
  if (!src || !out) return set_error(ASB_ERR_INVALID, "asb_model_create: null argument");
  *out = nullptr;
  if (hidden != kH) return set_error(ASB_ERR_UNSUPPORTED, "asb_model_create: hidden_size must be %d (got %d)", kH, hidden);
  if (cell != ASB_CELL_GRU && cell != ASB_CELL_LSTM)
    return set_error(ASB_ERR_INVALID, "asb_model_create: cell must be ASB_CELL_GRU or ASB_CELL_LSTM");
  if (vocab <= 0 || n_art <= 0 || embed_dim <= 0 || n_samples <= 0 || 2 * n_samples > kHeadPad)
    return set_error(ASB_ERR_INVALID, "asb_model_create: bad sizes (vocab %d, n_art %d, embed %d, n_samples %d <= %d)",
                     vocab, n_art, embed_dim, n_samples, kHeadPad / 2);
  if (precision != ASB_PRECISION_FAST && precision != ASB_PRECISION_FP32)
    return set_error(ASB_ERR_INVALID, "asb_model_create: bad precision %d", precision);
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  asb_model* m = new (std::nothrow) asb_model();
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_create: out of host memory");
  m->V = vocab; m->N = n_art; m->E = embed_dim; m->H = hidden; m->D = n_samples; m->G = cell;
  m->precision = precision;
  const int GH = m->G * kH;

  ArenaPlan p;
  const size_t o_tab0 = p.take(sizeof(float) * 2 * (size_t)vocab * GH);
  size_t o_frag[2][2], o_wt[2][2], o_bn[2][2];
  for (int l = 0; l < 2; ++l)
    for (int d = 0; d < 2; ++d) {
      o_frag[l][d] = p.take(sizeof(uint32_t) * (size_t)GH * kH / 2);
      o_wt[l][d] = p.take(sizeof(float) * (size_t)GH * kH);
      o_bn[l][d] = p.take(sizeof(float) * kH);
    }
  const size_t o_wih1 = p.take(sizeof(float) * 2 * (size_t)GH * 2 * kH);
  const size_t o_bih1 = p.take(sizeof(float) * 2 * GH);
  const size_t o_lin_w = p.take(sizeof(float) * kH * 2 * kH);
  const size_t o_lin_b = p.take(sizeof(float) * kH);
  const size_t N = n_art;
  const size_t o_ln0_w = p.take(sizeof(float) * N * kH), o_ln0_b = p.take(sizeof(float) * N * kH);
  const size_t o_fc1_w = p.take(sizeof(float) * N * kFc * kH), o_fc1_b = p.take(sizeof(float) * N * kFc);
  const size_t o_ln1_w = p.take(sizeof(float) * N * kFc), o_ln1_b = p.take(sizeof(float) * N * kFc);
  const size_t o_fc2_w = p.take(sizeof(float) * N * kFc * kFc), o_fc2_b = p.take(sizeof(float) * N * kFc);
  const size_t o_ln2_w = p.take(sizeof(float) * N * kFc), o_ln2_b = p.take(sizeof(float) * N * kFc);
  const size_t o_head_w = p.take(sizeof(float) * N * kHeadPad * kFc), o_head_b = p.take(sizeof(float) * N * kHeadPad);
  const bool umma = umma_decoder_supported(m);
  const size_t o_umma = umma ? p.take(sizeof(float) * umma_image_floats(n_art)) : 0;
  const bool umma16 = umma16_decoder_supported(m);   // both precisions (fp32: split fp16 operands)
  const size_t o_umma16 =
      umma16 ? p.take(sizeof(float) * umma16_image_floats(n_art, m->precision != ASB_PRECISION_FAST)) : 0;
  const bool rnnu = rnn_umma_supported(m);
  size_t o_whhu[2][2] = {{0, 0}, {0, 0}};
  if (rnnu)
    for (int l = 0; l < 2; ++l)
      for (int d = 0; d < 2; ++d) o_whhu[l][d] = p.take(whh_umma_bytes(m));
  const bool proj = proj_umma_supported(m);
  const size_t o_proj = proj ? p.take(sizeof(float) * proj_image_floats(m)) : 0;
  const bool proj16 = proj_umma16_supported(m);
  const size_t o_proj16 = proj16 ? p.take(sizeof(float) * proj16_image_floats(m)) : 0;

  cudaError_t e = cudaMalloc(&m->arena, p.total);
  if (e != cudaSuccess) {
    delete m;
    return check_cuda(e, "asb_model_create: cudaMalloc of the weight arena");
  }
  m->arena_bytes = p.total;
  e = cudaHostAlloc(reinterpret_cast<void**>(&m->tok_flag_host), 2 * sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable);
  if (e == cudaSuccess) {
    m->tok_flag_host[0] = m->tok_flag_host[1] = 0;
    e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&m->tok_flag_dev), m->tok_flag_host, 0);
  }
  if (e != cudaSuccess) {
    if (m->tok_flag_host) cudaFreeHost(m->tok_flag_host);
    cudaFree(m->arena);
    delete m;
    return check_cuda(e, "asb_model_create: mapped token-status word");
  }
  char* b = static_cast<char*>(m->arena);
  auto F = [&](size_t off) { return reinterpret_cast<float*>(b + off); };
  PackedWeights& w = m->w;
  w.tab0 = F(o_tab0);
  for (int l = 0; l < 2; ++l)
    for (int d = 0; d < 2; ++d) {
      w.whh_frag[l][d] = reinterpret_cast<uint32_t*>(b + o_frag[l][d]);
      w.whh_t[l][d] = F(o_wt[l][d]);
      w.bhh_n[l][d] = F(o_bn[l][d]);
    }
  w.wih1 = F(o_wih1); w.bih1 = F(o_bih1);
  w.lin_w = F(o_lin_w); w.lin_b = F(o_lin_b);
  w.ln0_w = F(o_ln0_w); w.ln0_b = F(o_ln0_b);
  w.fc1_w = F(o_fc1_w); w.fc1_b = F(o_fc1_b);
  w.ln1_w = F(o_ln1_w); w.ln1_b = F(o_ln1_b);
  w.fc2_w = F(o_fc2_w); w.fc2_b = F(o_fc2_b);
  w.ln2_w = F(o_ln2_w); w.ln2_b = F(o_ln2_b);
  w.head_w = F(o_head_w); w.head_b = F(o_head_b);
  w.umma_image = umma ? F(o_umma) : nullptr;
  w.umma16_image = umma16 ? F(o_umma16) : nullptr;
  w.proj_image = proj ? F(o_proj) : nullptr;
  w.proj16_image = proj16 ? F(o_proj16) : nullptr;
  for (int l = 0; l < 2; ++l)
    for (int d = 0; d < 2; ++d) w.whh_umma[l][d] = rnnu ? static_cast<void*>(b + o_whhu[l][d]) : nullptr;
  m->decoder_impl = umma16 ? 2 : 0;   // the fp16-operand tcgen05 decoder (split operands in fp32 precision)
  m->rnn_impl = rnnu ? 1 : 0;   // the tcgen05 kernel (W_hh in tensor memory): fast precision GRU / LSTM, fp32 precision GRU (split operands)

  int rc = pack_recurrent_weights(m, src, st);
  if (!rc) rc = pack_dense_weights(m, src, st);
  if (!rc && umma) rc = pack_umma_image(m, src, st);
  if (!rc && umma16) rc = pack_umma16_image(m, src, st);
  if (!rc && proj) rc = pack_proj_image(m, st);
  if (!rc && proj16) rc = pack_proj16_image(m, src, st);
  if (!rc && rnnu) rc = pack_whh_umma(m, src, st);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(st), "asb_model_create: packing kernels");
  if (rc) {
    cudaFreeHost(m->tok_flag_host);
    cudaFree(m->arena);
    delete m;
    return rc;
  }
  *out = m;
  return ASB_OK;
}

extern "C" void asb_model_destroy(asb_model* m) {
  if (!m) return;
  for (auto& set : m->pending)
    for (auto e : set) cudaEventDestroy(e);
  for (auto& set : m->free_sets)
    for (auto e : set) cudaEventDestroy(e);
  if (m->tok_flag_host) cudaFreeHost(m->tok_flag_host);
  if (m->ev_fork) cudaEventDestroy(m->ev_fork);
  if (m->ev_join) cudaEventDestroy(m->ev_join);
  cudaFree(m->arena);
  delete m;
}

extern "C" int asb_model_token_status(asb_model* m, int32_t* bad_out, int64_t* token_out, int reset) {
  if (!m || !bad_out) return set_error(ASB_ERR_INVALID, "asb_model_token_status: null argument");
  volatile int* f = m->tok_flag_host;
  *bad_out = f[0];
  if (token_out) *token_out = f[1];
  if (reset) f[0] = f[1] = 0;
  return ASB_OK;
}

extern "C" int32_t* asb_model_token_flag_ptr(asb_model* m) { return m ? m->tok_flag_dev : nullptr; }

extern "C" size_t asb_model_workspace_bytes(const asb_model* m, int B, int T) {
  if (!m || B <= 0 || T <= 0) return 0;
  return plan_workspace(m, nullptr, B, T).bytes;
}

extern "C" int asb_model_forward(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host, int B,
                                 float* out, void* workspace, size_t workspace_bytes, void* stream) {
  return asb_model_forward_resident(m, x, T_in, lengths_host, nullptr, B, out, workspace, workspace_bytes, stream);
}

static int forward_impl(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host,
                        const int32_t* lengths_dev, int B, float* out, void* workspace, size_t workspace_bytes,
                        void* stream, bool hidden_only) {
  if (!m || !x || !lengths_host || !out) return set_error(ASB_ERR_INVALID, "asb_model_forward: null argument");
  if (B <= 0 || T_in <= 0) return set_error(ASB_ERR_INVALID, "asb_model_forward: empty batch");
  // pack_padded_sequence(enforce_sorted=True) semantics, models.py:136
  for (int b = 0; b < B; ++b) {
    if (lengths_host[b] <= 0)
      return set_error(ASB_ERR_UNSORTED,
                       "Length of all samples has to be greater than 0, but found an element in 'lengths' that is <= 0");
    if (b > 0 && lengths_host[b] > lengths_host[b - 1])
      return set_error(ASB_ERR_UNSORTED,
                       "`lengths` array must be sorted in decreasing order when `enforce_sorted` is True.");
  }
  const int T = lengths_host[0];
  if (T > T_in) return set_error(ASB_ERR_INVALID, "asb_model_forward: lengths[0] = %d exceeds the sequence dimension %d", T, T_in);
  Workspace ws = plan_workspace(m, workspace, B, T);
  if (!workspace || workspace_bytes < ws.bytes)
    return set_error(ASB_ERR_WORKSPACE, "asb_model_forward: workspace %zu B < required %zu B", workspace_bytes, ws.bytes);
  if (hidden_only) ws.out1 = out;   // the second recurrent layer writes straight into the caller's buffer
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long M = (long long)B * T;
  if (M > 0x7fffffffLL / 64) return set_error(ASB_ERR_INVALID, "asb_model_forward: batch too large");

  int rc = ASB_OK;
  if (lengths_dev) {
    ws.lengths = const_cast<int*>(lengths_dev);  // caller keeps a device copy (CUDA-graph friendly)
  } else {
    rc = check_cuda(cudaMemcpyAsync(ws.lengths, lengths_host, sizeof(int) * (size_t)B, cudaMemcpyHostToDevice, st),
                    "asb_model_forward: lengths H2D");
    if (rc) return rc;
  }
  // padded steps of the recurrent outputs are zero (pad_packed_sequence, models.py:138): the tcgen05
  // recurrent kernel writes those zeros itself, the other kernels expect a cleared buffer
  const bool umma_rnn = m->rnn_impl == 1 && m->w.whh_umma[0][0];
  if (!umma_rnn) {
    rc = check_cuda(cudaMemsetAsync(ws.out0, 0, sizeof(float) * (size_t)M * 2 * kH, st), "asb_model_forward: memset");
    if (rc) return rc;
    rc = check_cuda(cudaMemsetAsync(ws.out1, 0, sizeof(float) * (size_t)M * 2 * kH, st), "asb_model_forward: memset");
    if (rc) return rc;
  }
  const long long* tok = reinterpret_cast<const long long*>(x);
  std::array<cudaEvent_t, ASB_N_STAGES + 1> ev{};
  const bool prof = m->profile;
  if (prof) {
    if (!m->free_sets.empty()) {
      ev = m->free_sets.back();
      m->free_sets.pop_back();
    } else {
      for (auto& e : ev)
        if ((rc = check_cuda(cudaEventCreate(&e), "asb_model_forward: cudaEventCreate"))) return rc;
    }
  }
  auto mark = [&](int k) {
    if (prof) cudaEventRecord(ev[k], st);
  };
  mark(0);
  // the tcgen05 recurrent kernel looks the embedding / layer-0 projection table up itself
  const bool fused_gather = umma_rnn && rnn_umma_reads_tokens(m, B, T);
  if (!fused_gather && (rc = launch_gather_gi0(m, tok, T_in, ws.gi, B, T, st))) return rc;
  mark(1);
  if (umma_rnn) {
    if ((rc = launch_rnn_umma(m, 0, ws.gi, fused_gather ? tok : nullptr, T_in, ws.lengths, ws.out0, B, T, st)))
      return rc;
  } else {
    if ((rc = launch_recurrent_layer(m, 0, ws.gi, ws.lengths, ws.out0, B, T, st))) return rc;
  }
  mark(2);
  if (m->w.proj16_image && m->decoder_impl == 2) {  // the tcgen05 kernels are selected together
    if ((rc = launch_proj_umma16(m, ws.out0, ws.gi, int(M), st))) return rc;
  } else if (m->w.proj_image && m->decoder_impl == 1) {
    if ((rc = launch_proj_umma(m, ws.out0, ws.gi, int(M), st))) return rc;
  } else {
    if ((rc = launch_linear(m, ws.out0, m->w.wih1, m->w.bih1, ws.gi, int(M), 2 * m->G * kH, 2 * kH, st))) return rc;
  }
  mark(3);
  if (umma_rnn) {
    if ((rc = launch_rnn_umma(m, 1, ws.gi, nullptr, 0, ws.lengths, ws.out1, B, T, st))) return rc;
  } else {
    if ((rc = launch_recurrent_layer(m, 1, ws.gi, ws.lengths, ws.out1, B, T, st))) return rc;
  }
  mark(4);
  if (hidden_only) {
    // nothing: `out` already holds the recurrent output
  } else if (m->decoder_impl == 2) {
    if ((rc = launch_decoder_umma16(m, ws.out1, out, int(M), st))) return rc;
  } else if (m->decoder_impl == 1) {
    if ((rc = launch_decoder_umma(m, ws.out1, out, int(M), st))) return rc;
  } else {
    if ((rc = launch_decoder(m, ws.out1, out, int(M), st))) return rc;
  }
  mark(5);
  if (prof) m->pending.push_back(ev);
  return ASB_OK;
}

extern "C" int asb_model_forward_resident(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host,
                                          const int32_t* lengths_dev, int B, float* out, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  return forward_impl(m, x, T_in, lengths_host, lengths_dev, B, out, workspace, workspace_bytes, stream, false);
}

extern "C" int asb_model_forward_hidden(asb_model* m, const int64_t* x, int T_in, const int32_t* lengths_host,
                                        const int32_t* lengths_dev, int B, float* hidden, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  return forward_impl(m, x, T_in, lengths_host, lengths_dev, B, hidden, workspace, workspace_bytes, stream, true);
}

// Decoder only: linear.0 + ReLU -> every predictor -> sigmoid on caller-supplied feature rows
// (models.py:140-145).  Used by SimpleArtSpeech (models.py:53-96), whose features are embedding rows.
extern "C" int asb_model_decode(asb_model* m, const float* features, int64_t rows, float* out, void* stream) {
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_decode: null model");
  if (rows < 0 || rows > 0x7fffffffLL / 64) return set_error(ASB_ERR_INVALID, "asb_model_decode: bad row count");
  if (rows == 0) return ASB_OK;
  if (!features || !out) return set_error(ASB_ERR_INVALID, "asb_model_decode: null pointer");
  if (reinterpret_cast<uintptr_t>(features) & 15)
    return set_error(ASB_ERR_INVALID, "asb_model_decode: features must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (m->decoder_impl == 2) return launch_decoder_umma16(m, features, out, int(rows), st);
  return m->decoder_impl == 1 ? launch_decoder_umma(m, features, out, int(rows), st)
                              : launch_decoder(m, features, out, int(rows), st);
}

extern "C" int asb_model_set_recurrent(asb_model* m, int impl) {
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_set_recurrent: null model");
  if (impl == 1 && !m->w.whh_umma[0][0])
    return set_error(ASB_ERR_UNSUPPORTED, "asb_model_set_recurrent: the tcgen05 recurrent kernel has no fp32-precision LSTM");
  if (impl != 0 && impl != 1) return set_error(ASB_ERR_INVALID, "asb_model_set_recurrent: impl must be 0 or 1");
  if (impl == 0 && m->G != ASB_CELL_GRU)
    return set_error(ASB_ERR_UNSUPPORTED, "asb_model_set_recurrent: the mma.sync kernel implements the GRU cell only");
  m->rnn_impl = impl;
  return ASB_OK;
}

extern "C" int asb_model_set_decoder(asb_model* m, int impl) {
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_set_decoder: null model");
  if (impl == 1 && !m->w.umma_image)
    return set_error(ASB_ERR_UNSUPPORTED, "asb_model_set_decoder: the tf32 tcgen05 decoder needs fast precision and n_samples = 50");
  if (impl == 2 && !m->w.umma16_image)
    return set_error(ASB_ERR_UNSUPPORTED, "asb_model_set_decoder: the tcgen05 decoder needs n_samples = 50");
  if (impl < 0 || impl > 2) return set_error(ASB_ERR_INVALID, "asb_model_set_decoder: impl must be 0, 1 or 2");
  m->decoder_impl = impl;
  return ASB_OK;
}

extern "C" int asb_model_profile_enable(asb_model* m, int enable) {
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_profile_enable: null model");
  m->profile = enable != 0;
  return ASB_OK;
}

extern "C" int asb_model_profile_read(asb_model* m, float* stage_ms_out, int32_t* stage_calls_out, int reset) {
  if (!m) return set_error(ASB_ERR_INVALID, "asb_model_profile_read: null model");
  for (auto& ev : m->pending) {
    int rc = check_cuda(cudaEventSynchronize(ev[ASB_N_STAGES]), "asb_model_profile_read: event sync");
    if (rc) return rc;
    for (int k = 0; k < ASB_N_STAGES; ++k) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ev[k], ev[k + 1]) == cudaSuccess) {
        m->stage_ms[k] += ms;
        m->stage_calls[k] += 1;
      }
    }
    m->free_sets.push_back(ev);
  }
  m->pending.clear();
  for (int k = 0; k < ASB_N_STAGES; ++k) {
    if (stage_ms_out) stage_ms_out[k] = float(m->stage_ms[k]);
    if (stage_calls_out) stage_calls_out[k] = m->stage_calls[k];
    if (reset) {
      m->stage_ms[k] = 0;
      m->stage_calls[k] = 0;
    }
  }
  return ASB_OK;
}
