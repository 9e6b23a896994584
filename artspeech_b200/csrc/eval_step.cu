// One evaluation step of run_test as ONE native call: forward -> P2CP -> tract variables of prediction
// and target -> length-masked accumulation (phoneme_to_articulation/encoder_decoder/evaluation.py:53-128
// and the frame driver phoneme_to_articulation/__init__.py:201-297, batched).
//
// The step is ~10 kernels of 20-150 us each.  Enqueued from Python they leave the device idle between
// launches (0.9 ms eager vs 0.48 ms of device time per 25 600-frame step); enqueued from here the host
// needs ~30 us per step, so an evaluation loop over resident batches is device-bound without a CUDA graph
// per batch, and several steps (three by default, evaluation.PipelinedEvaluator) can be kept in flight on their own streams from one host thread.
//
// The target's tract variables do not depend on the forward pass: they run on `side_stream` (forked
// here, joined before the accumulation), enqueued AFTER the forward's kernels - a full wave of
// tract-variable CTAs takes every register of every SM, so the recurrent kernels (the critical path)
// must be placed first.  Inside a stream capture the fork/join become graph edges.
#include "model.cuh"

extern "C" int asb_eval_step(asb_model* m, const int64_t* tokens, int T_in, const int32_t* lengths_host,
                             const int32_t* lengths_dev, int B, const float* targets,
                             const int32_t* art_index_host, const float* upper_incisor, int64_t upper_incisor_stride,
                             const asb_eval_buffers* buf, void* stream, void* side_stream) {
  using namespace asb;
  if (!m || !tokens || !lengths_host || !lengths_dev || !targets || !art_index_host || !buf)
    return set_error(ASB_ERR_INVALID, "asb_eval_step: null argument");
  if (!buf->outputs || !buf->p2cp || !buf->tv_pred || !buf->tv_pred_idx || !buf->tv_pred_poc || !buf->tv_target ||
      !buf->tv_target_idx || !buf->tv_target_poc || !buf->scratch || !buf->acc || !buf->workspace)
    return set_error(ASB_ERR_INVALID, "asb_eval_step: null buffer");
  if (B <= 0 || T_in <= 0) return set_error(ASB_ERR_INVALID, "asb_eval_step: empty batch");
  if (m->D != 50) return set_error(ASB_ERR_UNSUPPORTED, "asb_eval_step: the tract variables need 50-point contours");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaStream_t side = side_stream ? static_cast<cudaStream_t>(side_stream) : st;
  const int T = lengths_host[0];   // validated by the forward
  int rc;

  if (side != st) {
    if (!m->ev_fork) {
      if ((rc = check_cuda(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming), "asb_eval_step: event"))) return rc;
      if ((rc = check_cuda(cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming), "asb_eval_step: event"))) return rc;
    }
    if ((rc = check_cuda(cudaEventRecord(m->ev_fork, st), "asb_eval_step: fork"))) return rc;
  }
  // evaluation.py:57
  if ((rc = asb_model_forward_resident(m, tokens, T_in, lengths_host, lengths_dev, B, buf->outputs, buf->workspace,
                                       buf->workspace_bytes, stream)))
    return rc;
  const int64_t frames = int64_t(B) * T;
  // __init__.py:261 (target side), beside the forward
  if (side != st && (rc = check_cuda(cudaStreamWaitEvent(side, m->ev_fork, 0), "asb_eval_step: fork wait"))) return rc;
  if ((rc = asb_tract_variables_ex(targets, frames, m->N, art_index_host, upper_incisor, upper_incisor_stride, 0,
                                   buf->tv_target, buf->tv_target_idx, buf->tv_target_poc, side)))
    return rc;
  if (side != st && (rc = check_cuda(cudaEventRecord(m->ev_join, side), "asb_eval_step: join"))) return rc;
  // evaluation.py:72, all sentences at once
  if ((rc = asb_p2cp(buf->outputs, targets, buf->p2cp, frames * m->N, m->D, m->D, ASB_LAYOUT_PLANAR, stream))) return rc;
  // __init__.py:256 (prediction side)
  if ((rc = asb_tract_variables_ex(buf->outputs, frames, m->N, art_index_host, upper_incisor, upper_incisor_stride, 0,
                                   buf->tv_pred, buf->tv_pred_idx, buf->tv_pred_poc, stream)))
    return rc;
  if (side != st && (rc = check_cuda(cudaStreamWaitEvent(st, m->ev_join, 0), "asb_eval_step: join wait"))) return rc;
  // evaluation.py:68-84,142-156
  return asb_eval_accumulate_total(buf->p2cp, buf->tv_pred, buf->tv_target, lengths_dev, B, T, m->N, buf->scratch,
                                   buf->acc, buf->total, stream);
}
