// The fused training step of the example CNN as ONE host call: the ten kernel launches of forward + backward
// (cifar_example/sgd-torchad_nn-cifar.lua:166-183: model:forward, criterion:forward, criterion:backward, model:backward)
// are issued from C++, back to back, instead of through ten Python -> ctypes round trips with their argument marshalling
// and tensor allocations.  At B = 4096 a step is 0.23 ms of device time; the Python launch path cost 0.25 ms of host time.
#include "common.cuh"

using namespace tadnn;

#define STEP_TRY(call)          \
  do {                          \
    if (int _rc = (call)) return _rc; \
  } while (0)

extern "C" int tadnn_cnn_step(const tadnn_cnn_step_args* s, tadnn_stream_t stream) {
  TADNN_REQUIRE(s != nullptr && s->B > 0, "cnn_step: bad arguments");
  TADNN_REQUIRE(s->phase >= 1 && s->phase <= 3, "cnn_step: phase %d (1 = forward and backward down to dp1, 2 = conv1 backward, 3 = both)", s->phase);
  const int B = s->B;
  if (s->phase & 1) {
    TADNN_REQUIRE(s->xq && s->target && s->w1 && s->b1 && s->w2 && s->b2 && s->fw1 && s->fb1 && s->fw2 && s->fb2, "cnn_step: null input");
    TADNN_REQUIRE(s->w1x && s->w2p && s->Tt && s->dwp && s->p1 && s->idx1 && s->p2 && s->idx2 && s->hidden && s->logp && s->loss && s->da2w &&
                      s->dp1 && s->head_ws,
                  "cnn_step: null workspace");
    // ---- forward
    STEP_TRY(tadnn_cnn_pack_shadows_sliding(s->w1, s->b1, s->w2, s->w1x, s->w2p, s->Tt, stream));
    STEP_TRY(tadnn_conv1s_pool_forward(s->xq, s->w1x, s->p1, s->idx1, B, stream));
    STEP_TRY(tadnn_conv2_pool_forward(s->p1, s->w2p, s->b2, s->p2, s->idx2, nullptr, B, stream));
    tadnn_head_args h;
    memset(&h, 0, sizeof(h));
    h.p2 = s->p2; h.w1 = s->fw1; h.b1 = s->fb1; h.w2 = s->fw2; h.b2 = s->fb2;
    h.target = s->target;
    h.B = B; h.D = 64; h.H = 100; h.NC = 10;
    h.dropout_p = s->dropout_p; h.seed = s->seed; h.seed_dev = s->seed_dev;
    h.hidden = s->hidden; h.logp = s->logp; h.loss = s->loss;
    STEP_TRY(tadnn_head_forward(&h, s->head_ws, stream));
    if (s->forward_only) return 0;
    // ---- backward: head (+ pool2/ReLU routing, conv2 bias gradient), conv2 weight gradient, transposed conv2
    TADNN_REQUIRE(s->gw1 && s->gb1 && s->gw2 && s->gb2 && s->gfw1 && s->gfb1 && s->gfw2 && s->gfb2, "cnn_step: null gradient");
    h.gw1 = s->gfw1; h.gb1 = s->gfb1; h.gw2 = s->gfw2; h.gb2 = s->gfb2;
    h.grad_scale = s->grad_scale;
    h.pool_idx = s->idx2; h.dpre = s->da2w; h.gbias_pre = s->gb2;
    STEP_TRY(tadnn_head_backward(&h, stream));
    STEP_TRY(tadnn_conv2_wgrad(s->p1, s->da2w, s->dwp, B, stream));
    STEP_TRY(tadnn_conv2_wgrad_unpack(s->dwp, s->gw2, stream));
    tadnn_gemm_args g;
    memset(&g, 0, sizeof(g));
    g.M = B; g.N = 7 * 9 * 64; g.K = 576;
    g.A = s->da2w; g.lda = 576; g.a_mn_major = 0;
    g.B = s->Tt; g.ldb = 7 * 9 * 64; g.b_mn_major = 1;
    g.C = s->dp1; g.ldc = 81 * 64; g.c_fp32 = 0;
    g.alpha = 1.0f; g.tile_n = 256;
    g.kblock_mask = s->tt_mask; g.n_kblock_mask = s->n_tt_mask;
    STEP_TRY(tadnn_gemm_bf16(&g, stream));
  }
  if (s->phase & 2) {
    TADNN_REQUIRE(s->xq && s->dp1 && s->idx1 && s->gw1 && s->gb1, "cnn_step: null argument of the conv1 backward");
    STEP_TRY(tadnn_conv1s_pool_backward(s->xq, s->dp1, s->idx1, s->gw1, s->gb1, 7, B, stream));   // dp1 rows 7, 8 are never written
  }
  return 0;
}
