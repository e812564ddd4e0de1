/*
 * Deformation-network entry points of ed3dgs_b200 (included by ed3dgs_b200.h).
 *
 * Replaces the torch module scene/deformation.py:15-148 of the reference:
 *   get_temporal_embed (:53-67)  -> ed3dgs_temporal_embed_forward / _backward
 *   query_time (:72-88) + feature_out (:39-44) + deform (:90-106) + the five heads (:47-51),
 *   coarse and fine stage (:125-137)  -> ed3dgs_deform_forward / _backward
 *
 * The network input is cat([temporal embedding (identical for every Gaussian), per-Gaussian
 * embedding]); the temporal half is folded into a per-call bias, so per-Gaussian K is gemb_dim.
 * All matrices are nn.Linear layouts ([out, in], row-major, fp32). GEMMs run on the tensor
 * cores as 3xTF32 (error-compensated, fp32-grade) -- see csrc/tc_gemm.cuh.
 */
#ifndef ED3DGS_DEFORM_H_
#define ED3DGS_DEFORM_H_

#include <stddef.h>

#define ED3DGS_DEFORM_HEADS 5       /* pos, scales, rotations, opacity, rgb  (deformation.py:47-51) */
#define ED3DGS_DEFORM_MAX_DEPTH 8   /* Linear layers in feature_out (defor_depth) */

/* Output widths of the five heads: 3, 3, 4, 1, 48 (16 SH coefficients x RGB). */

typedef struct {
  int W;          /* net_width: multiple of 16, <= 128 */
  int depth;      /* number of Linear layers in feature_out: max(defor_depth, 1) */
  int temb_dim;   /* temporal_embedding_dim (256), multiple of 4 */
  int gemb_dim;   /* gaussian_embedding_dim (32), multiple of 4 */
  /* feature_out.{0,2,4,...}: weight [W, temb_dim+gemb_dim] for layer 0, [W, W] after; bias [W] */
  const float* feat_w[ED3DGS_DEFORM_MAX_DEPTH];
  const float* feat_b[ED3DGS_DEFORM_MAX_DEPTH];
  /* {pos,scales,rotations,opacity,rgb}_deform.1 : [W, W], [W];  .3 : [k, W], [k].
     head_w1[h] == NULL disables head h (no_ds / no_dr / no_do / no_dc; pos is always on). */
  const float* head_w1[ED3DGS_DEFORM_HEADS];
  const float* head_b1[ED3DGS_DEFORM_HEADS];
  const float* head_w2[ED3DGS_DEFORM_HEADS];
  const float* head_b2[ED3DGS_DEFORM_HEADS];
  /* Arithmetic of the tensor-core GEMMs. 0 (default): 3xTF32, every product as hi*hi + hi*lo + lo*hi of the
     operands' tf32 halves -- fp32-grade, the mode all parity bounds are stated for (outputs <= 2e-5, gradients
     <= 1e-4 relative). 1 ("tf32"): one tf32 MMA per product, what torch runs with allow_tf32 = True --
     outputs <= 2e-3, gradients <= 1e-2 relative (tests/test_deform_gpu.py), a third of the tensor work. */
  int precision;
} ed3dgs_deform_weights;

/* Same shapes as the weights; every non-NULL gradient is ACCUMULATED into (+=). */
typedef struct {
  float* feat_w[ED3DGS_DEFORM_MAX_DEPTH];
  float* feat_b[ED3DGS_DEFORM_MAX_DEPTH];
  float* head_w1[ED3DGS_DEFORM_HEADS];
  float* head_b1[ED3DGS_DEFORM_HEADS];
  float* head_w2[ED3DGS_DEFORM_HEADS];
  float* head_b2[ED3DGS_DEFORM_HEADS];
} ed3dgs_deform_grads;

/*
 * tau[dim] = the temporal embedding at time *t_dev (device scalar): bilinear resize of
 * weight[max_emb, dim] to n_rows rows (align_corners=True) then bilinear sampling with reflection
 * padding at y = (t - 0.5) * 2 (deformation.py:53-67). Backward accumulates (+=) into
 * d_weight[max_emb, dim] and d_t[1].
 */
int ed3dgs_temporal_embed_forward(const float* weight, int max_emb, int dim, int n_rows,
                                  const float* t_dev, float* tau, void* stream);
int ed3dgs_temporal_embed_backward(const float* weight, int max_emb, int dim, int n_rows,
                                   const float* t_dev, const float* d_tau, float* d_weight,
                                   float* d_t, void* stream);

/* Bytes of device workspace forward/backward need for N Gaussians (saved activations + scratch). */
size_t ed3dgs_deform_workspace_bytes(int W, int depth, int N);

/*
 * Multi-GPU overlap (no reference counterpart: the reference is single-GPU, train.py:171-187). The persistent
 * tcgen05 kernels of the deformation network launch one CTA per SM; `n_sms` > 0 makes every later launch leave that
 * many SMs free, so that an NCCL collective running on another stream (the all-reduce of the Gaussian-parameter
 * gradients, final before the network's backward starts) has SMs of its own. 0 restores the full grid. Kept per device:
 * it applies to the device that is current when it is called (and is read with the device current at the launch).
 * Returns the previous value.
 */
int ed3dgs_deform_set_sm_reserve(int n_sms);

/*
 * `cuda_event` (a cudaEvent_t, or NULL to switch it off) is recorded on the call's stream inside every later
 * ed3dgs_deform_backward right after the last kernel that writes d_embedding, i.e. before the weight-gradient launch of
 * the last stage: a multi-GPU caller makes its exchange stream wait for it and reduces the embedding gradient while that
 * launch runs. Kept per device: the event must belong to the device that is current at this call, and only backward
 * calls made with that device current record it. A failed record makes ed3dgs_deform_backward return ED3DGS_ERR_CUDA.
 * Set it back to NULL before destroying the event.
 */
void ed3dgs_deform_set_emb_ready_event(void* cuda_event);

/*
 * SMs left free by the weight-gradient launch of the LAST stage of every later ed3dgs_deform_backward (the launch that runs
 * after the event above), when larger than the general reserve: the embedding exchange has that one launch to finish in.
 * Per device, like the two settings above. Returns the previous value.
 */
int ed3dgs_deform_set_sm_reserve_last(int n_sms);

/*
 * Forward of both stages. coarse / fine may be NULL (no_coarse_deform / no_fine_deform).
 *   in[h], sub[h], out[h]: [N, k_h] for h = pos, scales, rotations, opacity, rgb.
 *   sub = in + delta_coarse * coef   (the reference's *_sub tuple), out = sub + delta_fine * coef.
 *   coef[h]: the annealing coefficients of deform() (:91-105): {scale, scale*coef_s, scale,
 *   scale*scale_o, scale_c}. Disabled heads copy their input through.
 * workspace keeps what backward needs; it must stay untouched until backward has run.
 */
int ed3dgs_deform_forward(const ed3dgs_deform_weights* coarse, const ed3dgs_deform_weights* fine,
                          int N, const float* embedding, const float* tau_c, const float* tau_f,
                          const float* const in[ED3DGS_DEFORM_HEADS],
                          float* const sub[ED3DGS_DEFORM_HEADS],
                          float* const out[ED3DGS_DEFORM_HEADS],
                          const float coef[ED3DGS_DEFORM_HEADS], void* workspace,
                          size_t workspace_bytes, void* stream);

/*
 * Backward of both stages.
 *   d_out[h]: dL/d out[h] (NULL = zero); d_sub[h]: extra dL/d sub[h] (NULL = zero). The rotation and rgb rows
 *   (h = 2, 4) are read as float4: those pointers must be 16-byte aligned (ED3DGS_ERR_ARG otherwise).
 *   Writes d_embedding[N, gemb_dim], d_tau_c[temb_dim], d_tau_f[temb_dim]; accumulates weight
 *   gradients into g_coarse / g_fine. The gradient w.r.t. in[h] is d_out[h] + d_sub[h]
 *   (identity path) and is left to the caller.
 */
int ed3dgs_deform_backward(const ed3dgs_deform_weights* coarse, const ed3dgs_deform_weights* fine,
                           const ed3dgs_deform_grads* g_coarse, const ed3dgs_deform_grads* g_fine,
                           int N, const float* embedding, const float* tau_c, const float* tau_f,
                           const float* const d_out[ED3DGS_DEFORM_HEADS],
                           const float* const d_sub[ED3DGS_DEFORM_HEADS],
                           const float coef[ED3DGS_DEFORM_HEADS], void* workspace,
                           size_t workspace_bytes, float* d_embedding, float* d_tau_c,
                           float* d_tau_f, void* stream);

#endif /* ED3DGS_DEFORM_H_ */
