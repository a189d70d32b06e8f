// One step of iterative SMPL fitting on the 2D re-projection loss (BASELINE config 3; loss block
// pose_model_3d.py:194-238, optimiser :393-401) with per-sample pose (and optionally shape) as the Adam
// variables.  Joint-only path: no vertices are touched.
//   fit_fused (joint_bwd_kernel with the residual formed inside: forward chain + projection + dL/dpx + loss partial sums +
//   hand-written adjoint, one warp per sample) -> loss_finish (mean, cap gate, Adam's lr_t) -> Adam (TF1 formula,
//   gradient gated by the cap like tf.minimum).  Round 1 ran pose_prep -> reproj_grad -> loss_finish -> joint_bwd -> Adam:
//   joint_bwd recomputes the forward chain anyway, so the two extra kernels (and their launches) were pure overhead.
// All launches are stream-ordered and allocation-free, so a caller can capture the step in a CUDA graph: with
// smplb_fit_io.step_dev the step counter lives on the device (loss_finish increments it and derives Adam's lr_t from
// it), so every replay of the captured step advances the bias correction like an eager step does.
#include "common.cuh"

namespace smplb {

int fit_fused_blocks(int B);
int32_t launch_fit_fused(const smplb_model* m, int B, const float* beta, const float* theta, const float* gt_px,
                         float px_coef, const float* cam_pos, const float* cam_rot, const float* focal,
                         const smplb_camcfg& cfg, float* px, float* g_px, double* partial, float* g_beta, float* g_theta,
                         const FitAdam* adam, cudaStream_t st);

constexpr int RG_THREADS = 256;

__device__ __forceinline__ float adam_lr_t(float lr, float b1, float b2, int t) {
  return lr * sqrtf(1.0f - powf(b2, (float)t)) / (1.0f - powf(b1, (float)t));
}

// g_px = 2 * scale / n_total * (px - gt)   (d/dpx of scale * mean((gt - px)^2), pose_model_3d.py:233-235)
// partial[block] = sum (px - gt)^2 over the block's elements (fixed tree order)
__global__ void __launch_bounds__(RG_THREADS) reproj_grad_kernel(int n, const float* __restrict__ px,
                                                                 const float* __restrict__ gt, float coef,
                                                                 float* __restrict__ g_px, double* __restrict__ partial) {
  __shared__ double red[RG_THREADS / 32];
  const int i = blockIdx.x * RG_THREADS + threadIdx.x;
  double acc = 0.0;
  if (i < n) {
    const float d = px[i] - gt[i];
    g_px[i] = coef * d;
    acc = (double)d * d;
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < RG_THREADS / 32; ++w) s += red[w];
    partial[blockIdx.x] = s;
  }
}

// loss[0] = scale * sum / n_total (this rank's share of the global mean), loss[1] = min(cap, loss[0]),
// loss[2] = gate (1 when the cap is inactive: tf.minimum passes the gradient, 0 otherwise),
// loss[3] = lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) of this step (tf.train.AdamOptimizer, pose_model_3d.py:393); t is
// the device counter *step_dev, incremented here, or step_host when there is none
__global__ void __launch_bounds__(256) loss_finish_kernel(int nblocks, const double* __restrict__ partial, float scale,
                                                          double inv_n_total, float cap, float* __restrict__ loss,
                                                          int* __restrict__ step_dev, int step_host, float lr, float b1,
                                                          float b2) {
  __shared__ double red[8];
  pdl_trigger();   // the next step's per-sample kernel may be scheduled (it waits for this grid before it reads anything)
  pdl_wait();      // the per-sample kernel of this step has completed: its partials are visible
  double acc = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 256) acc += partial[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += red[w];
    const float l = scale * (float)(s * inv_n_total);
    loss[0] = l;
    loss[1] = cap > 0.f ? fminf(cap, l) : l;
    loss[2] = (cap > 0.f && l > cap) ? 0.f : 1.f;
    int t = step_host;
    if (step_dev) {
      t = *step_dev + 1;
      *step_dev = t;
    }
    loss[3] = adam_lr_t(lr, b1, b2, t);
  }
}

// tf.train.AdamOptimizer (pose_model_3d.py:393) with the gradient multiplied by a device-side gate; lr_t is read from
// device memory (written by loss_finish_kernel / adam_tick_kernel), so a captured launch is not tied to one step
__global__ void adam_gated_kernel(int64_t n, float* __restrict__ x, const float* __restrict__ g,
                                  float* __restrict__ m, float* __restrict__ v, const float* __restrict__ gate,
                                  const float* __restrict__ lr_t_dev, float b1, float b2, float eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float lr_t = __ldg(lr_t_dev);
  const float gi = g[i] * (gate ? __ldg(gate) : 1.0f);
  const float mi = b1 * m[i] + (1.0f - b1) * gi;
  const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  x[i] = x[i] - lr_t * mi / (sqrtf(vi) + eps);
}

// step counter on the device: t = ++(*step_dev), lr_t_dev[0] = lr_t(t)
__global__ void adam_tick_kernel(int* __restrict__ step_dev, float lr, float b1, float b2, float* __restrict__ lr_t_dev) {
  const int t = *step_dev + 1;
  *step_dev = t;
  lr_t_dev[0] = adam_lr_t(lr, b1, b2, t);
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_adam_tick(int32_t* step_dev, float lr, float beta1, float beta2, float* lr_t_dev, void* stream) {
  if (!step_dev || !lr_t_dev) return fail(SMPLB_E_INVALID_ARG, "smplb_adam_tick: null pointer");
  adam_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev, lr, beta1, beta2, lr_t_dev);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_adam_step_dev(int64_t n, float* x, const float* g, float* m, float* v, const float* lr_t_dev,
                                       float beta1, float beta2, float eps, void* stream) {
  if (n < 0) return fail(SMPLB_E_INVALID_ARG, "smplb_adam_step_dev: n >= 0 required");
  if (n == 0) return SMPLB_OK;
  if (!x || !g || !m || !v || !lr_t_dev) return fail(SMPLB_E_INVALID_ARG, "smplb_adam_step_dev: null pointer");
  adam_gated_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, x, g, m, v, nullptr, lr_t_dev,
                                                                                   beta1, beta2, eps);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" size_t smplb_fit_scratch_bytes(int32_t B) {
  const size_t n = (size_t)(B > 0 ? B : 0) * NJ * 2;
  const size_t a = (n + RG_THREADS - 1) / RG_THREADS, b = (size_t)fit_fused_blocks(B > 0 ? B : 0);
  return (a > b ? a : b) * sizeof(double) + 256;
}

extern "C" int32_t smplb_fit_step(const smplb_model* m, const smplb_fit_io* io, void* stream) {
  if (!m || !io) return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: null model or io");
  if (int32_t rc = check_current_device(m->device, "smplb_fit_step")) return rc;
  if (io->B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_fit_step: negative batch");
  if (io->B == 0) return SMPLB_OK;
  if (!io->beta || !io->theta || !io->cam_pos || !io->cam_rot || !io->gt_px || !io->px || !io->g_px ||
      !io->g_beta || !io->g_theta || !io->m_theta || !io->v_theta || !io->loss || !io->scratch)
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: null pointer");
  if ((io->m_beta == nullptr) != (io->v_beta == nullptr))
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: m_beta and v_beta go together");
  if ((!io->step_dev && io->step < 1) || io->n_total < (int64_t)io->B * NJ * 2)
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: step >= 1 (or step_dev) and n_total >= B*48 required");
  // the cap of pose_model_3d.py:237 acts on the GLOBAL mean (tf.minimum of the batch loss); a shard only knows its
  // share, so the gate cannot be decided here when the batch is split over ranks
  if (io->cfg.reproj_cap > 0.f && io->n_total != (int64_t)io->B * NJ * 2)
    return fail(SMPLB_E_UNSUPPORTED, "smplb_fit_step: reproj_cap > 0 needs n_total == B*48 (the cap gates on the global loss; "
                                     "fit sharded batches uncapped or gate on the all-reduced loss in the caller)");
  cudaStream_t st = (cudaStream_t)stream;
  const int B = io->B;
  const int nblocks = fit_fused_blocks(B);
  double* partial = (double*)io->scratch;
  const double inv_n = 1.0 / (double)io->n_total;
  // g_px = 2 * scale / n_total * (px - gt)   (d/dpx of scale * mean((gt - px)^2), pose_model_3d.py:233-235)
  // With the cap off the gate is 1 whatever the loss is, so Adam runs inside the fused kernel on the gradient it holds in
  // registers (2 launches per step); with the cap on, the update has to wait for the global loss (3 - 4 launches).
  const bool adam_inside = !(io->cfg.reproj_cap > 0.f);
  FitAdam ad;
  ad.theta = io->theta; ad.m_theta = io->m_theta; ad.v_theta = io->v_theta;
  ad.beta = io->beta; ad.m_beta = io->m_beta; ad.v_beta = io->v_beta;
  ad.step_dev = io->step_dev; ad.step_host = io->step;
  ad.lr = io->lr; ad.b1 = io->beta1; ad.b2 = io->beta2; ad.eps = io->eps;
  int32_t rc = launch_fit_fused(m, B, io->beta, io->theta, io->gt_px, (float)(2.0 * io->cfg.reproj_scale * inv_n),
                                io->cam_pos, io->cam_rot, io->focal, io->cfg.cam, io->px, io->g_px, partial, io->g_beta,
                                io->g_theta, adam_inside ? &ad : nullptr, st);
  if (rc != SMPLB_OK) return rc;
  SMPLB_CUDA_CHECK(launch_pdl(loss_finish_kernel, dim3(1), dim3(256), 0, st, nblocks, (const double*)partial,
                              io->cfg.reproj_scale, inv_n, io->cfg.reproj_cap, io->loss, io->step_dev, io->step, io->lr,
                              io->beta1, io->beta2));
  if (!adam_inside) {
    const int64_t nt = (int64_t)B * NJ * 3;
    adam_gated_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(nt, io->theta, io->g_theta, io->m_theta, io->v_theta,
                                                                    io->loss + 2, io->loss + 3, io->beta1, io->beta2, io->eps);
    if (io->m_beta) {
      const int64_t nb = (int64_t)B * NB;
      adam_gated_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>(nb, io->beta, io->g_beta, io->m_beta, io->v_beta,
                                                                      io->loss + 2, io->loss + 3, io->beta1, io->beta2, io->eps);
    }
  }
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
