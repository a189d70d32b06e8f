// One step of iterative SMPL fitting on the 2D re-projection loss (BASELINE config 3; loss block
// pose_model_3d.py:194-238, optimiser :393-401) with per-sample pose (and optionally shape) as the Adam
// variables.  Joint-only path: no vertices are touched.
//   pose_prep (forward + projection) -> reproj_grad (dL/dpx, loss partial sums) -> loss_finish (mean, cap gate)
//   -> joint_bwd (hand-written adjoint) -> Adam (TF1 formula, gradient gated by the cap like tf.minimum)
// All launches are stream-ordered and allocation-free, so a caller can capture the step in a CUDA graph.
#include "common.cuh"

namespace smplb {

int32_t launch_joint_bwd_px(const smplb_model* m, int B, const float* beta, const float* theta, const float* g_px,
                            const float* cam_pos, const float* cam_rot, const float* focal, const smplb_camcfg& cfg,
                            float* g_beta, float* g_theta, cudaStream_t st);

constexpr int RG_THREADS = 256;

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
// loss[2] = gate (1 when the cap is inactive: tf.minimum passes the gradient, 0 otherwise)
__global__ void __launch_bounds__(256) loss_finish_kernel(int nblocks, const double* __restrict__ partial, float scale,
                                                          double inv_n_total, float cap, float* __restrict__ loss) {
  __shared__ double red[8];
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
  }
}

// tf.train.AdamOptimizer (pose_model_3d.py:393) with the gradient multiplied by a device-side gate
__global__ void adam_gated_kernel(int64_t n, float* __restrict__ x, const float* __restrict__ g,
                                  float* __restrict__ m, float* __restrict__ v, const float* __restrict__ gate,
                                  float lr_t, float b1, float b2, float eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gi = g[i] * (gate ? __ldg(gate) : 1.0f);
  const float mi = b1 * m[i] + (1.0f - b1) * gi;
  const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  x[i] = x[i] - lr_t * mi / (sqrtf(vi) + eps);
}

}  // namespace smplb

using namespace smplb;

extern "C" size_t smplb_fit_scratch_bytes(int32_t B) {
  const size_t n = (size_t)(B > 0 ? B : 0) * NJ * 2;
  return ((n + RG_THREADS - 1) / RG_THREADS) * sizeof(double) + 256;
}

extern "C" int32_t smplb_fit_step(const smplb_model* m, const smplb_fit_io* io, void* stream) {
  if (!m || !io) return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: null model or io");
  if (io->B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_fit_step: negative batch");
  if (io->B == 0) return SMPLB_OK;
  if (!io->beta || !io->theta || !io->cam_pos || !io->cam_rot || !io->gt_px || !io->px || !io->g_px ||
      !io->g_beta || !io->g_theta || !io->m_theta || !io->v_theta || !io->loss || !io->scratch)
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: null pointer");
  if ((io->m_beta == nullptr) != (io->v_beta == nullptr))
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: m_beta and v_beta go together");
  if (io->step < 1 || io->n_total < (int64_t)io->B * NJ * 2)
    return fail(SMPLB_E_INVALID_ARG, "smplb_fit_step: step >= 1 and n_total >= B*48 required");
  cudaStream_t st = (cudaStream_t)stream;
  const int B = io->B;
  PrepParams p;
  p.B = B;
  p.beta = io->beta;
  p.theta = io->theta;
  p.Jbar = m->Jbar;
  p.JS = m->JS;
  p.Ximg = nullptr;
  p.joints = nullptr;
  p.Rs = nullptr;
  p.A = nullptr;
  p.Aimg = nullptr;
  p.cam_pos = io->cam_pos;
  p.cam_rot = io->cam_rot;
  p.focal = io->focal;
  p.uv = nullptr;
  p.px = io->px;
  p.cfg = io->cfg.cam;
  p.tree = m->tree;
  int32_t rc = launch_pose_prep(p, st);
  if (rc != SMPLB_OK) return rc;
  const int n = B * NJ * 2;
  const int nblocks = (n + RG_THREADS - 1) / RG_THREADS;
  double* partial = (double*)io->scratch;
  const double inv_n = 1.0 / (double)io->n_total;
  reproj_grad_kernel<<<nblocks, RG_THREADS, 0, st>>>(n, io->px, io->gt_px, (float)(2.0 * io->cfg.reproj_scale * inv_n),
                                                    io->g_px, partial);
  loss_finish_kernel<<<1, 256, 0, st>>>(nblocks, partial, io->cfg.reproj_scale, inv_n, io->cfg.reproj_cap, io->loss);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  rc = launch_joint_bwd_px(m, B, io->beta, io->theta, io->g_px, io->cam_pos, io->cam_rot, io->focal, io->cfg.cam,
                           io->g_beta, io->g_theta, st);
  if (rc != SMPLB_OK) return rc;
  const float lr_t = io->lr * sqrtf(1.0f - powf(io->beta2, (float)io->step)) / (1.0f - powf(io->beta1, (float)io->step));
  const int64_t nt = (int64_t)B * NJ * 3;
  adam_gated_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(nt, io->theta, io->g_theta, io->m_theta, io->v_theta,
                                                                  io->loss + 2, lr_t, io->beta1, io->beta2, io->eps);
  if (io->m_beta) {
    const int64_t nb = (int64_t)B * NB;
    adam_gated_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>(nb, io->beta, io->g_beta, io->m_beta, io->v_beta,
                                                                    io->loss + 2, lr_t, io->beta1, io->beta2, io->eps);
  }
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
