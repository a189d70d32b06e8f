// Hand-written backward of the path (TF autodiff of batch_smpl.py:91-138 + project.py, driven by
// optimizer.minimize in pose_model_3d.py:393-400).  Adjoint derivation: SURVEY.md appendix A.
//
//   joint_bwd_kernel (K6)  warp per sample, lane = joint: projection adjoint, A -> G adjoint, reverse
//                          kinematic chain with warp shuffles, J = Jbar + JS beta adjoint, Rodrigues adjoint
//   vertex path            skinning adjoint (skin_bwd.cu) + transposed blend GEMM (blend_bwd.cu) feed K6 with
//                          g_A[B,24,12] and g_X[B,224] = [g_beta | g_pose_feature]
#include "geom.cuh"

namespace smplb {

struct JointBwdParams {
  int B;
  const float* beta;
  const float* theta;
  const float* Jbar;
  const float* JS;
  const float* JSt;        // smplb_model::JSt (joint-minor copy of Jbar / JS)
  const float* g_joints;   // [B,24,3] nullable
  const float* g_px;       // [B,24,2] nullable
  const float* cam_pos;
  const float* cam_rot;
  const float* focal;
  const float* g_A;        // [B,24,12] nullable (skinning adjoint)
  const float* g_X;        // [n_slices][B,KPAD] nullable (blend adjoint: 10 shape + 207 pose-feature columns)
  int n_slices;            // partial sums of the transposed blend GEMM, added here in slice order
  float* g_beta;
  float* g_theta;
  smplb_camcfg cfg;
  const int4* tree_tab;   // smplb_model::tree_tab: per lane {order, sub, up0, up1 | up2, up3, up4, -}
  int n_rounds;
  // fused fitting step (fit.cu): when gt_px is given the kernel forms the re-projection residual itself --
  // g_px = px_coef * (px - gt_px), px written to px_out, sum of squared residuals of the block to partial[blockIdx]
  // (fixed order) -- so pose_prep and reproj_grad need not run (this kernel recomputes the forward chain anyway)
  const float* gt_px;      // [B,24,2] nullable
  float px_coef;
  float* px_out;           // [B,24,2] nullable
  float* g_px_out;         // [B,24,2] nullable
  double* partial;         // [gridDim.x]
  // ... and, when the loss cap is off (the gate is 1 whatever the loss turns out to be), the TF1 Adam update of this
  // sample's theta (and beta) straight from the registers that hold the gradient: adam_m_theta != nullptr
  float* adam_theta;       // [B,72] updated in place (every lane has read its own theta before)
  float* adam_m_theta;
  float* adam_v_theta;
  float* adam_beta;        // [B,10] nullable
  float* adam_m_beta;
  float* adam_v_beta;
  const int* step_dev;     // t = *step_dev + 1 (incremented by loss_finish_kernel AFTER this kernel), else step_host
  int step_host;
  float lr, b1, b2, eps;
};

__device__ __forceinline__ float adam_lr_t_dev(float lr, float b1, float b2, int t) {
  return lr * sqrtf(1.0f - powf(b2, (float)t)) / (1.0f - powf(b1, (float)t));
}
__device__ __forceinline__ void adam_update(float* x, float* m, float* v, float g, float lr_t, float b1, float b2,
                                            float eps) {
  const float mi = b1 * *m + (1.0f - b1) * g;
  const float vi = b2 * *v + (1.0f - b2) * g * g;
  *m = mi;
  *v = vi;
  *x = *x - lr_t * mi / (sqrtf(vi) + eps);
}

constexpr int JB_WARPS = 4;

// One sample per warp.  Lane l holds joint order[l], the joints in depth-first PRE-order (Tree::order), so the subtree of
// a joint is the contiguous lane range [l, l + sub[l]).  That turns both chains into log-depth warp operations:
//   forward   G_i = L_root ... L_i with L_m = (R_m, J_m - J_parent(m)) by pointer jumping: round r composes every joint
//             with its 2^r-th ancestor (Tree::up[r]); ceil(log2(depth+1)) = 4 rounds for SMPL instead of 8 levels
//   backward  with the own gradients gGR_k, gGt_k of every joint (dL/dG_k when G_k is held independent), the chain rule
//             through G_k.R = G_i.R rel_k, G_k.t = G_i.R l_k + G_i.t for every descendant k of i collapses to
//               dL/dG_i.R = [ sum_k (gGR_k G_k.R^T + gGt_k G_k.t^T)  -  (sum_k gGt_k) G_i.t^T ] G_i.R ,  dL/dG_i.t = sum_k gGt_k
//             (k over the subtree of i, i included): 12 subtree sums per joint = range sums over contiguous lanes (a
//             doubling table of 5 rounds), instead of 8 levels of 45 shuffles
// Everything is fixed-order, so results are run-to-run identical.
__device__ __forceinline__ void joint_bwd_sample(const JointBwdParams& p, const int b, const int lane, double* sq_out) {
  const unsigned FULL = 0xffffffffu;
  const bool act = lane < NJ;
  const int4 tt0 = __ldg(p.tree_tab + lane * 2), tt1 = __ldg(p.tree_tab + lane * 2 + 1);
  const int j = tt0.x;          // joint of this lane (0 for idle lanes)
  const int sub = tt0.y;        // subtree size, itself included (1 for idle lanes)
  const int up[5] = {tt0.z, tt0.w, tt1.x, tt1.y, tt1.z};   // lane of the 2^r-th ancestor, -1 past the root
  const int par = up[0] >= 0 ? up[0] : 0;
  float lr_t = 0.f;
  if (p.adam_m_theta) lr_t = adam_lr_t_dev(p.lr, p.b1, p.b2, p.step_dev ? __ldg(p.step_dev) + 1 : p.step_host);

  // ---- recompute the forward quantities of this joint (pose_prep.cu) ----
  const float* th = p.theta + (size_t)b * (NJ * 3) + j * 3;
  const float tx = th[0], ty = th[1], tz = th[2];
  float R[9];
  rodrigues_smpl(tx, ty, tz, R);
  float beta[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) beta[k] = __ldg(p.beta + (size_t)b * NB + k);
  float J[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float acc = __ldg(p.JSt + (3 * NB + c) * 32 + j);
#pragma unroll
    for (int k = 0; k < NB; ++k) acc = fmaf(__ldg(p.JSt + (c * NB + k) * 32 + j), beta[k], acc);
    J[c] = acc;
  }
  float GR[9], Gt[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) GR[e] = R[e];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float pj = __shfl_sync(FULL, J[c], par);
    Gt[c] = up[0] >= 0 ? J[c] - pj : J[c];
  }
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    if (r < p.n_rounds) {
      const int a = up[r] >= 0 ? up[r] : 0;
      float aR[9], at[3];
#pragma unroll
      for (int e = 0; e < 9; ++e) aR[e] = __shfl_sync(FULL, GR[e], a);
#pragma unroll
      for (int c = 0; c < 3; ++c) at[c] = __shfl_sync(FULL, Gt[c], a);
      if (up[r] >= 0) {
        float nR[9], nt[3];
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
          for (int c = 0; c < 3; ++c)
            nR[rr * 3 + c] = aR[rr * 3 + 0] * GR[c] + aR[rr * 3 + 1] * GR[3 + c] + aR[rr * 3 + 2] * GR[6 + c];
          nt[rr] = aR[rr * 3 + 0] * Gt[0] + aR[rr * 3 + 1] * Gt[1] + aR[rr * 3 + 2] * Gt[2] + at[rr];
        }
#pragma unroll
        for (int e = 0; e < 9; ++e) GR[e] = nR[e];
#pragma unroll
        for (int c = 0; c < 3; ++c) Gt[c] = nt[c];
      }
    }
  }

  // ---- upstream gradients wrt J_transformed (posed joints) ----
  float gJt[3] = {0.f, 0.f, 0.f};
  if (act && p.g_joints) {
    const float* g = p.g_joints + ((size_t)b * NJ + j) * 3;
    gJt[0] = g[0]; gJt[1] = g[1]; gJt[2] = g[2];
  }
  const bool fused_px = p.gt_px != nullptr;
  float sq = 0.f;
  if (act && (p.g_px || fused_px)) {
    // px = uv / ss * H + off (pose_model_3d.py:229-231); camera held constant (:218-221)
    float2 gp = p.g_px ? reinterpret_cast<const float2*>(p.g_px)[(size_t)b * NJ + j] : make_float2(0.f, 0.f);
    const float sc = p.cfg.img_h / p.cfg.ss;
    const float* cp = p.cam_pos + (size_t)b * 3;
    const float* cr = p.cam_rot + (size_t)b * 3;
    const float f = p.focal ? __ldg(p.focal + b) : p.cfg.focal_const;
    float Rc[9], q[3], u, v, gq[3];
    rodrigues_cam(__ldg(cr), __ldg(cr + 1), __ldg(cr + 2), Rc);
    project_point(Rc, Gt[0], Gt[1], Gt[2], __ldg(cp), __ldg(cp + 1), __ldg(cp + 2), f, u, v, q);
    if (fused_px) {
      // the re-projection residual of the fitting step, formed here (pose_model_3d.py:229-235)
      float pxx, pxy;
      to_pixels(u, v, p.cfg, pxx, pxy);
      const float2 gt = reinterpret_cast<const float2*>(p.gt_px)[(size_t)b * NJ + j];
      const float dx = pxx - gt.x, dy = pxy - gt.y;
      gp = make_float2(p.px_coef * dx, p.px_coef * dy);
      sq = dx * dx + dy * dy;
      if (p.px_out) reinterpret_cast<float2*>(p.px_out)[(size_t)b * NJ + j] = make_float2(pxx, pxy);
      if (p.g_px_out) reinterpret_cast<float2*>(p.g_px_out)[(size_t)b * NJ + j] = gp;
    }
    project_bwd_q(q, f, gp.x * sc, gp.y * sc, gq);
#pragma unroll
    for (int r = 0; r < 3; ++r) gJt[r] += Rc[r * 3] * gq[0] + Rc[r * 3 + 1] * gq[1] + Rc[r * 3 + 2] * gq[2];
  }

  if (fused_px) {
    double sd = (double)sq;   // fixed-order warp tree -> the block's partial sum
    for (int o = 16; o > 0; o >>= 1) sd += __shfl_xor_sync(FULL, sd, o);
    if (lane == 0) *sq_out = sd;
  }

  // ---- own gradients: A.R = G.R, A.t = G.t - G.R J (batch_lbs.py:138-142), plus the posed-joint gradient on G.t ----
  float gJ[3];
  float S[12];   // the summands of the subtree sums: gGR_own G.R^T + gGt_own G.t^T (9) | gGt_own (3)
  float Fown[3], Qown[9];   // Qown = gGR_own G.R^T: the joint's own share of dL/dG_i.R G_i.R^T
  {
    float gAR[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, gAt[3] = {0.f, 0.f, 0.f};
    if (act && p.g_A) {
      const float4* g = reinterpret_cast<const float4*>(p.g_A + ((size_t)b * NJ + j) * 12);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float4 v = __ldg(g + r);
        gAR[r * 3] = v.x; gAR[r * 3 + 1] = v.y; gAR[r * 3 + 2] = v.z; gAt[r] = v.w;
      }
    }
    float gGR[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) gGR[r * 3 + c] = gAR[r * 3 + c] - gAt[r] * J[c];
      Fown[r] = act ? gAt[r] + gJt[r] : 0.f;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) gJ[c] = -(GR[c] * gAt[0] + GR[3 + c] * gAt[1] + GR[6 + c] * gAt[2]);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float v = gGR[r * 3] * GR[c * 3] + gGR[r * 3 + 1] * GR[c * 3 + 1] + gGR[r * 3 + 2] * GR[c * 3 + 2];
        Qown[r * 3 + c] = v;
        S[r * 3 + c] = act ? v + Fown[r] * Gt[c] : 0.f;
      }
      S[9 + r] = Fown[r];
    }
  }

  // ---- sums over the STRICT descendants = lanes [l + 1, l + sub): doubling table W_r[l] = S[l] + ... + S[l + 2^r - 1] built
  // in place; the range is cut into power-of-two blocks by the bits of its length.  No subtraction of prefixes: a leaf gets
  // an exact zero (as TF's autodiff does -- Adam would turn rounding noise on a zero gradient into steps of size lr) and
  // nothing cancels.
  float D[12];
#pragma unroll
  for (int e = 0; e < 12; ++e) D[e] = 0.f;
  {
    const int n = sub - 1;
    int pos = lane + 1;
#pragma unroll
    for (int r = 0; r < 5; ++r) {
      const bool take = (n >> r) & 1;
#pragma unroll
      for (int e = 0; e < 12; ++e) {
        const float blk = __shfl_sync(FULL, S[e], pos & 31);
        if (take) D[e] += blk;
      }
      if (take) pos += 1 << r;
      if (r < 4) {
#pragma unroll
        for (int e = 0; e < 12; ++e) S[e] += __shfl_down_sync(FULL, S[e], 1 << r);
      }
    }
  }

  // ---- dL/dG_i -> dL/dR_i and dL/dJ: G_i.R = G_p.R R_i ; G_i.t = G_p.R (J_i - J_p) + G_p.t (root: G_p = identity) ----
  float gR[9];
  {
    float pR[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) {
      const float v = __shfl_sync(FULL, GR[e], par);
      pR[e] = up[0] >= 0 ? v : ((e == 0 || e == 4 || e == 8) ? 1.f : 0.f);
    }
    float M[9], T[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        M[r * 3 + c] = Qown[r * 3 + c] + (D[r * 3 + c] - D[9 + r] * Gt[c]);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        T[r * 3 + c] = M[r * 3] * GR[c] + M[r * 3 + 1] * GR[3 + c] + M[r * 3 + 2] * GR[6 + c];   // dL/dG_i.R
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) gR[r * 3 + c] = pR[r] * T[c] + pR[3 + r] * T[3 + c] + pR[6 + r] * T[6 + c];   // G_p.R^T .
    // J_i enters G_i.t through G_p.R, and every child's G_c.t through -G_i.R
    const float gt0 = Fown[0] + D[9], gt1 = Fown[1] + D[10], gt2 = Fown[2] + D[11];   // dL/dG_i.t
#pragma unroll
    for (int c = 0; c < 3; ++c)
      gJ[c] += (pR[c] * gt0 + pR[3 + c] * gt1 + pR[6 + c] * gt2) -
               (GR[c] * D[9] + GR[3 + c] * D[10] + GR[6 + c] * D[11]);
  }

  // ---- pose feature adjoint (batch_smpl.py:109-110): phi = vec(R_1..23 - I) ----
  if (act && j >= 1 && p.g_X) {
    for (int sl = 0; sl < p.n_slices; ++sl) {
      const float* g = p.g_X + ((size_t)sl * p.B + b) * KPAD + NB + (j - 1) * 9;
#pragma unroll
      for (int e = 0; e < 9; ++e) gR[e] += __ldg(g + e);
    }
  }

  // ---- Rodrigues adjoint -> g_theta ----
  if (act) {
    float gth[3];
    rodrigues_smpl_bwd(tx, ty, tz, gR, gth);
    float* o = p.g_theta + (size_t)b * (NJ * 3) + j * 3;
    o[0] = gth[0]; o[1] = gth[1]; o[2] = gth[2];
    if (p.adam_m_theta) {
      const size_t at = (size_t)b * (NJ * 3) + j * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        adam_update(p.adam_theta + at + c, p.adam_m_theta + at + c, p.adam_v_theta + at + c, gth[c], lr_t, p.b1, p.b2, p.eps);
    }
  }

  // ---- J = Jbar + JS beta adjoint (+ shape-blend columns of g_X) -> g_beta ----
  float gb[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    float v = 0.f;
    if (act) {
#pragma unroll
      for (int c = 0; c < 3; ++c) v = fmaf(__ldg(p.JSt + (c * NB + k) * 32 + j), gJ[c], v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    gb[k] = v;
  }
  if (lane < NB) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < NB; ++k)
      if (k == lane) v = gb[k];
    if (p.g_X)
      for (int sl = 0; sl < p.n_slices; ++sl) v += __ldg(p.g_X + ((size_t)sl * p.B + b) * KPAD + lane);
    p.g_beta[(size_t)b * NB + lane] = v;
    if (p.adam_m_beta) {
      const size_t at = (size_t)b * NB + lane;
      adam_update(p.adam_beta + at, p.adam_m_beta + at, p.adam_v_beta + at, v, lr_t, p.b1, p.b2, p.eps);
    }
  }
}

__global__ void __launch_bounds__(JB_WARPS * 32, 7) joint_bwd_kernel(const JointBwdParams p) {
  __shared__ double sq_sum[JB_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * JB_WARPS + warp;
  pdl_wait();      // fitting step: the previous step's loss_finish_kernel (step counter, partials) has completed
  pdl_trigger();   // ... and this step's loss_finish_kernel may be scheduled now (it waits for this grid itself)
  if (p.gt_px) {
    // fused fitting step: every warp of the block takes part in the loss partial sum (no early exit)
    if (lane == 0) sq_sum[warp] = 0.0;
    __syncthreads();
  }
  if (b < p.B) joint_bwd_sample(p, b, lane, sq_sum + warp);
  if (p.gt_px) {
    __syncthreads();
    if (threadIdx.x == 0) {
      double sacc = 0.0;
      for (int w = 0; w < JB_WARPS; ++w) sacc += sq_sum[w];
      p.partial[blockIdx.x] = sacc;
    }
  }
}

// project() adjoint (project.py:7-48) wrt points and camera position / rotation.  Thread per camera
// group: loops over its P points so the per-camera sums have a fixed order.
__global__ void project_bwd_kernel(int N, int P, const float* __restrict__ pts, const float* __restrict__ cpos,
                                   const float* __restrict__ cang, const float* __restrict__ foc, int tiled,
                                   const float* __restrict__ g_uv, float* __restrict__ g_pts,
                                   float* __restrict__ g_cpos, float* __restrict__ g_cang,
                                   float* __restrict__ g_foc) {
  const int ncam = tiled ? N : N / P;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncam) return;
  const int np = tiled ? 1 : P;
  float Rc[9];
  const float ax = cang[3 * c], ay = cang[3 * c + 1], az = cang[3 * c + 2];
  rodrigues_cam(ax, ay, az, Rc);
  const float cx = cpos[3 * c], cy = cpos[3 * c + 1], cz = cpos[3 * c + 2], f = foc[c];
  float gc[3] = {0.f, 0.f, 0.f}, gRc[9], gf = 0.f;
#pragma unroll
  for (int e = 0; e < 9; ++e) gRc[e] = 0.f;
  for (int k = 0; k < np; ++k) {
    const int i = tiled ? c : c * P + k;
    float q[3], u, v, gq[3];
    project_point(Rc, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], cx, cy, cz, f, u, v, q);
    project_bwd_q(q, f, g_uv[2 * i], g_uv[2 * i + 1], gq);
    gf += (g_uv[2 * i] * q[0] + g_uv[2 * i + 1] * q[1]) / q[2];   // uv = f q_xy / q_z (project.py:35-46)
    const float d[3] = {pts[3 * i] - cx, pts[3 * i + 1] - cy, pts[3 * i + 2] - cz};
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float gp = Rc[r * 3] * gq[0] + Rc[r * 3 + 1] * gq[1] + Rc[r * 3 + 2] * gq[2];   // Rc gq
      if (g_pts) g_pts[3 * i + r] = gp;
      gc[r] -= gp;
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) gRc[r * 3 + cc] += d[r] * gq[cc];   // q_c = sum_r Rc[r][c] d_r
    }
  }
  if (g_cpos) {
    g_cpos[3 * c] = gc[0]; g_cpos[3 * c + 1] = gc[1]; g_cpos[3 * c + 2] = gc[2];
  }
  if (g_foc) g_foc[c] = gf;
  if (g_cang) {
    float gr[3];
    rodrigues_cam_bwd(ax, ay, az, gRc, gr);
    g_cang[3 * c] = gr[0]; g_cang[3 * c + 1] = gr[1]; g_cang[3 * c + 2] = gr[2];
  }
}

struct BwdScratch {
  float* g_A;     // [B,24,12]
  float* g_X;     // [n_slices][B,KPAD]
  float* g_img;   // bf16 split operand image of g_vposed (gimg16_off): [ceil(B/128)][BB_NKB][BB_ASTAGE x 16 bit]
  float* g_vtot;  // [B,NV3] g_verts + regressor^T g_kp (only touched when g_kp is given)
  size_t total;
};

static BwdScratch carve_bwd(void* base, int B) {
  BwdScratch w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* p = base ? (char*)base + off : nullptr;
    off += (bytes + 255) / 256 * 256;
    return (float*)p;
  };
  const size_t b = (size_t)(B > 0 ? B : 0);
  w.g_A = take(b * NJ * 12 * 4);
  w.g_X = take((size_t)blend_bwd_slices(B) * b * KPAD * 4);
  w.g_img = take(((b + BL_BM - 1) / BL_BM) * (size_t)BB_NKB * BB_ASTAGE * 2);
  w.g_vtot = take(b * NV3 * 4);
  w.total = off;
  return w;
}

size_t backward_scratch_bytes(int B) { return carve_bwd(nullptr, B).total; }

// The fused fitting step on the 2D re-projection loss (fit.cu): forward chain + projection + residual + backward chain in
// ONE per-sample kernel.  px / g_px are written for the caller; partial[i] = sum of squared pixel residuals of block i.
int fit_fused_blocks(int B) { return (B + JB_WARPS - 1) / JB_WARPS; }
int32_t launch_fit_fused(const smplb_model* m, int B, const float* beta, const float* theta, const float* gt_px,
                         float px_coef, const float* cam_pos, const float* cam_rot, const float* focal,
                         const smplb_camcfg& cfg, float* px, float* g_px, double* partial, float* g_beta, float* g_theta,
                         const FitAdam* adam, cudaStream_t st) {
  JointBwdParams p;
  p.adam_m_theta = nullptr;
  p.adam_m_beta = nullptr;
  if (adam) {
    p.adam_theta = adam->theta; p.adam_m_theta = adam->m_theta; p.adam_v_theta = adam->v_theta;
    p.adam_beta = adam->beta; p.adam_m_beta = adam->m_beta; p.adam_v_beta = adam->v_beta;
    p.step_dev = adam->step_dev; p.step_host = adam->step_host;
    p.lr = adam->lr; p.b1 = adam->b1; p.b2 = adam->b2; p.eps = adam->eps;
  }
  p.B = B;
  p.beta = beta;
  p.theta = theta;
  p.Jbar = m->Jbar;
  p.JS = m->JS;
  p.JSt = m->JSt;
  p.g_joints = nullptr;
  p.g_px = nullptr;
  p.cam_pos = cam_pos;
  p.cam_rot = cam_rot;
  p.focal = focal;
  p.g_A = nullptr;
  p.g_X = nullptr;
  p.n_slices = 0;
  p.g_beta = g_beta;
  p.g_theta = g_theta;
  p.cfg = cfg;
  p.tree_tab = m->tree_tab;
  p.n_rounds = m->tree.n_rounds;
  p.gt_px = gt_px;
  p.px_coef = px_coef;
  p.px_out = px;
  p.g_px_out = g_px;
  p.partial = partial;
  SMPLB_CUDA_CHECK(launch_pdl(joint_bwd_kernel, dim3(fit_fused_blocks(B)), dim3(JB_WARPS * 32), 0, st, p));
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_backward(const smplb_model* m, const smplb_backward_io* io, void* workspace,
                                  size_t workspace_bytes, int32_t flags, void* stream) {
  if (!m || !io) return fail(SMPLB_E_INVALID_ARG, "smplb_backward: null model or io");
  if (int32_t rc = check_current_device(m->device, "smplb_backward")) return rc;
  if (io->B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_backward: negative batch");
  if (io->B == 0) return SMPLB_OK;
  if (!io->beta || !io->theta || !io->g_beta || !io->g_theta)
    return fail(SMPLB_E_INVALID_ARG, "smplb_backward: beta/theta/g_beta/g_theta are required");
  if (io->g_px && (!io->cam_pos || !io->cam_rot))
    return fail(SMPLB_E_INVALID_ARG, "smplb_backward: g_px needs the camera of the forward call");
  const bool vertex_path = io->g_verts || io->g_kp;
  cudaStream_t st = (cudaStream_t)stream;
  const int B = io->B;
  float *g_A = nullptr, *g_X = nullptr;
  int n_slices = 0;
  if (vertex_path) {
    if (!(flags & SMPLB_F_KEEP_VPOSED) || (flags & SMPLB_F_NO_SKIN))
      return fail(SMPLB_E_INVALID_ARG, "smplb_backward: vertex gradients need the workspace of a forward call made with SMPLB_F_KEEP_VPOSED");
    if (!workspace || workspace_bytes < smplb_workspace_bytes(m, B, flags))
      return fail(SMPLB_E_INVALID_ARG, "smplb_backward: workspace too small (see smplb_workspace_bytes)");
    Workspace w = carve(workspace, B, flags);
    BwdScratch s = carve_bwd(w.bwd, B);
    const float* gy = io->g_verts;
    int32_t rc;
    if (io->g_kp) {   // adjoint of the keypoint regression folded into the vertex gradient
      rc = launch_gverts_total(m, B, io->g_verts, io->g_kp, s.g_vtot, st);
      if (rc != SMPLB_OK) return rc;
      gy = s.g_vtot;
    }
    timing_mark_bwd(0, st);
    // g_A = sum_v w (g_y (x) [v_posed;1]): tcgen05 reduction; SMPLB_F_SKIN_SIMT selects the CUDA-core verification kernel
    rc = (flags & SMPLB_F_SKIN_SIMT) ? launch_skin_bwd_gA(m, B, gy, w.vposed, s.g_A, st)
                                     : launch_skin_bwd_gA_tc(m, B, gy, w.vposed, s.g_A, st);
    if (rc != SMPLB_OK) return rc;
    timing_mark_bwd(1, st);
    rc = launch_skin_bwd_vposed(m, B, gy, w.Aimg, s.g_img, st);             // g_vp = T^T g_y  (operand image)
    if (rc != SMPLB_OK) return rc;
    timing_mark_bwd(2, st);
    rc = launch_blend_bwd(m, B, s.g_img, s.g_X, &n_slices, st);             // g_X  = g_vp . D^T
    if (rc != SMPLB_OK) return rc;
    timing_mark_bwd(3, st);
    g_A = s.g_A;
    g_X = s.g_X;
  }
  JointBwdParams p;
  p.B = B;
  p.beta = io->beta;
  p.theta = io->theta;
  p.Jbar = m->Jbar;
  p.JS = m->JS;
  p.JSt = m->JSt;
  p.g_joints = io->g_joints;
  p.g_px = io->g_px;
  p.cam_pos = io->cam_pos;
  p.cam_rot = io->cam_rot;
  p.focal = io->focal;
  p.g_A = g_A;
  p.g_X = g_X;
  p.n_slices = n_slices;
  p.g_beta = io->g_beta;
  p.g_theta = io->g_theta;
  p.cfg = io->cfg;
  p.tree_tab = m->tree_tab;
  p.n_rounds = m->tree.n_rounds;
  p.gt_px = nullptr; p.px_coef = 0.f; p.px_out = nullptr; p.g_px_out = nullptr; p.partial = nullptr;
  p.adam_m_theta = nullptr; p.adam_m_beta = nullptr;
  joint_bwd_kernel<<<(B + JB_WARPS - 1) / JB_WARPS, JB_WARPS * 32, 0, st>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  if (vertex_path) timing_mark_bwd(4, st);
  return SMPLB_OK;
}

extern "C" int32_t smplb_project_backward(int32_t N, int32_t P, const float* pts, const float* cams_pos,
                                          const float* cams_angles, const float* focal_lens, int32_t cam_tiled,
                                          const float* g_uv, float* g_pts, float* g_cams_pos, float* g_cams_angles,
                                          float* g_focal_lens, void* stream) {
  if (N < 0 || P < 1) return fail(SMPLB_E_BAD_SHAPE, "smplb_project_backward: N >= 0 and P >= 1 required");
  if (!cam_tiled && N % P != 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_project_backward: N must be a multiple of P");
  if (N == 0) return SMPLB_OK;
  if (!pts || !cams_pos || !cams_angles || !focal_lens || !g_uv)
    return fail(SMPLB_E_INVALID_ARG, "smplb_project_backward: null pointer");
  const int ncam = cam_tiled ? N : N / P;
  project_bwd_kernel<<<(ncam + 127) / 128, 128, 0, (cudaStream_t)stream>>>(N, P, pts, cams_pos, cams_angles,
                                                                           focal_lens, cam_tiled, g_uv, g_pts,
                                                                           g_cams_pos, g_cams_angles, g_focal_lens);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
