// K1 pose_prep: one warp per sample, lane = joint.
//   Rodrigues x24 (batch_lbs.py:38-57)  ->  207-d pose feature (batch_smpl.py:107-110) written as the
//   TF32 hi/lo split A-operand of the blend GEMM, beta in columns 0..9;
//   folded joint regression J = Jbar + JS beta (batch_smpl.py:91-101);
//   kinematic chain by depth level with warp shuffles (batch_lbs.py:84-144);
//   J_transformed, relative skinning transforms A[B,24,3,4];
//   optional perspective projection of the 24 posed joints + pixel rescale
//   (project.py:7-48, pose_model_3d.py:202-231).
// Also the small stand-alone operators of the reference API (rodrigues, global_rigid, project,
// orth_proj, reproj loss, Adam).
#include "geom.cuh"

namespace smplb {

constexpr int PREP_WARPS = 4;

__global__ void __launch_bounds__(PREP_WARPS * 32) pose_prep_kernel(const PrepParams p) {
  __shared__ float xs[PREP_WARPS][KPAD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * PREP_WARPS + warp;
  if (b >= p.B) return;
  const int j = lane < NJ ? lane : 0;  // lanes 24..31 shadow joint 0, never store
  const bool act = lane < NJ;

  const float* th = p.theta + (size_t)b * (NJ * 3) + j * 3;
  const float tx = th[0], ty = th[1], tz = th[2];
  float R[9];
  rodrigues_smpl(tx, ty, tz, R);

  // beta (broadcast loads) and folded joint regression
  float beta[NB];
#pragma unroll
  for (int k = 0; k < NB; ++k) beta[k] = __ldg(p.beta + (size_t)b * NB + k);
  if (p.status) {
    // the fp16 main parts of the split-operand GEMMs cannot carry |beta| > 65504 (theta only enters through rotations)
    bool bad_b = false;
#pragma unroll
    for (int k = 0; k < NB; ++k) bad_b |= !(fabsf(beta[k]) <= 65504.0f);
    const bool bad_t = !(fabsf(tx) <= 3.0e38f) || !(fabsf(ty) <= 3.0e38f) || !(fabsf(tz) <= 3.0e38f);
    const int f = (bad_b ? 1 : 0) | ((act && bad_t) ? 2 : 0);
    if (f && (lane == 0 || (f & 2))) atomicOr(p.status, f);
  }
  float J[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float acc = __ldg(p.Jbar + j * 3 + c);
#pragma unroll
    for (int k = 0; k < NB; ++k) acc = fmaf(__ldg(p.JS + (j * 3 + c) * NB + k), beta[k], acc);
    J[c] = acc;
  }

  // GEMM A-operand row: [beta(10) | vec(R_1..23 - I)(207) | 0(7)]
  if (p.Ximg) {
    float* x = xs[warp];
    if (lane < NB) x[lane] = beta[lane];
    if (lane < KPAD - KDIM) x[KDIM + lane] = 0.f;
    if (act && lane >= 1) {
      float* f = x + NB + (lane - 1) * 9;
#pragma unroll
      for (int e = 0; e < 9; ++e) f[e] = R[e] - ((e == 0 || e == 4 || e == 8) ? 1.0f : 0.0f);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < KPAD / 32; ++i) {
      const int k = lane + 32 * i;
      const float v = x[k];
      unsigned short* x16 = reinterpret_cast<unsigned short*>(p.Ximg);
      const unsigned short h = f16_rne_bits(v);
      x16[ximg16_off(b, k, 0)] = h;
      x16[ximg16_off(b, k, 1)] = f16_rne_bits(v - f16_bits_to_float(h));
      x16[ximg16_off(b, k, 2)] = bf16_rne_bits(v);
    }
  }
  if (p.Ximg2) {
    // operand image of the fused forward (ff_ximg_off): [Xh | Xl] per K16 step, sample-major inside the chunk half
    float* x = xs[warp];
    if (!p.Ximg) {
      if (lane < NB) x[lane] = beta[lane];
      if (lane < KPAD - KDIM) x[KDIM + lane] = 0.f;
      if (act && lane >= 1) {
        float* f = x + NB + (lane - 1) * 9;
#pragma unroll
        for (int e = 0; e < 9; ++e) f[e] = R[e] - ((e == 0 || e == 4 || e == 8) ? 1.0f : 0.0f);
      }
      __syncwarp();
    }
    unsigned short* x16 = reinterpret_cast<unsigned short*>(p.Ximg2);
#pragma unroll
    for (int i = 0; i < KPAD / 32; ++i) {
      const int k = lane + 32 * i;
      const float v = x[k];
      const unsigned short h = f16_rne_bits(v);
      x16[ff_ximg_off(b, k, 0)] = h;
      x16[ff_ximg_off(b, k, 1)] = f16_rne_bits(v - f16_bits_to_float(h));
    }
  }

  // kinematic chain: G = (GR, Gt).  Root: (R_0, J_0).  Level d: G_i = G_parent * [R_i | J_i - J_parent]
  float GR[9], Gt[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) GR[e] = R[e];
#pragma unroll
  for (int c = 0; c < 3; ++c) Gt[c] = J[c];
  const int par = p.tree.parent[j];
  const int dep = act ? p.tree.depth[j] : -1;
  for (int d = 1; d <= p.tree.max_depth; ++d) {
    float pR[9], pt[3], pJ[3];
#pragma unroll
    for (int e = 0; e < 9; ++e) pR[e] = __shfl_sync(0xffffffffu, GR[e], par);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      pt[c] = __shfl_sync(0xffffffffu, Gt[c], par);
      pJ[c] = __shfl_sync(0xffffffffu, J[c], par);
    }
    if (dep == d) {
      const float r0 = J[0] - pJ[0], r1 = J[1] - pJ[1], r2 = J[2] - pJ[2];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          GR[r * 3 + c] = pR[r * 3 + 0] * R[0 * 3 + c] + pR[r * 3 + 1] * R[1 * 3 + c] +
                          pR[r * 3 + 2] * R[2 * 3 + c];
        Gt[r] = pR[r * 3 + 0] * r0 + pR[r * 3 + 1] * r1 + pR[r * 3 + 2] * r2 + pt[r];
      }
    }
  }

  if (!act && p.Aimg2) {
    // pad joints 24..31 of the fused forward's A image: one 16-byte unit per (part, component)
    unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg2);
    for (int u = lane - NJ; u < 24; u += 32 - NJ)
      *reinterpret_cast<uint4*>(a16 + ff_aimg_off(b, NJ, u % 12, u / 12)) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (!act && p.Aimg) {
    // joints 24..31 of the padded K: the W columns are zero, these must be finite -> zero (workspace is not cleared).
    // Per (part, component e) the 8 pad joints are one 16-byte unit: 24 units per sample, spread over lanes 24..31.
    unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg);
    for (int u = lane - NJ; u < 24; u += 32 - NJ)
      *reinterpret_cast<uint4*>(a16 + aimg16_off(b, NJ, u % 12, u / 12)) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (act) {
    if (p.joints) {
      float* o = p.joints + ((size_t)b * NJ + j) * 3;
      o[0] = Gt[0]; o[1] = Gt[1]; o[2] = Gt[2];
    }
    if (p.Rs) {
      float* o = p.Rs + ((size_t)b * NJ + j) * 9;
#pragma unroll
      for (int e = 0; e < 9; ++e) o[e] = R[e];
    }
    if (p.A) {
      // A = G - [0 | G.R J] (batch_lbs.py:138-142): rows [R | t - R J]
      float4* o = reinterpret_cast<float4*>(p.A + ((size_t)b * NJ + j) * 12);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float tr = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
        o[r] = make_float4(GR[r * 3 + 0], GR[r * 3 + 1], GR[r * 3 + 2], tr);
      }
    }
    if (p.Aimg) {
      // skinning-GEMM B operand: component e of joint j of sample b in the 16-bit split format (common.cuh)
      unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float tr = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
        const float row[4] = {GR[r * 3 + 0], GR[r * 3 + 1], GR[r * 3 + 2], tr};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const unsigned short h = f16_rne_bits(row[c]);
          const float hf = f16_bits_to_float(h);
          a16[aimg16_off(b, j, r * 4 + c, 0)] = h;
          a16[aimg16_off(b, j, r * 4 + c, 1)] = f16_rne_bits(row[c] - hf);
        }
      }
    }
    if (p.Aimg2) {
      unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg2);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float tr = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
        const float row[4] = {GR[r * 3 + 0], GR[r * 3 + 1], GR[r * 3 + 2], tr};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const unsigned short h = f16_rne_bits(row[c]);
          a16[ff_aimg_off(b, j, r * 4 + c, 0)] = h;
          a16[ff_aimg_off(b, j, r * 4 + c, 1)] = f16_rne_bits(row[c] - f16_bits_to_float(h));
        }
      }
    }
    if (p.cam_pos) {
      const float* cp = p.cam_pos + (size_t)b * 3;
      const float* cr = p.cam_rot + (size_t)b * 3;
      const float f = p.focal ? __ldg(p.focal + b) : p.cfg.focal_const;
      float Rc[9], q[3], u, v;
      rodrigues_cam(__ldg(cr), __ldg(cr + 1), __ldg(cr + 2), Rc);
      project_point(Rc, Gt[0], Gt[1], Gt[2], __ldg(cp), __ldg(cp + 1), __ldg(cp + 2), f, u, v, q);
      if (p.uv) reinterpret_cast<float2*>(p.uv)[(size_t)b * NJ + j] = make_float2(u, v);
      if (p.px) {
        float x, y;
        to_pixels(u, v, p.cfg, x, y);
        reinterpret_cast<float2*>(p.px)[(size_t)b * NJ + j] = make_float2(x, y);
      }
    }
  }
}

int32_t launch_pose_prep(const PrepParams& p, cudaStream_t st) {
  if (p.B <= 0) return SMPLB_OK;
  const int grid = (p.B + PREP_WARPS - 1) / PREP_WARPS;
  pose_prep_kernel<<<grid, PREP_WARPS * 32, 0, st>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// ---------------------------------------------------------------------------------------------
// stand-alone operators of the reference API

__global__ void rodrigues_kernel(int N, const float* __restrict__ rv, float* __restrict__ out, int variant) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float R[9];
  if (variant == 0)
    rodrigues_smpl(rv[3 * i], rv[3 * i + 1], rv[3 * i + 2], R);
  else
    rodrigues_cam(rv[3 * i], rv[3 * i + 1], rv[3 * i + 2], R);
#pragma unroll
  for (int e = 0; e < 9; ++e) out[(size_t)i * 9 + e] = R[e];
}

// batch_global_rigid_transformation (batch_lbs.py:84-144), warp per sample, lane = joint.
__global__ void __launch_bounds__(128) global_rigid_kernel(int N, const float* __restrict__ Rs,
                                                           const float* __restrict__ Js, Tree tree,
                                                           int rotate_base, float* __restrict__ newJ,
                                                           float* __restrict__ A44) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + warp;
  if (b >= N) return;
  const int j = lane < NJ ? lane : 0;
  const bool act = lane < NJ;
  float R[9], J[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) R[e] = Rs[((size_t)b * NJ + j) * 9 + e];
#pragma unroll
  for (int c = 0; c < 3; ++c) J[c] = Js[((size_t)b * NJ + j) * 3 + c];
  if (rotate_base && j == 0) {  // Rs[:,0] @ diag(1,-1,-1)  (batch_lbs.py:102-106)
    R[1] = -R[1]; R[2] = -R[2]; R[4] = -R[4]; R[5] = -R[5]; R[7] = -R[7]; R[8] = -R[8];
  }
  float GR[9], Gt[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) GR[e] = R[e];
#pragma unroll
  for (int c = 0; c < 3; ++c) Gt[c] = J[c];
  const int par = tree.parent[j];
  const int dep = act ? tree.depth[j] : -1;
  for (int d = 1; d <= tree.max_depth; ++d) {
    float pR[9], pt[3], pJ[3];
#pragma unroll
    for (int e = 0; e < 9; ++e) pR[e] = __shfl_sync(0xffffffffu, GR[e], par);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      pt[c] = __shfl_sync(0xffffffffu, Gt[c], par);
      pJ[c] = __shfl_sync(0xffffffffu, J[c], par);
    }
    if (dep == d) {
      const float r0 = J[0] - pJ[0], r1 = J[1] - pJ[1], r2 = J[2] - pJ[2];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          GR[r * 3 + c] = pR[r * 3 + 0] * R[c] + pR[r * 3 + 1] * R[3 + c] + pR[r * 3 + 2] * R[6 + c];
        Gt[r] = pR[r * 3 + 0] * r0 + pR[r * 3 + 1] * r1 + pR[r * 3 + 2] * r2 + pt[r];
      }
    }
  }
  if (!act) return;
  if (newJ) {
    float* o = newJ + ((size_t)b * NJ + j) * 3;
    o[0] = Gt[0]; o[1] = Gt[1]; o[2] = Gt[2];
  }
  if (A44) {
    float* o = A44 + ((size_t)b * NJ + j) * 16;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      o[r * 4 + 0] = GR[r * 3 + 0];
      o[r * 4 + 1] = GR[r * 3 + 1];
      o[r * 4 + 2] = GR[r * 3 + 2];
      o[r * 4 + 3] = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
    }
    // bottom row of `results - init_bone`: [0,0,0,1] - [0,0,0,0]  (batch_lbs.py:138-142)
    o[12] = 0.f; o[13] = 0.f; o[14] = 0.f; o[15] = 1.f;
  }
}

__global__ void project_kernel(int N, int P, const float* __restrict__ pts, const float* __restrict__ cpos,
                               const float* __restrict__ cang, const float* __restrict__ foc, int tiled,
                               float* __restrict__ uv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int c = tiled ? i : i / P;
  float Rc[9], q[3], u, v;
  rodrigues_cam(cang[3 * c], cang[3 * c + 1], cang[3 * c + 2], Rc);
  project_point(Rc, pts[3 * i], pts[3 * i + 1], pts[3 * i + 2], cpos[3 * c], cpos[3 * c + 1],
                cpos[3 * c + 2], foc[c], u, v, q);
  reinterpret_cast<float2*>(uv)[i] = make_float2(u, v);
}

// batch_orth_proj_idrot (deps/tf_smpl/projection.py:10-26): camera[:,0] * (X[:,:,:2] + camera[:,1:])
__global__ void orth_proj_kernel(int N, int P, const float* __restrict__ X, const float* __restrict__ cam,
                                 float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * P) return;
  const int n = i / P;
  const float s = cam[3 * n], tx = cam[3 * n + 1], ty = cam[3 * n + 2];
  out[2 * (size_t)i] = s * (X[3 * (size_t)i] + tx);
  out[2 * (size_t)i + 1] = s * (X[3 * (size_t)i + 1] + ty);
}

// scale * mean((gt - pred)^2) over B*24*2 elements (pose_model_3d.py:233-237). Single block.
__global__ void __launch_bounds__(1024) reproj_loss_kernel(int n, const float* __restrict__ a,
                                                           const float* __restrict__ g, float scale,
                                                           float cap, float* __restrict__ out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float d = a[i] - g[i];
    acc += (double)d * d;
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) {
      const float l = scale * (float)(acc / (double)n);
      out[0] = l;
      out[1] = cap > 0.f ? fminf(cap, l) : l;
    }
  }
}

// tf.train.AdamOptimizer (pose_model_3d.py:393): lr_t = lr sqrt(1-b2^t)/(1-b1^t); x -= lr_t m/(sqrt(v)+eps)
__global__ void adam_kernel(int64_t n, float* __restrict__ x, const float* __restrict__ g,
                            float* __restrict__ m, float* __restrict__ v, float lr_t, float b1, float b2,
                            float eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float gi = g[i];
  const float mi = b1 * m[i] + (1.0f - b1) * gi;
  const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  x[i] = x[i] - lr_t * mi / (sqrtf(vi) + eps);
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_rodrigues(int32_t N, const float* rvecs, float* R_out, int32_t variant, void* stream) {
  if (N < 0 || (N > 0 && (!rvecs || !R_out))) return fail(SMPLB_E_INVALID_ARG, "smplb_rodrigues: bad argument");
  if (variant != 0 && variant != 1) return fail(SMPLB_E_INVALID_ARG, "smplb_rodrigues: variant must be 0 or 1");
  if (N == 0) return SMPLB_OK;
  rodrigues_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(N, rvecs, R_out, variant);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_global_rigid(int32_t N, const float* Rs, const float* Js, const int32_t* parents,
                                      int32_t rotate_base, float* new_J, float* A44, void* stream) {
  if (N < 0 || !parents || (N > 0 && (!Rs || !Js))) return fail(SMPLB_E_INVALID_ARG, "smplb_global_rigid: bad argument");
  Tree t;
  {
    int32_t rc = build_tree(parents, &t, "smplb_global_rigid");
    if (rc != SMPLB_OK) return rc;
  }
  if (N == 0) return SMPLB_OK;
  global_rigid_kernel<<<(N + 3) / 4, 128, 0, (cudaStream_t)stream>>>(N, Rs, Js, t, rotate_base, new_J, A44);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_project(int32_t N, int32_t P, const float* pts, const float* cams_pos,
                                 const float* cams_angles, const float* focal_lens, int32_t cam_tiled,
                                 float* uv, void* stream) {
  if (N < 0 || P < 1) return fail(SMPLB_E_BAD_SHAPE, "smplb_project: N >= 0 and P >= 1 required");
  if (!cam_tiled && N % P != 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_project: N must be a multiple of P");
  if (N > 0 && (!pts || !cams_pos || !cams_angles || !focal_lens || !uv))
    return fail(SMPLB_E_INVALID_ARG, "smplb_project: null pointer");
  if (N == 0) return SMPLB_OK;
  project_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(N, P, pts, cams_pos, cams_angles,
                                                                    focal_lens, cam_tiled, uv);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_orth_proj(int32_t N, int32_t P, const float* X, const float* camera, float* out,
                                   void* stream) {
  if (N < 0 || P < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_orth_proj: negative size");
  if ((int64_t)N * P == 0) return SMPLB_OK;
  if (!X || !camera || !out) return fail(SMPLB_E_INVALID_ARG, "smplb_orth_proj: null pointer");
  orth_proj_kernel<<<(N * P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(N, P, X, camera, out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_reproj_loss(int32_t B, const float* px_pred, const float* px_gt,
                                     const smplb_losscfg* cfg, float* loss_out, void* stream) {
  if (B <= 0 || !px_pred || !px_gt || !cfg || !loss_out) return fail(SMPLB_E_INVALID_ARG, "smplb_reproj_loss: bad argument");
  reproj_loss_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(B * NJ * 2, px_pred, px_gt, cfg->reproj_scale,
                                                          cfg->reproj_cap, loss_out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_adam_step(int64_t n, float* x, const float* g, float* m, float* v, int32_t t,
                                   float lr, float beta1, float beta2, float eps, void* stream) {
  if (n < 0 || t < 1) return fail(SMPLB_E_INVALID_ARG, "smplb_adam_step: n >= 0 and t >= 1 required");
  if (n == 0) return SMPLB_OK;
  if (!x || !g || !m || !v) return fail(SMPLB_E_INVALID_ARG, "smplb_adam_step: null pointer");
  const float lr_t = lr * sqrtf(1.0f - powf(beta2, (float)t)) / (1.0f - powf(beta1, (float)t));
  adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, x, g, m, v, lr_t, beta1,
                                                                              beta2, eps);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
