// K1 pose_prep: one warp per sample, lane = joint.
//   Rodrigues x24 (batch_lbs.py:38-57)  ->  207-d pose feature (batch_smpl.py:107-110) written as the
//   f16 / f16-residual split A-operand image of the blend GEMM (ximg16_off), beta in columns 0..9;
//   folded joint regression J = Jbar + JS beta (batch_smpl.py:91-101);
//   kinematic chain by pointer jumping with warp shuffles (batch_lbs.py:84-144);
//   J_transformed, relative skinning transforms A[B,24,3,4];
//   optional perspective projection of the 24 posed joints + pixel rescale
//   (project.py:7-48, pose_model_3d.py:202-231).
// Also the small stand-alone operators of the reference API (rodrigues, global_rigid, project,
// orth_proj, reproj loss, Adam).
#include "geom.cuh"

namespace smplb {

constexpr int PREP_WARPS = 8;          // samples per block: 8 x 16 bytes = one full line of the X operand image per k-chunk
constexpr int PREP_XPITCH = KPAD + 4;  // X rows in shared memory (pitch 228: the 8 samples' float4 reads hit distinct banks)
constexpr int PREP_STAGE = 2 * 4 * 13 * 8 / 2;   // floats per warp: the A operand image of one sample (1536 B, each of its 8
                                                 // (part, chunk) blocks padded by one 16-byte unit against bank conflicts) >= A (1152 B)
static_assert(PREP_STAGE >= NJ * 12, "staging buffer holds the largest per-joint output");
static_assert(BL_BM % PREP_WARPS == 0 && KPAD % 8 == 0, "a block's samples share one X tile");

// Per-joint outputs of one sample ([24][n] floats, n * 96 bytes per sample) leave through shared memory: every lane
// writes its joint's n floats, then the warp stores the sample's block with 16-byte coalesced stores -- lane = joint
// stores straight to global memory touch 24 sectors per instruction (9 instructions for Rs: 243 sector writes for 27
// sectors of data), and the store path, not the arithmetic, was what the kernel waited on.
template <int N>
__device__ __forceinline__ void store_staged(float* stage, float* dst, const float (&v)[N], int lane, bool act) {
  __syncwarp();
  if (act) {
    if (N % 4 == 0) {   // 16-byte stores: a 48-byte lane stride is conflict-free per quarter warp, 4-byte stores are 4-way
#pragma unroll
      for (int e = 0; e < N / 4; ++e)
        reinterpret_cast<float4*>(stage + lane * N)[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
    } else {
#pragma unroll
      for (int e = 0; e < N; ++e) stage[lane * N + e] = v[e];
    }
  }
  __syncwarp();
  constexpr int NQ = NJ * N / 4;   // 16-byte units of the block
  for (int q = lane; q < NQ; q += 32)
    reinterpret_cast<float4*>(dst)[q] = reinterpret_cast<const float4*>(stage)[q];
}

__global__ void __launch_bounds__(PREP_WARPS * 32) pose_prep_kernel(const PrepParams p) {
  __shared__ __align__(16) float xrow[PREP_WARPS][PREP_XPITCH];   // [beta | pose feature | 0] of the block's samples
  __shared__ __align__(16) float xs[PREP_WARPS][PREP_STAGE];      // per-warp staging of the outputs
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b0 = blockIdx.x * PREP_WARPS;
  const bool valid = b0 + warp < p.B;
  const int b = valid ? b0 + warp : p.B - 1;   // warps past the batch redo the last sample and store nothing
  const int j = lane < NJ ? lane : 0;  // lanes 24..31 shadow joint 0, never store
  const bool act = lane < NJ && valid;

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
    const int f = (bad_b ? 1 : 0) | ((lane < NJ && bad_t) ? 2 : 0);
    if (valid && f && (lane == 0 || (f & 2))) atomicOr(p.status, f);
  }
  float J[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float acc = __ldg(p.JSt + (3 * NB + c) * 32 + j);
#pragma unroll
    for (int k = 0; k < NB; ++k) acc = fmaf(__ldg(p.JSt + (c * NB + k) * 32 + j), beta[k], acc);
    J[c] = acc;
  }

  // GEMM A-operand row: [beta(10) | vec(R_1..23 - I)(207) | 0(7)]
  {
    float* x = xrow[warp];
    if (lane < NB) x[lane] = beta[lane];
    if (lane < KPAD - KDIM) x[KDIM + lane] = 0.f;
    if (lane >= 1 && lane < NJ) {
      float* f = x + NB + (lane - 1) * 9;
#pragma unroll
      for (int e = 0; e < 9; ++e) f[e] = R[e] - ((e == 0 || e == 4 || e == 8) ? 1.0f : 0.0f);
    }
  }
  __syncthreads();
  if (p.Ximg) {
    // 16-bit split image (ximg16_off): for one k-chunk of 8 the samples of a tile are 16 bytes apart, so thread
    // (chunk q, sample s) of the block converts 8 values and the 8 samples of a chunk leave as one 128-byte line per part
    const int s8 = threadIdx.x & 7, q = threadIdx.x >> 3;
    if (q < KPAD / 8 && b0 + s8 < p.B) {
      const float4 v0 = *reinterpret_cast<const float4*>(&xrow[s8][q * 8]);
      const float4 v1 = *reinterpret_cast<const float4*>(&xrow[s8][q * 8 + 4]);
      const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
      uint32_t h[4], l[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const unsigned short ha = f16_rne_bits(v[2 * i]), hb = f16_rne_bits(v[2 * i + 1]);
        h[i] = (uint32_t)ha | ((uint32_t)hb << 16);
        l[i] = (uint32_t)f16_rne_bits(v[2 * i] - f16_bits_to_float(ha)) |
               ((uint32_t)f16_rne_bits(v[2 * i + 1] - f16_bits_to_float(hb)) << 16);
      }
      unsigned short* x16 = reinterpret_cast<unsigned short*>(p.Ximg);
      *reinterpret_cast<uint4*>(x16 + ximg16_off(b0 + s8, q * 8, 0)) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(x16 + ximg16_off(b0 + s8, q * 8, 1)) = make_uint4(l[0], l[1], l[2], l[3]);
    }
  }
  if (p.Ximg2 && valid) {
    // operand image of the fused forward (ff_ximg_off): [Xh | Xl] per K16 step, sample-major inside the chunk half
    const float* x = xrow[warp];
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
  if (!valid) return;   // no block-wide barrier below

  // kinematic chain G_i = L_root ... L_i with L_m = (R_m, J_m - J_parent(m)) (batch_lbs.py:122-136) by pointer jumping:
  // round r composes every joint with its 2^r-th ancestor -- ceil(log2(depth + 1)) = 4 rounds for SMPL instead of one
  // pass per tree level (8)
  float GR[9], Gt[3];
  {
    const int4 t0 = __ldg(p.tree_tab + 64 + lane * 2), t1 = __ldg(p.tree_tab + 64 + lane * 2 + 1);
    const int up[5] = {t0.z, t0.w, t1.x, t1.y, t1.z};
    const int par = up[0] >= 0 ? up[0] : 0;
#pragma unroll
    for (int e = 0; e < 9; ++e) GR[e] = R[e];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float pj = __shfl_sync(0xffffffffu, J[c], par);
      Gt[c] = up[0] >= 0 ? J[c] - pj : J[c];
    }
#pragma unroll
    for (int r = 0; r < 5; ++r) {
      if (r < p.n_rounds) {
        const int a = up[r] >= 0 ? up[r] : 0;
        float aR[9], at[3];
#pragma unroll
        for (int e = 0; e < 9; ++e) aR[e] = __shfl_sync(0xffffffffu, GR[e], a);
#pragma unroll
        for (int c = 0; c < 3; ++c) at[c] = __shfl_sync(0xffffffffu, Gt[c], a);
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
  }

  if (!act && p.Aimg2) {
    // pad joints 24..31 of the fused forward's A image: one 16-byte unit per (part, component)
    unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg2);
    for (int u = lane - NJ; u < 24; u += 32 - NJ)
      *reinterpret_cast<uint4*>(a16 + ff_aimg_off(b, NJ, u % 12, u / 12)) = make_uint4(0u, 0u, 0u, 0u);
  }
  if (p.Aimg) {
    // skinning-GEMM B operand (aimg16_off): per part and chunk of 8 joints one sample owns 12 x 16 = 192 contiguous
    // bytes.  Every lane (joints 24..31 of the padded K are zero: the W columns are zero, the image must be finite)
    // puts its 2 x 12 halves into the staged [part][chunk][e][8] block; the warp then stores 96 16-byte units.
    unsigned short* st16 = reinterpret_cast<unsigned short*>(xs[warp]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float tr = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
      const float row[4] = {GR[r * 3 + 0], GR[r * 3 + 1], GR[r * 3 + 2], tr};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float val = lane < NJ ? row[c] : 0.f;
        const unsigned short h = f16_rne_bits(val);
        st16[(((lane >> 3)) * 13 + r * 4 + c) * 8 + (lane & 7)] = h;
        st16[((4 + (lane >> 3)) * 13 + r * 4 + c) * 8 + (lane & 7)] = f16_rne_bits(val - f16_bits_to_float(h));
      }
    }
    __syncwarp();
    unsigned short* a16 = reinterpret_cast<unsigned short*>(p.Aimg);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int u = lane + 32 * i, part = u / 48, chunk = (u % 48) / 12, e = u % 12;
      *reinterpret_cast<uint4*>(a16 + aimg16_off(b, chunk * 8, e, part)) =
          reinterpret_cast<const uint4*>(st16)[(part * 4 + chunk) * 13 + e];
    }
  }
  float uvv[2] = {0.f, 0.f}, pxv[2] = {0.f, 0.f};
  if (act) {
    if (p.joints && !p.staged) {
      float* o = p.joints + ((size_t)b * NJ + j) * 3;
      o[0] = Gt[0]; o[1] = Gt[1]; o[2] = Gt[2];
    }
    if (p.Rs && !p.staged) {
      float* o = p.Rs + ((size_t)b * NJ + j) * 9;
#pragma unroll
      for (int e = 0; e < 9; ++e) o[e] = R[e];
    }
    if (p.A && !p.staged) {
      // A = G - [0 | G.R J] (batch_lbs.py:138-142): rows [R | t - R J]
      float4* o = reinterpret_cast<float4*>(p.A + ((size_t)b * NJ + j) * 12);
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float tr = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
        o[r] = make_float4(GR[r * 3 + 0], GR[r * 3 + 1], GR[r * 3 + 2], tr);
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
      uvv[0] = u; uvv[1] = v;
      to_pixels(u, v, p.cfg, pxv[0], pxv[1]);
      if (!p.staged) {
        if (p.uv) reinterpret_cast<float2*>(p.uv)[(size_t)b * NJ + j] = make_float2(u, v);
        if (p.px) reinterpret_cast<float2*>(p.px)[(size_t)b * NJ + j] = make_float2(pxv[0], pxv[1]);
      }
    }
  }
  if (p.staged) {   // uniform over the warp
    float* stage = xs[warp];
    if (p.joints) store_staged(stage, p.joints + (size_t)b * NJ * 3, Gt, lane, act);
    if (p.Rs) store_staged(stage, p.Rs + (size_t)b * NJ * 9, R, lane, act);
    if (p.A) {
      // A = G - [0 | G.R J] (batch_lbs.py:138-142): rows [R | t - R J]
      float a[12];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        a[r * 4 + 0] = GR[r * 3 + 0]; a[r * 4 + 1] = GR[r * 3 + 1]; a[r * 4 + 2] = GR[r * 3 + 2];
        a[r * 4 + 3] = Gt[r] - (GR[r * 3 + 0] * J[0] + GR[r * 3 + 1] * J[1] + GR[r * 3 + 2] * J[2]);
      }
      store_staged(stage, p.A + (size_t)b * NJ * 12, a, lane, act);
    }
    if (p.cam_pos && p.uv) store_staged(stage, p.uv + (size_t)b * NJ * 2, uvv, lane, act);
    if (p.cam_pos && p.px) store_staged(stage, p.px + (size_t)b * NJ * 2, pxv, lane, act);
  }
}

int32_t launch_pose_prep(const PrepParams& p, cudaStream_t st) {
  if (p.B <= 0) return SMPLB_OK;
  const int grid = (p.B + PREP_WARPS - 1) / PREP_WARPS;
  PrepParams q = p;
  // the staged 16-byte stores need 16-byte aligned outputs (every per-sample block is a multiple of 96 bytes)
  q.staged = (((uintptr_t)p.joints | (uintptr_t)p.Rs | (uintptr_t)p.A | (uintptr_t)p.uv | (uintptr_t)p.px) & 15) == 0;
  pose_prep_kernel<<<grid, PREP_WARPS * 32, 0, st>>>(q);
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
