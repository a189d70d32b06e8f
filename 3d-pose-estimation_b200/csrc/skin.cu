// (1) CUDA-core skinning (verification path for skin_tc.cu, flag SMPLB_F_SKIN_SIMT): the same sums with the
//     compressed <=4-nonzero weight table, thread = vertex, A[b] of the sample in shared memory.
// (2) Keypoint regression joints = joint_regressor^T . verts (batch_smpl.py:134-138): the regressor is sparse
//     (cocoplus: ~32 vertices per keypoint), so this is a gather over the freshly written verts with a fixed
//     summation order (deterministic, no atomics).
#include "common.cuh"

namespace smplb {

__global__ void __launch_bounds__(SK_VT) skin_simt_kernel(int B, const float* __restrict__ vposed,
                                                          const float* __restrict__ A,
                                                          const uint32_t* __restrict__ skin_idx,
                                                          const float4* __restrict__ skin_w,
                                                          float* __restrict__ verts) {
  __shared__ float4 As[NJ * 3];
  const int b = blockIdx.x, vt = blockIdx.y, l = threadIdx.x;
  if (l < NJ * 3) As[l] = __ldg(reinterpret_cast<const float4*>(A + (size_t)b * NJ * 12) + l);
  __syncthreads();
  const int v = vt * SK_VT + l;
  if (v >= NV) return;
  const float* vp = vposed + (size_t)b * VPP + (size_t)vt * (3 * SK_VT);
  const float x = vp[l], y = vp[SK_VT + l], z = vp[2 * SK_VT + l];
  const uint32_t pk = __ldg(skin_idx + v);
  const float4 w = __ldg(skin_w + v);
  const int j0 = (pk & 0xff) * 3, j1 = ((pk >> 8) & 0xff) * 3, j2 = ((pk >> 16) & 0xff) * 3, j3 = (pk >> 24) * 3;
  float* o = verts + (size_t)b * NV3 + 3 * (size_t)v;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float4 a0 = As[j0 + r], a1 = As[j1 + r], a2 = As[j2 + r], a3 = As[j3 + r];
    const float t0 = w.x * a0.x + w.y * a1.x + w.z * a2.x + w.w * a3.x;
    const float t1 = w.x * a0.y + w.y * a1.y + w.z * a2.y + w.w * a3.y;
    const float t2 = w.x * a0.z + w.y * a1.z + w.z * a2.z + w.w * a3.z;
    const float t3 = w.x * a0.w + w.y * a1.w + w.z * a2.w + w.w * a3.w;
    o[r] = t0 * x + t1 * y + t2 * z + t3;
  }
}

int32_t launch_skin_simt(const smplb_model* m, int B, const float* vposed, const float* A, float* verts,
                         cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  if (!m->has_compressed)
    return fail(SMPLB_E_UNSUPPORTED, "SMPLB_F_SKIN_SIMT needs <= 4 non-zero LBS weights per vertex");
  dim3 grid(B, SK_NVT);
  skin_simt_kernel<<<grid, SK_VT, 0, st>>>(B, vposed, A, m->skin_idx, m->skin_w, verts);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// warp = sample; for each keypoint the lanes stride over its regressor entries, then a fixed-order
// butterfly reduction.
// Source rows: verts [B][NV3] with entry offset 3*vertex (mul = 3), or the compact keypoint-vertex rows
// [B][kpv_pitch] written by the skinning epilogue with entry offset = compact slot (mul = 1).  Same entry
// order either way, so both give bit-identical keypoints.
__global__ void __launch_bounds__(128) keypoints_kernel(int B, int NK, const float* __restrict__ verts, int pitch, int mul,
                                                        const int* __restrict__ kp_ptr,
                                                        const int* __restrict__ kp_vtx,
                                                        const float* __restrict__ kp_val,
                                                        float* __restrict__ kp) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 4 + warp;
  if (b >= B) return;
  const float* vb = verts + (size_t)b * pitch;

  for (int k = 0; k < NK; ++k) {
    const int e0 = __ldg(kp_ptr + k), e1 = __ldg(kp_ptr + k + 1);
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int e = e0 + lane; e < e1; e += 32) {
      const int v = __ldg(kp_vtx + e) * mul;
      const float r = __ldg(kp_val + e);
      sx = fmaf(r, vb[v], sx);
      sy = fmaf(r, vb[v + 1], sy);
      sz = fmaf(r, vb[v + 2], sz);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sx += __shfl_xor_sync(0xffffffffu, sx, o);
      sy += __shfl_xor_sync(0xffffffffu, sy, o);
      sz += __shfl_xor_sync(0xffffffffu, sz, o);
    }
    if (lane == 0) {
      float* o = kp + ((size_t)b * NK + k) * 3;
      o[0] = sx; o[1] = sy; o[2] = sz;
    }
  }
}

int32_t launch_keypoints(const smplb_model* m, int B, const float* verts, float* kp, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  keypoints_kernel<<<(B + 3) / 4, 128, 0, st>>>(B, m->NK, verts, NV3, 3, m->kp_ptr, m->kp_vtx, m->kp_val, kp);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// Keypoints from the compact rows kpv[group of 16 samples][slot][16] (skin_tc.cu): lane = sample (two groups per
// warp), every lane walks the regressor entries of its keypoints in order -> coalesced 64-byte row reads, uniform
// (broadcast) index/value loads, no cross-lane reduction; the warps of a block share 32 samples and split the
// keypoints -- one warp per keypoint (19 or 14 warps), so a small batch is not a chain of NK/4 regressor walks per
// warp (28 us at B = 1 with 4 warps; the kernel was a third of the B = 1024 step).  Same entry order as
// keypoints_kernel, plain sequential sum per keypoint.
__global__ void __launch_bounds__(640) keypoints_compact_kernel(int B, int NK, const float* __restrict__ kpv, int pitch,
                                                                const int* __restrict__ kp_ptr,
                                                                const int* __restrict__ kp_slot,
                                                                const float* __restrict__ kp_val,
                                                                float* __restrict__ kp) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 32 + lane;
  const bool ok = b < B;
  const int bc = ok ? b : B - 1;
  const float* row = kpv + (size_t)(bc / SK_S) * pitch * SK_S + (bc % SK_S);
  const int nw = blockDim.x >> 5;
  for (int k = warp; k < NK; k += nw) {
    const int e0 = __ldg(kp_ptr + k), e1 = __ldg(kp_ptr + k + 1);
    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll 16
    for (int e = e0; e < e1; ++e) {
      const float* p = row + (size_t)__ldg(kp_slot + e) * SK_S;
      const float r = __ldg(kp_val + e);
      sx = fmaf(r, __ldcs(p), sx);
      sy = fmaf(r, __ldcs(p + SK_S), sy);
      sz = fmaf(r, __ldcs(p + 2 * SK_S), sz);
    }
    if (ok) {
      float* o = kp + ((size_t)b * NK + k) * 3;
      o[0] = sx; o[1] = sy; o[2] = sz;
    }
  }
}

int32_t launch_keypoints_compact(const smplb_model* m, int B, const float* kpv, float* kp, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  const int nw = m->NK < 20 ? m->NK : 20;
  keypoints_compact_kernel<<<(B + 31) / 32, 32 * nw, 0, st>>>(B, m->NK, kpv, m->kpv_pitch, m->kp_ptr, m->kpc_slot, m->kp_val, kp);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
