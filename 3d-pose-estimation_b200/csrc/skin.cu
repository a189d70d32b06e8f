// K3 fused skinning pass (batch_smpl.py:120-138): linear blend skinning of v_posed with the per-sample
// relative joint transforms A, plus the sparse keypoint regression, in ONE pass over HBM:
//   read v_posed[b] (82.9 KB padded row) + A[b] (1152 B)  ->  write verts[b] (82 680 B) + kp[b].
// The reference materialises tile(W)[N,6890,24] and T[N,6890,4,4] (1.1 MB / sample); here W is the
// compressed <=4-nonzero table (20 B / vertex, L1/L2 resident) and T lives in registers.
//
// Mapping: one CTA = S consecutive samples x all 6890 vertices, walked in 27 tiles of 256 vertices;
// thread = vertex (its 4 weights / joint ids are loaded once per tile and reused for the S samples);
// A of the S samples sits in shared memory for the whole CTA; v_posed tiles arrive through a
// double-buffered cp.async pipeline; outputs are transposed through shared memory so that global
// stores are contiguous 8-byte-aligned runs (a verts row is only 8-byte aligned for odd b).
#include "common.cuh"

namespace smplb {

constexpr int SK_T = 256;                       // threads = vertices per tile
constexpr int SK_TILES = (NV + SK_T - 1) / SK_T;  // 27
constexpr int SK_TF = SK_T * 3;                 // floats per tile row (768)

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

template <int S>
struct SkinSmem {
  float4 A[S][NJ * 3];        // 3 float4 rows per joint
  float in[2][S][SK_TF];
  float out[S][SK_TF];
  float kp[S][NKMAX * 3];
};

template <int S>
__global__ void __launch_bounds__(SK_T) skin_kernel(int B, const float* __restrict__ vposed,
                                                    const float* __restrict__ A,
                                                    const uint32_t* __restrict__ skin_idx,
                                                    const float4* __restrict__ skin_w,
                                                    const int* __restrict__ kp_ptr,
                                                    const int* __restrict__ kp_col,
                                                    const float* __restrict__ kp_val, int NK,
                                                    float* __restrict__ verts, float* __restrict__ kp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SkinSmem<S>& sm = *reinterpret_cast<SkinSmem<S>*>(smem_raw);
  const int tid = threadIdx.x;
  const int b0 = blockIdx.x * S;
  const int ns = min(S, B - b0);

  auto prefetch = [&](int tile, int buf) {
    for (int i = tid; i < ns * (SK_TF / 4); i += SK_T) {
      const int s = i / (SK_TF / 4), q = i % (SK_TF / 4);
      cp_async16(&sm.in[buf][s][q * 4], vposed + (size_t)(b0 + s) * VPP + tile * SK_TF + q * 4);
    }
    cp_async_commit();
  };
  prefetch(0, 0);
  {
    const float4* Ag = reinterpret_cast<const float4*>(A + (size_t)b0 * NJ * 12);
    float4* As = &sm.A[0][0];
    for (int i = tid; i < ns * NJ * 3; i += SK_T) As[i] = Ag[i];
    float* kz = &sm.kp[0][0];
    for (int i = tid; i < S * NKMAX * 3; i += SK_T) kz[i] = 0.f;
  }

  for (int tile = 0; tile < SK_TILES; ++tile) {
    const int buf = tile & 1;
    const int v = tile * SK_T + tid;
    const bool valid = v < NV;
    const int vc = valid ? v : NV - 1;
    const uint32_t pk = __ldg(skin_idx + vc);
    const float4 w = __ldg(skin_w + vc);
    const int e0 = __ldg(kp_ptr + vc), e1 = valid ? __ldg(kp_ptr + vc + 1) : e0;
    if (tile + 1 < SK_TILES) {
      prefetch(tile + 1, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();  // tile `tile` landed; previous tile's out[] fully stored

    const int j0 = (pk & 0xff) * 3, j1 = ((pk >> 8) & 0xff) * 3, j2 = ((pk >> 16) & 0xff) * 3,
              j3 = (pk >> 24) * 3;
    if (valid) {
#pragma unroll
      for (int s = 0; s < S; ++s) {
        if (s < ns) {
          const float x = sm.in[buf][s][3 * tid], y = sm.in[buf][s][3 * tid + 1],
                      z = sm.in[buf][s][3 * tid + 2];
          const float4* As = sm.A[s];
          float o[3];
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const float4 a0 = As[j0 + r], a1 = As[j1 + r], a2 = As[j2 + r], a3 = As[j3 + r];
            // T row r = sum_k w_k A_{j_k} row r  (batch_smpl.py:125-127), then T . [v;1] (:128-130)
            const float t0 = w.x * a0.x + w.y * a1.x + w.z * a2.x + w.w * a3.x;
            const float t1 = w.x * a0.y + w.y * a1.y + w.z * a2.y + w.w * a3.y;
            const float t2 = w.x * a0.z + w.y * a1.z + w.z * a2.z + w.w * a3.z;
            const float t3 = w.x * a0.w + w.y * a1.w + w.z * a2.w + w.w * a3.w;
            o[r] = t0 * x + t1 * y + t2 * z + t3;
          }
          sm.out[s][3 * tid] = o[0];
          sm.out[s][3 * tid + 1] = o[1];
          sm.out[s][3 * tid + 2] = o[2];
          for (int e = e0; e < e1; ++e) {
            const int k = __ldg(kp_col + e);
            const float r = __ldg(kp_val + e);
            atomicAdd(&sm.kp[s][k * 3 + 0], r * o[0]);
            atomicAdd(&sm.kp[s][k * 3 + 1], r * o[1]);
            atomicAdd(&sm.kp[s][k * 3 + 2], r * o[2]);
          }
        }
      }
    }
    __syncthreads();
    if (verts) {
      const int nvalid = min(SK_TF, NV3 - tile * SK_TF);  // 768, or 702 on the last tile (even)
      for (int i = tid; i < ns * (SK_TF / 2); i += SK_T) {
        const int s = i / (SK_TF / 2), q = i % (SK_TF / 2);
        if (2 * q < nvalid) {
          const float2 val = *reinterpret_cast<const float2*>(&sm.out[s][2 * q]);
          float* dst = verts + (size_t)(b0 + s) * NV3 + tile * SK_TF + 2 * q;
          __stcs(reinterpret_cast<float2*>(dst), val);
        }
      }
    }
  }
  __syncthreads();
  if (kp) {
    for (int i = tid; i < ns * NK * 3; i += SK_T) {
      const int s = i / (NK * 3), r = i % (NK * 3);
      kp[(size_t)(b0 + s) * NK * 3 + r] = sm.kp[s][r];
    }
  }
}

template <int S>
static int32_t launch_skin_s(const smplb_model* m, int B, const float* vposed, const float* A, float* verts,
                             float* kp, cudaStream_t st) {
  const size_t smem = sizeof(SkinSmem<S>);
  static bool attr_done[64] = {};
  if (!attr_done[m->device & 63]) {
    SMPLB_CUDA_CHECK(cudaFuncSetAttribute(skin_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done[m->device & 63] = true;
  }
  const int grid = (B + S - 1) / S;
  skin_kernel<S><<<grid, SK_T, smem, st>>>(B, vposed, A, m->skin_idx, m->skin_w, m->kp_ptr, m->kp_col,
                                           m->kp_val, m->NK, verts, kp);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

int32_t launch_skin(const smplb_model* m, int B, const float* vposed, const float* A, float* verts, float* kp,
                    cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  // enough CTAs to fill 148 SMs a few times over; larger S amortises the per-tile weight fetch
  if (B >= 148 * 4 * 8) return launch_skin_s<8>(m, B, vposed, A, verts, kp, st);
  if (B >= 148 * 4 * 4) return launch_skin_s<4>(m, B, vposed, A, verts, kp, st);
  if (B >= 148 * 2 * 2) return launch_skin_s<2>(m, B, vposed, A, verts, kp, st);
  return launch_skin_s<1>(m, B, vposed, A, verts, kp, st);
}

}  // namespace smplb
