// Skinning adjoint, reduction half: the gradient of the relative joint transforms
//     g_A[b,j] = sum_v weights[v,j] * g_verts[b,v] (x) [v_posed[b,v]; 1]        (24 x 3x4 per sample)
// reduces 6890 vertices to 288 numbers per sample.  (The other half, g_vposed = T^T g_verts, is the MODE 1
// instantiation of skin_tc_kernel.)
//
// Mapping: lane = sample (32 samples per CTA); 16 warps = 8 joint triples x entry parity: a warp keeps the 36
// accumulators of its three joints in registers for the whole sweep over the 54 vertex tiles and takes every
// second (vertex, weight) pair of each (tile, joint) list; the two parities are combined in a fixed order at the
// end.  The lists come from a table built at model creation (staged in smem per tile), so the summation order is
// fixed (deterministic, no atomics).
// Tiles of g_verts / v_posed rows are staged through shared memory with cp.async (double buffered); the row
// pitch is odd so that lane = sample reads are bank-conflict free.
#include "common.cuh"

namespace smplb {

constexpr int GA_SAMPLES = 32;
constexpr int GA_JW = 8;                    // joint triples
constexpr int GA_WARPS = 16;                // two warps per joint triple: even / odd table entries
constexpr int GA_PITCH = 3 * SK_VT + 1;     // 385 floats
constexpr int GA_TILE = 3 * SK_VT;          // 384 floats of a row per vertex tile

constexpr int GA_MAXENT = 1024;             // (vertex, weight) pairs of one vertex tile staged in smem (128 x <= 8 nnz)

struct GaSmem {
  float gy[2][GA_SAMPLES][GA_PITCH];
  float vp[2][GA_SAMPLES][GA_PITCH];
  float2 ent[2][GA_MAXENT];                 // the tile's table entries: no dependent global load in the inner loop
  int ptr[SK_NVT * (NJ + 1)];               // the whole ent_ptr table (54 x 25), loaded once
};

__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}

__global__ void __launch_bounds__(GA_WARPS * 32, 1)
skin_bwd_gA_kernel(int B, const float* __restrict__ g_verts, const float* __restrict__ vposed,
                   const int* __restrict__ ent_ptr, const float2* __restrict__ ent, float* __restrict__ g_A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GaSmem& sm = *reinterpret_cast<GaSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int jw = warp & (GA_JW - 1), hf = warp >> 3;   // joint triple, entry parity
  const int b0 = blockIdx.x * GA_SAMPLES;
  const int nrow = min(GA_SAMPLES, B - b0);

  // warp w stages rows w, w+8, ...; lanes stride over the 384 floats of the row (no index division); the tile's
  // (vertex, weight) table goes along (ent_ptr rows are contiguous: [t][0..24])
  auto prefetch = [&](int t, int buf) {
    const int nvalid = min(GA_TILE, NV3 - t * GA_TILE);
    for (int s = warp; s < nrow; s += GA_WARPS) {
      const float* gsrc = g_verts + (size_t)(b0 + s) * NV3 + (size_t)t * GA_TILE;
      const float* vsrc = vposed + (size_t)(b0 + s) * VPP + (size_t)t * GA_TILE;
#pragma unroll 4
      for (int o = lane; o < GA_TILE; o += 32) {
        if (o < nvalid) cp_async4(&sm.gy[buf][s][o], gsrc + o);
        cp_async4(&sm.vp[buf][s][o], vsrc + o);
      }
    }
    const int e0 = sm.ptr[t * (NJ + 1)], e1 = sm.ptr[t * (NJ + 1) + NJ];
    if (e1 - e0 <= GA_MAXENT)
      for (int e = tid; e < e1 - e0; e += GA_WARPS * 32) cp_async8(&sm.ent[buf][e], ent + e0 + e);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  for (int i = tid; i < SK_NVT * (NJ + 1); i += GA_WARPS * 32) sm.ptr[i] = __ldg(ent_ptr + i);
  __syncthreads();

  float acc[3][12];
#pragma unroll
  for (int jj = 0; jj < 3; ++jj)
#pragma unroll
    for (int e = 0; e < 12; ++e) acc[jj][e] = 0.f;

  prefetch(0, 0);
  for (int t = 0; t < SK_NVT; ++t) {
    const int buf = t & 1;
    if (t + 1 < SK_NVT) {
      prefetch(t + 1, buf ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const float* gy = sm.gy[buf][lane];
    const float* vp = sm.vp[buf][lane];
    const int tb = sm.ptr[t * (NJ + 1)];
    const bool staged = sm.ptr[t * (NJ + 1) + NJ] - tb <= GA_MAXENT;   // dense models read the table from global
#pragma unroll
    for (int jj = 0; jj < 3; ++jj) {
      const int e0 = sm.ptr[t * (NJ + 1) + 3 * jw + jj] - (staged ? tb : 0);
      const int e1 = sm.ptr[t * (NJ + 1) + 3 * jw + jj + 1] - (staged ? tb : 0);
#pragma unroll 2
      for (int e = e0 + hf; e < e1; e += 2) {
        const float2 en = staged ? sm.ent[buf][e] : __ldg(ent + e);          // uniform across the warp
        const int v = __float_as_int(en.y);
        const float w = en.x;
        const float x = vp[v], y = vp[SK_VT + v], z = vp[2 * SK_VT + v];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const float a = w * gy[3 * v + r];
          acc[jj][4 * r + 0] = fmaf(a, x, acc[jj][4 * r + 0]);
          acc[jj][4 * r + 1] = fmaf(a, y, acc[jj][4 * r + 1]);
          acc[jj][4 * r + 2] = fmaf(a, z, acc[jj][4 * r + 2]);
          acc[jj][4 * r + 3] += a;
        }
      }
    }
    __syncthreads();
  }
  // combine the two entry-parity halves in a fixed order (even + odd): the odd warps hand their sums over through
  // the (now idle) staging buffer
  float* xch = &sm.gy[0][0][0];   // [8 triples][36][32 lanes]
  if (hf == 1) {
#pragma unroll
    for (int jj = 0; jj < 3; ++jj)
#pragma unroll
      for (int e = 0; e < 12; ++e) xch[(jw * 36 + jj * 12 + e) * 32 + lane] = acc[jj][e];
  }
  __syncthreads();
  if (hf == 0 && lane < nrow) {
    float4* o = reinterpret_cast<float4*>(g_A + ((size_t)(b0 + lane) * NJ + 3 * jw) * 12);
#pragma unroll
    for (int jj = 0; jj < 3; ++jj)
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const float* x = xch + (jw * 36 + jj * 12 + 4 * r) * 32 + lane;
        o[jj * 3 + r] = make_float4(acc[jj][4 * r] + x[0], acc[jj][4 * r + 1] + x[32], acc[jj][4 * r + 2] + x[64],
                                    acc[jj][4 * r + 3] + x[96]);
      }
  }
}

int32_t launch_skin_bwd_gA(const smplb_model* m, int B, const float* g_verts, const float* vposed, float* g_A,
                           cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  const size_t smem = sizeof(GaSmem);
  static KernelSetup setup;
  if (int32_t rc = setup.ensure(m->device, skin_bwd_gA_kernel, smem, nullptr)) return rc;
  skin_bwd_gA_kernel<<<(B + GA_SAMPLES - 1) / GA_SAMPLES, GA_WARPS * 32, smem, st>>>(B, g_verts, vposed, m->ent_ptr,
                                                                                     m->ent, g_A);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// g_verts_total = g_verts (or 0) + joint_regressor . g_kp  (adjoint of batch_smpl.py:134-138); only when the
// caller passes keypoint gradients.  One block per sample.
__global__ void __launch_bounds__(256) gverts_total_kernel(int B, int NK, const float* __restrict__ g_verts,
                                                           const float* __restrict__ g_kp,
                                                           const int* __restrict__ kp_ptr,
                                                           const int* __restrict__ kp_vtx,
                                                           const float* __restrict__ kp_val,
                                                           float* __restrict__ out) {
  const int b = blockIdx.x;
  float* ob = out + (size_t)b * NV3;
  for (int i = threadIdx.x; i < NV3; i += blockDim.x) ob[i] = g_verts ? g_verts[(size_t)b * NV3 + i] : 0.f;
  __syncthreads();
  if (threadIdx.x == 0) {   // a vertex may serve several keypoints: serial, fixed order
    for (int k = 0; k < NK; ++k) {
      const float gx = g_kp[((size_t)b * NK + k) * 3], gy = g_kp[((size_t)b * NK + k) * 3 + 1],
                  gz = g_kp[((size_t)b * NK + k) * 3 + 2];
      for (int e = kp_ptr[k]; e < kp_ptr[k + 1]; ++e) {
        const int v = kp_vtx[e];
        const float r = kp_val[e];
        ob[3 * v] += r * gx;
        ob[3 * v + 1] += r * gy;
        ob[3 * v + 2] += r * gz;
      }
    }
  }
}

int32_t launch_gverts_total(const smplb_model* m, int B, const float* g_verts, const float* g_kp, float* out,
                            cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  gverts_total_kernel<<<B, 256, 0, st>>>(B, m->NK, g_verts, g_kp, m->kp_ptr, m->kp_vtx, m->kp_val, out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
