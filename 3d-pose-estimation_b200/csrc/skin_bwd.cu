// Skinning adjoint, reduction half: the gradient of the relative joint transforms
//     g_A[b,j] = sum_v weights[v,j] * g_verts[b,v] (x) [v_posed[b,v]; 1]        (24 x 3x4 per sample)
// reduces 6890 vertices to 288 numbers per sample.  (The other half, g_vposed = T^T g_verts, is the MODE 1
// instantiation of skin_tc_kernel.)
//
// Mapping: lane = sample (32 samples per CTA); 16 COMPUTE warps = 8 joint triples x entry parity: a warp keeps the 36
// accumulators of its three joints in registers for the whole sweep over the 54 vertex tiles and takes every
// second (vertex, weight) pair of each (tile, joint) list; the two parities are combined in a fixed order at the
// end.  The lists come from a table built at model creation (staged in smem per tile), so the summation order is
// fixed (deterministic, no atomics).
// Tiles of g_verts / v_posed rows are staged through shared memory, double buffered, in a pitch that keeps the
// lane = sample reads of the compute warps bank-conflict free -- which rules out bulk copies (16-byte pitch) and, in
// round 1, meant 4-byte cp.async: 768 LDGSTS per (sample, tile) at 8 cycles each were 2/3 of the kernel (4.26 ms).
// Now 8 LOADER warps move the tiles through registers (48 coalesced 4-byte loads in flight per lane, lane = vertex) and
// re-pack them with 8-byte shared stores into the layout the compute warps want (below); loaders and compute warps hand the two buffers back and forth with named barriers (bar.arrive / bar.sync).
// A smem-bound tensor-core formulation (W^T . (g_verts (x) [v_posed;1]) as an MMA) was examined and dropped: forming
// the 12x-expanded operand costs 48 B of shared stores per vertex-sample and the MMA re-reads it at <= 128 B/clk
// (profiles/r02_mma_probe.txt), more shared-memory traffic than the 6 loads per (vertex, joint) of this kernel.
#include "common.cuh"

namespace smplb {

constexpr int GA_SAMPLES = 32;
constexpr int GA_JW = 8;                    // joint triples
constexpr int GA_CWARPS = 16;               // compute: two warps per joint triple (even / odd table entries)
constexpr int GA_LWARPS = 8;                // loaders: 4 samples each
constexpr int GA_THREADS = (GA_CWARPS + GA_LWARPS) * 32;
constexpr int GA_TILE = 3 * SK_VT;          // 384 floats of a row per vertex tile
// Staged tile: per sample one row of 128 vertices x 6 floats {g_x, g_y, g_z, x, y, z} (the loaders interleave the two
// sources), pitch 770 floats.  A compute lane (= sample) fetches an entry's six numbers with three 8-byte loads; with
// the pitch = 2 (mod 32) words the 16 lanes of a half-warp phase hit 16 distinct bank pairs: conflict-free.  The kernel
// is bound by the number of shared-memory INSTRUCTIONS (ncu: mio_throttle; ~0.5 LDS/STS per clock per SM), so the
// six 4-byte loads per entry of the first version (4.1 ms) became three.
constexpr int GA_VF = 6;
constexpr int GA_PITCH = GA_VF * SK_VT + 2; // 770 floats
constexpr int GA_MAXENT = 1024;             // (vertex, weight) pairs of one vertex tile staged in smem (128 x <= 8 nnz)

struct GaSmem {
  float tile[2][GA_SAMPLES][GA_PITCH];      // 2 x 98 560 B
  float2 ent[2][GA_MAXENT];                 // the tile's table entries: no dependent global load in the inner loop
  int ptr[SK_NVT * (NJ + 1)];               // the whole ent_ptr table (54 x 25), loaded once
};

__device__ __forceinline__ void ga_bar_sync(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(GA_THREADS) : "memory"); }
__device__ __forceinline__ void ga_bar_arrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(GA_THREADS) : "memory"); }
__device__ __forceinline__ float2 ldg_stream2(const float* p) {
  float2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
  return v;
}

__global__ void __launch_bounds__(GA_THREADS, 1)
skin_bwd_gA_kernel(int B, const float* __restrict__ g_verts, const float* __restrict__ vposed,
                   const int* __restrict__ ent_ptr, const float2* __restrict__ ent, float* __restrict__ g_A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GaSmem& sm = *reinterpret_cast<GaSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b0 = blockIdx.x * GA_SAMPLES;
  const int nrow = min(GA_SAMPLES, B - b0);

  for (int i = tid; i < SK_NVT * (NJ + 1); i += GA_THREADS) sm.ptr[i] = __ldg(ent_ptr + i);
  __syncthreads();

  // named barriers: 1 + buf = "buffer full" (loaders arrive, compute waits), 3 + buf = "buffer free" (compute arrives,
  // loaders wait)
  if (warp >= GA_CWARPS) {
    // ================= loaders =================
    const int lw = warp - GA_CWARPS, ltid = tid - GA_CWARPS * 32;
    for (int t = 0; t < SK_NVT; ++t) {
      const int buf = t & 1;
      if (t >= 2) ga_bar_sync(3 + buf);
      const int nvv = min(SK_VT, NV - t * SK_VT);   // vertices of this tile that exist (106 in the last one)
      // lane = vertex: per (sample, 32-vertex chunk) six 4-byte loads (coalesced: 384 / 128 contiguous bytes per
      // instruction) and three conflict-free 8-byte shared stores; two samples (48 loads) in flight per lane
#pragma unroll 1
      for (int sp = 0; sp < GA_SAMPLES / GA_LWARPS; sp += 2) {
        float v[2][4][6];
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int s = lw * (GA_SAMPLES / GA_LWARPS) + sp + q;
          const int sc = s < nrow ? s : nrow - 1;
          const float* gsrc = g_verts + (size_t)(b0 + sc) * NV3 + (size_t)t * GA_TILE;
          const float* vsrc = vposed + (size_t)(b0 + sc) * VPP + (size_t)t * GA_TILE;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int vv = 32 * c + lane;
            const bool ok = vv < nvv;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
              v[q][c][r] = ok ? __ldg(gsrc + 3 * vv + r) : 0.f;
              v[q][c][3 + r] = __ldg(vsrc + r * SK_VT + vv);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int s = lw * (GA_SAMPLES / GA_LWARPS) + sp + q;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float2* d = reinterpret_cast<float2*>(&sm.tile[buf][s][GA_VF * (32 * c + lane)]);
            d[0] = make_float2(v[q][c][0], v[q][c][1]);
            d[1] = make_float2(v[q][c][2], v[q][c][3]);
            d[2] = make_float2(v[q][c][4], v[q][c][5]);
          }
        }
      }
      const int e0 = sm.ptr[t * (NJ + 1)], e1 = sm.ptr[t * (NJ + 1) + NJ];
      if (e1 - e0 <= GA_MAXENT)
        for (int e = ltid; e < e1 - e0; e += GA_LWARPS * 32) sm.ent[buf][e] = __ldg(ent + e0 + e);
      ga_bar_arrive(1 + buf);
    }
  } else {
    // ================= compute =================
    const int jw = warp & (GA_JW - 1), hf = warp >> 3;   // joint triple, entry parity
    float acc[3][12];
#pragma unroll
    for (int jj = 0; jj < 3; ++jj)
#pragma unroll
      for (int e = 0; e < 12; ++e) acc[jj][e] = 0.f;

    for (int t = 0; t < SK_NVT; ++t) {
      const int buf = t & 1;
      ga_bar_sync(1 + buf);
      const float* row = sm.tile[buf][lane];
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
          const float2* pv = reinterpret_cast<const float2*>(row + GA_VF * v);
          const float2 p0 = pv[0], p1 = pv[1], p2 = pv[2];
          const float gyv[3] = {p0.x, p0.y, p1.x};
          const float x = p1.y, y = p2.x, z = p2.y;
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const float a = w * gyv[r];
            acc[jj][4 * r + 0] = fmaf(a, x, acc[jj][4 * r + 0]);
            acc[jj][4 * r + 1] = fmaf(a, y, acc[jj][4 * r + 1]);
            acc[jj][4 * r + 2] = fmaf(a, z, acc[jj][4 * r + 2]);
            acc[jj][4 * r + 3] += a;
          }
        }
      }
      if (t + 2 < SK_NVT) ga_bar_arrive(3 + buf);
    }
    // combine the two entry-parity halves in a fixed order (even + odd): the odd warps hand their sums over through
    // the (now idle) staging buffer.  Barrier 5: the 16 compute warps only.
    float* xch = &sm.tile[0][0][0];   // [8 triples][36][32 lanes]
    asm volatile("bar.sync 5, %0;" ::"n"(GA_CWARPS * 32) : "memory");   // every compute warp is done with the buffers
    if (hf == 1) {
#pragma unroll
      for (int jj = 0; jj < 3; ++jj)
#pragma unroll
        for (int e = 0; e < 12; ++e) xch[(jw * 36 + jj * 12 + e) * 32 + lane] = acc[jj][e];
    }
    asm volatile("bar.sync 5, %0;" ::"n"(GA_CWARPS * 32) : "memory");
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
}

int32_t launch_skin_bwd_gA(const smplb_model* m, int B, const float* g_verts, const float* vposed, float* g_A,
                           cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  const size_t smem = sizeof(GaSmem);
  static KernelSetup setup;
  if (int32_t rc = setup.ensure(m->device, skin_bwd_gA_kernel, smem, nullptr)) return rc;
  skin_bwd_gA_kernel<<<(B + GA_SAMPLES - 1) / GA_SAMPLES, GA_THREADS, smem, st>>>(B, g_verts, vposed, m->ent_ptr,
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
