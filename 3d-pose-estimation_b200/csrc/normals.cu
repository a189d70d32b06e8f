// Vertex normals of the posed meshes (deps/pose_3d/utils.py:92-127 normals_from_mesh), the
// first consumer of verts[B,6890,3] on the rendering side (utils.py:43).  One pass over HBM: a CTA keeps one
// sample's vertices (82 KB) and its normals (82 KB) in shared memory, so the kernel reads verts once and writes
// normals once = 165 360 algorithmic bytes per sample.
//
// The reference's graph is reproduced as written, including its two axis quirks:
//  * face normals are "normalised" over the TRIANGLE axis (`tf.nn.l2_normalize(face_normals, axis=1)` on a
//    [batch, faces, 3] tensor, utils.py:109-110): one scale per sample and coordinate, 1/sqrt(max(sum_t n_tc^2, eps));
//  * the outward test sums (v - mean) * n over the VERTEX axis (utils.py:119-121) and then counts positive /
//    negative entries among the three coordinates (:123-126): sign = +1 iff more coordinates are positive.
//  * vertex_faces pads the adjacency lists with the index n_vertices (utils.py:145-147), which is a VALID face
//    index whenever the mesh has more faces than vertices (SMPL: 13776 > 6890), so every vertex of less than
//    maximal order also receives (max_order - order) copies of face n_vertices' normal; an out-of-range pad
//    contributes zeros (the GPU tf.gather behaviour the docstring relies on).
#include "common.cuh"

namespace smplb {

constexpr int NM_THREADS = 1024;

__device__ __forceinline__ void face_cross(const float* __restrict__ v, const int* __restrict__ tri, int t, float n[3]) {
  const int a = 3 * __ldg(tri + 3 * t), b = 3 * __ldg(tri + 3 * t + 1), c = 3 * __ldg(tri + 3 * t + 2);
  const float e1x = v[b] - v[a], e1y = v[b + 1] - v[a + 1], e1z = v[b + 2] - v[a + 2];   // v2 - v1   (:107)
  const float e2x = v[c] - v[a], e2y = v[c + 1] - v[a + 1], e2z = v[c + 2] - v[a + 2];   // v3 - v1
  n[0] = e1y * e2z - e1z * e2y;
  n[1] = e1z * e2x - e1x * e2z;
  n[2] = e1x * e2y - e1y * e2x;
}

// fixed-order block sum of N values per thread; result broadcast to all threads.  Warp shuffle tree, one partial per
// warp in shared memory, then warp 0 alone adds the 32 partials (same tree) and publishes the totals -- letting every
// thread add the 32 partials itself cost 3 K cycles of broadcast shared-memory loads per call.
template <int N>
__device__ __forceinline__ void block_sum(float (&x)[N], float (*red)[8]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < N; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x[k] += __shfl_xor_sync(0xffffffffu, x[k], o);
  __syncthreads();   // red reusable
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < N; ++k) red[warp][k] = x[k];
  __syncthreads();
  if (warp == 0) {
    float t[N];
#pragma unroll
    for (int k = 0; k < N; ++k) {
      t[k] = red[lane][k];     // NM_THREADS / 32 == 32 partials
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
    }
    if (lane == 0)
#pragma unroll
      for (int k = 0; k < N; ++k) red[32][k] = t[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < N; ++k) x[k] = red[32][k];
}

// v2: one pass over the face CORNERS instead of a face pass plus a corner pass.  (b-a)x(c-a) is the same vector from
// any corner of the face, so the vertex thread takes its own position from registers and the two other vertices
// from one packed 4-byte table entry (mesh->vpair); the per-coordinate sum of squares over faces (:109-110) is one
// third of the sum over corners.  The unscaled vertex sums wait in shared memory for that scale.
__global__ void __launch_bounds__(NM_THREADS, 1)
normals_kernel(int B, int V, int T, int max_order, const int* __restrict__ tri, const int* __restrict__ vptr,
               const uint32_t* __restrict__ vpair, const uint4* __restrict__ vfan, const int* __restrict__ big_v, int n_big,
               const float* __restrict__ verts, float* __restrict__ normals) {
  extern __shared__ __align__(16) float nm_smem[];
  float* sv = nm_smem;              // [3V] vertices of the sample
  float* sn = nm_smem + 3 * V;      // [3V] vertex sums, then the (unsigned) normals
  unsigned char* so = reinterpret_cast<unsigned char*>(nm_smem + 6 * V);   // [V] min(vertex order, 255), loaded once
  for (int v = threadIdx.x; v < V; v += NM_THREADS) so[v] = (unsigned char)min(__ldg(vptr + v + 1) - __ldg(vptr + v), 255);
  __shared__ float red[NM_THREADS / 32 + 1][8];
  static_assert(NM_THREADS == 1024, "block_sum assumes 32 warps");
  const int n3 = 3 * V;
  if (blockIdx.x < B) {
    const float* src = verts + (size_t)blockIdx.x * n3;
#pragma unroll 8
    for (int i = threadIdx.x; i < n3; i += NM_THREADS) sv[i] = __ldcs(src + i);
  }
  __syncthreads();
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    float as[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // vertex sums (mean, :116-117) | sums of squared face normals
    float* acc = as;
    float* sq = as + 3;
    for (int v = threadIdx.x; v < V; v += NM_THREADS) {
      const float ax = sv[3 * v], ay = sv[3 * v + 1], az = sv[3 * v + 2];
      const uint4 fan = __ldg(vfan + v);
      if ((fan.x & 0xffffu) != 0xffffu) {
        // closed fan of 3..8 faces: ring of neighbours in one load, each neighbour read once
        const uint32_t w4[4] = {fan.x, fan.y, fan.z, fan.w};
        int qi[8];
#pragma unroll
        for (int k = 0; k < 4; ++k) { qi[2 * k] = (int)(w4[k] & 0xffffu); qi[2 * k + 1] = (int)(w4[k] >> 16); }
        int last = qi[2];
#pragma unroll
        for (int k = 3; k < 8; ++k) last = qi[k] != 0xffff ? qi[k] : last;
        // edge to the previous ring vertex carried in registers: one neighbour load and one subtraction per face
        float e1x = sv[3 * last] - ax, e1y = sv[3 * last + 1] - ay, e1z = sv[3 * last + 2] - az;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (qi[k] != 0xffff) {
            const float e2x = sv[3 * qi[k]] - ax, e2y = sv[3 * qi[k] + 1] - ay, e2z = sv[3 * qi[k] + 2] - az;
            const float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;   // :107
            s0 += nx; s1 += ny; s2 += nz;
            sq[0] += nx * nx; sq[1] += ny * ny; sq[2] += nz * nz;
            e1x = e2x; e1y = e2y; e1z = e2z;
          }
        }
        acc[0] += ax; acc[1] += ay; acc[2] += az;
        sn[3 * v] = s0; sn[3 * v + 1] = s1; sn[3 * v + 2] = s2;
        continue;
      }
      const int e0 = __ldg(vptr + v), e1 = __ldg(vptr + v + 1);
      if (e1 - e0 > NM_BIG_ORDER) continue;   // high-order vertex (e.g. a pole): a whole warp takes it below
      acc[0] += ax; acc[1] += ay; acc[2] += az;
      float s0 = 0.f, s1 = 0.f, s2 = 0.f;
      float qx = 0.f, qy = 0.f, qz = 0.f;   // previous face's second neighbour (fan order: this face's first)
#pragma unroll 4
      for (int e = e0; e < e1; ++e) {
        const uint32_t pr = __ldg(vpair + e);
        const int q = 3 * (int)((pr >> 15) & 0x7fffu);
        float px = qx, py = qy, pz = qz;
        if (!(pr >> 31)) {
          const int p = 3 * (int)(pr & 0x7fffu);
          px = sv[p]; py = sv[p + 1]; pz = sv[p + 2];
        }
        qx = sv[q]; qy = sv[q + 1]; qz = sv[q + 2];
        const float e1x = px - ax, e1y = py - ay, e1z = pz - az;
        const float e2x = qx - ax, e2y = qy - ay, e2z = qz - az;
        const float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;   // :107
        s0 += nx; s1 += ny; s2 += nz;
        sq[0] += nx * nx; sq[1] += ny * ny; sq[2] += nz * nz;
      }
      sn[3 * v] = s0; sn[3 * v + 1] = s1; sn[3 * v + 2] = s2;
    }
    // vertices with many incident faces: one warp each, lanes stride over the corners, fixed-order shuffle tree --
    // a single thread walking an 82-face pole would keep the other 1023 waiting at the barrier
    for (int i = threadIdx.x >> 5; i < n_big; i += NM_THREADS / 32) {
      const int v = __ldg(big_v + i), lane = threadIdx.x & 31;
      const float ax = sv[3 * v], ay = sv[3 * v + 1], az = sv[3 * v + 2];
      const int e0 = __ldg(vptr + v), e1 = __ldg(vptr + v + 1);
      float s0 = 0.f, s1 = 0.f, s2 = 0.f;
      for (int e = e0 + lane; e < e1; e += 32) {
        const uint32_t pr = __ldg(vpair + e);
        const int p = 3 * (int)(pr & 0x7fffu), q = 3 * (int)((pr >> 15) & 0x7fffu);
        const float e1x = sv[p] - ax, e1y = sv[p + 1] - ay, e1z = sv[p + 2] - az;
        const float e2x = sv[q] - ax, e2y = sv[q + 1] - ay, e2z = sv[q + 2] - az;
        const float nx = e1y * e2z - e1z * e2y, ny = e1z * e2x - e1x * e2z, nz = e1x * e2y - e1y * e2x;
        s0 += nx; s1 += ny; s2 += nz;
        sq[0] += nx * nx; sq[1] += ny * ny; sq[2] += nz * nz;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      }
      if (lane == 0) {
        sn[3 * v] = s0; sn[3 * v + 1] = s1; sn[3 * v + 2] = s2;
        acc[0] += ax; acc[1] += ay; acc[2] += az;
      }
    }
    block_sum<6>(as, red);
    const float mean[3] = {acc[0] / (float)V, acc[1] / (float)V, acc[2] / (float)V};     // :116-117
    // per-coordinate scale over the triangle axis (:109-110); every face was visited from its three corners
    const float third = 1.0f / 3.0f;
    const float sc[3] = {rsqrtf(fmaxf(sq[0] * third, 1e-12f)), rsqrtf(fmaxf(sq[1] * third, 1e-12f)),
                         rsqrtf(fmaxf(sq[2] * third, 1e-12f))};
    float pad[3] = {0.f, 0.f, 0.f};
    if (V < T) face_cross(sv, tri, V, pad);   // the pad index n_vertices is a real face (header)
    float sd[3] = {0.f, 0.f, 0.f};
    for (int v = threadIdx.x; v < V; v += NM_THREADS) {
      const int ord = so[v];
      const float np = (float)(max_order - (ord < 255 ? ord : __ldg(vptr + v + 1) - __ldg(vptr + v)));
      float s[3];
      s[0] = (sn[3 * v] + np * pad[0]) * sc[0];
      s[1] = (sn[3 * v + 1] + np * pad[1]) * sc[1];
      s[2] = (sn[3 * v + 2] + np * pad[2]) * sc[2];
      const float inv = rsqrtf(fmaxf(s[0] * s[0] + s[1] * s[1] + s[2] * s[2], 1e-12f));     // :112-114
      s[0] *= inv; s[1] *= inv; s[2] *= inv;
      sn[3 * v] = s[0]; sn[3 * v + 1] = s[1]; sn[3 * v + 2] = s[2];
      sd[0] += (sv[3 * v] - mean[0]) * s[0];                                               // :118-119
      sd[1] += (sv[3 * v + 1] - mean[1]) * s[1];
      sd[2] += (sv[3 * v + 2] - mean[2]) * s[2];
    }
    block_sum<3>(sd, red);   // also makes sn visible to every thread
    const int gt = (sd[0] > 0.f) + (sd[1] > 0.f) + (sd[2] > 0.f), lt = (sd[0] < 0.f) + (sd[1] < 0.f) + (sd[2] < 0.f);
    const float sign = gt > lt ? 1.0f : -1.0f;                                             // :121-126
    // write this sample's normals and, in the same sweep, fetch the next sample's vertices (sv is dead after the
    // reduction above): the load latency hides under the stores instead of opening the next iteration
    float* dst = normals + (size_t)b * n3;
    const int nb = b + gridDim.x;
    if (nb < B && (n3 & 1) == 0 && ((reinterpret_cast<uintptr_t>(verts) | reinterpret_cast<uintptr_t>(normals)) & 7) == 0) {
      // 8-byte accesses (a sample is 3V floats: 8-byte aligned for even V, never 16), eight loads in flight per
      // thread while the stores go out
      const float2* src = reinterpret_cast<const float2*>(verts + (size_t)nb * n3);
      float2* dst2 = reinterpret_cast<float2*>(dst);
      const float2* sn2 = reinterpret_cast<const float2*>(sn);
      float2* sv2 = reinterpret_cast<float2*>(sv);
      const int n2 = n3 >> 1;
      for (int i0 = threadIdx.x; i0 < n2; i0 += 8 * NM_THREADS) {
        float2 x[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          x[k] = i < n2 ? __ldcs(src + i) : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          if (i < n2) { const float2 t = sn2[i]; __stcs(dst2 + i, make_float2(t.x * sign, t.y * sign)); }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          if (i < n2) sv2[i] = x[k];
        }
      }
    } else if (nb < B) {
      const float* src = verts + (size_t)nb * n3;
      for (int i0 = threadIdx.x; i0 < n3; i0 += 8 * NM_THREADS) {
        float x[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          x[k] = i < n3 ? __ldcs(src + i) : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          if (i < n3) __stcs(dst + i, sn[i] * sign);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * NM_THREADS;
          if (i < n3) sv[i] = x[k];
        }
      }
    } else {
      for (int i = threadIdx.x; i < n3; i += NM_THREADS) __stcs(dst + i, sn[i] * sign);
    }
    __syncthreads();   // sn free for the next sample, sv complete
  }
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_mesh_normals(const smplb_mesh* m, int32_t B, const float* verts, float* normals, void* stream) {
  if (!m) return fail(SMPLB_E_INVALID_ARG, "smplb_mesh_normals: null mesh");
  if (B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_mesh_normals: negative batch");
  if (B == 0) return SMPLB_OK;
  if (!verts || !normals) return fail(SMPLB_E_INVALID_ARG, "smplb_mesh_normals: null pointer");
  if (m->T <= 0 || !m->vpair) return fail(SMPLB_E_UNSUPPORTED, "smplb_mesh_normals: mesh without triangles");
  const size_t smem = (size_t)6 * m->V * sizeof(float) + (((size_t)m->V + 15) & ~(size_t)15);
  if (smem > 220 * 1024) return fail(SMPLB_E_UNSUPPORTED, "smplb_mesh_normals: more than 9000 vertices per mesh");
  if (int32_t rc = check_current_device(m->device, "smplb_mesh_normals")) return rc;
  static KernelSetup setup;
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, normals_kernel, 220 * 1024, &n_sm)) return rc;
  const int grid = B < n_sm ? B : n_sm;
  normals_kernel<<<grid, NM_THREADS, smem, (cudaStream_t)stream>>>(B, m->V, m->T, m->max_order, m->tri, m->vptr, m->vpair,
                                                                   m->vfan, m->big_v, m->n_big, verts, normals);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
