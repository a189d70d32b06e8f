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

// fixed-order block sum of three values per thread; result broadcast to all threads
__device__ __forceinline__ void block_sum3(float x[3], float (*red)[3]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x[k] += __shfl_xor_sync(0xffffffffu, x[k], o);
  __syncthreads();   // red reusable
  if (lane == 0) { red[warp][0] = x[0]; red[warp][1] = x[1]; red[warp][2] = x[2]; }
  __syncthreads();
  float s[3] = {0.f, 0.f, 0.f};
#pragma unroll 8
  for (int w = 0; w < NM_THREADS / 32; ++w) { s[0] += red[w][0]; s[1] += red[w][1]; s[2] += red[w][2]; }
  x[0] = s[0]; x[1] = s[1]; x[2] = s[2];
}

// v2: one pass over the face CORNERS instead of a face pass plus a corner pass.  (b-a)x(c-a) is the same vector from
// any corner of the face, so the vertex thread takes its own position from registers and the two other vertices
// from one packed 4-byte table entry (mesh->vpair); the per-coordinate sum of squares over faces (:109-110) is one
// third of the sum over corners.  The unscaled vertex sums wait in shared memory for that scale.
__global__ void __launch_bounds__(NM_THREADS, 1)
normals_kernel(int B, int V, int T, int max_order, const int* __restrict__ tri, const int* __restrict__ vptr,
               const uint32_t* __restrict__ vpair, const int* __restrict__ big_v, int n_big,
               const float* __restrict__ verts, float* __restrict__ normals) {
  extern __shared__ __align__(16) float nm_smem[];
  float* sv = nm_smem;              // [3V] vertices of the sample
  float* sn = nm_smem + 3 * V;      // [3V] vertex sums, then the (unsigned) normals
  __shared__ float red[NM_THREADS / 32][3];
  const int n3 = 3 * V;
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const float* src = verts + (size_t)b * n3;
    __syncthreads();   // previous sample fully written out
    for (int i = threadIdx.x; i < n3; i += NM_THREADS) sv[i] = __ldcs(src + i);
    __syncthreads();
    float acc[3] = {0.f, 0.f, 0.f}, sq[3] = {0.f, 0.f, 0.f};
    for (int v = threadIdx.x; v < V; v += NM_THREADS) {
      const float ax = sv[3 * v], ay = sv[3 * v + 1], az = sv[3 * v + 2];
      const int e0 = __ldg(vptr + v), e1 = __ldg(vptr + v + 1);
      if (e1 - e0 > NM_BIG_ORDER) continue;   // high-order vertex (e.g. a pole): a whole warp takes it below
      acc[0] += ax; acc[1] += ay; acc[2] += az;
      float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll 4
      for (int e = e0; e < e1; ++e) {
        const uint32_t pr = __ldg(vpair + e);
        const int p = 3 * (int)(pr & 0xffffu), q = 3 * (int)(pr >> 16);
        const float e1x = sv[p] - ax, e1y = sv[p + 1] - ay, e1z = sv[p + 2] - az;
        const float e2x = sv[q] - ax, e2y = sv[q + 1] - ay, e2z = sv[q + 2] - az;
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
        const int p = 3 * (int)(pr & 0xffffu), q = 3 * (int)(pr >> 16);
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
    block_sum3(acc, red);
    const float mean[3] = {acc[0] / (float)V, acc[1] / (float)V, acc[2] / (float)V};     // :116-117
    block_sum3(sq, red);
    // per-coordinate scale over the triangle axis (:109-110); every face was visited from its three corners
    const float third = 1.0f / 3.0f;
    const float sc[3] = {rsqrtf(fmaxf(sq[0] * third, 1e-12f)), rsqrtf(fmaxf(sq[1] * third, 1e-12f)),
                         rsqrtf(fmaxf(sq[2] * third, 1e-12f))};
    float pad[3] = {0.f, 0.f, 0.f};
    if (V < T) face_cross(sv, tri, V, pad);   // the pad index n_vertices is a real face (header)
    float sd[3] = {0.f, 0.f, 0.f};
    for (int v = threadIdx.x; v < V; v += NM_THREADS) {
      const float np = (float)(max_order - (__ldg(vptr + v + 1) - __ldg(vptr + v)));
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
    block_sum3(sd, red);   // also makes sn visible to every thread
    const int gt = (sd[0] > 0.f) + (sd[1] > 0.f) + (sd[2] > 0.f), lt = (sd[0] < 0.f) + (sd[1] < 0.f) + (sd[2] < 0.f);
    const float sign = gt > lt ? 1.0f : -1.0f;                                             // :121-126
    float* dst = normals + (size_t)b * n3;
    for (int i = threadIdx.x; i < n3; i += NM_THREADS) __stcs(dst + i, sn[i] * sign);
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
  const size_t smem = (size_t)6 * m->V * sizeof(float);
  if (smem > 220 * 1024) return fail(SMPLB_E_UNSUPPORTED, "smplb_mesh_normals: more than 9386 vertices per mesh");
  static int n_sm[64] = {};
  const int dev = m->device & 63;
  if (!n_sm[dev]) {
    int v = 0;
    SMPLB_CUDA_CHECK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, m->device));
    SMPLB_CUDA_CHECK(cudaFuncSetAttribute(normals_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    n_sm[dev] = v;
  }
  const int grid = B < n_sm[dev] ? B : n_sm[dev];
  normals_kernel<<<grid, NM_THREADS, smem, (cudaStream_t)stream>>>(B, m->V, m->T, m->max_order, m->tri, m->vptr, m->vpair,
                                                                   m->big_v, m->n_big, verts, normals);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
