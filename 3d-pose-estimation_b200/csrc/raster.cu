// Triangle rasteriser + its gradient as CUDA kernels: the consumer of `verts`
// (pose_model_3d.py:319-332 -> tf_mesh_renderer).  Pure arithmetic, no graphics API.
//   forward : deps/tf_mesh_renderer/kernels/rasterize_triangles_impl.cc:75-155 (+ clears of
//             rasterize_triangles_op.cc:121-125): per pixel the barycentrics, triangle id and NDC depth of the
//             nearest triangle covering the pixel centre (16.16 fixed-point edge tests, both windings).
//   backward: kernels/rasterize_triangles_grad.cc:144-218: d f / d vertices(x, y) from d f / d barycentrics.
//
// The reference walks the triangles in order on one core; a pixel finally holds the triangle with the smallest
// z, ties going to the highest triangle id (`z > z_buffer` skips, so an equal z overwrites).  That final state is
// order independent, so the triangles run in parallel and resolve through ONE 64-bit atomicMin per covered pixel
// on the key (orderable z bits << 32 | ~id); a second pass decodes the winner and recomputes its barycentrics
// with the same float operations (no FMA contraction, the double-precision viewport transform of the
// reference, x86 float->int conversion semantics), so ids, barycentrics and depth are bit-identical to the
// reference's C++.  (Not reproduced: NaN depths of zero-area triangles, which the reference lets through.)
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace smplb {

constexpr int RS_BIG = 256;   // bounding boxes with more pixels than this go to the CTA-per-triangle kernel

// static_cast<int>(float) as x86-64 does it (cvttss2si): out of range / NaN -> INT_MIN
__device__ __forceinline__ int x86_f2i(float f) {
  return (f >= -2147483648.0f && f < 2147483648.0f) ? (int)f : (int)0x80000000;
}
__device__ __forceinline__ int clamped_min3(float a, float b, float c, int lo, int hi) {   // impl.cc:25-29
  return min(max(x86_f2i(floorf(fminf(fminf(a, b), c))), lo), hi);
}
__device__ __forceinline__ int clamped_max3(float a, float b, float c, int lo, int hi) {   // impl.cc:33-37
  return min(max(x86_f2i(ceilf(fmaxf(fmaxf(a, b), c))), lo), hi);
}
__device__ __forceinline__ int fixed_point(float f) { return x86_f2i(__fmul_rn(f, 65536.0f)); }   // impl.cc:41
__device__ __forceinline__ bool is_ccw(int v0x, int v0y, int v1x, int v1y, int px, int py) {       // impl.cc:45-56
  const int ex = v1x - v0x, ey = v1y - v0y, x = px - v0x, y = py - v0y;
  return (long long)ex * (long long)y >= (long long)ey * (long long)x;
}

struct TriSetup {
  float v0x, v0y, v1x, v1y, v2x, v2y, z0, z1, z2;
  int f0x, f0y, f1x, f1y, f2x, f2y;
  int left, right, bottom, top;
  bool safe;   // every fixed-point coordinate (vertices and pixel centres) below 2^29: no int32 wrap in the edge tests
};

// impl.cc:82-107.  (v + 1.0) * half is evaluated in double and rounded once, as the C++ does.
__device__ __forceinline__ void tri_setup(const float* __restrict__ ndc, const int* __restrict__ tri, int t, int W, int H,
                                          TriSetup& s) {
  const float half_w = (float)(0.5 * (double)W), half_h = (float)(0.5 * (double)H);
  const int i0 = 3 * __ldg(tri + 3 * t), i1 = 3 * __ldg(tri + 3 * t + 1), i2 = 3 * __ldg(tri + 3 * t + 2);
  auto vp = [](float v, float half) { return __double2float_rn(__dmul_rn(__dadd_rn((double)v, 1.0), (double)half)); };
  s.v0x = vp(ndc[i0], half_w); s.v0y = vp(ndc[i0 + 1], half_h); s.z0 = ndc[i0 + 2];
  s.v1x = vp(ndc[i1], half_w); s.v1y = vp(ndc[i1 + 1], half_h); s.z1 = ndc[i1 + 2];
  s.v2x = vp(ndc[i2], half_w); s.v2y = vp(ndc[i2 + 1], half_h); s.z2 = ndc[i2 + 2];
  s.left = clamped_min3(s.v0x, s.v1x, s.v2x, 0, W);
  s.right = clamped_max3(s.v0x, s.v1x, s.v2x, 0, W);
  s.bottom = clamped_min3(s.v0y, s.v1y, s.v2y, 0, H);
  s.top = clamped_max3(s.v0y, s.v1y, s.v2y, 0, H);
  s.f0x = fixed_point(s.v0x); s.f0y = fixed_point(s.v0y);
  s.f1x = fixed_point(s.v1x); s.f1y = fixed_point(s.v1y);
  s.f2x = fixed_point(s.v2x); s.f2y = fixed_point(s.v2y);
  auto small = [](int f) { return (uint32_t)f + (1u << 29) < (1u << 30); };   // -2^29 <= f < 2^29 (INT_MIN fails)
  s.safe = W <= 8192 && H <= 8192 && small(s.f0x) && small(s.f0y) && small(s.f1x) && small(s.f1y) && small(s.f2x) &&
           small(s.f2y);
  // The reference walks the box floor(min) .. ceil(max) of the float coordinates (above).  A pixel whose centre lies
  // strictly beyond the fixed-point extent of the triangle in x or in y cannot pass the inside test of a triangle
  // with non-zero fixed-point area (the union of "all edges >= 0" and "all edges <= 0" is then the closed triangle),
  // so the box is shrunk to the pixel centres inside the fixed-point extent -- for a body mesh at 320x240 most boxes
  // become empty or a single pixel.  Exactly collinear fixed-point triangles keep the reference box: their test
  // passes along the whole line.
  if (s.safe) {
    const long long area = (long long)(s.f1x - s.f0x) * (s.f2y - s.f0y) - (long long)(s.f1y - s.f0y) * (s.f2x - s.f0x);
    if (area != 0) {
      const int fx0 = min(min(s.f0x, s.f1x), s.f2x), fx1 = max(max(s.f0x, s.f1x), s.f2x);
      const int fy0 = min(min(s.f0y, s.f1y), s.f2y), fy1 = max(max(s.f0y, s.f1y), s.f2y);
      s.left = max(s.left, (fx0 - 32768 + 65535) >> 16);      // first ix with (ix << 16) + 32768 >= fx0
      s.right = min(s.right, ((fx1 - 32768) >> 16) + 1);      // one past the last ix with (ix << 16) + 32768 <= fx1
      s.bottom = max(s.bottom, (fy0 - 32768 + 65535) >> 16);
      s.top = min(s.top, ((fy1 - 32768) >> 16) + 1);
    }
  }
}

// impl.cc:112-124, the pixel-inside test: (all three edges counter-clockwise) or (all three clockwise), six is_ccw
// calls in the reference.  is_ccw(a, b, p) is E_ab(p) >= 0 with E_ab(p) = (b - a) x (p - a); the reversed call
// is_ccw(b, a, p) is E_ab(p) <= 0 -- the same integer, because moving the origin from a to b changes E by
// (b - a) x (b - a) = 0.  With every coordinate below 2^29 nothing wraps in int32 / int64, so three 64-bit edge
// functions decide both orientations exactly; otherwise the six calls are evaluated literally (wrap-around and all).
__device__ __forceinline__ bool tri_covers(const TriSetup& s, int ix, int iy) {
  if (s.safe) {
    const int fx = (ix << 16) + 32768, fy = (iy << 16) + 32768;   // == fixed_point(ix + 0.5f), exact
    const long long e01 = (long long)(s.f1x - s.f0x) * (fy - s.f0y) - (long long)(s.f1y - s.f0y) * (fx - s.f0x);
    const long long e12 = (long long)(s.f2x - s.f1x) * (fy - s.f1y) - (long long)(s.f2y - s.f1y) * (fx - s.f1x);
    const long long e20 = (long long)(s.f0x - s.f2x) * (fy - s.f2y) - (long long)(s.f0y - s.f2y) * (fx - s.f2x);
    return (e01 >= 0 && e12 >= 0 && e20 >= 0) || (e01 <= 0 && e12 <= 0 && e20 <= 0);
  }
  const int fx = fixed_point((float)ix + 0.5f), fy = fixed_point((float)iy + 0.5f);
  return (is_ccw(s.f0x, s.f0y, s.f1x, s.f1y, fx, fy) && is_ccw(s.f1x, s.f1y, s.f2x, s.f2y, fx, fy) &&
          is_ccw(s.f2x, s.f2y, s.f0x, s.f0y, fx, fy)) ||
         (is_ccw(s.f1x, s.f1y, s.f0x, s.f0y, fx, fy) && is_ccw(s.f2x, s.f2y, s.f1x, s.f1y, fx, fy) &&
          is_ccw(s.f0x, s.f0y, s.f2x, s.f2y, fx, fy));
}

// impl.cc:112-143 for one pixel: coverage test, barycentrics, depth, clip test.  No contraction anywhere.
__device__ __forceinline__ bool tri_eval(const TriSetup& s, int ix, int iy, float& b0, float& b1, float& b2, float& z) {
  if (!tri_covers(s, ix, iy)) return false;
  const float px = (float)ix + 0.5f, py = (float)iy + 0.5f;   // exact
  const float area2 = __fsub_rn(__fmul_rn(__fsub_rn(s.v2x, s.v0x), __fsub_rn(s.v1y, s.v0y)),
                                __fmul_rn(__fsub_rn(s.v2y, s.v0y), __fsub_rn(s.v1x, s.v0x)));
  b0 = __fdiv_rn(__fsub_rn(__fmul_rn(__fsub_rn(px, s.v1x), __fsub_rn(s.v2y, s.v1y)),
                           __fmul_rn(__fsub_rn(py, s.v1y), __fsub_rn(s.v2x, s.v1x))), area2);
  b1 = __fdiv_rn(__fsub_rn(__fmul_rn(__fsub_rn(px, s.v2x), __fsub_rn(s.v0y, s.v2y)),
                           __fmul_rn(__fsub_rn(py, s.v2y), __fsub_rn(s.v0x, s.v2x))), area2);
  b2 = __fsub_rn(__fsub_rn(1.0f, b0), b1);
  z = __fadd_rn(__fadd_rn(__fmul_rn(b0, s.z0), __fmul_rn(b1, s.z1)), __fmul_rn(b2, s.z2));
  return z >= -1.0f && z <= 1.0f;   // also rejects NaN (see header)
}

__device__ __forceinline__ unsigned long long pack_key(float z, int t) {
  if (z == 0.0f) z = 0.0f;   // -0 and +0 compare equal in the reference
  const uint32_t b = __float_as_uint(z);
  const uint32_t o = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return ((unsigned long long)o << 32) | (uint32_t)(~(uint32_t)t);
}
constexpr unsigned long long KEY_EMPTY = ~0ull;

// thread = (image, triangle): small bounding boxes are walked here, large ones queued
__global__ void __launch_bounds__(128) raster_tri_kernel(int T, int V, int W, int H, const float* __restrict__ ndc,
                                                         const int* __restrict__ tri, unsigned long long* __restrict__ keys,
                                                         int* __restrict__ big_list, int* __restrict__ big_count) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (t >= T) return;
  TriSetup s;
  tri_setup(ndc + (size_t)b * V * 3, tri, t, W, H, s);
  const int bw = s.right - s.left, bh = s.top - s.bottom;
  if (bw <= 0 || bh <= 0) return;
  if (bw * bh > RS_BIG) {
    big_list[atomicAdd(big_count, 1)] = b * T + t;
    return;
  }
  unsigned long long* k = keys + (size_t)b * W * H;
  // one flat loop over the box (the lanes of a warp have boxes of different shapes: a single trip count
  // reconverges better than two nested ones)
  int ix = s.left, iy = s.bottom;
  for (int i = bw * bh; i > 0; --i) {
    float b0, b1, b2, z;
    if (tri_eval(s, ix, iy, b0, b1, b2, z)) atomicMin(k + iy * W + ix, pack_key(z, t));
    if (++ix == s.right) { ix = s.left; ++iy; }
  }
}

// CTA = one queued large triangle, threads stride over its bounding box
__global__ void __launch_bounds__(256) raster_big_kernel(int T, int V, int W, int H, const float* __restrict__ ndc,
                                                         const int* __restrict__ tri, unsigned long long* __restrict__ keys,
                                                         const int* __restrict__ big_list, const int* __restrict__ big_count) {
  const int n = *big_count;
  for (int e = blockIdx.x; e < n; e += gridDim.x) {
    const int b = big_list[e] / T, t = big_list[e] % T;
    TriSetup s;
    tri_setup(ndc + (size_t)b * V * 3, tri, t, W, H, s);
    const int bw = s.right - s.left, np = bw * (s.top - s.bottom);
    unsigned long long* k = keys + (size_t)b * W * H;
    for (int i = threadIdx.x; i < np; i += blockDim.x) {
      const int ix = s.left + i % bw, iy = s.bottom + i / bw;
      float b0, b1, b2, z;
      if (tri_eval(s, ix, iy, b0, b1, b2, z)) atomicMin(k + iy * W + ix, pack_key(z, t));
    }
  }
}

// thread = pixel: decode the winner, recompute its barycentrics / depth, write the three outputs
__global__ void __launch_bounds__(256) raster_resolve_kernel(int B, int T, int V, int W, int H, const float* __restrict__ ndc,
                                                             const int* __restrict__ tri,
                                                             const unsigned long long* __restrict__ keys,
                                                             float* __restrict__ bary, int* __restrict__ ids,
                                                             float* __restrict__ zbuf) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * W * H) return;
  const int b = (int)(i / ((size_t)W * H)), p = (int)(i % ((size_t)W * H));
  const unsigned long long key = keys[i];
  float b0 = 0.f, b1 = 0.f, b2 = 0.f, z = 1.0f;
  int t = 0;
  if (key != KEY_EMPTY) {
    t = (int)(~(uint32_t)key);
    TriSetup s;
    tri_setup(ndc + (size_t)b * V * 3, tri, t, W, H, s);
    tri_eval(s, p % W, p / W, b0, b1, b2, z);
  }
  if (bary) {
    bary[3 * i] = b0;
    bary[3 * i + 1] = b1;
    bary[3 * i + 2] = b2;
  }
  if (ids) ids[i] = t;
  if (zbuf) zbuf[i] = z;
}

// same, thread = 4 consecutive pixels of one image (W*H % 4 == 0, 16-byte aligned buffers): 16-byte key loads and
// output stores, four independent key -> triangle -> vertex load chains in flight per thread
__global__ void __launch_bounds__(256) raster_resolve4_kernel(int B, int T, int V, int W, int H, const float* __restrict__ ndc,
                                                              const int* __restrict__ tri,
                                                              const unsigned long long* __restrict__ keys,
                                                              float* __restrict__ bary, int* __restrict__ ids,
                                                              float* __restrict__ zbuf) {
  const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x, i0 = 4 * q;
  if (i0 >= (size_t)B * W * H) return;
  const int b = (int)(i0 / ((size_t)W * H)), p0 = (int)(i0 % ((size_t)W * H));
  const uint4 ka = __ldcs(reinterpret_cast<const uint4*>(keys) + 2 * q), kb = __ldcs(reinterpret_cast<const uint4*>(keys) + 2 * q + 1);
  const uint32_t klo[4] = {ka.x, ka.z, kb.x, kb.z}, khi[4] = {ka.y, ka.w, kb.y, kb.w};
  float bc[12], z[4];
  int t[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    bc[3 * k] = bc[3 * k + 1] = bc[3 * k + 2] = 0.f;
    z[k] = 1.0f;
    t[k] = 0;
    if ((klo[k] & khi[k]) != 0xffffffffu) {
      t[k] = (int)(~klo[k]);
      TriSetup s;
      tri_setup(ndc + (size_t)b * V * 3, tri, t[k], W, H, s);
      const int p = p0 + k;
      tri_eval(s, p % W, p / W, bc[3 * k], bc[3 * k + 1], bc[3 * k + 2], z[k]);
    }
  }
  if (bary) {
    float4* o = reinterpret_cast<float4*>(bary) + 3 * q;
    __stcs(o, make_float4(bc[0], bc[1], bc[2], bc[3]));
    __stcs(o + 1, make_float4(bc[4], bc[5], bc[6], bc[7]));
    __stcs(o + 2, make_float4(bc[8], bc[9], bc[10], bc[11]));
  }
  if (ids) __stcs(reinterpret_cast<int4*>(ids) + q, make_int4(t[0], t[1], t[2], t[3]));
  if (zbuf) __stcs(reinterpret_cast<float4*>(zbuf) + q, make_float4(z[0], z[1], z[2], z[3]));
}

// ---- backward (rasterize_triangles_grad.cc:144-218) ------------------------------------------------------------
// The reference accumulates pixel by pixel into df_dvertices; here every (image, triangle) thread walks the pixels
// of its own bounding box in the same row-major order and keeps the six sums of its corners, then every
// (image, vertex) thread adds the partial sums of its incident corners in ascending triangle order: no atomics,
// run-to-run deterministic; differs from the reference only in float summation order (tolerance 1e-4, tests).
__global__ void __launch_bounds__(128) raster_bwd_tri_kernel(int T, int V, int W, int H, const float* __restrict__ ndc,
                                                             const int* __restrict__ tri, const float* __restrict__ bary,
                                                             const int* __restrict__ ids, const float* __restrict__ g_bary,
                                                             float* __restrict__ tri_grad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (t >= T) return;
  const float* nv = ndc + (size_t)b * V * 3;
  float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  TriSetup s;
  tri_setup(nv, tri, t, W, H, s);
  const int i0 = 3 * __ldg(tri + 3 * t), i1 = 3 * __ldg(tri + 3 * t + 1), i2 = 3 * __ldg(tri + 3 * t + 2);
  const float v0x = nv[i0], v0y = nv[i0 + 1], v1x = nv[i1], v1y = nv[i1 + 1], v2x = nv[i2], v2y = nv[i2 + 1];
  const float area = (v1x - v0x) * (v2y - v0y) - (v1y - v0y) * (v2x - v0x);   // grad.cc:183-184 (NDC)
  if (fabsf(area) >= 1e-13f) {                                                // grad.cc:186
    const float dx0 = v2y - v1y, dx1 = v0y - v2y, dx2 = v1y - v0y;            // grad.cc:198-203
    const float dy0 = v1x - v2x, dy1 = v2x - v0x, dy2 = v0x - v1x;
    const size_t base = (size_t)b * W * H;
    for (int iy = s.bottom; iy < s.top; ++iy)
      for (int ix = s.left; ix < s.right; ++ix) {
        const size_t p = base + (size_t)iy * W + ix;
        if (__ldg(ids + p) != t) continue;
        const float b0 = __ldg(bary + 3 * p), b1 = __ldg(bary + 3 * p + 1), b2 = __ldg(bary + 3 * p + 2);
        if (b0 + b1 + b2 < 0.9f) continue;                                     // grad.cc:155 (also: id 0 = background)
        const float d0 = __ldg(g_bary + 3 * p), d1 = __ldg(g_bary + 3 * p + 1), d2 = __ldg(g_bary + 3 * p + 2);
        const float gx = d0 * dx0 + d1 * dx1 + d2 * dx2, gy = d0 * dy0 + d1 * dy1 + d2 * dy2;   // grad.cc:207-210
        const float s0 = b0 / area, s1 = b1 / area, s2 = b2 / area;            // grad.cc:192-194
        acc[0] += s0 * gx; acc[1] += s0 * gy;                                  // grad.cc:211-216
        acc[2] += s1 * gx; acc[3] += s1 * gy;
        acc[4] += s2 * gx; acc[5] += s2 * gy;
      }
  }
  float* o = tri_grad + ((size_t)b * T + t) * 6;
#pragma unroll
  for (int k = 0; k < 6; ++k) o[k] = acc[k];
}

__global__ void __launch_bounds__(128) raster_bwd_vtx_kernel(int T, int V, const int* __restrict__ vptr,
                                                             const int* __restrict__ vent,
                                                             const float* __restrict__ tri_grad, float* __restrict__ g_ndc) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
  if (v >= V) return;
  float gx = 0.f, gy = 0.f;
  const float* tg = tri_grad + (size_t)b * T * 6;
  for (int e = __ldg(vptr + v); e < __ldg(vptr + v + 1); ++e) {
    const int tc = __ldg(vent + e);   // triangle * 3 + corner
    gx += tg[(size_t)tc * 2];
    gy += tg[(size_t)tc * 2 + 1];
  }
  float* o = g_ndc + ((size_t)b * V + v) * 3;
  o[0] = gx; o[1] = gy; o[2] = 0.f;   // no gradient through depth (grad.cc writes x, y only)
}

static size_t ru256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_mesh_create(smplb_mesh** out, int32_t device, const int32_t* triangles_host, int32_t T,
                                     int32_t V) {
  if (!out || (!triangles_host && T > 0)) return fail(SMPLB_E_INVALID_ARG, "smplb_mesh_create: null argument");
  if (T < 0 || V <= 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_mesh_create: need T >= 0 and V > 0");
  std::vector<int> vptr(V + 1, 0), vent((size_t)3 * T);
  for (int i = 0; i < 3 * T; ++i) {
    if (triangles_host[i] < 0 || triangles_host[i] >= V)
      return fail(SMPLB_E_INVALID_ARG, "smplb_mesh_create: triangle vertex index out of range");
    ++vptr[triangles_host[i] + 1];
  }
  for (int v = 0; v < V; ++v) vptr[v + 1] += vptr[v];
  std::vector<int> fill(vptr.begin(), vptr.end() - 1);
  for (int i = 0; i < 3 * T; ++i) vent[fill[triangles_host[i]]++] = i;   // i = triangle * 3 + corner, ascending
  int prev = 0;
  cudaGetDevice(&prev);
  SMPLB_CUDA_CHECK(cudaSetDevice(device));
  smplb_mesh* m = new smplb_mesh();
  m->device = device; m->T = T; m->V = V; m->tri = m->vptr = m->vent = nullptr; m->vpair = nullptr; m->vfan = nullptr; m->big_v = nullptr; m->n_big = 0;
  // vertex_faces_from_face_verts (deps/pose_3d/utils.py:130-149): adjacency lists padded to the maximum vertex order
  m->max_order = 0;
  for (int v = 0; v < V; ++v) m->max_order = std::max(m->max_order, vptr[v + 1] - vptr[v]);
  cudaError_t e = cudaMalloc((void**)&m->tri, sizeof(int) * (size_t)(3 * T > 0 ? 3 * T : 1));
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->vptr, sizeof(int) * (size_t)(V + 1));
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->vent, sizeof(int) * (size_t)(3 * T > 0 ? 3 * T : 1));
  if (e == cudaSuccess && T > 0) e = cudaMemcpy(m->tri, triangles_host, sizeof(int) * 3 * (size_t)T, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(m->vptr, vptr.data(), sizeof(int) * (size_t)(V + 1), cudaMemcpyHostToDevice);
  if (e == cudaSuccess && T > 0) e = cudaMemcpy(m->vent, vent.data(), sizeof(int) * 3 * (size_t)T, cudaMemcpyHostToDevice);
  {
    std::vector<int> big;
    for (int v = 0; v < V; ++v)
      if (vptr[v + 1] - vptr[v] > NM_BIG_ORDER) big.push_back(v);
    m->n_big = (int)big.size();
    if (e == cudaSuccess) e = cudaMalloc((void**)&m->big_v, sizeof(int) * (big.size() ? big.size() : 1));
    if (e == cudaSuccess && !big.empty()) e = cudaMemcpy(m->big_v, big.data(), sizeof(int) * big.size(), cudaMemcpyHostToDevice);
  }
  if (e == cudaSuccess && T > 0 && V <= 32768) {
    // (b - a) x (c - a) is the same vector from whichever corner it is taken: store, per incident face of a vertex,
    // the next two vertices (p, q) after this corner, 15 bits each, so that the normals kernel needs one 4-byte load
    // per face.  The faces of a vertex are put in FAN order (q of one = p of the next, as around any manifold
    // vertex); bit 31 marks an entry whose p is the previous entry's q, which the kernel then keeps in registers:
    // 3n + 6 shared-memory loads per vertex of order n instead of 6n + 3.
    std::vector<uint32_t> vpair((size_t)3 * T);
    std::vector<uint16_t> vfan((size_t)8 * V, (uint16_t)0xffff);
    std::vector<std::pair<int, int>> pq;
    std::vector<char> used;
    for (int v = 0; v < V; ++v) {
      const int e0 = vptr[v], n = vptr[v + 1] - vptr[v];
      pq.resize(n);
      used.assign(n, 0);
      for (int i = 0; i < n; ++i) {
        const int t = vent[e0 + i] / 3, k = vent[e0 + i] % 3;
        pq[i] = {triangles_host[3 * t + (k + 1) % 3], triangles_host[3 * t + (k + 2) % 3]};
      }
      int out = e0, last_q = -1, n_chains = 0;
      for (int done = 0; done < n; ++done) {
        int pick = -1;
        for (int i = 0; i < n && pick < 0; ++i)
          if (!used[i] && pq[i].first == last_q) pick = i;
        const bool chained = pick >= 0;
        n_chains += chained ? 0 : 1;
        if (!chained) {
          // start of a chain: prefer a face whose p is nobody's q (the open end of a boundary fan)
          for (int i = 0; i < n && pick < 0; ++i) {
            if (used[i]) continue;
            bool is_q = false;
            for (int j = 0; j < n; ++j) is_q |= (!used[j] && j != i && pq[j].second == pq[i].first);
            if (!is_q) pick = i;
          }
          for (int i = 0; i < n && pick < 0; ++i)
            if (!used[i]) pick = i;
        }
        used[pick] = 1;
        vpair[out++] = (uint32_t)pq[pick].first | ((uint32_t)pq[pick].second << 15) | (chained ? 0x80000000u : 0u);
        last_q = pq[pick].second;
      }
      // one closed fan of 3..8 faces (every interior vertex of a manifold mesh): the kernel walks the ring of
      // neighbours q_0 .. q_{n-1} from a single 16-byte load, the first face's p being q_{n-1}
      if (n >= 3 && n <= 8 && n_chains == 1 && (int)(vpair[e0] & 0x7fffu) == last_q)
        for (int i = 0; i < n; ++i) vfan[(size_t)8 * v + i] = (uint16_t)((vpair[e0 + i] >> 15) & 0x7fffu);
    }
    e = cudaMalloc((void**)&m->vfan, sizeof(uint16_t) * 8 * (size_t)V);
    if (e == cudaSuccess) e = cudaMemcpy(m->vfan, vfan.data(), sizeof(uint16_t) * 8 * (size_t)V, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc((void**)&m->vpair, sizeof(uint32_t) * 3 * (size_t)T);
    if (e == cudaSuccess) e = cudaMemcpy(m->vpair, vpair.data(), sizeof(uint32_t) * 3 * (size_t)T, cudaMemcpyHostToDevice);
  }
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    cudaFree(m->tri); cudaFree(m->vptr); cudaFree(m->vent); cudaFree(m->vpair); cudaFree(m->vfan); cudaFree(m->big_v);
    delete m;
    return fail(SMPLB_E_CUDA, std::string("smplb_mesh_create: ") + cudaGetErrorString(e));
  }
  *out = m;
  return SMPLB_OK;
}

extern "C" int32_t smplb_mesh_destroy(smplb_mesh* m) {
  if (!m) return SMPLB_OK;
  cudaFree(m->tri); cudaFree(m->vptr); cudaFree(m->vent); cudaFree(m->vpair); cudaFree(m->vfan); cudaFree(m->big_v);
  delete m;
  return SMPLB_OK;
}

// keys [B*W*H] u64 | big_count (256 B) | big_list [B*T] int | tri_grad [B*T*6] float
extern "C" size_t smplb_rasterize_workspace_bytes(const smplb_mesh* m, int32_t B, int32_t W, int32_t H) {
  if (!m || B <= 0 || W <= 0 || H <= 0) return 0;
  const size_t bt = (size_t)B * (size_t)(m->T > 0 ? m->T : 1);
  return ru256((size_t)B * W * H * 8) + 256 + ru256(bt * 4) + ru256(bt * 6 * 4);
}

static int32_t raster_check(const char* who, const smplb_mesh* m, int32_t B, int32_t W, int32_t H, const void* ws,
                            size_t ws_bytes) {
  if (!m) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": null mesh");
  if (int32_t rc = check_current_device(m->device, who)) return rc;
  if (B < 0) return fail(SMPLB_E_BAD_SHAPE, std::string(who) + ": negative batch");
  if (W <= 0 || H <= 0) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": image width and height must be > 0");
  if ((size_t)W * H > (1u << 30)) return fail(SMPLB_E_UNSUPPORTED, std::string(who) + ": image too large");
  if (B > 65535) return fail(SMPLB_E_UNSUPPORTED, std::string(who) + ": at most 65535 images per call");
  if (B > 0 && (!ws || ws_bytes < smplb_rasterize_workspace_bytes(m, B, W, H) || ((uintptr_t)ws & 255)))
    return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": workspace missing, too small or not 256-byte aligned");
  return SMPLB_OK;
}

extern "C" int32_t smplb_rasterize(const smplb_mesh* m, int32_t B, const float* ndc, int32_t W, int32_t H, float* bary,
                                   int32_t* ids, float* zbuf, void* workspace, size_t workspace_bytes, void* stream) {
  int32_t rc = raster_check("smplb_rasterize", m, B, W, H, workspace, workspace_bytes);
  if (rc != SMPLB_OK || B == 0) return rc;
  if (!ndc) return fail(SMPLB_E_INVALID_ARG, "smplb_rasterize: null vertices");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t npix = (size_t)B * W * H;
  char* w = (char*)workspace;
  unsigned long long* keys = (unsigned long long*)w;
  int* big_count = (int*)(w + ru256(npix * 8));
  int* big_list = (int*)(w + ru256(npix * 8) + 256);
  const size_t n4 = (npix + 3) / 4;
  // KEY_EMPTY is all ones: the key area is a byte fill (the copy engine's fill runs at write speed, above a
  // store kernel's), the big-triangle counter a second one
  SMPLB_CUDA_CHECK(cudaMemsetAsync(keys, 0xff, npix * 8, st));
  SMPLB_CUDA_CHECK(cudaMemsetAsync(big_count, 0, sizeof(int), st));
  if (m->T > 0) {
    raster_tri_kernel<<<dim3((m->T + 127) / 128, B), 128, 0, st>>>(m->T, m->V, W, H, ndc, m->tri, keys, big_list, big_count);
    raster_big_kernel<<<592, 256, 0, st>>>(m->T, m->V, W, H, ndc, m->tri, keys, big_list, big_count);
  }
  const bool vec = ((size_t)W * H) % 4 == 0 && (((uintptr_t)bary | (uintptr_t)ids | (uintptr_t)zbuf) & 15) == 0;
  if (vec)
    raster_resolve4_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(B, m->T, m->V, W, H, ndc, m->tri, keys, bary, ids, zbuf);
  else
    raster_resolve_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(B, m->T, m->V, W, H, ndc, m->tri, keys, bary, ids, zbuf);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_rasterize_backward(const smplb_mesh* m, int32_t B, const float* ndc, const float* bary,
                                            const int32_t* ids, const float* g_bary, int32_t W, int32_t H, float* g_ndc,
                                            void* workspace, size_t workspace_bytes, void* stream) {
  int32_t rc = raster_check("smplb_rasterize_backward", m, B, W, H, workspace, workspace_bytes);
  if (rc != SMPLB_OK || B == 0) return rc;
  if (!ndc || !bary || !ids || !g_bary || !g_ndc) return fail(SMPLB_E_INVALID_ARG, "smplb_rasterize_backward: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t npix = (size_t)B * W * H, bt = (size_t)B * (size_t)(m->T > 0 ? m->T : 1);
  float* tri_grad = (float*)((char*)workspace + ru256(npix * 8) + 256 + ru256(bt * 4));
  if (m->T > 0)
    raster_bwd_tri_kernel<<<dim3((m->T + 127) / 128, B), 128, 0, st>>>(m->T, m->V, W, H, ndc, m->tri, bary, ids, g_bary, tri_grad);
  raster_bwd_vtx_kernel<<<dim3((m->V + 127) / 128, B), 128, 0, st>>>(m->T, m->V, m->vptr, m->vent, tri_grad, g_ndc);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// ---- fused interpolation + diffuse Phong shading (the renderer path pose_3d drives: utils.py:20-58) ----------------
// rasterize_triangles.py:73-86 per vertex (clip transform, guarded perspective divide), then per pixel:
// winner decode (as raster_resolve_kernel) -> perspective-correct barycentrics (:125-144) -> interpolation of
// normals | positions | diffuse colours with the -1 background (:147-164, mesh_renderer.py:371-373) ->
// phong_shader without the specular term (mesh_renderer.py:84-118,158-162), rows reversed (:162).
// One pass: the per-pixel barycentrics / ids never go to HBM, the only output is RGBA (16 B / pixel).
namespace smplb {

__global__ void __launch_bounds__(256) clip_transform_kernel(int B, int V, const float* __restrict__ verts,
                                                             const float* __restrict__ clip_tf, float* __restrict__ ndc,
                                                             float* __restrict__ clip_w) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * V) return;
  const int b = (int)(i / V);
  const float* M = clip_tf + (size_t)b * 16;
  const float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
  float c[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) c[r] = M[4 * r] * x + M[4 * r + 1] * y + M[4 * r + 2] * z + M[4 * r + 3];
  const float sgn = c[3] > 0.f ? 1.f : (c[3] < 0.f ? -1.f : 0.f);
  const float w = fmaxf(fabsf(c[3]), 1e-6f) * sgn;                       // :81-84
  ndc[3 * i] = c[0] / w; ndc[3 * i + 1] = c[1] / w; ndc[3 * i + 2] = c[2] / w;   // :85-86
  clip_w[i] = w;
}

__global__ void __launch_bounds__(256) raster_shade_kernel(int B, int T, int V, int W, int H, int L,
                                                           const float* __restrict__ ndc, const float* __restrict__ clip_w,
                                                           const int* __restrict__ tri,
                                                           const unsigned long long* __restrict__ keys,
                                                           const float* __restrict__ verts, const float* __restrict__ normals,
                                                           const float* __restrict__ diffuse, const float* __restrict__ light_pos,
                                                           const float* __restrict__ light_int, const float* __restrict__ ambient,
                                                           float* __restrict__ rgba) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * W * H) return;
  const int b = (int)(i / ((size_t)W * H)), p = (int)(i % ((size_t)W * H));
  const int ix = p % W, iy = p / W;
  const unsigned long long key = __ldcs(keys + i);
  float bc[3] = {0.f, 0.f, 0.f}, z;
  int t = 0;
  if (key == KEY_EMPTY) {
    // background: zero barycentrics give alpha = 0, every attribute = the background value -1 (:161-164), hence
    // mask = 0 (:387) and the pixel is (0, 0, 0, 0) whatever the lights are (finite vertex attributes assumed)
    __stcs(reinterpret_cast<float4*>(rgba) + ((size_t)b * H + (H - 1 - iy)) * W + ix, make_float4(0.f, 0.f, 0.f, 0.f));
    return;
  }
  {
    t = (int)(~(uint32_t)key);
    TriSetup s;
    tri_setup(ndc + (size_t)b * V * 3, tri, t, W, H, s);
    tri_eval(s, ix, iy, bc[0], bc[1], bc[2], z);
  }
  float a9[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, alpha = 0.f;
  if (T > 0) {
    int vid[3];
    float wgt[3], fac = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      vid[k] = __ldg(tri + 3 * t + k);
      wgt[k] = bc[k] * (1.0f / clip_w[(size_t)b * V + vid[k]]);          // :125-135
      fac += wgt[k];
    }
    const float den = fmaxf(fac, 1e-6f);                                 // :139-144
    float csum = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float c = wgt[k] / den;
      csum += c;
      const size_t o = ((size_t)b * V + vid[k]) * 3;
#pragma unroll
      for (int e = 0; e < 3; ++e) {
        a9[e] += normals[o + e] * c;                                     // attribute order: mesh_renderer.py:361
        a9[3 + e] += verts[o + e] * c;
        a9[6 + e] += (diffuse ? diffuse[o + e] : 1.0f) * c;
      }
    }
    alpha = fminf(fmaxf(2.0f * csum, 0.f), 1.f);                         // :157-159
  }
#pragma unroll
  for (int e = 0; e < 9; ++e) a9[e] = alpha * a9[e] + (1.0f - alpha) * -1.0f;   // :161-164, background -1
  const float nn = rsqrtf(fmaxf(a9[0] * a9[0] + a9[1] * a9[1] + a9[2] * a9[2], 1e-12f));   // mesh_renderer.py:376
  const float n[3] = {a9[0] * nn, a9[1] * nn, a9[2] * nn};
  const float mask = (a9[6] >= 0.f || a9[7] >= 0.f || a9[8] >= 0.f) ? 1.f : 0.f;           // :387
  float col[3] = {0.f, 0.f, 0.f};
  if (ambient)
#pragma unroll
    for (int e = 0; e < 3; ++e) col[e] = ambient[3 * b + e] * a9[6 + e];                    // :92-94
  for (int l = 0; l < L; ++l) {
    const float* lp = light_pos + ((size_t)b * L + l) * 3;
    const float* li = light_int + ((size_t)b * L + l) * 3;
    const float d[3] = {lp[0] - a9[3], lp[1] - a9[4], lp[2] - a9[5]};
    const float dn = rsqrtf(fmaxf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], 1e-12f));        // :100-102
    const float ndl = fminf(fmaxf((n[0] * d[0] + n[1] * d[1] + n[2] * d[2]) * dn, 0.f), 1.f);   // :106-109
#pragma unroll
    for (int e = 0; e < 3; ++e) col[e] += a9[6 + e] * ndl * li[e];                          // :112-118
  }
  float4 o;
  o.x = mask > 0.5f ? col[0] : 0.f;                                                        // :158-161
  o.y = mask > 0.5f ? col[1] : 0.f;
  o.z = mask > 0.5f ? col[2] : 0.f;
  o.w = mask;
  reinterpret_cast<float4*>(rgba)[((size_t)b * H + (H - 1 - iy)) * W + ix] = o;            // :162 tf.reverse(axis=1)
}

}  // namespace smplb

// raster workspace | ndc [B,V,3] | clip_w [B,V]
extern "C" size_t smplb_render_workspace_bytes(const smplb_mesh* m, int32_t B, int32_t W, int32_t H) {
  if (!m || B <= 0) return 0;
  return smplb_rasterize_workspace_bytes(m, B, W, H) + ru256((size_t)B * m->V * 12) + ru256((size_t)B * m->V * 4);
}

extern "C" int32_t smplb_render_diffuse(const smplb_mesh* m, int32_t B, const float* verts, const float* normals,
                                        const float* diffuse, const float* clip_transforms, const float* light_positions,
                                        const float* light_intensities, int32_t L, const float* ambient, int32_t W,
                                        int32_t H, float* rgba, void* workspace, size_t workspace_bytes, void* stream) {
  if (!m) return fail(SMPLB_E_INVALID_ARG, "smplb_render_diffuse: null mesh");
  if (int32_t rc = check_current_device(m->device, "smplb_render_diffuse")) return rc;
  if (B < 0 || L < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_render_diffuse: negative batch or light count");
  if (W <= 0 || H <= 0) return fail(SMPLB_E_INVALID_ARG, "smplb_render_diffuse: image width and height must be > 0");
  if (B == 0) return SMPLB_OK;
  if (B > 65535) return fail(SMPLB_E_UNSUPPORTED, "smplb_render_diffuse: at most 65535 images per call");
  if (!verts || !normals || !clip_transforms || !rgba || (L > 0 && (!light_positions || !light_intensities)))
    return fail(SMPLB_E_INVALID_ARG, "smplb_render_diffuse: null pointer");
  if (!workspace || workspace_bytes < smplb_render_workspace_bytes(m, B, W, H) || ((uintptr_t)workspace & 255))
    return fail(SMPLB_E_INVALID_ARG, "smplb_render_diffuse: workspace missing, too small or not 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t rws = smplb_rasterize_workspace_bytes(m, B, W, H), npix = (size_t)B * W * H;
  char* w = (char*)workspace;
  float* ndc = (float*)(w + rws);
  float* clip_w = (float*)(w + rws + ru256((size_t)B * m->V * 12));
  unsigned long long* keys = (unsigned long long*)w;
  int* big_count = (int*)(w + ru256(npix * 8));
  int* big_list = (int*)(w + ru256(npix * 8) + 256);
  clip_transform_kernel<<<(unsigned)(((size_t)B * m->V + 255) / 256), 256, 0, st>>>(B, m->V, verts, clip_transforms, ndc, clip_w);
  SMPLB_CUDA_CHECK(cudaMemsetAsync(keys, 0xff, npix * 8, st));
  SMPLB_CUDA_CHECK(cudaMemsetAsync(big_count, 0, sizeof(int), st));
  if (m->T > 0) {
    raster_tri_kernel<<<dim3((m->T + 127) / 128, B), 128, 0, st>>>(m->T, m->V, W, H, ndc, m->tri, keys, big_list, big_count);
    raster_big_kernel<<<592, 256, 0, st>>>(m->T, m->V, W, H, ndc, m->tri, keys, big_list, big_count);
  }
  raster_shade_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(B, m->T, m->V, W, H, L, ndc, clip_w, m->tri, keys, verts,
                                                                     normals, diffuse, light_positions, light_intensities,
                                                                     ambient, rgba);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
