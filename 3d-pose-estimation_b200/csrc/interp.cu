// The element-wise graph of deps/tf_mesh_renderer/rasterize_triangles.py around the rasteriser op, forward and adjoint
// (SURVEY.md 8f N1), so that the differentiable renderer runs on kernels end to end:
//   smplb_clip_transform(_backward)   :73-86   homogeneous clip transform + guarded perspective divide
//   smplb_interpolate(_backward)      :119-164 perspective-correct barycentrics, attribute interpolation, alpha from the
//                                              barycentric sum, background compositing
// (the rasteriser between them: csrc/raster.cu; the shader after them: csrc/shade.cu).  TF differentiates these ops with
// its stock gradients; the adjoints below are those formulas.  The gather of vertex attributes turns into a scatter in
// the adjoint: float atomicAdd into zero-filled g_attr / g_clip_w (what TF's GPU gather gradient does too), so the LAST
// BITS of those two gradients depend on the order in which the pixels arrive -- the one place in this library that is
// not run-to-run bit-identical; g_bary and everything downstream of it (rasteriser adjoint) is.
#include "common.cuh"

namespace smplb {

constexpr float IT_W_MIN = 1e-6f;        // _MINIMUM_PERSPECTIVE_DIVIDE_THRESHOLD, rasterize_triangles.py:27
constexpr float IT_FACTOR_MIN = 1e-6f;   // _MINIMUM_REWEIGHTING_THRESHOLD, :26

__global__ void __launch_bounds__(256) clip_fwd_kernel(int B, int V, const float* __restrict__ verts,
                                                       const float* __restrict__ clip_tf, float* __restrict__ ndc,
                                                       float* __restrict__ clip_w) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * V) return;
  const float* M = clip_tf + (size_t)(i / V) * 16;
  const float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
  float c[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) c[r] = M[4 * r] * x + M[4 * r + 1] * y + M[4 * r + 2] * z + M[4 * r + 3];
  const float sgn = c[3] > 0.f ? 1.f : (c[3] < 0.f ? -1.f : 0.f);
  const float w = fmaxf(fabsf(c[3]), IT_W_MIN) * sgn;                    // :81-84
  ndc[3 * i] = c[0] / w; ndc[3 * i + 1] = c[1] / w; ndc[3 * i + 2] = c[2] / w;   // :85-86
  clip_w[i] = w;
}

// g_ndc [B,V,3], g_clip_w [B,V] (nullable) -> g_verts [B,V,3].  The clip matrices are constants.
__global__ void __launch_bounds__(256) clip_bwd_kernel(int B, int V, const float* __restrict__ verts,
                                                       const float* __restrict__ clip_tf, const float* __restrict__ g_ndc,
                                                       const float* __restrict__ g_clip_w, float* __restrict__ g_verts) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * V) return;
  const float* M = clip_tf + (size_t)(i / V) * 16;
  const float x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
  float c[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) c[r] = M[4 * r] * x + M[4 * r + 1] * y + M[4 * r + 2] * z + M[4 * r + 3];
  const float sgn = c[3] > 0.f ? 1.f : (c[3] < 0.f ? -1.f : 0.f);
  const float w = fmaxf(fabsf(c[3]), IT_W_MIN) * sgn;
  const float g0 = g_ndc[3 * i], g1 = g_ndc[3 * i + 1], g2 = g_ndc[3 * i + 2];
  float gc[4];
  gc[0] = g0 / w; gc[1] = g1 / w; gc[2] = g2 / w;
  // w = max(|c3|, eps) sign(c3): d w / d c3 = 1 where |c3| >= eps (the maximum passes its gradient at a tie), sign() is flat
  float gw = -(g0 * c[0] + g1 * c[1] + g2 * c[2]) / (w * w) + (g_clip_w ? g_clip_w[i] : 0.f);
  gc[3] = fabsf(c[3]) >= IT_W_MIN ? gw : 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) g_verts[3 * i + k] = M[k] * gc[0] + M[4 + k] * gc[1] + M[8 + k] * gc[2] + M[12 + k] * gc[3];
}

struct InterpParams {
  int B, V, C, H, W;
  const float* attr;      // [B,V,C]
  const float* clip_w;    // [B,V]
  const int* tri;         // [T,3]
  const float* bary;      // [B,H,W,3]
  const int* ids;         // [B,H,W]
  const float* bg;        // [C]
  float* out;             // [B,H,W,C]
  const float* g_out;
  float* g_attr;          // zero-filled, atomicAdd
  float* g_clip_w;        // zero-filled, atomicAdd
  float* g_bary;          // [B,H,W,3]
};

struct PixelW {
  int vid[3];
  float cwinv[3], weighted[3], corr[3], factor, den, csum, alpha;
};

__device__ __forceinline__ void pixel_weights(const InterpParams& p, int b, size_t pix, PixelW& q) {
  const int t = p.ids[pix];
  q.factor = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    q.vid[k] = __ldg(p.tri + 3 * t + k);
    q.cwinv[k] = 1.0f / p.clip_w[(size_t)b * p.V + q.vid[k]];            // :125-129
    q.weighted[k] = p.bary[3 * pix + k] * q.cwinv[k];                    // :133-135
    q.factor += q.weighted[k];
  }
  q.den = fmaxf(q.factor, IT_FACTOR_MIN);                                // :139-144
  q.csum = 0.f;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    q.corr[k] = q.weighted[k] / q.den;
    q.csum += q.corr[k];
  }
  q.alpha = fminf(fmaxf(2.0f * q.csum, 0.f), 1.f);                       // :157-159
}

__global__ void __launch_bounds__(256) interp_fwd_kernel(const InterpParams p) {
  const size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= (size_t)p.B * p.H * p.W) return;
  const int b = (int)(pix / ((size_t)p.H * p.W));
  PixelW q;
  pixel_weights(p, b, pix, q);
  const float* a0 = p.attr + ((size_t)b * p.V + q.vid[0]) * p.C;
  const float* a1 = p.attr + ((size_t)b * p.V + q.vid[1]) * p.C;
  const float* a2 = p.attr + ((size_t)b * p.V + q.vid[2]) * p.C;
  float* o = p.out + pix * p.C;
  for (int c = 0; c < p.C; ++c) {
    const float img = a0[c] * q.corr[0] + a1[c] * q.corr[1] + a2[c] * q.corr[2];      // :147-152
    o[c] = q.alpha * img + (1.0f - q.alpha) * p.bg[c];                               // :161-164
  }
}

__global__ void __launch_bounds__(256) interp_bwd_kernel(const InterpParams p) {
  const size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= (size_t)p.B * p.H * p.W) return;
  const int b = (int)(pix / ((size_t)p.H * p.W));
  PixelW q;
  pixel_weights(p, b, pix, q);
  const float* a[3];
  float* ga[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    a[k] = p.attr + ((size_t)b * p.V + q.vid[k]) * p.C;
    ga[k] = p.g_attr + ((size_t)b * p.V + q.vid[k]) * p.C;
  }
  const float* go = p.g_out + pix * p.C;
  float gcorr[3] = {0.f, 0.f, 0.f}, galpha = 0.f;
  for (int c = 0; c < p.C; ++c) {
    const float g = go[c];
    const float img = a[0][c] * q.corr[0] + a[1][c] * q.corr[1] + a[2][c] * q.corr[2];
    galpha += g * (img - p.bg[c]);
    const float gimg = q.alpha * g;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      gcorr[k] += gimg * a[k][c];
      if (gimg != 0.f && q.corr[k] != 0.f) atomicAdd(ga[k] + c, gimg * q.corr[k]);
    }
  }
  const float two = 2.0f * q.csum;
  const float gsum = (two >= 0.f && two <= 1.f) ? 2.0f * galpha : 0.f;   // clip to [0,1]
  float gden = 0.f, gw[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    gcorr[k] += gsum;
    gw[k] = gcorr[k] / q.den;
    gden -= gcorr[k] * q.weighted[k] / (q.den * q.den);
  }
  const float gfac = q.factor >= IT_FACTOR_MIN ? gden : 0.f;             // maximum(factor, eps)
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    gw[k] += gfac;
    p.g_bary[3 * pix + k] = gw[k] * q.cwinv[k];
    const float gcw = -(gw[k] * p.bary[3 * pix + k]) * q.cwinv[k] * q.cwinv[k];   // d (1 / w) = -1 / w^2
    if (gcw != 0.f) atomicAdd(p.g_clip_w + (size_t)b * p.V + q.vid[k], gcw);
  }
}

static int32_t interp_check(const char* who, const smplb_mesh* m, int B, int C, int H, int W) {
  if (!m) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": null mesh");
  if (int32_t rc = check_current_device(m->device, who)) return rc;
  if (B < 0 || C <= 0 || H <= 0 || W <= 0) return fail(SMPLB_E_BAD_SHAPE, std::string(who) + ": need B >= 0 and C, H, W > 0");
  if (m->T <= 0) return fail(SMPLB_E_UNSUPPORTED, std::string(who) + ": mesh without triangles");
  return SMPLB_OK;
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_clip_transform(int32_t B, int32_t V, const float* verts, const float* clip_transforms, float* ndc,
                                        float* clip_w, void* stream) {
  if (B < 0 || V < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_clip_transform: negative size");
  if ((size_t)B * V == 0) return SMPLB_OK;
  if (!verts || !clip_transforms || !ndc || !clip_w) return fail(SMPLB_E_INVALID_ARG, "smplb_clip_transform: null pointer");
  clip_fwd_kernel<<<(unsigned)(((size_t)B * V + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, V, verts, clip_transforms, ndc,
                                                                                            clip_w);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_clip_transform_backward(int32_t B, int32_t V, const float* verts, const float* clip_transforms,
                                                 const float* g_ndc, const float* g_clip_w, float* g_verts, void* stream) {
  if (B < 0 || V < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_clip_transform_backward: negative size");
  if ((size_t)B * V == 0) return SMPLB_OK;
  if (!verts || !clip_transforms || !g_ndc || !g_verts)
    return fail(SMPLB_E_INVALID_ARG, "smplb_clip_transform_backward: null pointer");
  clip_bwd_kernel<<<(unsigned)(((size_t)B * V + 255) / 256), 256, 0, (cudaStream_t)stream>>>(B, V, verts, clip_transforms,
                                                                                            g_ndc, g_clip_w, g_verts);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_interpolate(const smplb_mesh* m, int32_t B, int32_t C, const float* attributes, const float* clip_w,
                                     const float* bary, const int32_t* ids, const float* background, int32_t W, int32_t H,
                                     float* images, void* stream) {
  if (int32_t rc = interp_check("smplb_interpolate", m, B, C, H, W)) return rc;
  if (B == 0) return SMPLB_OK;
  if (!attributes || !clip_w || !bary || !ids || !background || !images)
    return fail(SMPLB_E_INVALID_ARG, "smplb_interpolate: null pointer");
  InterpParams p = {};
  p.B = B; p.V = m->V; p.C = C; p.H = H; p.W = W;
  p.attr = attributes; p.clip_w = clip_w; p.tri = m->tri; p.bary = bary; p.ids = ids; p.bg = background; p.out = images;
  const size_t npix = (size_t)B * H * W;
  interp_fwd_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_interpolate_backward(const smplb_mesh* m, int32_t B, int32_t C, const float* attributes,
                                              const float* clip_w, const float* bary, const int32_t* ids,
                                              const float* background, int32_t W, int32_t H, const float* g_images,
                                              float* g_attributes, float* g_clip_w, float* g_bary, void* stream) {
  if (int32_t rc = interp_check("smplb_interpolate_backward", m, B, C, H, W)) return rc;
  if (B == 0) return SMPLB_OK;
  if (!attributes || !clip_w || !bary || !ids || !background || !g_images || !g_attributes || !g_clip_w || !g_bary)
    return fail(SMPLB_E_INVALID_ARG, "smplb_interpolate_backward: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  SMPLB_CUDA_CHECK(cudaMemsetAsync(g_attributes, 0, (size_t)B * m->V * C * sizeof(float), st));
  SMPLB_CUDA_CHECK(cudaMemsetAsync(g_clip_w, 0, (size_t)B * m->V * sizeof(float), st));
  InterpParams p = {};
  p.B = B; p.V = m->V; p.C = C; p.H = H; p.W = W;
  p.attr = attributes; p.clip_w = clip_w; p.tri = m->tri; p.bary = bary; p.ids = ids; p.bg = background;
  p.g_out = g_images; p.g_attr = g_attributes; p.g_clip_w = g_clip_w; p.g_bary = g_bary;
  const size_t npix = (size_t)B * H * W;
  interp_bwd_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
