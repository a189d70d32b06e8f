// Phong shading with the specular term and its adjoint (deps/tf_mesh_renderer/mesh_renderer.py:23-162).
// Thread = pixel, all lights in registers.  The only cross-pixel coupling is the reference's l2_normalize of
// reflection . camera over the PIXEL axis (:134-136): one sum of squares per (image, light) in the forward pass, one more
// sum (g_u . rdc) in the backward pass; both are fixed-order (block tree, then one finishing block per image).
//   forward   phong_stats_kernel (specular only) -> phong_finish_kernel -> phong_shade_kernel
//   backward  phong_bwd_stats_kernel (specular only) -> phong_finish_kernel -> phong_bwd_kernel
#include "common.cuh"

namespace smplb {

constexpr int PH_THREADS = 256;
constexpr int PH_LMAX = 8;
constexpr float PH_EPS = 1e-12f;   // tf.nn.l2_normalize epsilon

struct PhongParams {
  smplb_phong_io io;
  int P;            // pixels per image
  int nblk;         // blocks per image
  float* partial;   // [B][L][nblk]
  float* S;         // [B][L]  sum over pixels of rdc^2
  float* Tsum;      // [B][L]  sum over pixels of g_u * rdc (backward)
  const float* g_rgba;
  float* rgba;
  float* g_attr;
  float* g_alphas;
};

// y = x / sqrt(max(|x|^2, eps)); returns 1 / sqrt(max(|x|^2, eps))
__device__ __forceinline__ float l2n(const float x[3], float y[3], float& ss) {
  ss = x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
  const float inv = rsqrtf(fmaxf(ss, PH_EPS));
  y[0] = x[0] * inv; y[1] = x[1] * inv; y[2] = x[2] * inv;
  return inv;
}
// adjoint of l2n: gx = inv * (gy - y (y.gy)) when |x|^2 > eps, else inv * gy (the clamp holds the denominator constant)
__device__ __forceinline__ void l2n_bwd(const float y[3], float inv, float ss, const float gy[3], float gx[3]) {
  const float d = ss > PH_EPS ? y[0] * gy[0] + y[1] * gy[1] + y[2] * gy[2] : 0.f;
  gx[0] = inv * (gy[0] - y[0] * d); gx[1] = inv * (gy[1] - y[1] * d); gx[2] = inv * (gy[2] - y[2] * d);
}

// Per-pixel state shared by all passes.
struct Pixel {
  float nraw[3], n[3], ninv, nss;   // normal before / after the optional normalisation
  float pos[3];
  float tc[3], tcinv, tcss;         // direction to the camera
};

__device__ __forceinline__ void load_pixel(const smplb_phong_io& io, const float* a, const float* cam, bool spec, Pixel& px) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    px.nraw[c] = a[io.off_normal + c];
    px.pos[c] = a[io.off_position + c];
  }
  if (io.normalize_normals) {
    px.ninv = l2n(px.nraw, px.n, px.nss);
  } else {
#pragma unroll
    for (int c = 0; c < 3; ++c) px.n[c] = px.nraw[c];
    px.ninv = 1.f; px.nss = 1.f;
  }
  if (spec) {
    const float d[3] = {cam[0] - px.pos[0], cam[1] - px.pos[1], cam[2] - px.pos[2]};
    px.tcinv = l2n(d, px.tc, px.tcss);
  }
}

// One light at one pixel: everything the forward needs, kept for the adjoint.
struct LightTerm {
  float tl[3], tlinv, tlss;   // direction to the light
  float d, ndl;               // n . tl before / after the clip to [0,1]
  float mr[3], minv, mss;     // mirror reflection direction
  float rdc;                  // to_camera . mirror
};

__device__ __forceinline__ void light_term(const Pixel& px, const float* lp, bool spec, LightTerm& t) {
  const float dl[3] = {lp[0] - px.pos[0], lp[1] - px.pos[1], lp[2] - px.pos[2]};
  t.tlinv = l2n(dl, t.tl, t.tlss);                                                       // :100-102
  t.d = px.n[0] * t.tl[0] + px.n[1] * t.tl[1] + px.n[2] * t.tl[2];
  t.ndl = fminf(fmaxf(t.d, 0.f), 1.f);                                                   // :106-109
  if (spec) {
    const float m[3] = {2.f * t.ndl * px.n[0] - t.tl[0], 2.f * t.ndl * px.n[1] - t.tl[1], 2.f * t.ndl * px.n[2] - t.tl[2]};
    t.minv = l2n(m, t.mr, t.mss);                                                        // :123-126
    t.rdc = px.tc[0] * t.mr[0] + px.tc[1] * t.mr[1] + px.tc[2] * t.mr[2];                // :129-132
  }
}

// fixed-order block sums of L values -> partial[(b * L + l) * nblk + blockIdx.x]
__device__ __forceinline__ void block_partials(const float (&v)[PH_LMAX], int L, float* partial, int b, int nblk) {
  __shared__ float red[PH_THREADS / 32][PH_LMAX];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int l = 0; l < L; ++l) {
    float x = v[l];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) red[warp][l] = x;
  }
  __syncthreads();
  if (threadIdx.x < L) {
    float s = 0.f;
    for (int w = 0; w < PH_THREADS / 32; ++w) s += red[w][threadIdx.x];
    partial[((size_t)b * L + threadIdx.x) * nblk + blockIdx.x] = s;
  }
}

// S[b][l] = sum over pixels of rdc^2
__global__ void __launch_bounds__(PH_THREADS) phong_stats_kernel(const PhongParams p) {
  const smplb_phong_io& io = p.io;
  const int b = blockIdx.y, i = blockIdx.x * PH_THREADS + threadIdx.x;
  float v[PH_LMAX];
#pragma unroll
  for (int l = 0; l < PH_LMAX; ++l) v[l] = 0.f;
  if (i < p.P) {
    Pixel px;
    load_pixel(io, io.attr + ((size_t)b * p.P + i) * io.C, io.camera + b * 3, true, px);
    for (int l = 0; l < io.L; ++l) {
      LightTerm t;
      light_term(px, io.light_positions + ((size_t)b * io.L + l) * 3, true, t);
      v[l] = t.rdc * t.rdc;
    }
  }
  block_partials(v, io.L, p.partial, b, p.nblk);
}

// one block per (image, light): fixed-order sum of the block partials
__global__ void __launch_bounds__(PH_THREADS) phong_finish_kernel(int nblk, const float* __restrict__ partial,
                                                                   float* __restrict__ out) {
  __shared__ float red[PH_THREADS / 32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblk; i += PH_THREADS) acc += partial[(size_t)blockIdx.x * nblk + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < PH_THREADS / 32; ++w) s += red[w];
    out[blockIdx.x] = s;
  }
}

// specular weight of one light at one pixel: r = where(ndl != 0, clip(rdc / ||rdc||_pixels, 0, 1), 0) (:134-143)
__device__ __forceinline__ float spec_base(const LightTerm& t, float S, float& u, float& ninv) {
  ninv = rsqrtf(fmaxf(S, PH_EPS));
  u = t.rdc * ninv;
  return t.ndl != 0.f ? fminf(fmaxf(u, 0.f), 1.f) : 0.f;
}

__device__ __forceinline__ float pixel_alpha(const smplb_phong_io& io, const float* a, size_t pix) {
  if (io.alphas) return io.alphas[pix];
  const float* d = a + io.off_diffuse;
  return (d[0] >= 0.f || d[1] >= 0.f || d[2] >= 0.f) ? 1.f : 0.f;                       // mesh_renderer.py:387
}

__global__ void __launch_bounds__(PH_THREADS) phong_shade_kernel(const PhongParams p) {
  const smplb_phong_io& io = p.io;
  const int b = blockIdx.y, i = blockIdx.x * PH_THREADS + threadIdx.x;
  if (i >= p.P) return;
  const bool spec = io.off_specular >= 0;
  const float* a = io.attr + ((size_t)b * p.P + i) * io.C;
  Pixel px;
  load_pixel(io, a, spec ? io.camera + b * 3 : nullptr, spec, px);
  float D[3], Sc[3] = {0.f, 0.f, 0.f}, out[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < 3; ++c) D[c] = a[io.off_diffuse + c];
  float shin = 0.f;
  if (spec) {
#pragma unroll
    for (int c = 0; c < 3; ++c) Sc[c] = a[io.off_specular + c];
    shin = io.off_shininess >= 0 ? a[io.off_shininess] : io.shininess[b];
  }
  if (io.ambient) {
#pragma unroll
    for (int c = 0; c < 3; ++c) out[c] = io.ambient[b * 3 + c] * D[c];                   // :92-94
  }
  for (int l = 0; l < io.L; ++l) {
    LightTerm t;
    light_term(px, io.light_positions + ((size_t)b * io.L + l) * 3, spec, t);
    const float* I = io.light_intensities + ((size_t)b * io.L + l) * 3;
    float sp = 0.f;
    if (spec) {
      float u, ninv;
      sp = powf(spec_base(t, p.S[b * io.L + l], u, ninv), shin);                         // :148-152
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) out[c] += D[c] * t.ndl * I[c] + Sc[c] * sp * I[c];       // :112-118, :153-157
  }
  const float alpha = pixel_alpha(io, a, (size_t)b * p.P + i);
  const int y = i / io.W, x = i - y * io.W;
  float4* dst = reinterpret_cast<float4*>(p.rgba) + ((size_t)b * p.P + (size_t)(io.H - 1 - y) * io.W + x);   // :162
  const bool valid = alpha > 0.5f;                                                       // :160-161
  *dst = make_float4(valid ? out[0] : 0.f, valid ? out[1] : 0.f, valid ? out[2] : 0.f, alpha);
}

// gradient reaching u = rdc / ||rdc|| of light l at this pixel
__device__ __forceinline__ float grad_u(const LightTerm& t, float S, float shin, float gspec, float& r, float& ninv) {
  float u;
  r = spec_base(t, S, u, ninv);
  if (t.ndl == 0.f || u < 0.f || u > 1.f || shin == 0.f) return 0.f;   // where / clip pass nothing; d pow(r, 0) / dr = 0
  return gspec * shin * powf(r, shin - 1.f);
}

__device__ __forceinline__ void load_grad(const PhongParams& p, int b, int i, float alpha, float g[3], float& galpha) {
  const smplb_phong_io& io = p.io;
  const int y = i / io.W, x = i - y * io.W;
  const float4 v = reinterpret_cast<const float4*>(p.g_rgba)[(size_t)b * p.P + (size_t)(io.H - 1 - y) * io.W + x];
  const bool valid = alpha > 0.5f;
  g[0] = valid ? v.x : 0.f; g[1] = valid ? v.y : 0.f; g[2] = valid ? v.z : 0.f;
  galpha = v.w;
}

// Tsum[b][l] = sum over pixels of g_u * rdc
__global__ void __launch_bounds__(PH_THREADS) phong_bwd_stats_kernel(const PhongParams p) {
  const smplb_phong_io& io = p.io;
  const int b = blockIdx.y, i = blockIdx.x * PH_THREADS + threadIdx.x;
  float v[PH_LMAX];
#pragma unroll
  for (int l = 0; l < PH_LMAX; ++l) v[l] = 0.f;
  if (i < p.P) {
    const float* a = io.attr + ((size_t)b * p.P + i) * io.C;
    Pixel px;
    load_pixel(io, a, io.camera + b * 3, true, px);
    float g[3], galpha;
    load_grad(p, b, i, pixel_alpha(io, a, (size_t)b * p.P + i), g, galpha);
    const float shin = io.off_shininess >= 0 ? a[io.off_shininess] : io.shininess[b];
    for (int l = 0; l < io.L; ++l) {
      LightTerm t;
      light_term(px, io.light_positions + ((size_t)b * io.L + l) * 3, true, t);
      const float* I = io.light_intensities + ((size_t)b * io.L + l) * 3;
      const float gspec = g[0] * a[io.off_specular] * I[0] + g[1] * a[io.off_specular + 1] * I[1] +
                          g[2] * a[io.off_specular + 2] * I[2];
      float r, ninv;
      v[l] = grad_u(t, p.S[b * io.L + l], shin, gspec, r, ninv) * t.rdc;
    }
  }
  block_partials(v, io.L, p.partial, b, p.nblk);
}

__global__ void __launch_bounds__(PH_THREADS) phong_bwd_kernel(const PhongParams p) {
  const smplb_phong_io& io = p.io;
  const int b = blockIdx.y, i = blockIdx.x * PH_THREADS + threadIdx.x;
  if (i >= p.P) return;
  const bool spec = io.off_specular >= 0;
  const float* a = io.attr + ((size_t)b * p.P + i) * io.C;
  float* ga = p.g_attr + ((size_t)b * p.P + i) * io.C;
  Pixel px;
  load_pixel(io, a, spec ? io.camera + b * 3 : nullptr, spec, px);
  float D[3], Sc[3] = {0.f, 0.f, 0.f}, g[3], galpha;
#pragma unroll
  for (int c = 0; c < 3; ++c) D[c] = a[io.off_diffuse + c];
  float shin = 0.f;
  if (spec) {
#pragma unroll
    for (int c = 0; c < 3; ++c) Sc[c] = a[io.off_specular + c];
    shin = io.off_shininess >= 0 ? a[io.off_shininess] : io.shininess[b];
  }
  load_grad(p, b, i, pixel_alpha(io, a, (size_t)b * p.P + i), g, galpha);
  if (p.g_alphas) p.g_alphas[(size_t)b * p.P + i] = galpha;

  float gD[3] = {0.f, 0.f, 0.f}, gS[3] = {0.f, 0.f, 0.f}, gn[3] = {0.f, 0.f, 0.f}, gpos[3] = {0.f, 0.f, 0.f},
        gtc[3] = {0.f, 0.f, 0.f}, gshin = 0.f;
  if (io.ambient) {
#pragma unroll
    for (int c = 0; c < 3; ++c) gD[c] = g[c] * io.ambient[b * 3 + c];
  }
  for (int l = 0; l < io.L; ++l) {
    LightTerm t;
    light_term(px, io.light_positions + ((size_t)b * io.L + l) * 3, spec, t);
    const float* I = io.light_intensities + ((size_t)b * io.L + l) * 3;
    float gndl = 0.f, gtl[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      gD[c] += g[c] * t.ndl * I[c];
      gndl += g[c] * D[c] * I[c];
    }
    if (spec) {
      const float gspec = g[0] * Sc[0] * I[0] + g[1] * Sc[1] * I[1] + g[2] * Sc[2] * I[2];
      const float S = p.S[b * io.L + l];
      float r, ninv;
      const float gu = grad_u(t, S, shin, gspec, r, ninv);
      const float sp = powf(r, shin);
#pragma unroll
      for (int c = 0; c < 3; ++c) gS[c] += g[c] * sp * I[c];
      if (io.off_shininess >= 0 && r > 0.f) gshin += gspec * sp * logf(r);   // d r^s / ds = r^s ln r (0 at r = 0)
      // u = rdc / sqrt(max(S, eps)), S = sum over pixels of rdc^2
      const float grdc = gu * ninv - (S > PH_EPS ? t.rdc * p.Tsum[b * io.L + l] * ninv * ninv * ninv : 0.f);
      float gm[3], gmraw[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        gtc[c] += grdc * t.mr[c];
        gm[c] = grdc * px.tc[c];
      }
      l2n_bwd(t.mr, t.minv, t.mss, gm, gmraw);     // mirror = l2n(2 ndl n - tl)
      gndl += 2.f * (px.n[0] * gmraw[0] + px.n[1] * gmraw[1] + px.n[2] * gmraw[2]);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        gn[c] += 2.f * t.ndl * gmraw[c];
        gtl[c] -= gmraw[c];
      }
    }
    const float gd = (t.d >= 0.f && t.d <= 1.f) ? gndl : 0.f;   // clip to [0,1]
    float gdl[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      gn[c] += gd * t.tl[c];
      gtl[c] += gd * px.n[c];
    }
    l2n_bwd(t.tl, t.tlinv, t.tlss, gtl, gdl);      // tl = l2n(light - pos)
#pragma unroll
    for (int c = 0; c < 3; ++c) gpos[c] -= gdl[c];
  }
  if (spec) {
    float gdc[3];
    l2n_bwd(px.tc, px.tcinv, px.tcss, gtc, gdc);   // tc = l2n(camera - pos)
#pragma unroll
    for (int c = 0; c < 3; ++c) gpos[c] -= gdc[c];
  }
  float gnraw[3] = {gn[0], gn[1], gn[2]};
  if (io.normalize_normals) l2n_bwd(px.n, px.ninv, px.nss, gn, gnraw);
  for (int c = 0; c < io.C; ++c) ga[c] = 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    ga[io.off_normal + c] = gnraw[c];
    ga[io.off_position + c] = gpos[c];
    ga[io.off_diffuse + c] = gD[c];
    if (spec) ga[io.off_specular + c] = gS[c];
  }
  if (spec && io.off_shininess >= 0) ga[io.off_shininess] = gshin;
}

static int32_t phong_check(const smplb_phong_io* io, const char* who) {
  if (!io) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": null io");
  if (io->B < 0 || io->H <= 0 || io->W <= 0 || io->C < 9 || io->L < 0 || io->L > PH_LMAX)
    return fail(SMPLB_E_BAD_SHAPE, std::string(who) + ": need B >= 0, H, W > 0, C >= 9, 0 <= L <= 8");
  const int offs[5] = {io->off_normal, io->off_position, io->off_diffuse, io->off_specular, io->off_shininess};
  const int wid[5] = {3, 3, 3, 3, 1};
  for (int k = 0; k < 5; ++k) {
    if (k < 3 && offs[k] < 0) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": normal / position / diffuse offsets are required");
    if (offs[k] >= 0 && offs[k] + wid[k] > io->C) return fail(SMPLB_E_BAD_SHAPE, std::string(who) + ": channel offset outside the attribute image");
  }
  if (!io->attr || (io->L > 0 && (!io->light_positions || !io->light_intensities)))
    return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": null pointer");
  if (io->off_specular >= 0 && (!io->camera || (io->off_shininess < 0 && !io->shininess)))
    return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": a specular term needs the camera position and shininess");
  if ((int64_t)io->H * io->W > (int64_t)1 << 30) return fail(SMPLB_E_BAD_SHAPE, std::string(who) + ": image too large");
  return SMPLB_OK;
}

static size_t phong_carve(const smplb_phong_io* io, void* scratch, PhongParams* p) {
  const int P = io->H * io->W, nblk = (P + PH_THREADS - 1) / PH_THREADS;
  const size_t nS = ((size_t)(io->B > 0 ? io->B : 0) * PH_LMAX * 4 + 255) / 256 * 256;
  if (p) {
    p->io = *io;
    p->P = P;
    p->nblk = nblk;
    p->S = (float*)scratch;
    p->Tsum = (float*)((char*)scratch + nS);
    p->partial = (float*)((char*)scratch + 2 * nS);
  }
  return 2 * nS + (size_t)(io->B > 0 ? io->B : 0) * PH_LMAX * nblk * 4 + 256;
}

}  // namespace smplb

using namespace smplb;

extern "C" size_t smplb_phong_scratch_bytes(int32_t B, int32_t H, int32_t W, int32_t L) {
  smplb_phong_io io = {};
  io.B = B; io.H = H > 0 ? H : 1; io.W = W > 0 ? W : 1; io.L = L;
  return phong_carve(&io, nullptr, nullptr);
}

extern "C" int32_t smplb_phong_shade(const smplb_phong_io* io, float* rgba, void* scratch, size_t scratch_bytes,
                                     void* stream) {
  if (int32_t rc = phong_check(io, "smplb_phong_shade")) return rc;
  if (io->B == 0) return SMPLB_OK;
  if (!rgba || !scratch) return fail(SMPLB_E_INVALID_ARG, "smplb_phong_shade: null pointer");
  PhongParams p = {};
  if (scratch_bytes < phong_carve(io, scratch, &p)) return fail(SMPLB_E_INVALID_ARG, "smplb_phong_shade: scratch too small");
  p.rgba = rgba;
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid(p.nblk, io->B);
  if (io->off_specular >= 0 && io->L > 0) {
    phong_stats_kernel<<<grid, PH_THREADS, 0, st>>>(p);
    phong_finish_kernel<<<io->B * io->L, PH_THREADS, 0, st>>>(p.nblk, p.partial, p.S);
  }
  phong_shade_kernel<<<grid, PH_THREADS, 0, st>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_phong_shade_backward(const smplb_phong_io* io, const float* g_rgba, float* g_attr,
                                              float* g_alphas, void* scratch, size_t scratch_bytes, void* stream) {
  if (int32_t rc = phong_check(io, "smplb_phong_shade_backward")) return rc;
  if (io->B == 0) return SMPLB_OK;
  if (!g_rgba || !g_attr || !scratch) return fail(SMPLB_E_INVALID_ARG, "smplb_phong_shade_backward: null pointer");
  PhongParams p = {};
  if (scratch_bytes < phong_carve(io, scratch, &p))
    return fail(SMPLB_E_INVALID_ARG, "smplb_phong_shade_backward: scratch too small");
  p.g_rgba = g_rgba;
  p.g_attr = g_attr;
  p.g_alphas = g_alphas;
  cudaStream_t st = (cudaStream_t)stream;
  const dim3 grid(p.nblk, io->B);
  if (io->off_specular >= 0 && io->L > 0) {
    phong_bwd_stats_kernel<<<grid, PH_THREADS, 0, st>>>(p);
    phong_finish_kernel<<<io->B * io->L, PH_THREADS, 0, st>>>(p.nblk, p.partial, p.Tsum);
  }
  phong_bwd_kernel<<<grid, PH_THREADS, 0, st>>>(p);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
