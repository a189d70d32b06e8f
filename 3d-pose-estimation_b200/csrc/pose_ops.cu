// Per-sample operators either side of the SMPL path:
//   N2  rotate_global_pose / add_axis_angle_rotations  deps/pose_3d/utils.py:152-199  (GT pose pre-processing,
//       called at pose_model_3d.py:371-372 right before SMPL)
//   N3  camera-angle (ray) loss + camera regularisers  deps/pose_3d/pose_model_3d.py:256-292 with
//       utils.get_2d_points_in_3d (utils.py:75-89): value and gradient wrt camera position / rotation / focal.
// Latency-class kernels: thread = sample, a few hundred bytes each; the loss sums are reduced in a fixed
// order (block tree, then one finishing block) so the scalars are run-to-run identical.
#include "common.cuh"
#include "geom.cuh"

namespace smplb {

// utils.py:170-199: rotation c*n equivalent to applying rv2 then rv1 (half-angle / quaternion product).
__device__ __forceinline__ void add_axis_angle(const float r1[3], const float r2[3], float out[3]) {
  const float a = sqrtf(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
  const float b = sqrtf(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
  if (b == 0.0f) { out[0] = r1[0]; out[1] = r1[1]; out[2] = r1[2]; return; }   // :198 (applied last, wins)
  if (a == 0.0f) { out[0] = r2[0]; out[1] = r2[1]; out[2] = r2[2]; return; }   // :197
  const float l[3] = {r1[0] / a, r1[1] / a, r1[2] / a}, m[3] = {r2[0] / b, r2[1] / b, r2[2] / b};
  float sa, ca, sb, cb;
  sincosf(a * 0.5f, &sa, &ca);
  sincosf(b * 0.5f, &sb, &cb);
  const float ldm = l[0] * m[0] + l[1] * m[1] + l[2] * m[2];
  const float cos_half_c = ca * cb - sa * sb * ldm;                                       // :186-188
  const float cr[3] = {l[1] * m[2] - l[2] * m[1], l[2] * m[0] - l[0] * m[2], l[0] * m[1] - l[1] * m[0]};
  // :189-195.  The reference takes half_c = acos(cos_half_c) and divides by sin(half_c); |sin_half_c_n| IS sin(half_c)
  // (n is a unit vector, half_c in [0, pi]), so half_c = atan2(|s|, cos_half_c) is the same function -- evaluated this
  // way it keeps full float precision where acos loses half of its digits (composed angles near 0 and 2*pi).
  float sn[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) sn[k] = sa * cb * l[k] + (ca * sb * m[k] + sa * sb * cr[k]);
  const float sh = sqrtf(sn[0] * sn[0] + sn[1] * sn[1] + sn[2] * sn[2]);
  const float half_c = atan2f(sh, cos_half_c);
#pragma unroll
  for (int k = 0; k < 3; ++k) out[k] = sn[k] / sh * half_c * 2.0f;   // 0/0 = NaN for a composed identity, like :194
}

__global__ void add_axis_angle_kernel(int N, const float* __restrict__ rv1, const float* __restrict__ rv2,
                                      float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float a[3] = {rv1[3 * i], rv1[3 * i + 1], rv1[3 * i + 2]}, b[3] = {rv2[3 * i], rv2[3 * i + 1], rv2[3 * i + 2]};
  float o[3];
  add_axis_angle(a, b, o);
  out[3 * i] = o[0]; out[3 * i + 1] = o[1]; out[3 * i + 2] = o[2];
}

// utils.py:152-167: global rotation <- (turn_y(zrot + pi/2) o turn_x(-pi/2)) o theta[:3]; the other 69 copied.
__global__ void rotate_global_pose_kernel(int B, const float* __restrict__ thetas, const float* __restrict__ zrot,
                                          float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * 24) return;
  const int b = i / 24, j = i % 24;
  const float* t = thetas + (size_t)b * 72 + 3 * j;
  float o[3] = {t[0], t[1], t[2]};
  if (j == 0) {
    const float tx[3] = {-1.5707963267948966f, 0.f, 0.f};
    const float ty[3] = {0.f, zrot[b] + 1.5707963267948966f, 0.f};
    float both[3];
    add_axis_angle(ty, tx, both);
    add_axis_angle(both, t, o);
  }
  float* d = out + (size_t)b * 72 + 3 * j;
  d[0] = o[0]; d[1] = o[1]; d[2] = o[2];
}

// ---- N3 ---------------------------------------------------------------------------------------------------
// per sample: sum_j angle^2 (pose_model_3d.py:260-268), the three limit terms (:276-281), the rotation-norm term
// (:285-289); optional gradients with the global scales already applied (scale / element count).
__global__ void __launch_bounds__(128) cam_losses_kernel(int B, const float* __restrict__ j3d, const float* __restrict__ p2d,
                                                         const float* __restrict__ cpos, const float* __restrict__ crot,
                                                         const float* __restrict__ cf, const float* __restrict__ f_raw,
                                                         smplb_camcfg cfg, float w_angle, float w_neg, float w_rot,
                                                         float* __restrict__ partial /*[gridDim][3]*/,
                                                         float* __restrict__ g_pos, float* __restrict__ g_rot,
                                                         float* __restrict__ g_f, float* __restrict__ g_fraw) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  float s_ang = 0.f, s_neg = 0.f, s_rot = 0.f;
  if (b < B) {
    const float ax = crot[3 * b], ay = crot[3 * b + 1], az = crot[3 * b + 2];
    const float c[3] = {cpos[3 * b], cpos[3 * b + 1], cpos[3 * b + 2]};
    const float f = cf[b];
    float R[9], gR[9], gc[3] = {0.f, 0.f, 0.f}, gf = 0.f;
    rodrigues_cam(ax, ay, az, R);
#pragma unroll
    for (int e = 0; e < 9; ++e) gR[e] = 0.f;
    for (int j = 0; j < 24; ++j) {
      const float* p = j3d + ((size_t)b * 24 + j) * 3;
      const float d[3] = {p[0] - c[0], p[1] - c[1], p[2] - c[2]};                       // :260
      const float nd = sqrtf(fmaxf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2], 1e-12f));   // l2_normalize eps
      const float u[3] = {d[0] / nd, d[1] / nd, d[2] / nd};
      const float* q2 = p2d + ((size_t)b * 24 + j) * 2;
      const float q[3] = {cfg.ss * (q2[0] - 0.5f * cfg.img_w) / cfg.img_h,               // utils.py:77
                          cfg.ss * (q2[1] - 0.5f * cfg.img_h) / cfg.img_h, f};
      float rq[3];
#pragma unroll
      for (int r = 0; r < 3; ++r) rq[r] = R[3 * r] * q[0] + R[3 * r + 1] * q[1] + R[3 * r + 2] * q[2];   // :84-85
      const float nq = sqrtf(fmaxf(rq[0] * rq[0] + rq[1] * rq[1] + rq[2] * rq[2], 1e-12f));
      const float w[3] = {rq[0] / nq, rq[1] / nq, rq[2] / nq};
      const float dot = u[0] * w[0] + u[1] * w[1] + u[2] * w[2];                         // :265
      const float lo = -1.0f + 1e-6f, hi = 1.0f - 1e-6f;
      // :267-268 acos(clip(dot)).  Inside the clip range the angle is taken as atan2(|u x w|, u.w) -- the same
      // function of two unit vectors, without acos' loss of digits at dot -> 1 (where a fitted camera lives);
      // outside, the clipped constants exactly as the float32 graph has them.
      const float cx = u[1] * w[2] - u[2] * w[1], cy = u[2] * w[0] - u[0] * w[2], cz = u[0] * w[1] - u[1] * w[0];
      const float sn = sqrtf(cx * cx + cy * cy + cz * cz);
      const bool inside = dot >= lo && dot <= hi;
      const float ang = inside ? atan2f(sn, dot) : acosf(dot > hi ? hi : lo);
      s_ang += ang * ang;
      if (g_pos && inside) {
        const float gd = -2.0f * w_angle * ang / sn;                                     // d acos(x)/dx = -1/sin
        // through u = d / |d|
        const float ugw = u[0] * w[0] + u[1] * w[1] + u[2] * w[2];
#pragma unroll
        for (int r = 0; r < 3; ++r) gc[r] -= gd * (w[r] - u[r] * ugw) / nd;
        // through w = Rq / |Rq|
        float grq[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) grq[r] = gd * (u[r] - w[r] * ugw) / nq;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
#pragma unroll
          for (int k = 0; k < 3; ++k) gR[3 * r + k] += grq[r] * q[k];
          gf += R[3 * r + 2] * grq[r];
        }
      }
    }
    // camera regularisers
    const float lim0 = -c[2] - 0.3f, lim1 = c[2] + 10.0f;                                // :276-278
    float gz = 0.f, gfr = 0.f;
    if (lim0 < 0.f) { s_neg += lim0 * lim0; gz -= 2.0f * w_neg * lim0; }
    if (lim1 < 0.f) { s_neg += lim1 * lim1; gz += 2.0f * w_neg * lim1; }
    if (f_raw) {
      const float l2 = f_raw[b];
      if (l2 < 0.f) { s_neg += l2 * l2; gfr = 2.0f * w_neg * l2; }
    }
    const float nr = sqrtf(ax * ax + ay * ay + az * az);                                 // :285-286
    float grn = 0.f;
    if (nr > 0.39269908169872414f) { s_rot += nr * nr; grn = 2.0f * w_rot * nr; }        // pi * 0.125, :288
    if (g_pos) {
      g_pos[3 * b] = gc[0]; g_pos[3 * b + 1] = gc[1]; g_pos[3 * b + 2] = gc[2] + gz;
      float gr[3];
      rodrigues_cam_bwd(ax, ay, az, gR, gr);
      const float inv = nr > 0.f ? grn / nr : 0.f;
      g_rot[3 * b] = gr[0] + inv * ax; g_rot[3 * b + 1] = gr[1] + inv * ay; g_rot[3 * b + 2] = gr[2] + inv * az;
      if (g_f) g_f[b] = gf;
      if (g_fraw) g_fraw[b] = gfr;
    }
  }
  __shared__ float red[3][128];
  red[0][threadIdx.x] = s_ang; red[1][threadIdx.x] = s_neg; red[2][threadIdx.x] = s_rot;
  __syncthreads();
  for (int o = 64; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o)
      for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x < 3) partial[blockIdx.x * 3 + threadIdx.x] = red[threadIdx.x][0];
}

// loss_out = {cam_angle_loss, cam_neg_loss, cam_rot_loss, sum}
__global__ void cam_losses_finish_kernel(int nblocks, const float* __restrict__ partial, float w_angle, float w_neg,
                                         float w_rot, float* __restrict__ loss_out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float s[3] = {0.f, 0.f, 0.f};
  for (int i = 0; i < nblocks; ++i)
    for (int k = 0; k < 3; ++k) s[k] += partial[i * 3 + k];
  loss_out[0] = w_angle * s[0];
  loss_out[1] = w_neg * s[1];
  loss_out[2] = w_rot * s[2];
  loss_out[3] = loss_out[0] + loss_out[1] + loss_out[2];
}

// ---- 3D losses (pose_model_3d.py:183-192): weight * mean((pred - target)^2), value and gradient ----------------
// Grid-stride over the elements, per-thread sequential partial -> fixed-order block tree -> one partial per block;
// the finishing block adds the MSE_BLOCKS partials in index order: run-to-run identical.
constexpr int MSE_BLOCKS = 1184, MSE_THREADS = 256;
__global__ void __launch_bounds__(MSE_THREADS) mse_grad_kernel(long long n, const float* __restrict__ pred,
                                                               const float* __restrict__ target, float gscale,
                                                               float* __restrict__ g_out, float* __restrict__ partial) {
  float acc = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float d = __ldcs(pred + i) - __ldcs(target + i);
    acc = fmaf(d, d, acc);
    if (g_out) __stcs(g_out + i, gscale * d);
  }
  __shared__ float red[MSE_THREADS];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = MSE_THREADS / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = red[0];
}
__global__ void mse_finish_kernel(const float* __restrict__ partial, float lscale, float* __restrict__ loss_out,
                                  int accumulate) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float s = 0.f;
  for (int i = 0; i < MSE_BLOCKS; ++i) s += partial[i];
  loss_out[0] = (accumulate ? loss_out[0] : 0.f) + lscale * s;
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_add_axis_angle(int32_t N, const float* rv1, const float* rv2, float* out, void* stream) {
  if (N < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_add_axis_angle: negative count");
  if (N == 0) return SMPLB_OK;
  if (!rv1 || !rv2 || !out) return fail(SMPLB_E_INVALID_ARG, "smplb_add_axis_angle: null pointer");
  add_axis_angle_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(N, rv1, rv2, out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" int32_t smplb_rotate_global_pose(int32_t B, const float* thetas, const float* zrot, float* out, void* stream) {
  if (B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_rotate_global_pose: negative batch");
  if (B == 0) return SMPLB_OK;
  if (!thetas || !zrot || !out) return fail(SMPLB_E_INVALID_ARG, "smplb_rotate_global_pose: null pointer");
  rotate_global_pose_kernel<<<(B * 24 + 127) / 128, 128, 0, (cudaStream_t)stream>>>(B, thetas, zrot, out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" size_t smplb_cam_losses_scratch_bytes(int32_t B) { return (size_t)((B > 0 ? B : 0) + 127) / 128 * 3 * sizeof(float) + 16; }

extern "C" int32_t smplb_cam_losses(int32_t B, const float* gt_joints3d, const float* gt_joints2d, const float* cam_pos,
                                    const float* cam_rot, const float* cam_f, const float* f_raw,
                                    const smplb_camcfg* cfg, float angle_scale, float* loss_out, float* g_cam_pos,
                                    float* g_cam_rot, float* g_cam_f, float* g_f_raw, void* scratch, void* stream) {
  if (B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_cam_losses: negative batch");
  if (!cfg || !loss_out) return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses: null cfg or loss_out");
  if (B == 0) return SMPLB_OK;
  if (!gt_joints3d || !gt_joints2d || !cam_pos || !cam_rot || !cam_f || !scratch)
    return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses: null pointer");
  if ((g_cam_pos == nullptr) != (g_cam_rot == nullptr))
    return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses: g_cam_pos and g_cam_rot go together");
  const int nb = (B + 127) / 128;
  // tf.losses.mean_squared_error: weight * mean over all elements (B*24 angles, B*3 limits, B norms)
  const float w_angle = angle_scale / (float)((long long)B * 24), w_neg = 1.0f / (float)((long long)B * 3),
              w_rot = 10.0f / (float)B;
  cudaStream_t st = (cudaStream_t)stream;
  cam_losses_kernel<<<nb, 128, 0, st>>>(B, gt_joints3d, gt_joints2d, cam_pos, cam_rot, cam_f, f_raw, *cfg, w_angle, w_neg,
                                       w_rot, (float*)scratch, g_cam_pos, g_cam_rot, g_cam_f, g_f_raw);
  cam_losses_finish_kernel<<<1, 32, 0, st>>>(nb, (const float*)scratch, w_angle, w_neg, w_rot, loss_out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

extern "C" size_t smplb_mse_scratch_bytes(void) { return MSE_BLOCKS * sizeof(float); }

extern "C" int32_t smplb_mse_grad(int64_t n, const float* pred, const float* target, float weight, int64_t n_total,
                                  float* g_out, float* loss_out, int32_t accumulate, void* scratch, void* stream) {
  if (n < 0 || n_total <= 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_mse_grad: n >= 0 and n_total > 0 required");
  if (!pred || !target || !loss_out || !scratch) return fail(SMPLB_E_INVALID_ARG, "smplb_mse_grad: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const float inv = weight / (float)n_total;
  mse_grad_kernel<<<MSE_BLOCKS, MSE_THREADS, 0, st>>>((long long)n, pred, target, 2.0f * inv, g_out, (float*)scratch);
  mse_finish_kernel<<<1, 32, 0, st>>>((const float*)scratch, inv, loss_out, accumulate);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

// ---- allocation-free variant of smplb_cam_losses (no counterpart in the reference; added
// for callers that cannot allocate device memory: CUDA graph capture, tight serving loops) ----------------
// The per-block partial sums live in a file-scope device array instead of a caller-provided scratch buffer.  Every GPU
// that loads the library holds the array (SMPLB_CAM_LOSSES_STATIC_MAX_B / 128 blocks x 3 floats + 16 B = 6 160 B) until
// exit.  All calls on one device share it: calls that overlap on the GPU (different streams) race.
#define SMPLB_CAM_STATIC_FLOATS ((SMPLB_CAM_LOSSES_STATIC_MAX_B / 128) * 3 + 4)
static __device__ __align__(256) float g_cam_losses_scratch[SMPLB_CAM_STATIC_FLOATS];

extern "C" int32_t smplb_cam_losses_static(int32_t B, const float* gt_joints3d, const float* gt_joints2d,
                                           const float* cam_pos, const float* cam_rot, const float* cam_f,
                                           const float* f_raw, const smplb_camcfg* cfg, float angle_scale,
                                           float* loss_out, float* g_cam_pos, float* g_cam_rot, float* g_cam_f,
                                           float* g_f_raw, void* stream) {
  if (B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_cam_losses_static: negative batch");
  if (B > SMPLB_CAM_LOSSES_STATIC_MAX_B)
    return fail(SMPLB_E_UNSUPPORTED, "smplb_cam_losses_static: B exceeds SMPLB_CAM_LOSSES_STATIC_MAX_B (" +
                                         std::to_string(SMPLB_CAM_LOSSES_STATIC_MAX_B) +
                                         "): the built-in scratch array is fixed; use smplb_cam_losses with a caller-owned scratch");
  if (!cfg || !loss_out) return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses_static: null cfg or loss_out");
  if (B == 0) return SMPLB_OK;
  if (!gt_joints3d || !gt_joints2d || !cam_pos || !cam_rot || !cam_f)
    return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses_static: null pointer");
  if ((g_cam_pos == nullptr) != (g_cam_rot == nullptr))
    return fail(SMPLB_E_INVALID_ARG, "smplb_cam_losses_static: g_cam_pos and g_cam_rot go together");
  // In host code the array's name is only a stand-in: the device address is looked up once per device, with that
  // device current (cudaGetSymbolAddress answers for the current device), and cached.
  static std::mutex mu;
  static float* addr[64] = {};
  int dev = -1;
  SMPLB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(SMPLB_E_UNSUPPORTED, "smplb_cam_losses_static: device index beyond 63");
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
      void* p = nullptr;
      SMPLB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_cam_losses_scratch));
      addr[dev] = (float*)p;
    }
    scratch = addr[dev];
  }
  const int nb = (B + 127) / 128;
  const float w_angle = angle_scale / (float)((long long)B * 24), w_neg = 1.0f / (float)((long long)B * 3),
              w_rot = 10.0f / (float)B;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's partial sums: clear the part this call uses (stream-ordered, no allocation)
  SMPLB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, (size_t)nb * 3 * sizeof(float), st));
  cam_losses_kernel<<<nb, 128, 0, st>>>(B, gt_joints3d, gt_joints2d, cam_pos, cam_rot, cam_f, f_raw, *cfg, w_angle, w_neg,
                                       w_rot, scratch, g_cam_pos, g_cam_rot, g_cam_f, g_f_raw);
  cam_losses_finish_kernel<<<1, 32, 0, st>>>(nb, scratch, w_angle, w_neg, w_rot, loss_out);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}
