// K2 (verification variant): blend-shape contraction v_posed = [beta | pose_feature] . [shapedirs; posedirs]
// + v_template (batch_smpl.py:91-95, :113-115) on CUDA cores in plain fp32 FFMA.  Bring-up / cross-check
// path for the tcgen05 3xTF32 kernel in blend_tc.cu (flag SMPLB_F_BLEND_SIMT); not the fast path.
#include "common.cuh"

namespace smplb {

constexpr int BS_TM = 64, BS_TN = 64, BS_TK = 32;

__global__ void __launch_bounds__(256) blend_simt_kernel(int B, const float* __restrict__ Ximg,
                                                         const float* __restrict__ D32,
                                                         const float* __restrict__ vt,
                                                         float* __restrict__ vposed) {
  __shared__ float Xs[BS_TK][BS_TM + 1];
  __shared__ float Ds[BS_TK][BS_TN];
  const int n0 = blockIdx.x * BS_TN, b0 = blockIdx.y * BS_TM;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // tx -> 4 columns, ty -> 4 samples
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < KPAD; k0 += BS_TK) {
    for (int i = threadIdx.x; i < BS_TM * BS_TK; i += 256) {
      const int r = i / BS_TK, k = i % BS_TK;
      const int b = b0 + r;
      float v = 0.f;
      if (b < B) {   // X = f16(X) + f16(X - f16(X)): good to 2^-22 relative (2^-25 absolute below 0.12)
        const unsigned short* x16 = reinterpret_cast<const unsigned short*>(Ximg);
        v = f16_bits_to_float(x16[ximg16_off(b, k0 + k, 0)]) + f16_bits_to_float(x16[ximg16_off(b, k0 + k, 1)]);
      }
      Xs[k][r] = v;
    }
    for (int i = threadIdx.x; i < BS_TK * BS_TN; i += 256) {
      const int k = i / BS_TN, n = i % BS_TN;
      Ds[k][n] = D32[(size_t)(k0 + k) * VPP + n0 + n];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BS_TK; ++k) {
      float a[4], d[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = Xs[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) d[j] = Ds[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], d[j], acc[i][j]);
    }
    __syncthreads();
  }
  const float4 t = *reinterpret_cast<const float4*>(vt + n0 + tx * 4);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + ty * 4 + i;
    if (b < B)
      *reinterpret_cast<float4*>(vposed + (size_t)b * VPP + n0 + tx * 4) =
          make_float4(acc[i][0] + t.x, acc[i][1] + t.y, acc[i][2] + t.z, acc[i][3] + t.w);
  }
}

int32_t launch_blend_simt(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  dim3 grid(VPP / BS_TN, (B + BS_TM - 1) / BS_TM);
  blend_simt_kernel<<<grid, 256, 0, st>>>(B, Ximg, m->D32, m->v_template_pad, vposed);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
