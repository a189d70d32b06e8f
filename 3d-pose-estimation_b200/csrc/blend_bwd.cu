// K5 transposed blend-shape GEMM (adjoint of batch_smpl.py:91-95, :113-115):
//     g_X[B, 224] = g_vposed[B, 20736] . D^T[20736, 224]     ->  [g_beta (10) | g_pose_feature (207) | pad]
// on tcgen05 in 3xTF32 like the forward GEMM.  The contraction runs over the 20736 internal v_posed columns
// (648 k-blocks of 32); the output is tiny.  A operand = the hi/lo image written by skin_tc_kernel<1>,
// B operand = the hi/lo image of D^T built at model creation; both are streamed with 1-D bulk copies.
// Work unit = (128-sample M tile, K slice); slices give small batches enough CTAs, their partial sums are
// written to g_X[slice] and added in a fixed order by joint_bwd_kernel.
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int BB_STAGES = 2;
constexpr uint32_t BB_STAGE_BYTES = (BB_GSTAGE + BB_DSTAGE) * 4;
constexpr int BB_THREADS = 192;
constexpr int BB_ACC_STRIDE = 256;   // TMEM columns between the two accumulators (224 used)

struct BbSmem {
  float g[BB_STAGES][BB_GSTAGE];
  float d[BB_STAGES][BB_DSTAGE];
  uint64_t full[BB_STAGES], empty[BB_STAGES], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(BB_THREADS, 1)
blend_bwd_kernel(int B, int n_mtiles, int n_slices, const float* __restrict__ Gimg, const float* __restrict__ DTimg,
                 float* __restrict__ g_X) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BbSmem& sm = *reinterpret_cast<BbSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = n_mtiles * n_slices;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < BB_STAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&sm.acc_full[a], 1);
      mbar_init(&sm.acc_empty[a], 4);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int mt = u / n_slices, sl = u % n_slices;
        const int kb0 = (int)((long long)sl * BB_NKB / n_slices), kb1 = (int)((long long)(sl + 1) * BB_NKB / n_slices);
        const float* gs = Gimg + (size_t)mt * ((size_t)BB_NKB * BB_GSTAGE);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % BB_STAGES;
          mbar_wait(&sm.empty[s], ((it / BB_STAGES) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.full[s], BB_STAGE_BYTES);
          bulk_g2s(sm.g[s], gs + (size_t)kb * BB_GSTAGE, BB_GSTAGE * 4, &sm.full[s]);
          bulk_g2s(sm.d[s], DTimg + (size_t)kb * BB_DSTAGE, BB_DSTAGE * 4, &sm.full[s]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(2, BL_BM, KPAD);
      constexpr uint32_t G_LBO = BL_BM * 16, D_LBO = KPAD * 16, SBO = 128;
      constexpr uint32_t G_HL = (BB_KB / 4) * G_LBO, D_HL = (BB_KB / 4) * D_LBO;
      uint32_t it = 0, ul = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
        const int sl = u % n_slices;
        const int kb0 = (int)((long long)sl * BB_NKB / n_slices), kb1 = (int)((long long)(sl + 1) * BB_NKB / n_slices);
        const int a = ul & 1;
        mbar_wait(&sm.acc_empty[a], ((ul >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + (uint32_t)a * BB_ACC_STRIDE;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % BB_STAGES;
          mbar_wait(&sm.full[s], (it / BB_STAGES) & 1);
          tc_fence_after();
          const uint32_t gb = smem_u32(sm.g[s]), db = smem_u32(sm.d[s]);
#pragma unroll
          for (int ks = 0; ks < BB_KB / 8; ++ks) {
            const uint64_t gh = umma_desc_kmajor_noswz(gb + ks * 2 * G_LBO, G_LBO, SBO);
            const uint64_t gl = umma_desc_kmajor_noswz(gb + G_HL + ks * 2 * G_LBO, G_LBO, SBO);
            const uint64_t dh = umma_desc_kmajor_noswz(db + ks * 2 * D_LBO, D_LBO, SBO);
            const uint64_t dl = umma_desc_kmajor_noswz(db + D_HL + ks * 2 * D_LBO, D_LBO, SBO);
            umma_tf32_ss(dcol, gh, dh, idesc, (kb != kb0) || (ks != 0));
            umma_tf32_ss(dcol, gl, dh, idesc, 1);
            umma_tf32_ss(dcol, gh, dl, idesc, 1);
          }
          umma_commit(&sm.empty[s]);
          if (kb == kb1 - 1) umma_commit(&sm.acc_full[a]);
        }
      }
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    uint32_t ul = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
      const int mt = u / n_slices, sl = u % n_slices;
      const int a = ul & 1;
      mbar_wait(&sm.acc_full[a], (ul >> 1) & 1);
      tc_fence_after();
      const int b = mt * BL_BM + q * 32 + lane;
      float* orow = g_X + ((size_t)sl * B + (b < B ? b : 0)) * KPAD;
      const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)a * BB_ACC_STRIDE;
#pragma unroll 1
      for (int c0 = 0; c0 < KPAD; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tbase + c0, r);
        tmem_ld_wait();
        if (b < B) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(orow + c0 + j) =
                make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                            __uint_as_float(r[j + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.acc_empty[a]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// enough (M tile, K slice) units for ~2 waves of CTAs, at least 8 k-blocks per slice
int blend_bwd_slices(int B, int n_sm) {
  const int n_mtiles = (B + BL_BM - 1) / BL_BM;
  int s = (2 * n_sm + n_mtiles - 1) / n_mtiles;
  if (n_mtiles >= 2 * n_sm) s = 1;
  if (s > BB_NKB / 8) s = BB_NKB / 8;
  if (s < 1) s = 1;
  return s;
}

int32_t launch_blend_bwd(const smplb_model* m, int B, const float* g_img, float* g_X, int* n_slices_out,
                         cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static KernelSetup setup;
  const size_t smem = sizeof(BbSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, blend_bwd_kernel, smem, &n_sm)) return rc;
  const int n_mtiles = (B + BL_BM - 1) / BL_BM;
  const int n_slices = blend_bwd_slices(B, 148);   // fixed constant: the workspace size must not depend on the device
  const int n_units = n_mtiles * n_slices;
  const int grid = n_units < n_sm ? n_units : n_sm;
  blend_bwd_kernel<<<grid, BB_THREADS, smem, st>>>(B, n_mtiles, n_slices, g_img, m->DTimg, g_X);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  *n_slices_out = n_slices;
  return SMPLB_OK;
}

}  // namespace smplb
