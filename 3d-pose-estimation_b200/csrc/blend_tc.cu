// K2 blend-shape GEMM on the 5th-generation tensor cores (batch_smpl.py:91-95, :113-115):
//     v_posed[B, 20670] = [beta | pose_feature][B, 217] . [shapedirs; posedirs][217, 20670] + v_template
// in 3xTF32 so the result stays fp32-faithful: with X = Xhi + Xlo and D = Dhi + Dlo (each part exactly
// representable in TF32), acc += Xhi.Dhi + Xlo.Dhi + Xhi.Dlo (the dropped Xlo.Dlo term is ~2^-22).
//
// v1 organisation (one CTA per SM, persistent over 128 x 256 output tiles, cta_group::1):
//   warp 0    producer : one 1-D bulk copy (TMA engine) per operand per k-block into a 2-stage smem ring;
//                        the operands already sit in global memory as ready-made smem images (common.cuh)
//   warp 1    MMA      : one elected thread issues tcgen05.mma.kind::tf32 (M=128, N=256, K=8) x 3 products
//                        x 28 k-steps per tile into one of two 256-column TMEM accumulators
//   warps 2-5 epilogue : tcgen05.ld the accumulator (lane = sample row), add v_template, store v_posed
// smem ring <-> MMA and TMEM <-> epilogue are mbarrier pipelines, so the epilogue of tile i overlaps the
// MMAs of tile i+1.
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int TC_STAGES = 2;
constexpr int TC_X_STAGE_FLOATS = 2 * (BL_KB / 4) * BL_BM * 4;   // hi+lo : 8192 floats = 32 KB
constexpr int TC_D_STAGE_FLOATS = 2 * (BL_KB / 4) * BL_BN * 4;   // hi+lo : 16384 floats = 64 KB
constexpr uint32_t TC_STAGE_BYTES = (TC_X_STAGE_FLOATS + TC_D_STAGE_FLOATS) * 4;
constexpr int TC_THREADS = 192;
constexpr int TC_TMEM_COLS = 512;

struct TcSmem {
  float x[TC_STAGES][TC_X_STAGE_FLOATS];
  float d[TC_STAGES][TC_D_STAGE_FLOATS];
  uint64_t full[TC_STAGES];
  uint64_t empty[TC_STAGES];
  uint64_t acc_full[2];
  uint64_t acc_empty[2];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(TC_THREADS, 1)
blend_tc_kernel(int B, int n_mtiles, const float* __restrict__ Ximg, const float* __restrict__ Dimg,
                const float* __restrict__ vt, float* __restrict__ vposed) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = n_mtiles * BL_NT;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < TC_STAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&sm.acc_full[a], 1);
      mbar_init(&sm.acc_empty[a], 4);   // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, TC_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= producer =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const int mt = t / BL_NT, nt = t % BL_NT;
        const float* xs = Ximg + (size_t)mt * BL_XTILE;
        const float* ds = Dimg + (size_t)nt * BL_DTILE;
        for (int kb = 0; kb < BL_NKB; ++kb, ++it) {
          const int s = it % TC_STAGES;
          const uint32_t ph = (it / TC_STAGES) & 1;
          mbar_wait(&sm.empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&sm.full[s], TC_STAGE_BYTES);
          bulk_g2s(sm.x[s], xs + (size_t)kb * TC_X_STAGE_FLOATS, TC_X_STAGE_FLOATS * 4, &sm.full[s]);
          bulk_g2s(sm.d[s], ds + (size_t)kb * TC_D_STAGE_FLOATS, TC_D_STAGE_FLOATS * 4, &sm.full[s]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(2, BL_BM, BL_BN);
      constexpr uint32_t X_LBO = BL_BM * 16, D_LBO = BL_BN * 16, SBO = 128;   // bytes
      constexpr uint32_t X_HL = (BL_KB / 4) * BL_BM * 16, D_HL = (BL_KB / 4) * BL_BN * 16;   // hi -> lo offset
      uint32_t it = 0, tl = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
        const int a = tl & 1;
        mbar_wait(&sm.acc_empty[a], ((tl >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + (uint32_t)a * BL_BN;
        for (int kb = 0; kb < BL_NKB; ++kb, ++it) {
          const int s = it % TC_STAGES;
          mbar_wait(&sm.full[s], (it / TC_STAGES) & 1);
          tc_fence_after();
          const uint32_t xb = smem_u32(sm.x[s]), db = smem_u32(sm.d[s]);
#pragma unroll
          for (int ks = 0; ks < BL_KB / 8; ++ks) {
            const uint64_t xh = umma_desc_kmajor_noswz(xb + ks * 2 * X_LBO, X_LBO, SBO);
            const uint64_t xl = umma_desc_kmajor_noswz(xb + X_HL + ks * 2 * X_LBO, X_LBO, SBO);
            const uint64_t dh = umma_desc_kmajor_noswz(db + ks * 2 * D_LBO, D_LBO, SBO);
            const uint64_t dl = umma_desc_kmajor_noswz(db + D_HL + ks * 2 * D_LBO, D_LBO, SBO);
            umma_tf32_ss(dcol, xh, dh, idesc, (kb | ks) != 0);
            umma_tf32_ss(dcol, xl, dh, idesc, 1);
            umma_tf32_ss(dcol, xh, dl, idesc, 1);
          }
          umma_commit(&sm.empty[s]);                         // stage free once these MMAs have read it
          if (kb == BL_NKB - 1) umma_commit(&sm.acc_full[a]);  // accumulator complete
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5) =================
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    uint32_t tl = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
      const int mt = t / BL_NT, nt = t % BL_NT;
      const int a = tl & 1;
      mbar_wait(&sm.acc_full[a], (tl >> 1) & 1);
      tc_fence_after();
      const int b = mt * BL_BM + q * 32 + lane;
      float* orow = vposed + (size_t)b * VPP + (size_t)nt * BL_BN;
      const float* vrow = vt + (size_t)nt * BL_BN;
      const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)a * BL_BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BL_BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tbase + c0, r);
        tmem_ld_wait();
        if (b < B) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 tv = __ldg(reinterpret_cast<const float4*>(vrow + c0 + j));
            float4 o;
            o.x = __uint_as_float(r[j]) + tv.x;
            o.y = __uint_as_float(r[j + 1]) + tv.y;
            o.z = __uint_as_float(r[j + 2]) + tv.z;
            o.w = __uint_as_float(r[j + 3]) + tv.w;
            *reinterpret_cast<float4*>(orow + c0 + j) = o;
          }
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
    tmem_dealloc(tmem, TC_TMEM_COLS);
  }
}

void free_tmap_cache(smplb_model* m) { (void)m; }

int32_t launch_blend_tc(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static int n_sm[64] = {};
  const int dev = m->device & 63;
  const size_t smem = sizeof(TcSmem);
  if (!n_sm[dev]) {
    int v = 0;
    SMPLB_CUDA_CHECK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, m->device));
    SMPLB_CUDA_CHECK(cudaFuncSetAttribute(blend_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    n_sm[dev] = v;
  }
  const int n_mtiles = (B + BL_BM - 1) / BL_BM;
  const int n_tiles = n_mtiles * BL_NT;
  const int grid = n_tiles < n_sm[dev] ? n_tiles : n_sm[dev];
  blend_tc_kernel<<<grid, TC_THREADS, smem, st>>>(B, n_mtiles, Ximg, m->Dimg, m->v_template_pad, vposed);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
