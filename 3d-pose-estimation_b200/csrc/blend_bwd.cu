// K5 transposed blend-shape GEMM (adjoint of batch_smpl.py:91-95, :113-115):
//     g_X[B, 224] = g_vposed[B, 20736] . D^T[20736, 224]     ->  [g_beta (10) | g_pose_feature (207) | pad]
// on tcgen05 from bf16 split operands (common.cuh: g1.D1 + g2.D1 + g1.D2, three kind::f16 MMAs per K16 step, fp32
// accumulation in TMEM; the 3xTF32 version of round 1 took 2.97 ms at B = 65536: 4 B operands at half the tensor rate,
// one CTA per 128-sample tile pulling 88 KB of operands per 1344 tensor cycles through a 2-stage ring).
// Organisation like blend_tc.cu: CTA PAIRS (cluster of 2, cta_group::2, M = 256 samples, N = 224), each CTA stages its
// own 128 sample rows (A operand, MN-major image written by skin_tc_kernel<1>, 32 KB per 64-column k-block) and one
// 112-row half of the D^T k-block (28 KB), 3-stage ring of 1-D bulk copies; the leader issues for both, commits are
// multicast.  The contraction runs over the 20736 internal v_posed columns (324 k-blocks of 64); work unit = (pair
// tile of 256 samples, K slice); slices balance the 74 clusters and give small batches enough CTAs, their partial
// sums go to g_X[slice] and are added in slice order by joint_bwd_kernel (fixed order: deterministic).
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int BB_STAGES = 3;
constexpr uint32_t BB_A_BYTES = BB_ASTAGE * 2;        // 32768: [g1 | g2], each [16 sample groups][8 column groups][128 B]
constexpr uint32_t BB_D_BYTES = 2 * BB_DPART * 2;     // 28672: [D1 | D2] of this CTA's 112 rows
constexpr uint32_t BB_STAGE_BYTES = BB_A_BYTES + BB_D_BYTES;
constexpr int BB_THREADS = 192;
constexpr int BB_ACC_STRIDE = 256;   // TMEM columns between the two accumulators (224 used)

struct BbSmem {
  unsigned char a[BB_STAGES][BB_A_BYTES];
  unsigned char d[BB_STAGES][BB_D_BYTES];
  uint64_t full[BB_STAGES];        // local bulk copies landed
  uint64_t peer_full[BB_STAGES];   // (leader) the peer's copies landed
  uint64_t empty[BB_STAGES];       // MMAs that read the stage completed (multicast commit)
  uint64_t acc_full[2];            // accumulator complete (multicast commit)
  uint64_t acc_empty[2];           // (leader) both CTAs' epilogue warps drained the accumulator
  uint32_t tmem_base;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(BB_THREADS, 1)
blend_bwd_kernel(int B, int n_mtiles, int n_slices, const unsigned short* __restrict__ Gimg,
                 const unsigned short* __restrict__ DTimg, float* __restrict__ g_X) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BbSmem& sm = *reinterpret_cast<BbSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int n_units = ((n_mtiles + 1) / 2) * n_slices;
  const int cl = blockIdx.x >> 1, n_cl = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < BB_STAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.peer_full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&sm.acc_full[a], 1);
      mbar_init(&sm.acc_empty[a], 8);   // one arrival per epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_pair(&sm.tmem_base, 512);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= producer (both CTAs) =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = cl; u < n_units; u += n_cl) {
        const int pt = u / n_slices, sl = u % n_slices;
        const int kb0 = (int)((long long)sl * BB_NKB / n_slices), kb1 = (int)((long long)(sl + 1) * BB_NKB / n_slices);
        int mt = pt * 2 + (int)rank;
        if (mt >= n_mtiles) mt = n_mtiles - 1;   // odd tile count: the peer recomputes a valid tile, stores nothing
        const unsigned short* gs = Gimg + (size_t)mt * BB_NKB * BB_ASTAGE;
        const unsigned short* ds = DTimg + (size_t)rank * (2 * BB_DPART);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % BB_STAGES;
          mbar_wait(&sm.empty[s], ((it / BB_STAGES) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.full[s], BB_STAGE_BYTES);
          bulk_g2s(sm.a[s], gs + (size_t)kb * BB_ASTAGE, BB_A_BYTES, &sm.full[s]);
          bulk_g2s(sm.d[s], ds + (size_t)kb * BB_DSTAGE, BB_D_BYTES, &sm.full[s]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // Whole warp runs the loops (uniform control flow); one elected lane issues the tcgen05 / arrive ops.
    const uint32_t issue = elect_one() ? 1u : 0u;
    const uint32_t full0 = smem_u32(&sm.full[0]), pfull0 = smem_u32(&sm.peer_full[0]);
    if (rank != 0) {
      // ================= peer: relay stage-full to the leader =================
      uint32_t s = 0, ph = 0;
      for (int u = cl; u < n_units; u += n_cl) {
        const int sl = u % n_slices;
        const int kb0 = (int)((long long)sl * BB_NKB / n_slices), kb1 = (int)((long long)(sl + 1) * BB_NKB / n_slices);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_spin(full0 + s * 8, ph);
          if (issue) mbar_arrive_remote(&sm.peer_full[s], 0);
          if (++s == BB_STAGES) { s = 0; ph ^= 1; }
        }
      }
    } else {
      // ================= MMA issuer (leader CTA) =================
      // A: MN-major bf16 (a core matrix = 8 columns x 8 samples), LBO = 128 B between column groups, SBO = 1024 B
      // between sample groups; B: K-major bf16, LBO = 112 rows x 16 B between column groups, SBO = 128 B between row groups
      constexpr uint32_t idesc = umma_idesc_ex(1, 1, 1, 0, 2 * BL_BM, KPAD);
      constexpr uint32_t A_LBO = 128, A_SBO = (BB_KB / 8) * 128, D_LBO = BB_NH * 16, D_SBO = 128;
      constexpr uint32_t A_PART = BB_APART * 2 / 16, D_PART = BB_DPART * 2 / 16;   // descriptor units (16 B)
      constexpr uint32_t A_KS = 2 * A_LBO / 16, D_KS = 2 * D_LBO / 16;              // per K16 step
      const uint64_t atempl = umma_desc_kmajor_noswz(0, A_LBO, A_SBO);
      const uint64_t dtempl = umma_desc_kmajor_noswz(0, D_LBO, D_SBO);
      const uint32_t empty0 = smem_u32(&sm.empty[0]);
      const uint32_t ab0 = smem_u32(sm.a[0]) >> 4, db0 = smem_u32(sm.d[0]) >> 4;
      uint32_t s = 0, ph = 0, ul = 0;
      for (int u = cl; u < n_units; u += n_cl, ++ul) {
        const int sl = u % n_slices;
        const int kb0 = (int)((long long)sl * BB_NKB / n_slices), kb1 = (int)((long long)(sl + 1) * BB_NKB / n_slices);
        const uint32_t a = ul & 1;
        mbar_wait_spin(smem_u32(&sm.acc_empty[a]), ((ul >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + a * BB_ACC_STRIDE;
#pragma unroll 1
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_spin(full0 + s * 8, ph);
          mbar_wait_spin(pfull0 + s * 8, ph);
          tc_fence_after();
          const uint64_t ad = atempl + (uint64_t)(ab0 + s * (BB_A_BYTES / 16));
          const uint64_t dd = dtempl + (uint64_t)(db0 + s * (BB_D_BYTES / 16));
#pragma unroll
          for (int ks = 0; ks < BB_KB / 16; ++ks) {
            const uint64_t a1 = ad + (uint64_t)(ks * A_KS), d1 = dd + (uint64_t)(ks * D_KS);
            umma_f16_ss_pair(dcol, a1, d1, idesc, (kb != kb0 || ks != 0) ? 1u : 0u, issue);              // g1.D1
            umma_f16_ss_pair(dcol, a1 + (uint64_t)A_PART, d1, idesc, 1u, issue);                          // g2.D1
            umma_f16_ss_pair(dcol, a1, d1 + (uint64_t)D_PART, idesc, 1u, issue);                          // g1.D2
          }
          umma_commit_pair(empty0 + s * 8, 3, issue);                                        // stage free in both CTAs
          if (kb == kb1 - 1) umma_commit_pair(smem_u32(&sm.acc_full[a]), 3, issue);          // accumulator complete
          if (++s == BB_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5, both CTAs): thread = sample row =================
    const int q = warp & 3;
    uint32_t ul = 0;
    for (int u = cl; u < n_units; u += n_cl, ++ul) {
      const int pt = u / n_slices, sl = u % n_slices;
      const int a = ul & 1;
      const int mt = pt * 2 + (int)rank;
      mbar_wait(&sm.acc_full[a], (ul >> 1) & 1);
      tc_fence_after();
      const int b = mt * BL_BM + q * 32 + lane;
      const bool ok = mt < n_mtiles && b < B;
      float* orow = g_X + ((size_t)sl * B + (ok ? b : 0)) * KPAD;
      const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)a * BB_ACC_STRIDE;
#pragma unroll 1
      for (int c0 = 0; c0 < KPAD; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tbase + c0, r);
        tmem_ld_wait();
        if (ok) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(orow + c0 + j) =
                make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                            __uint_as_float(r[j + 3]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(&sm.acc_empty[a], 0);
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem, 512);
  }
}

// K slices per pair tile of 256 samples: enough units for the 74 clusters of a B200 with little tail (the slice count
// sizes the g_X scratch, so it depends on B only, never on the device); at least 6 k-blocks per slice.
int blend_bwd_slices(int B) {
  const int n_pt = ((B + BL_BM - 1) / BL_BM + 1) / 2;
  if (n_pt <= 0) return 1;
  const int n_cl = 74;
  int best = 1;
  double best_cost = 1e30;
  for (int s = 1; s <= BB_NKB / 6; ++s) {
    const int units = n_pt * s;
    const int rounds = (units + n_cl - 1) / n_cl;
    // time ~ rounds x (k-blocks per slice + the accumulator hand-over / epilogue of a unit, ~8 k-blocks' worth)
    const double cost = rounds * ((double)BB_NKB / s + 8.0);
    if (cost < best_cost * 0.98) { best_cost = cost; best = s; }   // prefer fewer slices unless 2 % better
  }
  return best;
}

int32_t launch_blend_bwd(const smplb_model* m, int B, const float* g_img, float* g_X, int* n_slices_out,
                         cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static KernelSetup setup;
  const size_t smem = sizeof(BbSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, blend_bwd_kernel, smem, &n_sm)) return rc;
  const int n_mtiles = (B + BL_BM - 1) / BL_BM;
  const int n_slices = blend_bwd_slices(B);
  const int n_units = ((n_mtiles + 1) / 2) * n_slices;
  const int n_cl = n_units < n_sm / 2 ? n_units : n_sm / 2;
  blend_bwd_kernel<<<2 * n_cl, BB_THREADS, smem, st>>>(B, n_mtiles, n_slices,
                                                      reinterpret_cast<const unsigned short*>(g_img),
                                                      reinterpret_cast<const unsigned short*>(m->DTimg), g_X);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  *n_slices_out = n_slices;
  return SMPLB_OK;
}

}  // namespace smplb
