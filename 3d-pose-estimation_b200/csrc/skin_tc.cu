// K3 fused skinning pass (batch_smpl.py:120-132), the HBM-bound kernel of the forward:
//     T[b,v]   = sum_j weights[v,j] * A[b,j]              (:122-127, tile(W) x A never materialised)
//     verts[b,v] = T[b,v][:3,:3] . v_posed[b,v] + T[b,v][:3,3]   (:128-132)
// One pass over HBM: read v_posed (82 944 B / sample, internal SoA tile layout) + the A operand image
// (2 304 B / sample, L2 resident across the 54 vertex tiles), write verts (82 680 B / sample).
//
// The weight blend is the reference's own matmul W[6890,24] x A[24,12]; doing it per thread from shared
// memory costs 192 B of smem gather per vertex-sample and was 4x off the HBM roofline (profiles/r01), so
// it runs on the tensor cores instead, in 3xTF32 (fp32-faithful):
//   MMA: D[128 vertices, 240 = 20 samples x 12] += W_tile[128, 24] . A_group[240, 24]^T   (9 x M128 N240 K8)
//   accumulator in TMEM: lane = vertex, 12 columns per sample = the 3x4 blended transform of that vertex.
// CTA = one vertex tile (W tile resident in smem) x a chunk of sample groups, persistent over units:
//   warp 0    producer : bulk copies (TMA engine): A group image + the 20 v_posed tile rows of the group
//   warp 1    MMA      : tcgen05.mma.kind::tf32 into one of two 240-column TMEM accumulators
//   warps 2-5 epilogue : tcgen05.ld T, apply to v_posed (thread = vertex), transpose through smem so the
//                        verts rows are written as contiguous 8-byte-aligned runs (streaming stores)
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int SKT_THREADS = 192;
constexpr int SKT_TMEM_COLS = 512;
constexpr uint32_t SKT_VP_ROW_BYTES = 3 * SK_VT * 4;   // 1536

struct SktSmem {
  float w[SK_WTILE];                 // 24 576 B  W tile image (hi|lo)
  float a[2][SK_AGRP];               // 46 080 B  per stage: A group image (hi|lo)
  float vp[2][SK_S][3 * SK_VT];      // 30 720 B  per stage: v_posed tile rows of the group
  float out[SK_S][3 * SK_VT];        // 30 720 B  verts staging (AoS)
  uint64_t full[2], empty[2], acc_full[2], acc_empty[2], w_full, w_empty;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(SKT_THREADS, 1)
skin_tc_kernel(int B, int n_groups, int n_chunks, const float* __restrict__ vposed,
               const float* __restrict__ Aimg, const float* __restrict__ Wimg, float* __restrict__ verts) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SktSmem& sm = *reinterpret_cast<SktSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = SK_NVT * n_chunks;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 5);       // MMA commit + 4 epilogue warps
      mbar_init(&sm.acc_full[s], 1);
      mbar_init(&sm.acc_empty[s], 4);
    }
    mbar_init(&sm.w_full, 1);
    mbar_init(&sm.w_empty, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, SKT_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= producer =================
    if (lane == 0) {
      uint32_t it = 0, ul = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
        const int vt = u % SK_NVT, ch = u / SK_NVT;
        const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
        mbar_wait(&sm.w_empty, (ul & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.w_full, SK_WTILE * 4);
        bulk_g2s(sm.w, Wimg + (size_t)vt * SK_WTILE, SK_WTILE * 4, &sm.w_full);
        for (int g = g0; g < g1; ++g, ++it) {
          const int s = it & 1;
          mbar_wait(&sm.empty[s], ((it >> 1) & 1) ^ 1);
          const int ns = min(SK_S, B - g * SK_S);
          mbar_arrive_expect_tx(&sm.full[s], SK_AGRP * 4 + ns * SKT_VP_ROW_BYTES);
          bulk_g2s(sm.a[s], Aimg + (size_t)g * SK_AGRP, SK_AGRP * 4, &sm.full[s]);
          const float* src = vposed + (size_t)g * SK_S * VPP + (size_t)vt * (3 * SK_VT);
          for (int i = 0; i < ns; ++i) bulk_g2s(sm.vp[s][i], src + (size_t)i * VPP, SKT_VP_ROW_BYTES, &sm.full[s]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(2, SK_VT, SK_N);
      constexpr uint32_t W_LBO = SK_VT * 16, A_LBO = SK_N * 16, SBO = 128;
      constexpr uint32_t W_HL = SK_KCH * W_LBO, A_HL = SK_KCH * A_LBO;
      const uint32_t wb = smem_u32(sm.w);
      uint32_t it = 0, ul = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
        const int ch = u / SK_NVT;
        const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
        mbar_wait(&sm.w_full, ul & 1);
        for (int g = g0; g < g1; ++g, ++it) {
          const int s = it & 1;
          const uint32_t ph = (it >> 1) & 1;
          mbar_wait(&sm.acc_empty[s], ph ^ 1);
          mbar_wait(&sm.full[s], ph);
          tc_fence_after();
          const uint32_t ab = smem_u32(sm.a[s]);
          const uint32_t dcol = tmem + (uint32_t)s * SK_N;
#pragma unroll
          for (int ks = 0; ks < NJ / 8; ++ks) {
            const uint64_t wh = umma_desc_kmajor_noswz(wb + ks * 2 * W_LBO, W_LBO, SBO);
            const uint64_t wl = umma_desc_kmajor_noswz(wb + W_HL + ks * 2 * W_LBO, W_LBO, SBO);
            const uint64_t ah = umma_desc_kmajor_noswz(ab + ks * 2 * A_LBO, A_LBO, SBO);
            const uint64_t al = umma_desc_kmajor_noswz(ab + A_HL + ks * 2 * A_LBO, A_LBO, SBO);
            umma_tf32_ss(dcol, wh, ah, idesc, ks != 0);
            umma_tf32_ss(dcol, wl, ah, idesc, 1);
            umma_tf32_ss(dcol, wh, al, idesc, 1);
          }
          umma_commit(&sm.empty[s]);
          umma_commit(&sm.acc_full[s]);
        }
        umma_commit(&sm.w_empty);   // W tile may be replaced once every MMA of this unit has read it
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5): thread = vertex lane =================
    const int q = warp & 3;
    const int l4 = q * 32 + lane;                 // TMEM lane = vertex within the tile
    const int et = (warp - 2) * 32 + lane;        // 0..127 index for the cooperative copy-out
    uint32_t it = 0;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int vt = u % SK_NVT, ch = u / SK_NVT;
      const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
      const int nvalid = min(3 * SK_VT, NV3 - vt * 3 * SK_VT);   // floats of this tile that exist in verts
      for (int g = g0; g < g1; ++g, ++it) {
        const int s = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        mbar_wait(&sm.full[s], ph);       // v_posed rows of this group have landed
        mbar_wait(&sm.acc_full[s], ph);   // blended transforms are in TMEM
        tc_fence_after();
        const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)s * SK_N;
#pragma unroll 1
        for (int bt = 0; bt < SK_S / 4; ++bt) {
          uint32_t r0[16], r1[16], r2[16];
          tmem_ld16(tbase + bt * 48, r0);
          tmem_ld16(tbase + bt * 48 + 16, r1);
          tmem_ld16(tbase + bt * 48 + 32, r2);
          tmem_ld_wait();
          float T[48];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            T[e] = __uint_as_float(r0[e]);
            T[16 + e] = __uint_as_float(r1[e]);
            T[32 + e] = __uint_as_float(r2[e]);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int i = bt * 4 + k;
            const float x = sm.vp[s][i][l4], y = sm.vp[s][i][SK_VT + l4], z = sm.vp[s][i][2 * SK_VT + l4];
            const float* t = T + 12 * k;
#pragma unroll
            for (int r = 0; r < 3; ++r)
              sm.out[i][3 * l4 + r] = fmaf(t[4 * r], x, fmaf(t[4 * r + 1], y, fmaf(t[4 * r + 2], z, t[4 * r + 3])));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&sm.acc_empty[s]);
          mbar_arrive(&sm.empty[s]);
        }
        named_bar_sync(1, 128);   // staging complete
        const int ns = min(SK_S, B - g * SK_S);
        float* dst = verts + (size_t)g * SK_S * NV3 + (size_t)vt * (3 * SK_VT);
        for (int idx = et; idx < ns * (3 * SK_VT / 2); idx += 128) {
          const int i = idx / (3 * SK_VT / 2), p = idx % (3 * SK_VT / 2);
          if (2 * p < nvalid) {
            const float2 v = *reinterpret_cast<const float2*>(&sm.out[i][2 * p]);
            __stcs(reinterpret_cast<float2*>(dst + (size_t)i * NV3 + 2 * p), v);
          }
        }
        named_bar_sync(1, 128);   // staging reusable
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, SKT_TMEM_COLS);
  }
}

int32_t launch_skin_tc(const smplb_model* m, int B, const float* vposed, const float* Aimg, float* verts,
                       cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static int n_sm[64] = {};
  const int dev = m->device & 63;
  const size_t smem = sizeof(SktSmem);
  if (!n_sm[dev]) {
    int v = 0;
    SMPLB_CUDA_CHECK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, m->device));
    SMPLB_CUDA_CHECK(cudaFuncSetAttribute(skin_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    n_sm[dev] = v;
  }
  const int n_groups = (B + SK_S - 1) / SK_S;
  // units = 54 vertex tiles x n_chunks sample chunks; aim at ~3 full waves of CTAs
  int n_chunks = (3 * n_sm[dev]) / SK_NVT;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > n_groups) n_chunks = n_groups;
  const int n_units = SK_NVT * n_chunks;
  const int grid = n_units < n_sm[dev] ? n_units : n_sm[dev];
  skin_tc_kernel<<<grid, SKT_THREADS, smem, st>>>(B, n_groups, n_chunks, vposed, Aimg, m->Wimg, verts);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
