// K3 fused skinning pass (batch_smpl.py:120-132), the HBM-bound kernel of the forward:
//     T[b,v]   = sum_j weights[v,j] * A[b,j]              (:122-127, tile(W) x A never materialised)
//     verts[b,v] = T[b,v][:3,:3] . v_posed[b,v] + T[b,v][:3,3]   (:128-132)
// One pass over HBM: read v_posed (82 944 B / sample, internal SoA tile layout) + the A operand image
// (1 536 B / sample, L2 resident across the 54 vertex tiles), write verts (82 680 B / sample).
//
// The weight blend is the reference's own matmul W[6890,24] x A[24,12]; doing it per thread from shared
// memory costs 192 B of smem gather per vertex-sample and was 4x off the HBM roofline (profiles/r01), so
// it runs on the tensor cores instead, fp32-faithful from 16-bit split operands like the blend GEMM (common.cuh:
// f16 main parts + f16 residuals, three products; 3xTF32 = 9 K8 MMAs before, and the tensor work matters even in
// this HBM-bound kernel because the whole step runs at the board's power cap):
//   MMA: D[128 vertices, 192 = 16 samples x 12] += W_tile[128, 32] . A_group[192, 32]^T   (6 x M128 N192 K16, K padded)
//   accumulator in TMEM: lane = vertex, 12 columns per sample = the 3x4 blended transform of that vertex.
// CTA = one vertex tile (W tile resident in smem) x a chunk of sample groups, persistent over units:
//   warp 0    producer : bulk copies (TMA engine): A group image (2-stage ring) + the 16 v_posed tile rows of
//                        the group (5-stage ring = the HBM prefetch depth)
//   warp 1    MMA      : tcgen05.mma.kind::f16 into one of two 192-column TMEM accumulators
//   warps 2-9 epilogue : tcgen05.ld T, apply to v_posed (thread = vertex), transpose through smem so the
//                        verts rows are written as contiguous 8-byte-aligned runs (streaming stores)
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int SKT_EPI_WARPS = 8;
constexpr int SKT_THREADS = 96 + 32 * SKT_EPI_WARPS;   // row-producer warp + MMA warp + epilogue warps + operand-producer warp
constexpr int SKT_TMEM_COLS = 512;
constexpr int SKT_ASTAGES = 2;   // A-image ring (released by the MMA)
constexpr int SKT_VSTAGES = 5;   // v_posed ring (released by the epilogue): HBM prefetch depth
constexpr uint32_t SKT_VP_ROW_BYTES = 3 * SK_VT * 4;   // 1536
constexpr int SKT_OUT_PITCH = 3 * SK_VT + 2;
constexpr int SKT_GY_PITCH = 3 * SK_VT + 4;   // MODE 1: g_verts row of a tile + 16 bytes (rows of odd samples are 8 bytes off a 16-byte boundary)

struct SktSmem {
  float w[SK_WTILE];                          // 16 384 B  W tile image (2 parts)
  float a[SKT_ASTAGES][SK_AGRP];              // 24 576 B  per stage: A group image (2 parts)
  union {
    struct {
      float vp[SKT_VSTAGES][SK_S][3 * SK_VT];   // 24 576 B  per stage: v_posed tile rows of the group
      float out[SK_S][SKT_OUT_PITCH];           // 24 704 B  verts staging (AoS); pitch 386: the keypoint-vertex
                                                //           copy reads 16 samples x 2 slots per warp conflict-free
    };
    float gy[SKT_VSTAGES][SK_S][SKT_GY_PITCH];  // MODE 1: 24 832 B per stage: g_verts tile rows of the group
  };
  uint64_t afull[SKT_ASTAGES], aempty[SKT_ASTAGES], vfull[SKT_VSTAGES], vempty[SKT_VSTAGES];
  uint64_t acc_full[2], acc_empty[2], w_full, w_empty;
  uint32_t tmem_base;
};

// MODE 0: forward (above).  MODE 1: adjoint wrt v_posed, g_vposed[b,v] = T[b,v][:3,:3]^T g_verts[b,v]: same
// weight-blend MMA; the epilogue reads g_verts straight from global (stride-3 loads, fetched ONE GROUP AHEAD so
// the DRAM latency hides behind the previous group) and writes the result from registers as the bf16 split
// MN-major A-operand image of the transposed blend GEMM (blend_bwd.cu; 16-byte stores, 512 contiguous bytes per
// warp); `vposed` = g_verts [B,6890,3], `verts` = that image.
template <int MODE>
__global__ void __launch_bounds__(SKT_THREADS, 1)
skin_tc_kernel(int B, int n_groups, int n_chunks, const float* __restrict__ vposed,
               const float* __restrict__ Aimg, const float* __restrict__ Wimg, float* __restrict__ verts,
               float* __restrict__ kpv, int kpv_pitch, const int* __restrict__ kpc_tile_ptr,
               const int* __restrict__ kpc_src, int gy_ring) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SktSmem& sm = *reinterpret_cast<SktSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_units = SK_NVT * n_chunks;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < SKT_ASTAGES; ++s) {
      mbar_init(&sm.afull[s], 1);
      mbar_init(&sm.aempty[s], 1);              // MMA commit
    }
    for (int s = 0; s < SKT_VSTAGES; ++s) {
      mbar_init(&sm.vfull[s], 1);
      mbar_init(&sm.vempty[s], SKT_EPI_WARPS);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sm.acc_full[s], 1);
      mbar_init(&sm.acc_empty[s], SKT_EPI_WARPS);
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
    // ================= row producer: the 16 tile rows of every sample group, one bulk copy per lane =================
    // (one thread issuing the 16 copies and their addresses was the bottleneck of the MODE 1 kernel: ~2500 cycles per group)
    uint32_t itv = 0;   // ring counter (MODE 1: units of the last tile do not use the ring)
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int vt = u % SK_NVT, ch = u / SK_NVT;
      const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
      if (MODE == 1 && !(gy_ring && vt != SK_NVT - 1)) continue;
      for (int g = g0; g < g1; ++g, ++itv) {
        const int sv = itv % SKT_VSTAGES;
        const int ns = min(SK_S, B - g * SK_S);
        if (lane == 0) mbar_wait(&sm.vempty[sv], ((itv / SKT_VSTAGES) & 1) ^ 1);
        __syncwarp();
        if (MODE == 0) {
          if (lane == 0) mbar_arrive_expect_tx(&sm.vfull[sv], ns * SKT_VP_ROW_BYTES);
          __syncwarp();
          if (lane < ns)
            bulk_g2s(sm.vp[sv][lane], vposed + ((size_t)g * SK_S + lane) * VPP + (size_t)vt * (3 * SK_VT), SKT_VP_ROW_BYTES,
                     &sm.vfull[sv]);
        } else {
          // g_verts rows [B][20670] are only 8-byte aligned (82 680 = 16 * 5167 + 8): a row segment that starts 8 bytes
          // off a 16-byte boundary is copied from 8 bytes earlier (valid memory: the previous tile or sample) and the
          // epilogue skips those two floats.  The last tile (1272 bytes per row, not a multiple of 16) is read with
          // plain loads instead.
          const char* r = reinterpret_cast<const char*>(vposed + ((size_t)g * SK_S + (lane < ns ? lane : 0)) * NV3 +
                                                        (size_t)vt * (3 * SK_VT));
          const uint32_t mis = (uint32_t)((uintptr_t)r & 15u);   // 0 or 8
          const uint32_t n_mis = __popc(__ballot_sync(0xffffffffu, lane < ns && mis != 0));
          if (lane == 0) mbar_arrive_expect_tx(&sm.vfull[sv], ns * SKT_VP_ROW_BYTES + 16 * n_mis);
          __syncwarp();
          if (lane < ns) bulk_g2s(sm.gy[sv][lane], r - mis, SKT_VP_ROW_BYTES + 2 * mis, &sm.vfull[sv]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 2 + SKT_EPI_WARPS) {
    // ================= operand producer: W tile per unit, A group image per sample group (2-stage ring) =================
    if (lane == 0) {
      uint32_t it = 0, ul = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
        const int vt = u % SK_NVT, ch = u / SK_NVT;
        const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
        mbar_wait(&sm.w_empty, (ul & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.w_full, SK_WTILE * 4);
        bulk_g2s(sm.w, Wimg + (size_t)vt * SK_WTILE, SK_WTILE * 4, &sm.w_full);
        for (int g = g0; g < g1; ++g, ++it) {
          const int sa = it % SKT_ASTAGES;
          mbar_wait(&sm.aempty[sa], ((it / SKT_ASTAGES) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.afull[sa], SK_AGRP * 4);
          bulk_g2s(sm.a[sa], Aimg + (size_t)g * SK_AGRP, SK_AGRP * 4, &sm.afull[sa]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issuer =================
    if (lane == 0) {
      constexpr uint32_t idesc_h = umma_idesc(0, SK_VT, SK_N);   // kind::f16, f16 operands
      constexpr uint32_t W_LBO = SK_VT * 16, A_LBO = SK_N * 16, SBO = 128;
      constexpr uint32_t W_PART = SK_WPART * 2, A_PART = SK_APART * 2;   // bytes per operand part
      const uint32_t wb = smem_u32(sm.w);
      uint32_t it = 0, ul = 0;
      for (int u = blockIdx.x; u < n_units; u += gridDim.x, ++ul) {
        const int ch = u / SK_NVT;
        const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
        mbar_wait(&sm.w_full, ul & 1);
        for (int g = g0; g < g1; ++g, ++it) {
          const int sa = it % SKT_ASTAGES, ac = it & 1;
          mbar_wait(&sm.acc_empty[ac], ((it >> 1) & 1) ^ 1);
          mbar_wait(&sm.afull[sa], (it / SKT_ASTAGES) & 1);
          tc_fence_after();
          const uint32_t ab = smem_u32(sm.a[sa]);
          const uint32_t dcol = tmem + (uint32_t)ac * SK_N;
#pragma unroll
          for (int ks = 0; ks < SK_KPAD / 16; ++ks) {   // f16(W).f16(A) + f16(Wl).f16(A) + f16(W).f16(Al)
            const uint64_t w0 = umma_desc_kmajor_noswz(wb + ks * 2 * W_LBO, W_LBO, SBO);
            const uint64_t w1 = umma_desc_kmajor_noswz(wb + W_PART + ks * 2 * W_LBO, W_LBO, SBO);
            const uint64_t a0 = umma_desc_kmajor_noswz(ab + ks * 2 * A_LBO, A_LBO, SBO);
            const uint64_t a1 = umma_desc_kmajor_noswz(ab + A_PART + ks * 2 * A_LBO, A_LBO, SBO);
            umma_f16_ss(dcol, w0, a0, idesc_h, ks != 0);
            umma_f16_ss(dcol, w1, a0, idesc_h, 1);
            umma_f16_ss(dcol, w0, a1, idesc_h, 1);
          }
          umma_commit(&sm.aempty[sa]);
          umma_commit(&sm.acc_full[ac]);
        }
        umma_commit(&sm.w_empty);   // W tile may be replaced once every MMA of this unit has read it
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..9): thread = vertex lane, two warps per TMEM lane quarter ====
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;             // samples [8*half, 8*half+8) of the group
    const int l4 = q * 32 + lane;                 // TMEM lane = vertex within the tile
    const int ew = warp - 2;                      // 0..7: copy-out rows 2*ew, 2*ew+1
    uint32_t it = 0, itv = 0;                     // itv: MODE 1 g_verts ring counter (units of the last tile skip the ring)
    // MODE 1: float offset of this thread's vertex inside the staged rows of even / odd samples (a row starts
    // 0 or 2 floats into its 16-byte unit: NV3 = 2 (mod 4), groups start at even samples)
    const int gy_sh0 = (int)(((uintptr_t)vposed >> 2) & 3);
    const int gy_off0 = gy_sh0 + 3 * l4, gy_off1 = ((gy_sh0 + 2) & 3) + 3 * l4;
    // MODE 1: 16-bit-element offset of this thread's column inside the g_vposed image (gimg16_off without the sample terms)
    const size_t img_col = (size_t)(l4 >> 6) * BB_ASTAGE + (size_t)(((l4 & 63) >> 3) * 64 + (l4 & 7) * 8);
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
      const int vt = u % SK_NVT, ch = u / SK_NVT;
      const int g0 = (int)((long long)ch * n_groups / n_chunks), g1 = (int)((long long)(ch + 1) * n_groups / n_chunks);
      const int nvalid2 = min(3 * SK_VT, NV3 - vt * 3 * SK_VT) / 2;   // float2 of this tile that exist in verts
      // compact keypoint-vertex slots of this tile: thread -> (sample i = et & 15, slot ks0 + (et >> 4) + 16 p);
      // the source indices of the first 4 passes (64 slots) are fetched once per unit
      int ks0 = 0, ks1 = 0, ksrc[4] = {-1, -1, -1, -1};
      const int et = (int)threadIdx.x - 64;
      if (MODE == 0 && kpv) {
        ks0 = __ldg(kpc_tile_ptr + vt);
        ks1 = __ldg(kpc_tile_ptr + vt + 1);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const int sl = ks0 + (et >> 4) + 16 * p;
          if (sl < ks1) ksrc[p] = __ldg(kpc_src + sl);
        }
      }
      for (int g = g0; g < g1; ++g, ++it) {
        const int sv = it % SKT_VSTAGES, ac = it & 1;
        const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)ac * SK_N + half * (SK_S / 2) * 12;
        if (MODE == 1) {
          // ---- adjoint wrt v_posed ----
          const int v = vt * SK_VT + l4;
          const bool vok = v < NV;
          // g_verts of this thread's vertex: from the bulk-copy ring (5 groups of prefetch; stride-3 smem reads are
          // conflict-free), or with plain loads for the last tile / an unaligned buffer (1/54 of the work)
          const bool ring = gy_ring && vt != SK_NVT - 1;
          int sr = 0;
          if (ring) {
            sr = itv % SKT_VSTAGES;
            mbar_wait_spin(smem_u32(&sm.vfull[sr]), (itv / SKT_VSTAGES) & 1);
            ++itv;
          }
          mbar_wait_spin(smem_u32(&sm.acc_full[ac]), (it >> 1) & 1);
          tc_fence_after();
          float res[3][SK_S / 2];   // g_vposed of this vertex: [component][sample of this warp's half]
#pragma unroll
          for (int bt = 0; bt < SK_S / 8; ++bt) {
            uint32_t r0[16], r1[16], r2[16];
            tmem_ld16(tbase + bt * 48, r0);
            tmem_ld16(tbase + bt * 48 + 16, r1);
            tmem_ld16(tbase + bt * 48 + 32, r2);
            float gy[4][3];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int i = half * (SK_S / 2) + bt * 4 + k;
              const int b = g * SK_S + i;
              if (ring) {
                const float* row = sm.gy[sr][i] + ((k & 1) ? gy_off1 : gy_off0);
#pragma unroll
                for (int r = 0; r < 3; ++r) gy[k][r] = row[r];
              } else {
                const float* src = vposed + (size_t)(b < B ? b : B - 1) * NV3 + 3 * (size_t)(vok ? v : 0);
#pragma unroll
                for (int r = 0; r < 3; ++r) gy[k][r] = vok ? __ldg(src + r) : 0.f;
              }
            }
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              // rows of T for sample k: r0 = T[0..15], r1 = T[16..31], r2 = T[32..47]; sample k uses T[12k .. 12k+11]
              float t[12];
#pragma unroll
              for (int e = 0; e < 12; ++e) {
                const int idx = 12 * k + e;
                t[e] = __uint_as_float(idx < 16 ? r0[idx & 15] : (idx < 32 ? r1[idx & 15] : r2[idx & 15]));
              }
#pragma unroll
              for (int c = 0; c < 3; ++c)
                res[c][bt * 4 + k] = fmaf(t[c], gy[k][0], fmaf(t[4 + c], gy[k][1], t[8 + c] * gy[k][2]));   // (T.R^T gy)[c]
            }
          }
          if (ring) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.vempty[sr]);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.acc_empty[ac]);
          // Straight from registers into the A-operand image of the transposed blend GEMM (gimg16_off, MN-major): the 8
          // samples of this thread are one sample group of the image, i.e. 16 contiguous bytes per (column, part); the 32
          // lanes of a warp (consecutive vertices = consecutive columns of one component) cover 512 contiguous bytes.
          // bf16 split: g1 = bf16(g), g2 = bf16(g - g1).  (Round 1 staged a TF32 hi/lo image through shared memory:
          // twice the bytes, two block barriers per group.)
          {
            const int b0 = g * SK_S + half * (SK_S / 2);          // first sample of this thread's group (multiple of 8)
            if (b0 < B) {
              // gimg16_off(b0, vt * 384 + c * 128 + l4, part): the component moves two k-blocks, the part BB_APART
              unsigned short* img = reinterpret_cast<unsigned short*>(verts) +
                                    ((size_t)(b0 / BL_BM) * BB_NKB + (size_t)vt * 6) * BB_ASTAGE +
                                    (size_t)((b0 % BL_BM) / 8) * ((BB_KB / 8) * 64) + img_col;
#pragma unroll
              for (int c = 0; c < 3; ++c) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int k2 = 0; k2 < 4; ++k2) {   // one cvt.rn.bf16x2.f32 per pair and part
                  const float x0 = res[c][2 * k2], x1 = res[c][2 * k2 + 1];
                  hi[k2] = pack_bf16x2(x0, x1);
                  lo[k2] = pack_bf16x2(x0 - __uint_as_float(hi[k2] << 16), x1 - __uint_as_float(hi[k2] & 0xffff0000u));
                }
                unsigned short* dst = img + (size_t)c * (2 * BB_ASTAGE);
                __stcs(reinterpret_cast<uint4*>(dst), make_uint4(hi[0], hi[1], hi[2], hi[3]));
                __stcs(reinterpret_cast<uint4*>(dst + BB_APART), make_uint4(lo[0], lo[1], lo[2], lo[3]));
              }
            }
          }
          continue;
        }
        mbar_wait_spin(smem_u32(&sm.vfull[sv]), (it / SKT_VSTAGES) & 1);   // v_posed rows of this group have landed
        mbar_wait_spin(smem_u32(&sm.acc_full[ac]), (it >> 1) & 1);         // blended transforms are in TMEM
        tc_fence_after();
#pragma unroll
        for (int bt = 0; bt < SK_S / 8; ++bt) {
          uint32_t r0[16], r1[16], r2[16];
          tmem_ld16(tbase + bt * 48, r0);
          tmem_ld16(tbase + bt * 48 + 16, r1);
          tmem_ld16(tbase + bt * 48 + 32, r2);
          float xs[4], ys[4], zs[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float* row = sm.vp[sv][half * (SK_S / 2) + bt * 4 + k];
            xs[k] = row[l4]; ys[k] = row[SK_VT + l4]; zs[k] = row[2 * SK_VT + l4];
          }
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
            float* orow = sm.out[half * (SK_S / 2) + bt * 4 + k] + 3 * l4;
            const float* t = T + 12 * k;
#pragma unroll
            for (int r = 0; r < 3; ++r)
              orow[r] = fmaf(t[4 * r], xs[k], fmaf(t[4 * r + 1], ys[k], fmaf(t[4 * r + 2], zs[k], t[4 * r + 3])));
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&sm.acc_empty[ac]);
          mbar_arrive(&sm.vempty[sv]);
        }
        named_bar_sync(1, 32 * SKT_EPI_WARPS);   // staging complete
        const int ns = min(SK_S, B - g * SK_S);
        // vertices the keypoint regressor touches -> compact rows kpv[group][slot][16 samples]: one contiguous
        // (slots x 64 B) run per (tile, group), each warp store = 128 contiguous bytes
        if (ks1 > ks0) {
          const int i = et & 15;
          float* kd = kpv + ((size_t)g * kpv_pitch + ks0 + (et >> 4)) * SK_S + i;
#pragma unroll
          for (int p = 0; p < 4; ++p)
            if (ks0 + (et >> 4) + 16 * p < ks1) kd[p * 16 * SK_S] = ksrc[p] >= 0 ? sm.out[i][ksrc[p]] : 0.f;
          for (int sl = ks0 + (et >> 4) + 64; sl < ks1; sl += 16) {   // tiles with more than 64 slots (dense regressors)
            const int src = __ldg(kpc_src + sl);
            kpv[((size_t)g * kpv_pitch + sl) * SK_S + i] = src >= 0 ? sm.out[i][src] : 0.f;
          }
        }
        float* dst = verts + (size_t)g * SK_S * NV3 + (size_t)vt * (3 * SK_VT);
#pragma unroll
        for (int rr = 0; rr < SK_S / SKT_EPI_WARPS; ++rr) {
          const int i = ew * (SK_S / SKT_EPI_WARPS) + rr;
          if (i < ns && verts) {
            float2 v[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) v[k] = *reinterpret_cast<const float2*>(&sm.out[i][2 * (lane + 32 * k)]);
#pragma unroll
            for (int k = 0; k < 6; ++k)
              if (lane + 32 * k < nvalid2) __stcs(reinterpret_cast<float2*>(dst + (size_t)i * NV3) + lane + 32 * k, v[k]);
          }
        }
        named_bar_sync(1, 32 * SKT_EPI_WARPS);   // staging reusable
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

template <int MODE>
static int32_t launch_skin_tc_mode(const smplb_model* m, int B, const float* in, const float* Aimg, float* out,
                                   float* kpv, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static KernelSetup setup;   // one per MODE instantiation
  const size_t smem = sizeof(SktSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, skin_tc_kernel<MODE>, smem, &n_sm)) return rc;
  const int n_groups = (B + SK_S - 1) / SK_S;
  // units = 54 vertex tiles x n_chunks sample chunks over a whole number of CTA waves
  // 7 waves when every unit still covers >= 32 sample groups (tail of the last wave 1 % instead of 3 %: 2.12 -> 2.10 ms
  // at B = 65536), else 3: every unit reloads the 24 KB W tile and refills its pipeline
  int n_chunks = (7 * n_sm) / SK_NVT;
  if (n_groups < 32 * n_chunks) n_chunks = (3 * n_sm) / SK_NVT;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > n_groups) n_chunks = n_groups;
  const int n_units = SK_NVT * n_chunks;
  const int grid = n_units < n_sm ? n_units : n_sm;
  // MODE 1: g_verts rows go through the bulk-copy ring when the buffer is 8-byte aligned (any contiguous fp32
  // [B,6890,3] tensor whose base is; a view that starts at an odd float falls back to plain loads)
  const int gy_ring = (MODE == 1 && (((uintptr_t)in) & 7u) == 0) ? 1 : 0;
  skin_tc_kernel<MODE><<<grid, SKT_THREADS, smem, st>>>(B, n_groups, n_chunks, in, Aimg, m->Wimg, out, kpv, m->kpv_pitch,
                                                        m->kpc_tile_ptr, m->kpc_src, gy_ring);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

int32_t launch_skin_tc(const smplb_model* m, int B, const float* vposed, const float* Aimg, float* verts, float* kpv,
                       cudaStream_t st) {
  if (!verts && !kpv) return SMPLB_OK;
  return launch_skin_tc_mode<0>(m, B, vposed, Aimg, verts, kpv, st);
}

// g_verts [B,6890,3] -> bf16 split operand image of g_vposed (gimg16_off)
int32_t launch_skin_bwd_vposed(const smplb_model* m, int B, const float* g_verts, const float* Aimg, float* g_img,
                               cudaStream_t st) {
  return launch_skin_tc_mode<1>(m, B, g_verts, Aimg, g_img, nullptr, st);
}

}  // namespace smplb
