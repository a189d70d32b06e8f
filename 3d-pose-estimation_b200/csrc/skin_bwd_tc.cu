// Skinning adjoint, reduction half, on the tensor cores (VERDICT r1 "next" 2(i)):
//     g_A[b, j, (r,c)] = sum_v weights[v,j] * P[b, v, (r,c)],      P[b,v,(r,c)] = g_verts[b,v,r] * [v_posed[b,v]; 1][c]
// i.e. per group of 8 samples the GEMM  D[(e,s) 96 rows, 32 joints] += P[(e,s), 128 vertices] . W[32 joints, 128 vertices]^T
// accumulated over the 54 vertex tiles IN TMEM (no partial sums, fixed order: deterministic).
//   * A operand = P, formed in shared memory by the compute warps (thread = vertex, 8 samples): MN-MAJOR bf16 image, so a
//     thread's 8 samples of one component e are one 16-byte store (rows (e,s) = e * 8 + s); bf16 split P = P1 + P2 like
//     the transposed blend GEMM (gradient magnitudes ~1/(B*20670) leave the fp16 range): W1.P1 + W2.P1 + W1.P2, error 2^-16.
//     Rows 96..127 of the M = 128 MMA read whatever follows the 12 row groups; they only reach TMEM lanes 96..127, which
//     nobody reads.
//   * B operand = LBS weights of the tile, K-major bf16 split image built at model creation (dense weights supported).
//   * a CTA owns a PASS of up to 64 samples (8 groups x 64 TMEM columns) and sweeps the 54 tiles; g_verts / v_posed
//     tile rows stream through a 3-stage bulk-copy ring, the W tile through a 2-stage ring.
//   warp 0      producer (rows: one bulk copy per lane; W tile per vertex tile)
//   warp 1      MMA issue: 8 x (M128 N64 K16 + M128 N32 K16) per (tile, group)
//   warps 2-9   two warpgroups that alternate over the (tile, group) granules, each with its own P buffer; warpgroup 0
//               also runs the epilogue of a pass (TMEM -> g_A)
// The CUDA-core kernel (skin_bwd.cu, 4.1 ms at B = 65536, bound by shared-memory wavefronts of its lane = sample
// gather) stays as the verification path.
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int GT_S = 8;                        // samples per group
constexpr int GT_MAXG = 8;                     // groups per pass: 8 x 64 = 512 TMEM columns
constexpr int GT_RST = 4;                      // row-ring stages (what fits beside the two 48 KB P buffers)
constexpr int GT_GPITCH = 3 * SK_VT + 4;       // floats per staged g_verts row (+16 B: rows of odd samples start 8 B off)
constexpr uint32_t GT_PPART = 12 * 2048;       // bytes per part of a P granule: 12 row groups (e) x 16 k-groups x 128 B
constexpr uint32_t GT_WPART = 16 * 32 * 16;    // bytes per part of a W tile image: 16 k-groups x 32 rows x 16 B (the two parts
                                               // are interleaved per k-group: rows 0..31 = W1, rows 32..63 = W2)
constexpr int GT_THREADS = 64 + 256;

struct GtSmem {
  unsigned char p[2][2 * GT_PPART];            // 2 x 49 152 B (MMA rows 96..127 of the last part read 8 KB into `w`)
  unsigned char w[2][2 * GT_WPART];            // 2 x 16 384 B
  float gy[GT_RST][GT_S][GT_GPITCH];           // 3 x 12 416 B
  float vp[GT_RST][GT_S][3 * SK_VT];           // 3 x 12 288 B
  uint64_t rfull[GT_RST], rempty[GT_RST], pfull[2], pempty[2], wfull[2], wempty[2], acc_full, acc_empty;
  uint32_t tmem_base;
};

// 16-bit-element offset of (vertex v, joint j) in the W image of this kernel: [tile][k-group 16][part 2][row 32][8], i.e.
// one K-major operand of 64 rows [W1; W2] per tile (and its first 32 rows alone are the operand W1)
__host__ __device__ inline size_t gt_wimg_off(int v, int j, int part) {
  return (size_t)(v / SK_VT) * (2 * GT_WPART / 2) + (size_t)(((v % SK_VT) / 8) * 64 + part * 32 + j) * 8 + (size_t)(v & 7);
}

__global__ void __launch_bounds__(GT_THREADS, 1)
skin_bwd_gA_tc_kernel(int B, int G, int n_pass, const float* __restrict__ g_verts, const float* __restrict__ vposed,
                      const unsigned short* __restrict__ Wimg3, float* __restrict__ g_A, int g_ring) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  GtSmem& sm = *reinterpret_cast<GtSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int PS = G * GT_S;   // samples per pass

  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_RST; ++s) {
      mbar_init(&sm.rfull[s], 1);
      mbar_init(&sm.rempty[s], 4);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sm.pfull[s], 4);
      mbar_init(&sm.pempty[s], 1);
      mbar_init(&sm.wfull[s], 1);
      mbar_init(&sm.wempty[s], 1);
    }
    mbar_init(&sm.acc_full, 1);
    mbar_init(&sm.acc_empty, 4);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= producer =================
    uint32_t n = 0, wn = 0;
    for (int ps = blockIdx.x; ps < n_pass; ps += gridDim.x) {
      const int b0 = ps * PS;
      const int ng = min(G, (B - b0 + GT_S - 1) / GT_S);
      for (int vt = 0; vt < SK_NVT; ++vt, ++wn) {
        if (lane == 0) {
          const uint32_t wb = wn & 1;
          mbar_wait(&sm.wempty[wb], ((wn >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.wfull[wb], 2 * GT_WPART);
          bulk_g2s(sm.w[wb], Wimg3 + (size_t)vt * GT_WPART, 2 * GT_WPART, &sm.wfull[wb]);
        }
        const bool gr = g_ring && vt != SK_NVT - 1;   // last tile: 1272 B per row, read with plain loads
        for (int gi = 0; gi < ng; ++gi, ++n) {
          const uint32_t s = n % GT_RST;
          const int bg = b0 + gi * GT_S;
          const int ns = min(GT_S, B - bg);
          if (lane == 0) mbar_wait(&sm.rempty[s], ((n / GT_RST) & 1) ^ 1);
          __syncwarp();
          const char* r = reinterpret_cast<const char*>(g_verts + ((size_t)bg + (lane < ns ? lane : 0)) * NV3 +
                                                        (size_t)vt * (3 * SK_VT));
          const uint32_t mis = (uint32_t)((uintptr_t)r & 15u);   // 0 or 8 (rows are 82 680 B apart)
          const uint32_t n_mis = __popc(__ballot_sync(0xffffffffu, lane < ns && mis != 0));
          if (lane == 0)
            mbar_arrive_expect_tx(&sm.rfull[s], (uint32_t)ns * (3 * SK_VT * 4) * (gr ? 2u : 1u) + (gr ? 16u * n_mis : 0u));
          __syncwarp();
          if (lane < ns) {
            if (gr) bulk_g2s(sm.gy[s][lane], r - mis, 3 * SK_VT * 4 + 2 * mis, &sm.rfull[s]);
            bulk_g2s(sm.vp[s][lane], vposed + ((size_t)bg + lane) * VPP + (size_t)vt * (3 * SK_VT), 3 * SK_VT * 4, &sm.rfull[s]);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================= MMA issue =================
    if (lane == 0) {
      // bf16 x bf16, A MN-major, B K-major, M 128.  The three split products W1.P1 + W2.P1 + W1.P2 are TWO MMAs per K16
      // step: P1 against the 64-row operand [W1; W2] (columns 0..31 and 32..63 of the group's accumulator, added in the
      // epilogue) and P2 against W1 -- P1 is read from shared memory once instead of twice (the kernel is bound by
      // shared-memory bandwidth: TMA writes + LDS reads + P stores + MMA operand reads).
      constexpr uint32_t idesc64 = umma_idesc_ex(1, 1, 1, 0, 128, 64), idesc32 = umma_idesc_ex(1, 1, 1, 0, 128, 32);
      const uint64_t a_t = umma_desc_kmajor_noswz(0, 128, 2048);       // P: LBO = k-group stride, SBO = row-group stride
      const uint64_t b_t = umma_desc_kmajor_noswz(0, 64 * 16, 128);    // W: LBO = k-group stride, SBO = 8-row group stride
      uint32_t n = 0, wn = 0, pn = 0;
      for (int ps = blockIdx.x; ps < n_pass; ps += gridDim.x, ++pn) {
        const int b0 = ps * PS;
        const int ng = min(G, (B - b0 + GT_S - 1) / GT_S);
        if (pn > 0) mbar_wait(&sm.acc_empty, (pn - 1) & 1);
        tc_fence_after();
        for (int vt = 0; vt < SK_NVT; ++vt, ++wn) {
          const uint32_t wb = wn & 1;
          mbar_wait(&sm.wfull[wb], (wn >> 1) & 1);
          const uint32_t wa = smem_u32(sm.w[wb]) >> 4;
          for (int gi = 0; gi < ng; ++gi, ++n) {
            const uint32_t pb = n & 1;
            mbar_wait(&sm.pfull[pb], (n >> 1) & 1);
            tc_fence_after();
            const uint32_t pa = smem_u32(sm.p[pb]) >> 4;
            const uint32_t dcol = tmem + (uint32_t)gi * 64;
#pragma unroll
            for (int ks = 0; ks < SK_VT / 16; ++ks) {
              const uint64_t p1 = a_t + (uint64_t)(pa + ks * (256 / 16)), p2 = p1 + (uint64_t)(GT_PPART / 16);
              const uint64_t w12 = b_t + (uint64_t)(wa + ks * (2048 / 16));
              umma_f16_ss(dcol, p1, w12, idesc64, (vt != 0 || ks != 0) ? 1u : 0u);   // P1.[W1; W2]
              umma_f16_ss(dcol, p2, w12, idesc32, 1u);                                // P2.W1
            }
            umma_commit(&sm.pempty[pb]);
          }
          umma_commit(&sm.wempty[wb]);
        }
        umma_commit(&sm.acc_full);
      }
    }
    __syncwarp();
  } else {
    // ================= compute warps: form P; warpgroup 0 also writes g_A at the end of a pass =================
    const int wg = (warp - 2) >> 2;                 // warpgroup: granules n = wg (mod 2), P buffer wg
    const int k = ((warp - 2) & 3) * 32 + lane;     // vertex inside the tile
    const int q = warp & 3;                         // TMEM lane quarter of this warp (epilogue)
    const int gy_sh0 = (int)(((uintptr_t)g_verts >> 2) & 3);
    unsigned char* pbuf = sm.p[wg] + (k >> 3) * 128 + (k & 7) * 16;
    uint32_t n = 0, pn = 0;
    for (int ps = blockIdx.x; ps < n_pass; ps += gridDim.x, ++pn) {
      const int b0 = ps * PS;
      const int ng = min(G, (B - b0 + GT_S - 1) / GT_S);
      for (int vt = 0; vt < SK_NVT; ++vt) {
        const bool gr = g_ring && vt != SK_NVT - 1;
        const int v = vt * SK_VT + k;
        const bool vok = v < NV;
        for (int gi = 0; gi < ng; ++gi, ++n) {
          if ((int)(n & 1) != wg) continue;
          const uint32_t s = n % GT_RST;
          const int bg = b0 + gi * GT_S;
          const int ns = min(GT_S, B - bg);
          mbar_wait_spin(smem_u32(&sm.rfull[s]), (n / GT_RST) & 1);
          float g[GT_S][3], x[GT_S][3];
#pragma unroll
          for (int i = 0; i < GT_S; ++i) {
            if (i < ns && vok) {
              if (gr) {
                const float* row = sm.gy[s][i] + ((gy_sh0 + 2 * ((bg + i) & 1)) & 3) + 3 * k;
#pragma unroll
                for (int r = 0; r < 3; ++r) g[i][r] = row[r];
              } else {
                const float* src = g_verts + (size_t)(bg + i) * NV3 + 3 * (size_t)v;
#pragma unroll
                for (int r = 0; r < 3; ++r) g[i][r] = __ldg(src + r);
              }
#pragma unroll
              for (int c = 0; c < 3; ++c) x[i][c] = sm.vp[s][i][c * SK_VT + k];
            } else {
#pragma unroll
              for (int r = 0; r < 3; ++r) { g[i][r] = 0.f; x[i][r] = 0.f; }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.rempty[s]);
          mbar_wait_spin(smem_u32(&sm.pempty[wg]), ((n >> 1) & 1) ^ 1);
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t hi[4], lo[4];
#pragma unroll
              for (int i2 = 0; i2 < 4; ++i2) {
                const float v0 = c < 3 ? g[2 * i2][r] * x[2 * i2][c] : g[2 * i2][r];
                const float v1 = c < 3 ? g[2 * i2 + 1][r] * x[2 * i2 + 1][c] : g[2 * i2 + 1][r];
                hi[i2] = pack_bf16x2(v0, v1);
                lo[i2] = pack_bf16x2(v0 - __uint_as_float(hi[i2] << 16), v1 - __uint_as_float(hi[i2] & 0xffff0000u));
              }
              unsigned char* d = pbuf + (r * 4 + c) * 2048;
              *reinterpret_cast<uint4*>(d) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
              *reinterpret_cast<uint4*>(d + GT_PPART) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&sm.pfull[wg]);
        }
      }
      if (wg == 0) {
        // ---- epilogue of the pass: lane L of the accumulator = (e = L / 8, sample = L % 8), 32 joint columns per group ----
        mbar_wait_spin(smem_u32(&sm.acc_full), pn & 1);
        tc_fence_after();
        const int L = q * 32 + lane, e = L >> 3, si = L & 7;
        for (int gi = 0; gi < ng; ++gi) {
          uint32_t r[32], r2[32];
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)gi * 64, r);
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)gi * 64 + 32, r2);
          tmem_ld_wait();
          const int b = b0 + gi * GT_S + si;
          if (e < 12 && b < B) {
            float* o = g_A + (size_t)b * (NJ * 12) + e;
#pragma unroll
            for (int j = 0; j < NJ; ++j) o[j * 12] = __uint_as_float(r[j]) + __uint_as_float(r2[j]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.acc_empty);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// host: W image of the kernel (bf16 split, K-major), built at model creation
void build_wimg3(const float* weights, unsigned short* out) {
  for (int v = 0; v < NV; ++v)
    for (int j = 0; j < NJ; ++j) {
      const float x = weights[(size_t)v * NJ + j];
      const unsigned short h = bf16_rne_bits(x);
      out[gt_wimg_off(v, j, 0)] = h;
      out[gt_wimg_off(v, j, 1)] = bf16_rne_bits(x - bf16_bits_to_float(h));
    }
}
size_t wimg3_elems() { return (size_t)SK_NVT * 2 * (GT_WPART / 2); }

int32_t launch_skin_bwd_gA_tc(const smplb_model* m, int B, const float* g_verts, const float* vposed, float* g_A,
                              cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static KernelSetup setup;
  const size_t smem = sizeof(GtSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, skin_bwd_gA_tc_kernel, smem, &n_sm)) return rc;
  // groups per pass: the choice with the fewest (waves x groups), i.e. the least wave quantisation
  const int n_groups = (B + GT_S - 1) / GT_S;
  int best_g = 1;
  long long best = 1ll << 60;
  for (int g = 1; g <= GT_MAXG; ++g) {
    const int np = (n_groups + g - 1) / g;
    const long long cost = (long long)((np + n_sm - 1) / n_sm) * (g * 54 + 6);   // + a pass's fixed cost
    if (cost < best) { best = cost; best_g = g; }
  }
  const int n_pass = (n_groups + best_g - 1) / best_g;
  const int grid = n_pass < n_sm ? n_pass : n_sm;
  const int g_ring = (((uintptr_t)g_verts) & 7u) == 0 ? 1 : 0;
  skin_bwd_gA_tc_kernel<<<grid, GT_THREADS, smem, st>>>(B, best_g, n_pass, g_verts, vposed,
                                                       reinterpret_cast<const unsigned short*>(m->Wimg3), g_A, g_ring);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
