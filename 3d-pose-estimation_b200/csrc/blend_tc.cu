// K2 blend-shape GEMM on the 5th-generation tensor cores (batch_smpl.py:91-95, :113-115):
//     v_posed[B, 20670] = [beta | pose_feature][B, 217] . [shapedirs; posedirs][217, 20670] + v_template
// fp32-faithful on the tensor cores from 16-bit operands (operand images and error analysis: common.cuh):
//     acc += f16(X).f16(D') + f16(X - f16(X)).f16(D') + f16(X).f16(D' - f16(D')),   D' = 2^e D (the epilogue multiplies by 2^-e)
// three kind::f16 MMAs (K = 16) per 16-element k-block, fp32 accumulation in TMEM.  History of this kernel at
// B = 65536: 3xTF32, 6 MMA slots and 8 B per element: 2.50 ms (tensor pipe 83 % busy);  TF32 main product + two
// bf16 corrections, 4 slots, 8 B: 2.41 ms, bound by the EPILOGUE (an L2 round trip for v_template in every
// 32-column block), 2.08 ms with that load hoisted above the accumulator wait, then at the ~43 B/clk/SM the L2
// delivers (ncu: 19.1 GB of operands in 2.9 M cycles);  keeping the TF32 part of X resident in smem cut the
// operand stream by 25 % but left room for only 4 x 24 KB stages and became latency-bound (2.28 ms, rejected);
// fp16 main product (same 11 significant bits as TF32): 3 slots, 6 B per element, 8 x 24 KB stages: 1.67 ms;
// the ring depth, not the L2 bandwidth, is what that version waited on (ring bytes / consumption rate = 3000
// cycles of cover; the X-stationary variants traded ring for stream one for one), so the second product now
// reuses the f16(D) part (f16(Xl) instead of bf16(Xl).bf16(f16 D)): D slabs 12 -> 8 KB, 10 x 20 KB stages: 1.55 ms, tensor pipe
// 71 % active, the pair asking L2 for 53 B/clk/SM at the full tensor rate against the ~43 it delivers;  D stored scaled by a
// power of two so that its residual is a normal fp16 number too: the bf16(X) part goes, X slabs 12 -> 8 KB (41 B/clk/SM),
// 12 x 16 KB stages.
//
// v2 organisation: CTA PAIRS (cluster of 2 = one TPC, tcgen05 cta_group::2), persistent over 256 x 256
// output tiles.  v1 (one CTA per 128 x 256 tile, 2 x 96 KB stages) pulled 688 KB of operands from L2 per
// tile = 64 B/clk/SM at the tensor-pipe rate, more than the L2 delivers (~43 B/clk/SM): tensor pipe 50 %
// busy.  A pair shares the D tile (each CTA stages one 128-column half), so a CTA reads 287 KB per
// 128 x 256 outputs, and the smaller stages (20 KB) make a 10-deep ring.
//   warp 0    producer : per k-block one 1-D bulk copy (TMA engine) of its 128 X rows and one of its D half
//                        into a 10-stage smem ring; the operands sit in global memory as ready-made smem
//                        images (common.cuh)
//   warp 1    leader CTA: one thread issues tcgen05.mma.cta_group::2.kind::f16 (M=256, N=256, K=16) x 3
//                        products x 14 k-blocks per tile into one of two 256-column TMEM accumulators (each
//                        CTA holds its 128 rows); commits are multicast to both CTAs' barriers.
//             peer CTA : one thread relays "my stage is full" to the leader's peer_full barrier
//   warps 2-5 epilogue : tcgen05.ld the accumulator (lane = sample row), transpose 32 x 32 blocks through
//                        smem so that every store instruction writes four full 128-byte lines, + v_template
// smem ring <-> MMA and TMEM <-> epilogue are mbarrier pipelines, so the epilogue of tile i overlaps the
// MMAs of tile i+1.
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int TC_STAGES = 12;
constexpr int TC_X_STAGE_FLOATS = BL_SLAB;   // one X slab: 2 parts x 4 KB
constexpr int TC_D_STAGE_FLOATS = BL_DSLAB;  // one D half slab: 2 parts x 4 KB
constexpr uint32_t TC_STAGE_BYTES = (TC_X_STAGE_FLOATS + TC_D_STAGE_FLOATS) * 4;
constexpr int TC_THREADS = 192;
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_EP_PITCH = 36;   // floats per staged row (32 + 4: conflict-free 16-byte accesses both ways)

struct TcSmem {
  float x[TC_STAGES][TC_X_STAGE_FLOATS];
  float d[TC_STAGES][TC_D_STAGE_FLOATS];
  float ep[4][32 * TC_EP_PITCH];
  uint64_t full[TC_STAGES];        // local bulk copies landed
  uint64_t peer_full[TC_STAGES];   // (leader) the peer's copies landed
  uint64_t empty[TC_STAGES];       // MMAs that read the stage completed (multicast commit)
  uint64_t acc_full[2];            // accumulator complete (multicast commit)
  uint64_t acc_empty[2];           // (leader) both CTAs' epilogue warps drained the accumulator
  uint32_t tmem_base;
};

// One 32-row x 32-column block of the accumulator: registers (thread = row) -> smem -> coalesced stores.
__device__ __forceinline__ void tc_store_block(const uint32_t (&r)[32], float* stage, int lane, const float4 tv,
                                               float* obase, int rows_valid, float sc) {
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(stage + lane * TC_EP_PITCH + j * 4) =
        make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                    __uint_as_float(r[4 * j + 3]));
  __syncwarp();
  const int c4 = (lane & 7) * 4, rsub = lane >> 3;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int row = it * 4 + rsub;
    float4 v = *reinterpret_cast<const float4*>(stage + row * TC_EP_PITCH + c4);
    v.x = fmaf(v.x, sc, tv.x);   // sc = 2^-e undoes the scale of the D image (exact), + v_template
    v.y = fmaf(v.y, sc, tv.y);
    v.z = fmaf(v.z, sc, tv.z);
    v.w = fmaf(v.w, sc, tv.w);
    if (row < rows_valid) __stcs(reinterpret_cast<float4*>(obase + (size_t)row * VPP + c4), v);
  }
  __syncwarp();
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_THREADS, 1)
blend_tc_kernel(int B, int n_mtiles, const float* __restrict__ Ximg, const float* __restrict__ Dimg,
                const float* __restrict__ vt, float* __restrict__ vposed, float d_scale_inv) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int n_tiles = ((n_mtiles + 1) / 2) * BL_NT;   // 256-row x 256-column pair tiles
  const int cl = blockIdx.x >> 1, n_cl = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < TC_STAGES; ++s) {
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
  if (warp == 1) tmem_alloc_pair(&sm.tmem_base, TC_TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= producer (both CTAs) =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = cl; t < n_tiles; t += n_cl) {
        const int mt2 = t / BL_NT, nt = t % BL_NT;
        int mt = mt2 * 2 + (int)rank;
        if (mt >= n_mtiles) mt = n_mtiles - 1;   // odd tile count: the peer recomputes a valid tile, stores nothing
        const float* xs = Ximg + (size_t)mt * BL_XTILE;
        const float* ds = Dimg + (size_t)nt * BL_DTILE + (size_t)rank * TC_D_STAGE_FLOATS;
        for (int kb = 0; kb < BL_NKB; ++kb, ++it) {
          const int s = it % TC_STAGES;
          const uint32_t ph = (it / TC_STAGES) & 1;
          mbar_wait(&sm.empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&sm.full[s], TC_STAGE_BYTES);
          bulk_g2s(sm.x[s], xs + (size_t)kb * TC_X_STAGE_FLOATS, TC_X_STAGE_FLOATS * 4, &sm.full[s]);
          bulk_g2s(sm.d[s], ds + (size_t)kb * 2 * TC_D_STAGE_FLOATS, TC_D_STAGE_FLOATS * 4, &sm.full[s]);
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
      for (int t = cl; t < n_tiles; t += n_cl)
        for (int kb = 0; kb < BL_NKB; ++kb) {
          mbar_wait_spin(full0 + s * 8, ph);
          if (issue) mbar_arrive_remote(&sm.peer_full[s], 0);
          if (++s == TC_STAGES) { s = 0; ph ^= 1; }
        }
    } else {
      // ================= MMA issuer (leader CTA) =================
      constexpr uint32_t idesc_h = umma_idesc(0, 2 * BL_BM, BL_BN);   // kind::f16, f16 operands
      constexpr uint32_t LBO = BL_BM * 16, SBO = 128;                 // bytes; X and D halves are both 128 rows
      constexpr uint32_t PART = BL_PART16 * 2 / 16;                   // one 4 KB part in descriptor units
      static_assert(BL_BM == BL_BH, "one descriptor template for both operands");
      const uint64_t dtempl = umma_desc_kmajor_noswz(0, LBO, SBO);
      const uint32_t empty0 = smem_u32(&sm.empty[0]);
      const uint32_t xb0 = smem_u32(sm.x[0]) >> 4, db0 = smem_u32(sm.d[0]) >> 4;
      uint32_t s = 0, ph = 0, tl = 0;
      for (int t = cl; t < n_tiles; t += n_cl, ++tl) {
        const uint32_t a = tl & 1;
        mbar_wait_spin(smem_u32(&sm.acc_empty[a]), ((tl >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem + a * BL_BN;
#pragma unroll 1
        for (int kb = 0; kb < BL_NKB; ++kb) {
          mbar_wait_spin(full0 + s * 8, ph);
          mbar_wait_spin(pfull0 + s * 8, ph);
          tc_fence_after();
          const uint64_t xd = dtempl + (uint64_t)(xb0 + s * (TC_X_STAGE_FLOATS * 4 / 16));
          const uint64_t dd = dtempl + (uint64_t)(db0 + s * (TC_D_STAGE_FLOATS * 4 / 16));
          umma_f16_ss_pair(dcol, xd, dd, idesc_h, kb != 0, issue);                                  // f16(X).f16(D)
          umma_f16_ss_pair(dcol, xd + (uint64_t)PART, dd, idesc_h, 1, issue);                       // f16(Xl).f16(D)
          umma_f16_ss_pair(dcol, xd, dd + (uint64_t)PART, idesc_h, 1, issue);                       // f16(X).f16(Dl)
          umma_commit_pair(empty0 + s * 8, 3, issue);                                        // stage free in both CTAs
          if (kb == BL_NKB - 1) umma_commit_pair(smem_u32(&sm.acc_full[a]), 3, issue);       // accumulator complete
          if (++s == TC_STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 2..5, both CTAs) =================
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    float* stage = sm.ep[warp - 2];
    uint32_t tl = 0;
    for (int t = cl; t < n_tiles; t += n_cl, ++tl) {
      const int mt2 = t / BL_NT, nt = t % BL_NT;
      const int a = tl & 1;
      const int row0 = (mt2 * 2 + (int)rank) * BL_BM + q * 32;
      const int rows_valid = B - row0;   // <= 0: nothing to store (padding tile)
      float* obase = vposed + (size_t)row0 * VPP + (size_t)nt * BL_BN;
      // v_template of this N tile (thread: 4 columns of each 32-column block), fetched BEFORE the accumulator wait:
      // a load inside the store loop put an L2 round trip on the critical path of every block (ncu: half of the
      // epilogue warps' samples), which bounds the kernel once the MMAs of a tile take less time than its epilogue
      float4 tv[BL_BN / 32];
#pragma unroll
      for (int j = 0; j < BL_BN / 32; ++j)
        tv[j] = __ldg(reinterpret_cast<const float4*>(vt + (size_t)nt * BL_BN + j * 32 + (lane & 7) * 4));
      mbar_wait(&sm.acc_full[a], (tl >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)a * BL_BN;
      uint32_t ra[32], rb[32];
      tmem_ld32(tbase, ra);
#pragma unroll
      for (int c0 = 0; c0 < BL_BN; c0 += 64) {
        tmem_ld_wait();
        tmem_ld32(tbase + c0 + 32, rb);
        tc_store_block(ra, stage, lane, tv[c0 / 32], obase + c0, rows_valid, d_scale_inv);
        tmem_ld_wait();
        if (c0 + 64 < BL_BN) {
          tmem_ld32(tbase + c0 + 64, ra);
        } else {
          // every TMEM read of this accumulator has completed: hand it back before the last stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(&sm.acc_empty[a], 0);
        }
        tc_store_block(rb, stage, lane, tv[c0 / 32 + 1], obase + c0 + 32, rows_valid, d_scale_inv);
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem, TC_TMEM_COLS);
  }
}

int32_t launch_blend_tc(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st) {
  if (B <= 0) return SMPLB_OK;
  static KernelSetup setup;
  const size_t smem = sizeof(TcSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, blend_tc_kernel, smem, &n_sm)) return rc;
  const int n_mtiles = (B + BL_BM - 1) / BL_BM;
  const int n_tiles = ((n_mtiles + 1) / 2) * BL_NT;
  const int n_cl = n_tiles < n_sm / 2 ? n_tiles : n_sm / 2;
  const int grid = 2 * n_cl;   // CTA pairs
  blend_tc_kernel<<<grid, TC_THREADS, smem, st>>>(B, n_mtiles, Ximg, m->Dimg, m->v_template_pad, vposed, m->d_scale_inv);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
