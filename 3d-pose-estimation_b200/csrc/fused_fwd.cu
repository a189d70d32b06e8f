// FUSED forward (north_star item 4; batch_smpl.py:91-95, :113-115, :120-132): blend-shape GEMM + weight blend + skinning
// in ONE kernel -- v_posed never reaches HBM.
//     v_posed[b,v,c] = v_template[v,c] + sum_k [beta|phi][b,k] D[k,(v,c)]          (tcgen05, accumulator in TMEM)
//     T[b,v]         = sum_j weights[v,j] A[b,j]                                     (tcgen05, accumulator in TMEM)
//     verts[b,v]     = T[b,v][:3,:3] v_posed[b,v] + T[b,v][:3,3]                     (epilogue, thread = vertex)
// The unfused pair (blend_tc_kernel -> skin_tc_kernel<0>) moves v_posed through HBM once each way (10.9 GB at
// B = 65536) and stays as the graded LBS kernel; this kernel writes only verts.
//
// Layout (operand images and the TMEM lane argument: common.cuh "FUSED forward"):
//   * D-STATIONARY, transposed GEMM: a CTA pair (cluster of 2, cta_group::2, M = 128) owns a 128-vertex tile, 64
//     vertices per CTA; each CTA keeps the D rows of its vertices resident in shared memory for x, y, z (3 x 2 parts x
//     28 KB = 168 KB).  N = 96 samples per blend MMA (the A operand is re-read from shared memory by every MMA at
//     <= 128 B/clk -- profiles/r02_mma_probe.txt -- so N must be large: ~24 cycles per M128 N96 K16 MMA).
//   * per 96-sample chunk: 14 K16 steps x 3 components x 3 split products = 126 blend MMAs into three 48-column
//     accumulators (double buffered), then 12 groups of 8 samples: 6 MMAs (M128 N96 K16) for T into a 48-column
//     accumulator (4 in flight); blend and weight-blend MMAs are issued by two warps with their own operand rings.
//   STATUS (profiles/r02_fused_forward.md): parity-green, 3.7 - 4.4 ms at B = 65536 against 3.5 ms for the unfused pair,
//   so smplb_forward uses it only on request (SMPLB_F_FUSED).  With 168 KB of D resident only 52 KB are left for the X
//   and A-image rings, whose round trip (bulk copy from L2 + peer relay + MMA completion + multicast commit) was
//   measured at 1800 - 2200 cycles: X ring alone 1.68 ms, A ring alone 1.17 ms with every MMA, TMEM load and store removed.
//   * TMEM lane = vertex for both accumulators (lanes 64..127 repeat the vertices for the second half of the columns;
//     a T group is 4 samples of each chunk half, so v_posed and T of a sample sit in the same lane).
//   warp 0      X producer: one 3 KB bulk copy per K16 step (5-stage ring)
//   warp 1      leader CTA: blend MMA issue; peer CTA: relays "my stage landed" to the leader's barriers
//   warp 2      A-image producer (6 KB per group, 5-stage ring); D / W tile loads when the vertex tile changes
//   warps 3-10  epilogue: tcgen05.ld T and v_posed (thread = vertex, 4 samples per warp and group, two sets of warps
//               alternate over the groups), apply, 12-byte-stride stores; keypoint vertices also go to the compact rows
//   warp 11     leader CTA: weight-blend MMA issue; peer CTA: relay of the A ring
// Schedule: 54 clusters own one vertex tile each and sweep the first m chunks in step (X / A images are then shared
// through L2); the other 20 clusters split the remaining chunks of all tiles.
#include "tc_common.cuh"

namespace smplb {

using namespace tc;

constexpr int FF_XST = 5;   // X ring: 3 KB per K16 step
constexpr int FF_AST = 5;   // A-image ring: 6 KB per group, ~450 cycles per group, ~1500 cycles of L2 latency -> >= 4 in flight
constexpr int FF_EPI_WARPS = 8;
constexpr int FF_THREADS = 128 + 32 * FF_EPI_WARPS;   // X producer, blend issuer, A/D producer, 8 epilogue warps, T issuer
constexpr uint32_t FF_D_BYTES = FF_DTILE * 2;   // 172 032
constexpr uint32_t FF_W_BYTES = FF_WHT * 2;     // 8 192
constexpr uint32_t FF_X_BYTES = FF_XKS * 2;     // 3 072
constexpr uint32_t FF_A_BYTES = FF_AGH * 2;     // 6 144
constexpr uint32_t FF_ACC_B = 3 * (FF_CH / 2);  // 144 TMEM columns of one blend accumulator set (3 components x 48)
constexpr uint32_t FF_TCOL0 = 2 * FF_ACC_B;     // 288: first column of the T accumulators
constexpr uint32_t FF_ACC_T = 48;               // TMEM columns of one T accumulator (N = 96 over two lane halves)
constexpr int FF_TBUF = 4;                      // T accumulators in flight: the T round trip (MMA -> multicast commit ->
                                                // tcgen05.ld in both CTAs -> remote arrive -> next MMA) takes > 1000 cycles,
                                                // a group's tensor work ~400; with 2 accumulators the kernel ran at 1/3.6 of
                                                // its tensor floor.  288 + 4 x 48 = 480 of the 512 columns.

struct FfSmem {
  unsigned char d[FF_D_BYTES];
  unsigned char w[FF_W_BYTES];
  unsigned char x[FF_XST][FF_X_BYTES];
  unsigned char a[FF_AST][FF_A_BYTES];
  // full barriers: the leader's count the local producer AND the peer's relay (one wait per stage on the MMA thread)
  uint64_t dfull, dempty;
  uint64_t xfull[FF_XST], xempty[FF_XST];
  uint64_t afull[FF_AST], aempty[FF_AST];
  uint64_t bfull[2], bempty[2], tfull[FF_TBUF], tempty[FF_TBUF];
  uint32_t tmem_base;
};

struct FfSched {
  int n_ch;      // chunks of 128 samples
  int n_main;    // clusters that own one vertex tile each (0 or 54)
  int m_main;    // chunks [0, m_main) of every tile go to its main cluster
  int n_cl;      // clusters launched
};
__device__ __forceinline__ int ff_n_items(const FfSched& s, int cl) {
  if (cl < s.n_main) return s.m_main;
  const long long t = cl - s.n_main, Tn = s.n_cl - s.n_main, L = (long long)SK_NVT * (s.n_ch - s.m_main);
  return (int)((t + 1) * L / Tn - t * L / Tn);
}
__device__ __forceinline__ void ff_item(const FfSched& s, int cl, int idx, int& vt, int& ch) {
  if (cl < s.n_main) {
    vt = cl;
    ch = idx;
    return;
  }
  const long long t = cl - s.n_main, Tn = s.n_cl - s.n_main, tw = s.n_ch - s.m_main, L = (long long)SK_NVT * tw;
  const long long I = t * L / Tn + idx;
  vt = (int)(I / tw);
  ch = s.m_main + (int)(I % tw);
}

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&r)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld2(uint32_t taddr, uint32_t (&r)[2]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FF_THREADS, 1)
fused_fwd_kernel(int B, FfSched sched, const unsigned short* __restrict__ Ximg2, const unsigned short* __restrict__ Aimg2,
                 const unsigned short* __restrict__ Dimg2, const unsigned short* __restrict__ Wimg2,
                 const float* __restrict__ vt_plain, float* __restrict__ verts, float* __restrict__ kpv, int kpv_pitch,
                 const int* __restrict__ kpc_slot_of) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  FfSmem& sm = *reinterpret_cast<FfSmem*>(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rk = cluster_ctarank();
  const int cl = blockIdx.x >> 1;
  const int n_items = ff_n_items(sched, cl);

  if (threadIdx.x == 0) {
    const uint32_t nfull = rk == 0 ? 2 : 1;   // leader: own producer + the peer's relay
    mbar_init(&sm.dfull, nfull);
    mbar_init(&sm.dempty, 2);   // both issuing warps of the leader commit at a tile change
    for (int s = 0; s < FF_XST; ++s) {
      mbar_init(&sm.xfull[s], nfull);
      mbar_init(&sm.xempty[s], 1);
    }
    for (int s = 0; s < FF_AST; ++s) {
      mbar_init(&sm.afull[s], nfull);
      mbar_init(&sm.aempty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&sm.bfull[s], 1);
      mbar_init(&sm.bempty[s], 2 * FF_EPI_WARPS);   // epilogue warps of both CTAs
    }
    for (int s = 0; s < FF_TBUF; ++s) {
      mbar_init(&sm.tfull[s], 1);
      mbar_init(&sm.tempty[s], FF_EPI_WARPS);       // the four warps of one epilogue set in both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_pair(&sm.tmem_base, 512);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 0) {
    // ================= X producer (both CTAs: the 64 samples of this CTA's chunk half) =================
    if (lane == 0) {
      uint32_t xit = 0;
      for (int il = 0; il < n_items; ++il) {
        int vt, ch;
        ff_item(sched, cl, il, vt, ch);
        const unsigned short* src = Ximg2 + (size_t)(ch * 2 + (int)rk) * FF_XHALF;
        for (int ks = 0; ks < FF_NKS; ++ks, ++xit) {
          const int s = xit % FF_XST;
          mbar_wait(&sm.xempty[s], ((xit / FF_XST) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.xfull[s], FF_X_BYTES);
          bulk_g2s(sm.x[s], src + (size_t)ks * FF_XKS, FF_X_BYTES, &sm.xfull[s]);
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ================= A-image producer; D / W loads at a tile change (both CTAs) =================
    uint32_t ait = 0, dk = 0;
    int cur_vt = -1;
    for (int il = 0; il < n_items; ++il) {
      int vt, ch;
      ff_item(sched, cl, il, vt, ch);
      if (vt != cur_vt) {
        cur_vt = vt;
        const int ht = vt * 2 + (int)rk;
        if (lane == 0) {
          if (dk > 0) mbar_wait(&sm.dempty, (dk - 1) & 1);   // every MMA that read the old tile has completed
          mbar_arrive_expect_tx(&sm.dfull, FF_D_BYTES + FF_W_BYTES);
        }
        __syncwarp();
        if (lane < 6)
          bulk_g2s(sm.d + (size_t)lane * (FF_DPART * 2), Dimg2 + (size_t)ht * FF_DTILE + (size_t)lane * FF_DPART,
                   FF_DPART * 2, &sm.dfull);
        else if (lane == 6)
          bulk_g2s(sm.w, Wimg2 + (size_t)ht * FF_WHT, FF_W_BYTES, &sm.dfull);
        ++dk;
      }
      if (lane == 0) {
        const unsigned short* src = Aimg2 + ((size_t)ch * FF_NG * 2 + rk) * FF_AGH;
        for (int g = 0; g < FF_NG; ++g, ++ait) {
          const int s = ait % FF_AST;
          mbar_wait(&sm.aempty[s], ((ait / FF_AST) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.afull[s], FF_A_BYTES);
          bulk_g2s(sm.a[s], src + (size_t)g * 2 * FF_AGH, FF_A_BYTES, &sm.afull[s]);
        }
      } else {
        ait += FF_NG;
      }
      __syncwarp();
    }
  } else if (warp == 1 || warp == 3 + FF_EPI_WARPS) {
    // ================= MMA issue (leader) / stage relay (peer) =================
    // Two issuing warps with their own rings, so that a stall of one stream does not hold up the other (with one
    // issuer the T groups waited behind blend K steps that waited for X slabs: 4x the tensor floor):
    //   warp 1   blend GEMM: per chunk 14 K16 steps x 9 MMAs from the X ring into blend accumulator il & 1
    //   warp 11  weight blend: per chunk 12 groups x 6 MMAs from the A ring into T accumulator (12 il + g) % 4
    // Whole warp runs the loops (uniform control flow); one elected lane issues the tcgen05 / arrive ops.
    const uint32_t issue = elect_one() ? 1u : 0u;
    const bool leader = rk == 0;
    const bool blend_role = warp == 1;
    constexpr uint32_t idesc_b = umma_idesc(0, 128, FF_CH);        // kind::f16, f16 operands, M 128 (pair), N 96
    constexpr uint32_t idesc_t = umma_idesc(0, 128, 2 * FF_AROWS);  // N 96
    const uint64_t d_t = umma_desc_kmajor_noswz(0, FF_VH * 16, 128);      // D and W: 64 rows per CTA
    const uint64_t x_t = umma_desc_kmajor_noswz(0, FF_SH * 16, 128);      // X: 48 rows per CTA
    const uint64_t a_t = umma_desc_kmajor_noswz(0, FF_AROWS * 16, 128);   // A image: 48 rows per CTA
    const uint32_t db = smem_u32(sm.d) >> 4, wb = smem_u32(sm.w) >> 4, xb = smem_u32(sm.x[0]) >> 4, ab = smem_u32(sm.a[0]) >> 4;
    uint32_t xit = 0, ait = 0, tg = 0, dk = 0;

    auto stage_wait = [&](uint64_t* full, uint32_t ph, bool relay) {
      mbar_wait_spin(smem_u32(full), ph);                             // leader: local data AND the peer's relay have arrived
      if (!leader && relay && issue) mbar_arrive_remote(full, 0);     // peer: my half has landed
    };
    int cur_vt = -1;
    for (int il = 0; il < n_items; ++il) {
      int vt, ch;
      ff_item(sched, cl, il, vt, ch);
      if (vt != cur_vt) {
        // tile change: this issuer's MMAs on the old tile are complete -> (both issuers) the loaders may overwrite D / W
        if (cur_vt >= 0 && leader) umma_commit_pair(smem_u32(&sm.dempty), 3, issue);
        stage_wait(&sm.dfull, dk & 1, blend_role);   // the peer's blend warp relays the tile load once
        ++dk;
        cur_vt = vt;
      }
      if (blend_role) {
        const uint32_t bb = (uint32_t)il & 1;
        if (leader) mbar_wait_spin(smem_u32(&sm.bempty[bb]), (((uint32_t)il >> 1) & 1) ^ 1);
#pragma unroll 1
        for (int ks = 0; ks < FF_NKS; ++ks, ++xit) {
          const uint32_t s = xit % FF_XST;
          stage_wait(&sm.xfull[s], (xit / FF_XST) & 1, true);
          if (!leader) continue;
          tc_fence_after();
          const uint64_t xh = x_t + (uint64_t)(xb + s * (FF_X_BYTES / 16)), xl = xh + (uint64_t)(FF_XPART * 2 / 16);
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const uint32_t dcol = tmem + bb * FF_ACC_B + (uint32_t)c * (FF_CH / 2);
            const uint64_t dh = d_t + (uint64_t)(db + (uint32_t)(c * 2) * (FF_DPART * 2 / 16) + (uint32_t)ks * (2 * FF_VH * 16 / 16));
            const uint64_t dl = dh + (uint64_t)(FF_DPART * 2 / 16);
            umma_f16_ss_pair(dcol, dh, xh, idesc_b, ks != 0 ? 1u : 0u, issue);   // Dh.Xh
            umma_f16_ss_pair(dcol, dh, xl, idesc_b, 1u, issue);                  // Dh.Xl
            umma_f16_ss_pair(dcol, dl, xh, idesc_b, 1u, issue);                  // Dl.Xh
          }
          umma_commit_pair(smem_u32(&sm.xempty[s]), 3, issue);
          if (ks == FF_NKS - 1) umma_commit_pair(smem_u32(&sm.bfull[bb]), 3, issue);
        }
      } else {
#pragma unroll 1
        for (int g = 0; g < FF_NG; ++g, ++tg, ++ait) {
          const uint32_t tb = tg % FF_TBUF;
          if (leader) mbar_wait_spin(smem_u32(&sm.tempty[tb]), ((tg / FF_TBUF) & 1) ^ 1);
          const uint32_t s = ait % FF_AST;
          stage_wait(&sm.afull[s], (ait / FF_AST) & 1, true);
          if (!leader) continue;
          tc_fence_after();
          const uint32_t tcol = tmem + FF_TCOL0 + tb * FF_ACC_T;
#pragma unroll
          for (int ks = 0; ks < SK_KPAD / 16; ++ks) {
            const uint64_t wh = d_t + (uint64_t)(wb + (uint32_t)ks * (2 * FF_VH * 16 / 16)), wl = wh + (uint64_t)(FF_WPART * 2 / 16);
            const uint64_t ah = a_t + (uint64_t)(ab + s * (FF_A_BYTES / 16) + (uint32_t)ks * (2 * FF_AROWS * 16 / 16));
            const uint64_t al = ah + (uint64_t)(FF_APART * 2 / 16);
            umma_f16_ss_pair(tcol, wh, ah, idesc_t, ks != 0 ? 1u : 0u, issue);   // Wh.Ah
            umma_f16_ss_pair(tcol, wl, ah, idesc_t, 1u, issue);                  // Wl.Ah
            umma_f16_ss_pair(tcol, wh, al, idesc_t, 1u, issue);                  // Wh.Al
          }
          umma_commit_pair(smem_u32(&sm.aempty[s]), 3, issue);
          umma_commit_pair(smem_u32(&sm.tfull[tb]), 3, issue);
        }
      }
    }
    __syncwarp();
  } else {
    // ================= epilogue (warps 3..10, both CTAs) =================
    // Two sets of four warps (one warp per TMEM lane quarter): set 0 takes the even T groups (accumulator 0), set 1 the
    // odd ones (accumulator 1), each warp all 4 samples of its chunk half -- a group's fixed costs (barrier waits,
    // arrivals, address set-up) are paid once per 4 samples and a warp has two group periods for one group.
    const int ew = warp - 3;
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    const int set = ew >> 2;             // parity of the groups this warp handles = its T accumulator
    const int hl = q >> 1;               // chunk half held by this lane quarter
    const int r = (q & 1) * 32 + lane;   // vertex inside the CTA's 64
    const uint32_t lsel = (uint32_t)(q * 32) << 16;
    const uint32_t tcol0 = tmem + lsel + FF_TCOL0 + (uint32_t)set * FF_ACC_T;   // this set uses T accumulators set and set + 2
    float* const verts_r = verts;
    float* const kpv_r = kpv;
    const int Bn = B;
    uint32_t tu = 0;                     // uses of this set's T accumulators so far
    for (int il = 0; il < n_items; ++il) {
      int vt, ch;
      ff_item(sched, cl, il, vt, ch);
      const int v = vt * SK_VT + (int)rk * FF_VH + r;
      const bool vok = v < NV;
      float tv[3] = {0.f, 0.f, 0.f};
      int slot[3] = {-1, -1, -1};
      if (vok) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          tv[c] = __ldg(vt_plain + 3 * v + c);
          if (kpv_r) slot[c] = __ldg(kpc_slot_of + 3 * v + c);
        }
      }
      const bool any_slot = (slot[0] & slot[1] & slot[2]) >= 0 || slot[0] >= 0 || slot[1] >= 0 || slot[2] >= 0;
      const int v0 = vt * SK_VT + (int)rk * FF_VH + (q & 1) * 32;   // first vertex of this warp
      const int nvf = 3 * max(0, min(32, NV - v0));                 // floats of the warp's 96 that exist
      const uint32_t bb = (uint32_t)il & 1;
      const uint32_t pcol = tmem + lsel + bb * FF_ACC_B;
      mbar_wait_spin(smem_u32(&sm.bfull[bb]), ((uint32_t)il >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int g = set; g < FF_NG; g += 2, ++tu) {
        const uint32_t tbuf = (uint32_t)set + 2 * (tu & 1);
        const uint32_t tcol = tcol0 + 2 * (tu & 1) * FF_ACC_T;
        mbar_wait_spin(smem_u32(&sm.tfull[tbuf]), (tu >> 1) & 1);
        tc_fence_after();
        uint32_t t0[16], t1[16], t2[16], px[4], py[4], pz[4];
        tmem_ld16(tcol, t0);
        tmem_ld16(tcol + 16, t1);
        tmem_ld16(tcol + 32, t2);
        tmem_ld4(pcol + 4 * g, px);
        tmem_ld4(pcol + 4 * g + FF_CH / 2, py);
        tmem_ld4(pcol + 4 * g + 2 * (FF_CH / 2), pz);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(&sm.tempty[tbuf], 0);
        const int b0 = ch * FF_CH + hl * FF_SH + 4 * g;
        float y[4][3];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float x0 = __uint_as_float(px[i]) + tv[0], x1 = __uint_as_float(py[i]) + tv[1], x2 = __uint_as_float(pz[i]) + tv[2];
          float t[12];
#pragma unroll
          for (int e = 0; e < 12; ++e) {
            const int idx = 12 * i + e;
            t[e] = __uint_as_float(idx < 16 ? t0[idx & 15] : (idx < 32 ? t1[idx & 15] : t2[idx & 15]));
          }
#pragma unroll
          for (int rr = 0; rr < 3; ++rr)
            y[i][rr] = fmaf(t[4 * rr], x0, fmaf(t[4 * rr + 1], x1, fmaf(t[4 * rr + 2], x2, t[4 * rr + 3])));
        }
        // straight from registers: 12-byte lane stride, the three stores of a sample fill the same 384 bytes (the L2 merges
        // them into whole sectors); a transposition through smem (as in the unfused kernel) cost more epilogue latency
        // than it saved: 4.49 -> 3.65 ms
        if (verts_r && vok) {
          float* dst = verts_r + (size_t)b0 * NV3 + (size_t)v * 3;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (b0 + i < Bn) {
#pragma unroll
              for (int c = 0; c < 3; ++c) __stcs(dst + (size_t)i * NV3 + c, y[i][c]);
            }
          }
        }
        if (any_slot) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int b = b0 + i;
            if (b < Bn) {
#pragma unroll
              for (int c = 0; c < 3; ++c)
                if (slot[c] >= 0) kpv_r[((size_t)(b / SK_S) * kpv_pitch + slot[c]) * SK_S + (b % SK_S)] = y[i][c];
            }
          }
        }
      }
      // every TMEM read of this blend accumulator set has completed (the last group's tcgen05.wait::ld)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(&sm.bempty[bb], 0);
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem, 512);
  }
}

int32_t launch_fused_fwd(const smplb_model* m, int B, const float* Ximg2, const float* Aimg2, float* verts, float* kpv,
                         cudaStream_t st) {
  if (B <= 0 || (!verts && !kpv)) return SMPLB_OK;
  static KernelSetup setup;
  const size_t smem = sizeof(FfSmem);
  int n_sm = 0;
  if (int32_t rc = setup.ensure(m->device, fused_fwd_kernel, smem, &n_sm)) return rc;
  FfSched s;
  s.n_ch = (B + FF_CH - 1) / FF_CH;
  const int G = n_sm / 2;
  const long long total = (long long)SK_NVT * s.n_ch;
  if (G > SK_NVT && s.n_ch >= 8) {
    // 54 clusters own a vertex tile each and sweep chunks [0, m) in step; the other G - 54 share the rest
    s.n_main = SK_NVT;
    s.m_main = (int)(total / G);
    s.n_cl = G;
    if (s.m_main >= s.n_ch) { s.n_main = 0; s.m_main = 0; }
  } else {
    s.n_main = 0;
    s.m_main = 0;
    s.n_cl = (int)(total < G ? total : G);
  }
  fused_fwd_kernel<<<2 * s.n_cl, FF_THREADS, smem, st>>>(B, s, reinterpret_cast<const unsigned short*>(Ximg2),
                                                        reinterpret_cast<const unsigned short*>(Aimg2),
                                                        reinterpret_cast<const unsigned short*>(m->Dimg2),
                                                        reinterpret_cast<const unsigned short*>(m->Wimg2), m->vt_plain, verts,
                                                        kpv, m->kpv_pitch, m->kpc_slot_of);
  SMPLB_CUDA_CHECK(cudaGetLastError());
  return SMPLB_OK;
}

}  // namespace smplb
