// Shared declarations of the smplb200 library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/smplb200.h"

#define NV SMPLB_NV
#define NV3 SMPLB_NV3
#define NJ SMPLB_NJ
#define NB SMPLB_NB
#define NP SMPLB_NP
#define KDIM SMPLB_K
#define KPAD SMPLB_KPAD
#define VPP SMPLB_VP_PITCH      // v_posed row pitch in floats (81 * 256)
#define NKMAX 32                // keypoint-regressor columns supported (cocoplus 19 / lsp 14)

// ---- operand images of the blend-shape GEMM (blend_tc.cu) -------------------------------------
// Both operands are kept in global memory in exactly the byte order the tcgen05 MMA reads them from
// shared memory (K-major, no swizzle: 16-byte K-chunks, rows contiguous inside a chunk), so that one
// 1-D bulk copy moves one pipeline stage.  X (samples) : [m_tile][k_block][hi|lo][chunk][row 0..127][4]
//                                          D (columns) : [n_tile][k_block][hi|lo][chunk][row 0..255][4]
#define BL_BM 128               // samples per M tile
#define BL_BN 256               // v_posed columns per N tile (VPP = 81 * 256)
#define BL_KB 32                // K elements per pipeline stage (k-block) = 8 chunks of 4
#define BL_NKB (KPAD / BL_KB)   // 7 k-blocks
#define BL_NT (VPP / BL_BN)     // 81 N tiles
#define BL_XTILE (BL_NKB * 2 * (BL_KB / 4) * BL_BM * 4)   // floats per X m-tile image (57344)
#define BL_DTILE (BL_NKB * 2 * (BL_KB / 4) * BL_BN * 4)   // floats per D n-tile image (114688)

namespace smplb {

void set_error(const std::string& msg);
int32_t fail(int32_t code, const std::string& msg);

#define SMPLB_CUDA_CHECK(expr)                                                              \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return smplb::fail(SMPLB_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// Kinematic tree, passed by value to the kernels (batch_smpl.py:54, batch_lbs.py:122-127).
struct Tree {
  int parent[NJ];
  int depth[NJ];
  int max_depth;
};

}  // namespace smplb

// Device-resident model constants (SMPL.__init__, batch_smpl.py:17-69) plus derived tables.
struct smplb_model {
  int device;
  int NK;
  smplb::Tree tree;
  // reference-layout constants
  float* v_template_pad;  // [VPP]   v_template flattened (3v+c), zero padded
  // blend-shape basis D = [shapedirs; posedirs] : K=217 rows, padded to KPAD x VPP
  float* D32;             // [KPAD][VPP] fp32 (SIMT verification path, backward)
  float* Dimg;            // [BL_NT][BL_DTILE] TF32 hi/lo operand image of D (tcgen05 B operand)
  // folded joint regression: J = Jbar + JS * beta  (batch_smpl.py:91-101)
  float* Jbar;            // [24*3]
  float* JS;              // [24*3*10]
  // compressed LBS weights (<= 4 non-zeros per vertex), sorted by joint index
  uint32_t* skin_idx;     // [NV] 4 x u8 joint ids
  float4* skin_w;         // [NV]
  // keypoint regressor, CSR by vertex (joint_regressor, batch_smpl.py:63-69)
  int* kp_ptr;            // [NV+1]
  int* kp_col;            // [nnz] keypoint id
  float* kp_val;          // [nnz]
  int kp_nnz;
  // J_regressor CSR by joint (for d(J)/d(beta) in backward: JS is enough) — not needed
  void* tmap_cache;       // host-side cache of TMA descriptors (blend_tc.cu)
};

namespace smplb {

struct PrepParams {
  int B;
  const float* beta;
  const float* theta;
  const float* Jbar;
  const float* JS;
  float* Ximg;    // [ceil(B/128)][BL_XTILE] TF32 hi/lo operand image of [beta | pose feature], nullable
  float* joints;  // [B][24][3] nullable
  float* Rs;      // [B][24][9] nullable
  float* A;       // [B][24][12] nullable
  const float* cam_pos;
  const float* cam_rot;
  const float* focal;
  float* uv;
  float* px;
  smplb_camcfg cfg;
  Tree tree;
};

int32_t launch_pose_prep(const PrepParams& p, cudaStream_t st);
int32_t launch_blend_simt(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st);
int32_t launch_blend_tc(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st);

// float offset of element (row b, k) of the X operand image; hl = 0 (hi) / 1 (lo)
__host__ __device__ inline size_t ximg_off(int b, int k, int hl) {
  return (size_t)(b / BL_BM) * BL_XTILE +
         (size_t)((((k / BL_KB) * 2 + hl) * (BL_KB / 4) + (k % BL_KB) / 4) * BL_BM + (b % BL_BM)) * 4 + (k & 3);
}
// float offset of element (column n, k) of the D operand image
__host__ __device__ inline size_t dimg_off(int n, int k, int hl) {
  return (size_t)(n / BL_BN) * BL_DTILE +
         (size_t)((((k / BL_KB) * 2 + hl) * (BL_KB / 4) + (k % BL_KB) / 4) * BL_BN + (n % BL_BN)) * 4 + (k & 3);
}
int32_t launch_skin(const smplb_model* m, int B, const float* vposed, const float* A, float* verts,
                    float* kp, cudaStream_t st);

// TF32 helpers --------------------------------------------------------------------------------
__host__ __device__ inline float tf32_rna_host(float x) {
  // round-to-nearest (ties away) to 10 explicit mantissa bits, like cvt.rna.tf32.f32
  union { float f; uint32_t u; } c;
  c.f = x;
  if ((c.u & 0x7f800000u) == 0x7f800000u) return x;
  c.u = (c.u + 0x1000u) & 0xffffe000u;
  return c.f;
}

}  // namespace smplb
