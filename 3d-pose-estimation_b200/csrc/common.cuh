// Shared declarations of the smplb200 library (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <mutex>
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
//                                          D (columns) : [n_tile][k_block][half][hi|lo][chunk][row 0..127][4]
// (the two CTAs of a pair each hold one 128-column half of the 256-column N tile: cta_group::2 MMA)
#define BL_BM 128               // samples per M tile
#define BL_BN 256               // v_posed columns per N tile (VPP = 81 * 256)
#define BL_BH 128               // columns of the N tile held by each CTA of the pair
#define BL_KB 16                // K elements per pipeline stage (k-block) = 4 chunks of 4
#define BL_NKB (KPAD / BL_KB)   // 14 k-blocks
#define BL_NT (VPP / BL_BN)     // 81 N tiles
#define BL_SLAB (2 * BL_KB * BL_BM / 2)   // floats per (k-block, 128 rows) operand slab: 2 sixteen-bit parts = 8 KB
#define BL_XTILE (BL_NKB * BL_SLAB)       // floats per X m-tile image (28672)
#define BL_DSLAB (2 * BL_KB * BL_BH / 2)  // floats per (k-block, 128 columns) D slab: 2 sixteen-bit parts = 8 KB
#define BL_DTILE (BL_NKB * 2 * BL_DSLAB)  // floats per D n-tile image (57344): [k-block][column half] slabs

namespace smplb {

void set_error(const std::string& msg);
int32_t fail(int32_t code, const std::string& msg);

#define SMPLB_CUDA_CHECK(expr)                                                              \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return smplb::fail(SMPLB_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// The kernels launch on the CURRENT device: every entry point that takes a model / mesh handle checks that it is the
// handle's device (SMPLB_E_INVALID_ARG otherwise) instead of silently configuring or launching on another GPU.
int32_t check_current_device(int device, const char* who);

// One-time, thread-safe, per-device kernel setup (dynamic shared memory opt-in + SM count).  A launcher keeps one
// `static KernelSetup` per kernel; std::call_once makes concurrent first calls from several host threads safe, and the
// attribute is set on the device the handle lives on (the caller has verified that it is the current one).
struct KernelSetup {
  std::once_flag once[64];
  int n_sm[64] = {};
  cudaError_t err[64] = {};
  template <class Kernel>
  int32_t ensure(int device, Kernel kernel, size_t smem_bytes, int* n_sm_out) {
    const int d = device & 63;
    std::call_once(once[d], [&] {
      int v = 0;
      cudaError_t e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device);
      if (e == cudaSuccess && smem_bytes > 48 * 1024)
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
      n_sm[d] = v;
      err[d] = e;
    });
    if (err[d] != cudaSuccess)
      return fail(SMPLB_E_CUDA, std::string("kernel setup: ") + cudaGetErrorString(err[d]));
    if (n_sm_out) *n_sm_out = n_sm[d];
    return SMPLB_OK;
  }
};

// Programmatic dependent launch (PDL): a kernel launched with launch_pdl() may be scheduled while its predecessor on the
// stream is still running; it must call pdl_wait() before it touches anything the predecessor wrote (or overwrites
// anything the predecessor reads) and may call pdl_trigger() to let ITS successor be scheduled early.  Both are no-ops
// for a kernel launched the ordinary way.  Used where a chain of short kernels is bound by launch gaps (the fitting step).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at;
  at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = &at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Kinematic tree, passed by value to the kernels (batch_smpl.py:54, batch_lbs.py:122-127).
struct Tree {
  int parent[NJ];
  int depth[NJ];
  int child[NJ][3];   // -1 padded (SMPL: at most 3 children per joint)
  int max_depth;
  // depth-first pre-order view, used by the per-sample adjoint kernel (backward.cu): lane l holds joint order[l], its
  // subtree is the lane range [l, l + sub[l]), up[r][l] is the lane of its 2^r-th ancestor (-1 past the root)
  int order[NJ];
  int rank[NJ];
  int sub[NJ];
  int up[5][NJ];
  int upj[5][NJ];     // the same in joint order: upj[r][j] = 2^r-th ancestor joint of joint j, -1 past the root
  int n_rounds;       // ceil(log2(max_depth + 1)): pointer-jumping rounds of the forward chain
};
int32_t build_tree(const int32_t* parents, Tree* out, const char* who);

// Adam state handed to the fused fitting kernel (backward.cu) when the update can run inside it (fit.cu)
struct FitAdam {
  float *theta, *m_theta, *v_theta;
  float *beta, *m_beta, *v_beta;   // m_beta == nullptr: beta is not fitted
  const int* step_dev;
  int step_host;
  float lr, b1, b2, eps;
};

}  // namespace smplb

// Device-resident model constants (SMPL.__init__, batch_smpl.py:17-69) plus derived tables.
struct smplb_model {
  int device;
  int NK;
  smplb::Tree tree;
  int4* tree_tab;   // device: [32 lanes][2] = {order, sub, up0, up1 | up2, up3, up4, 0} (idle lanes: 0, 1, -1 ...), then the
                    // same table with lane = joint (pose_prep.cu): {j, 1, upj0, upj1 | upj2, upj3, upj4, 0}
  // reference-layout constants
  float* v_template_pad;  // [VPP]   v_template flattened (3v+c), zero padded
  // blend-shape basis D = [shapedirs; posedirs] : K=217 rows, padded to KPAD x VPP
  float* D32;             // [KPAD][VPP] fp32 (SIMT verification path, backward)
  float* Dimg;            // [BL_NT][BL_DTILE] fp16 split operand image of 2^e D (tcgen05 B operand, dimg16_off)
  float d_scale_inv;      // 2^-e: the blend GEMM's epilogue multiplies the accumulator by it (exact)
  // folded joint regression: J = Jbar + JS * beta  (batch_smpl.py:91-101)
  float* Jbar;            // [24*3]
  float* JS;              // [24*3*10]
  float* JSt;             // [33][32] joint-minor copy: row c*10+k = JS[.][c][k], rows 30..32 = Jbar[.][c]; lanes that hold
                          // joints read one 128-byte line per row (the per-sample kernels)
  // compressed LBS weights (<= 4 non-zeros per vertex), sorted by joint index (SIMT skinning path);
  // has_compressed == 0 when some vertex has more than 4 non-zero weights (dense models: TC path only)
  uint32_t* skin_idx;     // [NV] 4 x u8 joint ids
  float4* skin_w;         // [NV]
  int has_compressed;
  float* Wimg;            // [SK_NVT][SK_WTILE] f16/bf16 operand image of the dense LBS weights (wimg16_off)
  // backward tables
  int* ent_ptr;           // [SK_NVT][NJ+1] (vertex, weight) pairs of every (vertex tile, joint): skin_bwd.cu
  float2* ent;            // [nnz] {weight, vertex index inside the tile as int bits}
  float* DTimg;           // [BB_NKB][BB_DSTAGE x 16 bit] bf16 split image of D^T (blend_bwd.cu B operand, dtimg16_off)
  // keypoint regressor (joint_regressor, batch_smpl.py:63-69), CSR by keypoint, fixed summation order
  int* kp_ptr;            // [NK+1]
  int* kp_vtx;            // [nnz] vertex id
  float* kp_val;          // [nnz]
  int kp_nnz;
  // compact keypoint-vertex rows (skin_tc.cu -> keypoints): the skinning epilogue copies the vertices the
  // regressor touches into kpv[b][kpv_pitch] (runs per vertex tile, padded to whole 32-byte sectors), so the
  // keypoint pass reads ~8 KB/sample of contiguous data instead of gathering 608 sectors from verts
  int kpv_pitch;          // floats per sample (multiple of 8)
  int* kpc_tile_ptr;      // [SK_NVT+1] first compact slot of every vertex tile
  int* kpc_src;           // [kpv_pitch] float index 3*lv+c inside the tile's staged verts row, -1 = pad
  int* kpc_slot;          // [nnz] compact slot of the vertex of every regressor entry
  float* Wimg3;           // bf16 split K-major image of the LBS weights for the tensor-core g_A reduction (skin_bwd_tc.cu)
  // fused forward (fused_fwd.cu)
  float* Dimg2;           // [FF_NHT][FF_DTILE x 16 bit] D rows of every half tile, f16 split (ff_dimg_off)
  float* Wimg2;           // [FF_NHT][FF_WHT x 16 bit] LBS weights, f16 split (ff_wimg_off)
  float* vt_plain;        // [NV3] v_template in the reference order 3v+c
  int* kpc_slot_of;       // [NV3] compact keypoint slot of (vertex, component), -1 = not a keypoint vertex
};

#define NM_BIG_ORDER 16
// Triangle mesh handle (raster.cu, normals.cu)
struct smplb_mesh {
  int device;
  int T, V;
  int max_order;   // largest number of faces around a vertex
  int* tri;        // [T][3]
  int* vptr;       // [V+1] CSR vertex -> incident (triangle, corner), ascending triangle id
  int* vent;       // [3T]  triangle * 3 + corner
  int* big_v;      // [n_big] vertices with more than NM_BIG_ORDER incident faces (handled by a whole warp in normals.cu)
  int n_big;
  uint4* vfan;     // [V] closed fans of order 3..8: the ring of neighbours as 8 x uint16 (0xffff = unused); else all ones
  uint32_t* vpair; // [3T]  per CSR entry: the two other vertices of the face, in winding order after this corner
                   //       (lo 16 bits = next, hi 16 bits = next-next); only when V <= 65536, else NULL
};

namespace smplb {

struct PrepParams {
  int B;
  const float* beta;
  const float* theta;
  const float* Jbar;
  const float* JS;
  const float* JSt;   // smplb_model::JSt
  float* Ximg;    // [ceil(B/128)][BL_XTILE] fp16 split operand image of [beta | pose feature] (ximg16_off), nullable
  float* joints;  // [B][24][3] nullable
  float* Rs;      // [B][24][9] nullable
  float* A;       // [B][24][12] nullable
  float* Aimg;    // [ceil(B/20)][SK_AGRP] f16/bf16 operand image of A for the skinning GEMM (aimg16_off), nullable
  float* Ximg2;   // fused forward: [ceil(B/96)][2][FF_XHALF] f16 split image of [beta | pose feature] (ff_ximg_off), nullable
  float* Aimg2;   // fused forward: [ceil(B/96)][12][2][FF_AGH] f16 split image of A (ff_aimg_off), nullable
  const float* cam_pos;
  const float* cam_rot;
  const float* focal;
  float* uv;
  float* px;
  smplb_camcfg cfg;
  const int4* tree_tab;   // smplb_model::tree_tab
  int n_rounds;           // Tree::n_rounds
  int staged;             // set by launch_pose_prep: per-joint outputs leave through shared memory (16-byte stores)
  int* status;    // nullable: input-range flags (see Workspace)
};

// Carving of the caller-owned workspace (capi.cu)
struct Workspace {
  int* status;     // [1] bit 0: |beta| > 65504 or not finite, bit 1: theta not finite (set by pose_prep, smplb_forward_status)
  float* Ximg;
  float* A;
  float* Aimg;
  float* vposed;
  float* verts;    // scratch verts when only keypoints are wanted (SMPLB_F_SCRATCH_VERTS)
  float* joints;   // scratch J_transformed when the caller passes none
  float* bwd;      // backward scratch (backward.cu)
  float* kpv;      // compact keypoint-vertex rows [B][kpv_pitch] (last: earlier offsets do not depend on it)
  float* Ximg2;    // fused forward operand images (SMPLB_F_FUSED)
  float* Aimg2;
  size_t total;
};
Workspace carve(void* base, int B, int flags, int kpv_pitch = 0);
size_t backward_scratch_bytes(int B);

int32_t launch_pose_prep(const PrepParams& p, cudaStream_t st);
int32_t launch_blend_simt(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st);
int32_t launch_blend_tc(const smplb_model* m, int B, const float* Ximg, float* vposed, cudaStream_t st);

// ---- internal v_posed layout ----------------------------------------------------------------------
// v_posed is an internal buffer (blend GEMM -> skinning), so its column order is ours to choose: vertices
// are grouped in tiles of 128 (= TMEM lanes of the skinning kernel) and stored SoA inside a tile, i.e.
// column(v, c) = (v / 128) * 384 + c * 128 + v % 128.  54 tiles * 384 = 20736 = VPP; pad columns are 0.
#define SK_VT 128                       // vertices per skinning tile
#define SK_NVT ((NV + SK_VT - 1) / SK_VT)   // 54
__host__ __device__ inline int vp_col(int v, int c) { return (v / SK_VT) * (3 * SK_VT) + c * SK_VT + (v % SK_VT); }

// ---- operand images of the skinning GEMM (skin_tc.cu): T[v, (s,e)] = sum_j W[v,j] A[s][j][e] -------
#define SK_S 16                         // samples per MMA group: N = 12 * 16 = 192
#define SK_N (12 * SK_S)
#define SK_KPAD 32                      // joints padded to two K = 16 steps (pad columns are zero in both images)
#define SK_KCH (SK_KPAD / 8)            // 4 K-chunks of 8 sixteen-bit elements
#define SK_WPART (SK_KCH * SK_VT * 8)   // 16-bit elements per part of a W tile image (4096 = 8 KB)
#define SK_APART (SK_KCH * SK_N * 8)    // 16-bit elements per part of an A group image (6144 = 12 KB)
#define SK_WTILE (2 * SK_WPART / 2)     // floats per W tile image  [part 2][chunk][128 rows][8 x 16 bit]  (4096)
#define SK_AGRP (2 * SK_APART / 2)      // floats per A group image [part 2][chunk][192 rows][8 x 16 bit]  (6144)
// Both operands of the skinning GEMM T = W.A are split into an fp16 main part and the fp16 residual:
//     W.A ~= f16(W).f16(A) + f16(Wl).f16(A) + f16(W).f16(Al),     Wl = W - f16(W), Al = A - f16(A)
// (weights in [0,1], rotations in [-1,1], translations in metres: a residual is a normal fp16 number when the value
// is >= 0.25 and otherwise has an absolute error <= 2^-25, which is what counts against the vertex scale).
// parts: W [f16(W) | f16(Wl)],  A [f16(A) | f16(Al)].
// 16-bit-element offset of component e (3x4 row-major) of joint j of sample b in the A operand image
__host__ __device__ inline size_t aimg16_off(int b, int j, int e, int part) {
  return (size_t)(b / SK_S) * (SK_AGRP * 2) + (size_t)part * SK_APART +
         (size_t)(((j / 8) * SK_N + (b % SK_S) * 12 + e) * 8 + (j & 7));
}
// 16-bit-element offset of (vertex v, joint j) in the W operand image
__host__ __device__ inline size_t wimg16_off(int v, int j, int part) {
  return (size_t)(v / SK_VT) * (SK_WTILE * 2) + (size_t)part * SK_WPART + (size_t)(((j / 8) * SK_VT + (v % SK_VT)) * 8 + (j & 7));
}

// ---- operand images of the transposed blend GEMM (blend_bwd.cu): g_X[b,k] = sum_n g_vposed[b,n] D[k,n] ----
// 16-bit split operands like the forward GEMMs, but bf16 parts (gradient magnitudes, ~1/(B*20670) for a mean loss,
// leave the fp16 range; tcgen05 kind::f16 needs both operands of one MMA in the same format -- probed,
// profiles/r02_mma_probe.txt):   g.D ~= g1.D1 + g2.D1 + g1.D2,   x1 = bf16(x), x2 = bf16(x - x1)   (error ~2^-16)
// A operand = g_vposed, rows = samples, K = the 20736 internal columns, stored MN-MAJOR: 8 consecutive samples of one
//   column are 16 contiguous bytes, which is what a thread of skin_tc_kernel<1> holds (thread = vertex, 8 samples) --
//   it writes the image straight from registers, no transposition through shared memory:
//   [m_tile of 128 samples][k_block of 64 columns][part 2][sample group m/8: 16][column group k/8: 8][k%8][m%8]
// B operand = D^T, rows = k 0..223 split in two halves of 112 (one per CTA of the pair), K-major:
//   [k_block][half 2][part 2][column group: 8][row group r/8: 14][r%8][k%8]
#define BB_KB 64                                      // internal columns per pipeline stage (4 K16 steps)
#define BB_NKB (VPP / BB_KB)                          // 324 k-blocks
#define BB_APART (BL_BM * BB_KB)                      // 16-bit elements per part of an A stage (16 KB)
#define BB_ASTAGE (2 * BB_APART)                      // elements per A stage: [g1 | g2] (32 KB)
#define BB_NH (KPAD / 2)                              // 112 rows of D^T per CTA of the pair
#define BB_DPART (BB_NH * BB_KB)                      // elements per part of a D^T half stage (14 KB)
#define BB_DSTAGE (2 * 2 * BB_DPART)                  // elements per k-block: [half][D1 | D2] (56 KB)
__host__ __device__ inline size_t gimg16_off(int b, int n, int part) {
  return ((size_t)(b / BL_BM) * BB_NKB + (size_t)(n / BB_KB)) * BB_ASTAGE + (size_t)part * BB_APART +
         (size_t)(((b % BL_BM) / 8) * (BB_KB / 8) + (n % BB_KB) / 8) * 64 + (size_t)((n & 7) * 8 + (b & 7));
}
__host__ __device__ inline size_t dtimg16_off(int k, int n, int part) {
  const int half = k / BB_NH, r = k % BB_NH;
  return ((size_t)((n / BB_KB) * 2 + half) * 2 + (size_t)part) * BB_DPART +
         (size_t)(((n % BB_KB) / 8) * (BB_NH / 8) + r / 8) * 64 + (size_t)((r & 7) * 8 + (n & 7));
}

// Operand images of the blend GEMM (blend_tc.cu).  The product is evaluated fp32-faithfully from fp16 operands:
//     X.D ~= 2^-e [ f16(X).f16(D')  +  f16(Xl).f16(D')  +  f16(X).f16(Dl') ],   D' = 2^e D, Xl = X - f16(X), Dl' = D' - f16(D')
// (kind::f16 MMAs, K = 16 each, fp32 accumulation).  fp16 carries the same 11 significant bits as TF32 at twice the
// tensor rate and half the bytes; the two correction products are 2^-11 of the main one.  Xl (inputs are O(1)) stays an
// fp16 number: normal for |X| >= 0.12, below that its absolute error is <= 2^-25, which is what matters against the vertex
// scale.  The residual of SMPL's bases (~5e-6) would sit below the fp16 normal range, so the model is stored scaled by the
// power of two that puts max|D| just under 2^15 (smplb_model::d_scale_inv = 2^-e, applied exactly in the epilogue): both
// parts of D' are then normal fp16 numbers down to entries 2^-17 of the largest one, and all three products read the
// same two X parts -- the kernel is bound by the operand stream from L2 (~43 B/clk/SM delivered), and the bf16(X) part
// the first version of this round needed for a bf16(Dl) partner was a third of the X bytes.
// Error ~2^-21 of the blend offsets (tests/test_blend_split.py).  3 MMA slots, 4 B per X element and 4 B per D element
// against 6 slots and 8 B for 3xTF32 (Xh.Dh + Xl.Dh + Xh.Dl).
// Per 16-element k-block a 128-row slab is made of 4 KB parts, each 2 K-chunks x 128 rows x 16 B:
//     X (8 KB): f16(X) | f16(Xl)          D (8 KB): f16(D') | f16(Dl')
#define BL_PART16 (BL_KB * BL_BM)   // 16-bit elements per part (2048)
// 16-bit-element offset of part `part` of element (row b, k) of the X operand image
__host__ __device__ inline size_t ximg16_off(int b, int k, int part) {
  return ((size_t)(b / BL_BM) * BL_XTILE + (size_t)(k / BL_KB) * BL_SLAB) * 2 + (size_t)part * BL_PART16 +
         (size_t)(((k % BL_KB) / 8) * BL_BM + (b % BL_BM)) * 8 + (k & 7);
}
// the same for element (column n, k) of the D operand image: slabs ordered [n-tile][k-block][column half]
__host__ __device__ inline size_t dimg16_off(int n, int k, int part) {
  const int half = (n % BL_BN) / BL_BH;
  return ((size_t)(n / BL_BN) * BL_DTILE + (size_t)((k / BL_KB) * 2 + half) * BL_DSLAB) * 2 + (size_t)part * BL_PART16 +
         (size_t)(((k % BL_KB) / 8) * BL_BH + (n % BL_BH)) * 8 + (k & 7);
}
// fp32 -> bf16 bits, round to nearest even (finite inputs)
__host__ __device__ inline unsigned short bf16_rne_bits(float x) {
  union { float f; uint32_t u; } c;
  c.f = x;
  c.u += 0x7FFFu + ((c.u >> 16) & 1u);
  return (unsigned short)(c.u >> 16);
}
__host__ __device__ inline float bf16_bits_to_float(unsigned short h) {
  union { float f; uint32_t u; } c;
  c.u = (uint32_t)h << 16;
  return c.f;
}
// fp32 <-> fp16 bits, round to nearest even (subnormals included)
__host__ __device__ inline unsigned short f16_rne_bits(float x) { return __half_as_ushort(__float2half_rn(x)); }
__host__ __device__ inline float f16_bits_to_float(unsigned short h) { return __half2float(__ushort_as_half(h)); }
// ---- operand images of the FUSED forward (fused_fwd.cu): blend GEMM + weight blend + skinning in one kernel ----
// D-stationary, transposed: a CTA pair (cta_group::2, M = 128) owns one 128-vertex tile, 64 vertices per CTA, and keeps
// the rows of D = [shapedirs; posedirs] of its vertices resident in shared memory for all three components; samples
// stream through in chunks of 96 (48 per CTA: in pair mode the B operand is split between the CTAs).  With M = 128 the
// TMEM of each CTA holds rows = its 64 vertices in lanes 0..63 for the first half of the N columns and again in lanes
// 64..127 for the second half (profiles/r02_mma_probe.txt), so v_posed[vertex, sample] (N = 96 samples) and the blended
// transforms T[vertex, (sample, 3x4)] (N = 96: 8 samples) meet in the same lane when the 8 samples of a T group are 4
// of the first and 4 of the second half of the chunk.  All operands f16 main part + f16 residual (X.D ~= Xh.Dh + Xl.Dh
// + Xh.Dl; the residual of D, ~2^-12 |D|, is a subnormal fp16 number with absolute error <= 2^-25: ~2^-18 |D|).
// All images K-major, no swizzle: element (row, k) at (k/8) * rows*8 + (row/8)*64 + (row%8)*8 + k%8   (16-bit units)
#define FF_VH 64                          // vertices per CTA
#define FF_NHT (2 * SK_NVT)               // 108 half tiles
#define FF_CH 96                          // samples per chunk (N of the blend MMAs; leaves TMEM for 4 T accumulators)
#define FF_SH 48                          // samples per CTA of the pair
#define FF_NKS (KPAD / 16)                // 14 K16 steps
#define FF_DPART (FF_VH * KPAD)           // elements per (component, part) block of a half tile (28 672 B)
#define FF_DTILE (6 * FF_DPART)           // per half tile: [component 3][part 2]   (172 032 B)
#define FF_XPART (FF_SH * 16)             // elements per part of a (chunk, half, k-step) slab (1 536 B)
#define FF_XKS (2 * FF_XPART)             // per (chunk, half, k-step): [Xh | Xl]   (3 072 B)
#define FF_XHALF (FF_NKS * FF_XKS)        // per (chunk, half)
#define FF_G 8                            // samples per T group: 4 of each half of the chunk
#define FF_NG (FF_CH / FF_G)              // 12 groups per chunk
#define FF_AROWS 48                       // rows (4 samples x 12) per (group, half)
#define FF_APART (FF_AROWS * SK_KPAD)     // elements per part (3 072 B)
#define FF_AGH (2 * FF_APART)             // per (chunk, group, half): [Ah | Al]   (6 144 B)
#define FF_WPART (FF_VH * SK_KPAD)        // elements per part of a half tile of W (4 096 B)
#define FF_WHT (2 * FF_WPART)             // per half tile: [Wh | Wl]   (8 192 B)
__host__ __device__ inline size_t ff_kmajor(int rows, int row, int k) {
  return (size_t)(k / 8) * (rows * 8) + (size_t)(row / 8) * 64 + (size_t)((row & 7) * 8 + (k & 7));
}
__host__ __device__ inline size_t ff_dimg_off(int v, int c, int k, int part) {
  return (size_t)(v / FF_VH) * FF_DTILE + (size_t)(c * 2 + part) * FF_DPART + ff_kmajor(FF_VH, v % FF_VH, k);
}
__host__ __device__ inline size_t ff_wimg_off(int v, int j, int part) {
  return (size_t)(v / FF_VH) * FF_WHT + (size_t)part * FF_WPART + ff_kmajor(FF_VH, v % FF_VH, j);
}
__host__ __device__ inline size_t ff_ximg_off(int b, int k, int part) {
  const int ch = b / FF_CH, h = (b % FF_CH) / FF_SH, s = b % FF_SH;
  return ((size_t)(ch * 2 + h) * FF_NKS + (size_t)(k / 16)) * FF_XKS + (size_t)part * FF_XPART + ff_kmajor(FF_SH, s, k % 16);
}
__host__ __device__ inline size_t ff_aimg_off(int b, int j, int e, int part) {
  const int ch = b / FF_CH, h = (b % FF_CH) / FF_SH, s = b % FF_SH;
  return ((size_t)(ch * FF_NG + s / 4) * 2 + (size_t)h) * FF_AGH + (size_t)part * FF_APART +
         ff_kmajor(FF_AROWS, (s % 4) * 12 + e, j);
}
int32_t launch_fused_fwd(const smplb_model* m, int B, const float* Ximg2, const float* Aimg2, float* verts, float* kpv,
                         cudaStream_t st);

int32_t launch_skin_simt(const smplb_model* m, int B, const float* vposed, const float* A, float* verts,
                         cudaStream_t st);
// verts and kpv are each nullable (kpv: compact keypoint-vertex rows, see smplb_model)
int32_t launch_skin_tc(const smplb_model* m, int B, const float* vposed, const float* Aimg, float* verts, float* kpv,
                       cudaStream_t st);
int32_t launch_skin_bwd_vposed(const smplb_model* m, int B, const float* g_verts, const float* Aimg, float* g_img,
                               cudaStream_t st);
int32_t launch_skin_bwd_gA(const smplb_model* m, int B, const float* g_verts, const float* vposed, float* g_A,
                           cudaStream_t st);
void build_wimg3(const float* weights, unsigned short* out);   // skin_bwd_tc.cu (host)
size_t wimg3_elems();
int32_t launch_skin_bwd_gA_tc(const smplb_model* m, int B, const float* g_verts, const float* vposed, float* g_A,
                              cudaStream_t st);
int32_t launch_gverts_total(const smplb_model* m, int B, const float* g_verts, const float* g_kp, float* out,
                            cudaStream_t st);
// g_X partial sums: [n_slices][B][KPAD]; returns the slice count through *n_slices
int32_t launch_blend_bwd(const smplb_model* m, int B, const float* g_img, float* g_X, int* n_slices, cudaStream_t st);
int blend_bwd_slices(int B);
void timing_mark_bwd(int slot, cudaStream_t st);   // capi.cu: backward per-kernel event hooks (no-op unless enabled)
int32_t launch_keypoints(const smplb_model* m, int B, const float* verts, float* kp, cudaStream_t st);
int32_t launch_keypoints_compact(const smplb_model* m, int B, const float* kpv, float* kp, cudaStream_t st);

// TF32 helpers --------------------------------------------------------------------------------
__device__ __forceinline__ float tf32_rna_dev(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__host__ __device__ inline float tf32_rna_host(float x) {
  // round-to-nearest (ties away) to 10 explicit mantissa bits, like cvt.rna.tf32.f32
  union { float f; uint32_t u; } c;
  c.f = x;
  if ((c.u & 0x7f800000u) == 0x7f800000u) return x;
  c.u = (c.u + 0x1000u) & 0xffffe000u;
  return c.f;
}

}  // namespace smplb
