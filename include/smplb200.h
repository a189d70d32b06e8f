/*
 * smplb200 — C ABI of the B200-native (sm_100a) batched SMPL + projection hot path.
 *
 * The reference (blzq/3D-pose-estimation) reaches this path through plain Python call
 * signatures inside a TF1 graph, not through an FFI.  Each entry point below
 * names the reference interface it replaces (paths relative to the reference root).  The
 * Python drop-ins (`tf_smpl.batch_smpl.SMPL`, `tf_perspective_projection.project.project`,
 * …) bind these with ctypes; tensors are only carriers (raw device pointers).
 *
 * Conventions
 *   - every function returns an int32 status: SMPLB_OK or a negative SMPLB_E_* code;
 *     `smplb_last_error()` returns a thread-local message for the last failing call;
 *   - all `const float*` / `float*` batch arguments are DEVICE pointers (fp32, contiguous)
 *     unless the name ends in `_host`; model-constant arguments of smplb_model_create are
 *     HOST pointers and are copied;
 *   - launches are asynchronous on the given stream (a `cudaStream_t` passed as void*);
 *     no hidden device synchronisation, no allocation on the hot calls; the caller owns the
 *     workspace (`smplb_workspace_bytes`);
 *   - a model handle is immutable after creation: concurrent calls from several host threads /
 *     streams are legal when each call uses its own workspace;
 *   - the kernels launch on the CURRENT CUDA device of the calling thread: it must be the device
 *     the model / mesh handle was created on (cudaSetDevice first), else SMPLB_E_INVALID_ARG.
 */
#ifndef SMPLB200_H_
#define SMPLB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMPLB_NV 6890      /* vertices            (batch_smpl.py:30)  */
#define SMPLB_NV3 20670    /* 3*NV                                    */
#define SMPLB_NJ 24        /* joints              (config.py:7)       */
#define SMPLB_NB 10        /* betas               (batch_smpl.py:31)  */
#define SMPLB_NP 207       /* pose-feature width  (batch_smpl.py:109) */
#define SMPLB_K 217        /* NB + NP: blend-shape contraction length */
#define SMPLB_KPAD 224     /* K padded to the MMA k-block             */
#define SMPLB_VP_PITCH 20736 /* row pitch (floats) of the internal v_posed buffer: 81*256 */

enum {
  SMPLB_OK = 0,
  SMPLB_E_INVALID_ARG = -1,
  SMPLB_E_BAD_SHAPE = -2,
  SMPLB_E_CUDA = -3,
  SMPLB_E_NOT_FINITE = -4,
  SMPLB_E_UNSUPPORTED = -5
};

/* flags for smplb_forward / smplb_workspace_bytes */
enum {
  SMPLB_F_BLEND_SIMT = 1 << 0,   /* blend-shape contraction on CUDA cores (fp32 FFMA) instead of
                                    the tcgen05 split-operand kernel — verification / bring-up path   */
  SMPLB_F_KEEP_VPOSED = 1 << 1,  /* keep v_posed in the workspace for smplb_backward           */
  SMPLB_F_NO_SKIN = 1 << 2,      /* get_skin=False and no keypoints wanted: joints-only path   */
  SMPLB_F_SKIN_SIMT = 1 << 3,    /* skinning on CUDA cores with the compressed <=4-nonzero weights instead
                                    of the tcgen05 split-operand weight blend — verification path       */
  SMPLB_F_SCRATCH_VERTS = 1 << 4, /* SIMT skinning path only: keypoints wanted but no verts buffer given, the workspace holds verts (the tensor-core path needs no verts for keypoints) */
  SMPLB_F_FUSED = 1 << 5         /* one kernel for blend GEMM + weight blend + skinning: v_posed stays in TMEM and never
                                    reaches HBM (north_star item 4).  Not combinable with SMPLB_F_KEEP_VPOSED (the
                                    backward needs v_posed), SMPLB_F_BLEND_SIMT or SMPLB_F_SKIN_SIMT */
};

typedef struct smplb_model smplb_model;

/* Camera / image constants of the loss block: deps/pose_3d/config.py:9-12, pose_model_3d.py:41-42 */
typedef struct {
  float img_w;       /* 320 */
  float img_h;       /* 240 */
  float ss;          /* 0.024 sensor size */
  float focal_const; /* 0.05; used when the focal pointer is NULL (pose_model_3d.py:209-210) */
} smplb_camcfg;

/* I/O of one forward call.  Replaces SMPL.__call__ (deps/tf_smpl/batch_smpl.py:71-143) and,
 * when cam_pos != NULL, the projection of the 24 posed joints + pixel rescale
 * (deps/tf_perspective_projection/project.py:7-48, deps/pose_3d/pose_model_3d.py:202-231). */
typedef struct {
  int32_t B;
  const float* beta;   /* [B,10]                                                       */
  const float* theta;  /* [B,72] axis-angle, global rotation first                      */
  float* verts;        /* [B,6890,3]   nullable                                         */
  float* joints;       /* [B,24,3]     J_transformed (batch_smpl.py:118), nullable      */
  float* kp;           /* [B,NK,3]     cocoplus/lsp keypoints (batch_smpl.py:134-138)   */
  float* Rs;           /* [B,24,3,3]   nullable                                         */
  float* A;            /* [B,24,3,4]   top three rows of the relative transforms
                                       (batch_lbs.py:138-142), nullable                 */
  const float* cam_pos;   /* [B,3] nullable: no projection                              */
  const float* cam_rot;   /* [B,3] axis-angle                                           */
  const float* focal;     /* [B]   nullable -> cfg.focal_const                          */
  float* uv;              /* [B,24,2] project() output, nullable                        */
  float* px;              /* [B,24,2] pixels, nullable                                  */
  smplb_camcfg cfg;
} smplb_forward_io;

/* Loss configuration of the re-projection block (pose_model_3d.py:194-254, config.py:17-23). */
typedef struct {
  smplb_camcfg cam;
  float reproj_scale;   /* 1.0*ss                       (config.py:22)  */
  float reproj_cap;     /* 1/24, <=0 disables the cap   (pose_model_3d.py:237) */
  float joint_scale;    /* 2.0                          (config.py:18)  */
  float mesh_scale;     /* 1.0                          (config.py:17)  */
} smplb_losscfg;

const char* smplb_last_error(void);
const char* smplb_version(void);

/* SMPL.__init__ (deps/tf_smpl/batch_smpl.py:17-69).  All pointers HOST, reference layouts:
 *   v_template [6890*3]; shapedirs [10*20670] (row k = blend shape k, column 3v+c, :34-37);
 *   posedirs [207*20670] (:46-51); J_regressor [6890*24] dense (:40-43); weights [6890*24]
 *   (:57-60); kp_regressor [6890*NK] dense (:63-69), NK = 19 or 14; parents [24]
 *   (kintree_table[0] as int32, entry 0 ignored, :54).
 * Dense LBS weights are supported (the tensor-core skinning contracts all 24 joints); only the
 * SMPLB_F_SKIN_SIMT verification path needs at most 4 non-zeros per vertex (true of SMPL).
 * Builds the derived constants on `device`. */
int32_t smplb_model_create(smplb_model** out, int32_t device,
                           const float* v_template_host, const float* shapedirs_host,
                           const float* posedirs_host, const float* J_regressor_host,
                           const float* weights_host, const float* kp_regressor_host,
                           int32_t NK, const int32_t* parents_host);
int32_t smplb_model_destroy(smplb_model* m);
int32_t smplb_model_nk(const smplb_model* m);

/* Caller-owned scratch for one forward (+ backward when SMPLB_F_KEEP_VPOSED) call of batch B. */
size_t smplb_workspace_bytes(const smplb_model* m, int32_t B, int32_t flags);

/* SMPL.__call__ (+ optional projection).  Kernels: pose_prep -> blend GEMM -> skinning.
 * The two GEMMs run on tcgen05 from 16-bit split operands (fp16 main part + fp16 residuals in the forward, bf16 pairs in the backward, fp32 accumulation,
 * fp32-faithful: ~2e-7 of the vertex scale): |beta| and the model constants must stay inside the fp16 range
 * (<= 65504; smplb_model_create rejects a model outside it with SMPLB_E_UNSUPPORTED). */
int32_t smplb_forward(const smplb_model* m, const smplb_forward_io* io, void* workspace,
                      size_t workspace_bytes, int32_t flags, void* stream);

/* Input-range check of the LAST smplb_forward call that used `workspace` (stream-ordered after it; synchronises the
 * stream): SMPLB_E_NOT_FINITE when a beta entry was not finite or beyond the fp16 range (|beta| > 65504: the fp16 main
 * part of the split-operand GEMMs cannot carry it -- the vertices of that call are invalid), or a theta entry was not
 * finite; SMPLB_OK otherwise.  The forward itself never synchronises, so the check is a separate, optional call. */
int32_t smplb_forward_status(const void* workspace, void* stream);

/* Same call with HOST buffers (pinned or pageable): H2D of beta/theta/camera, the kernels, D2H of
 * every non-NULL output, then a stream synchronise.  `io` holds HOST pointers here; `dev_io`
 * holds caller-owned DEVICE staging buffers of the same shapes.  This is the end-to-end entry
 * the bench times ("e2e"). */
int32_t smplb_forward_host(const smplb_model* m, const smplb_forward_io* io_host,
                           const smplb_forward_io* dev_io, void* workspace,
                           size_t workspace_bytes, int32_t flags, void* stream);

/* batch_rodrigues (deps/tf_smpl/batch_lbs.py:38-57) when variant==0;
 * rodrigues_batch (deps/tf_perspective_projection/project.py:51-81) when variant==1. */
int32_t smplb_rodrigues(int32_t N, const float* rvecs, float* R_out, int32_t variant, void* stream);

/* batch_global_rigid_transformation (deps/tf_smpl/batch_lbs.py:84-144): Rs[N,24,3,3], Js[N,24,3]
 * -> new_J[N,24,3], A[N,24,4,4]. */
int32_t smplb_global_rigid(int32_t N, const float* Rs, const float* Js, const int32_t* parents_host,
                           int32_t rotate_base, float* new_J, float* A44, void* stream);

/* project (deps/tf_perspective_projection/project.py:7-48): pts[N,3] -> uv[N,2].
 * cam_tiled==1: cams_* are per point ([N,3],[N,3],[N]) exactly like the reference call;
 * cam_tiled==0: cams_* are per group of P consecutive points ([N/P,...]). */
int32_t smplb_project(int32_t N, int32_t P, const float* pts, const float* cams_pos,
                      const float* cams_angles, const float* focal_lens, int32_t cam_tiled,
                      float* uv, void* stream);

/* batch_orth_proj_idrot (deps/tf_smpl/projection.py:10-26): X[N,P,3], camera[N,3] -> [N,P,2]. */
int32_t smplb_orth_proj(int32_t N, int32_t P, const float* X, const float* camera, float* out,
                        void* stream);

/* Re-projection loss block (deps/pose_3d/pose_model_3d.py:218-238): px_pred/px_gt [B,24,2] ->
 * loss_out[0] = scale * mean((px_gt-px_pred)^2) (uncapped), loss_out[1] = capped value. */
int32_t smplb_reproj_loss(int32_t B, const float* px_pred, const float* px_gt,
                          const smplb_losscfg* cfg, float* loss_out, void* stream);

/* Backward of smplb_forward (TF autodiff of batch_smpl.py:91-138, pose_model_3d.py:393-400;
 * adjoint in SURVEY.md appendix A).  Upstream gradients may be NULL (treated as zero).
 * g_px is the gradient wrt the pixel coordinates produced by the fused projection (camera
 * treated as constant, pose_model_3d.py:218-221).  Requires the workspace of the matching
 * forward call made with SMPLB_F_KEEP_VPOSED. */
typedef struct {
  int32_t B;
  const float* beta;
  const float* theta;
  const float* g_verts;   /* [B,6890,3] nullable */
  const float* g_joints;  /* [B,24,3]   nullable */
  const float* g_kp;      /* [B,NK,3]   nullable */
  const float* g_px;      /* [B,24,2]   nullable */
  const float* cam_pos;   /* needed when g_px != NULL */
  const float* cam_rot;
  const float* focal;
  smplb_camcfg cfg;
  float* g_beta;          /* [B,10] */
  float* g_theta;         /* [B,72] */
} smplb_backward_io;

int32_t smplb_backward(const smplb_model* m, const smplb_backward_io* io, void* workspace,
                       size_t workspace_bytes, int32_t flags, void* stream);

/* Backward of project() wrt points and camera (pose_model_3d.py:243-245 path; project.py:30-42 is differentiable in
 * the focal length too): g_uv[N,2] -> g_pts[N,3], g_cam_pos / g_cam_angles ([N,3] tiled or [N/P,3]),
 * g_focal_lens ([N] tiled or [N/P]); every output nullable. */
int32_t smplb_project_backward(int32_t N, int32_t P, const float* pts, const float* cams_pos,
                               const float* cams_angles, const float* focal_lens,
                               int32_t cam_tiled, const float* g_uv, float* g_pts,
                               float* g_cams_pos, float* g_cams_angles, float* g_focal_lens, void* stream);

/* tf.train.AdamOptimizer update (pose_model_3d.py:393): eps outside the bias correction.  `t` (1-based) is a host
 * value baked into the launch: do not capture this call in a CUDA graph (use the _dev pair below). */
int32_t smplb_adam_step(int64_t n, float* x, const float* g, float* m, float* v, int32_t t,
                        float lr, float beta1, float beta2, float eps, void* stream);
/* The same update with the step counter on the DEVICE, so that a captured step advances Adam's bias correction on
 * every graph replay: smplb_adam_tick does t = ++(*step_dev) and writes lr_t(t) = lr sqrt(1-b2^t)/(1-b1^t) to
 * lr_t_dev[0] (once per optimiser step); smplb_adam_step_dev applies the update to one variable with that lr_t. */
int32_t smplb_adam_tick(int32_t* step_dev, float lr, float beta1, float beta2, float* lr_t_dev, void* stream);
int32_t smplb_adam_step_dev(int64_t n, float* x, const float* g, float* m, float* v, const float* lr_t_dev,
                            float beta1, float beta2, float eps, void* stream);

/* One Adam step of iterative fitting on the 2D re-projection loss with per-sample pose (and, when m_beta /
 * v_beta are given, shape) as the variables: the loss block of deps/pose_3d/pose_model_3d.py:194-238 driven by
 * the optimiser of :393-401, joint-only path (no vertices).  theta / beta / m / v are updated in place.
 * loss[0] = scale * sum((px - gt_px)^2) / n_total (this shard's share of the global mean; n_total = global
 * B*24*2 so that the shares of several GPUs add up under an all-reduce), loss[1] = min(cap, loss[0]),
 * loss[2] = 1 when the cap is inactive else 0 (the gradient gate of tf.minimum, :237), loss[3] = Adam's lr_t of this
 * step.  The cap acts on the GLOBAL mean, so reproj_cap > 0 requires n_total == B*48 (SMPLB_E_UNSUPPORTED otherwise).
 * Two launches when the cap is off (forward chain + projection + residual + adjoint + the Adam update of the sample in
 * one per-sample kernel, then the loss / step-counter kernel), four when it is on (Adam waits for the global loss).
 * Stream-ordered, allocation-free: capturable in a CUDA graph when step_dev is given (the step counter then lives on
 * the device and every replay advances the bias correction; with step_dev == NULL the host value `step` is used and
 * a captured launch would freeze it). */
typedef struct {
  int32_t B;
  float* beta;            /* [B,10]  updated when m_beta != NULL */
  float* theta;           /* [B,72]  updated */
  const float* cam_pos;   /* [B,3] */
  const float* cam_rot;   /* [B,3] */
  const float* focal;     /* [B] nullable -> cfg.cam.focal_const */
  const float* gt_px;     /* [B,24,2] target 2D joints in pixels */
  float* px;              /* [B,24,2] out: current projection */
  float* g_px;            /* [B,24,2] scratch */
  float* g_beta;          /* [B,10]  out */
  float* g_theta;         /* [B,72]  out */
  float* m_theta;         /* Adam moments, [B,72] */
  float* v_theta;
  float* m_beta;          /* nullable pair */
  float* v_beta;
  float* loss;            /* [4] device */
  void* scratch;          /* smplb_fit_scratch_bytes(B) device bytes */
  smplb_losscfg cfg;
  int64_t n_total;
  int32_t step;           /* 1-based; ignored when step_dev != NULL */
  float lr, beta1, beta2, eps;
  int32_t* step_dev;      /* device, nullable: steps done so far (start at 0); incremented by the call */
} smplb_fit_io;
size_t smplb_fit_scratch_bytes(int32_t B);
int32_t smplb_fit_step(const smplb_model* m, const smplb_fit_io* io, void* stream);

/* ---- Triangle rasteriser: deps/tf_mesh_renderer ---------------------------------------
 * Replaces the custom TF ops RasterizeTriangles (kernels/rasterize_triangles_op.cc:26-147, implementation
 * kernels/rasterize_triangles_impl.cc:75-155) and RasterizeTrianglesGrad (kernels/rasterize_triangles_grad.cc:
 * 52-224), which the reference loads with tf.load_op_library (rasterize_triangles.py:21-23) and calls once per
 * image through tf.map_fn (:98-117).  Here one call rasterises a whole batch.
 * A mesh handle owns the device copy of the triangle list ([T,3] vertex ids, clockwise = front facing) and the
 * vertex -> corner table of the gradient; it is immutable, so any number of streams may share it. */
typedef struct smplb_mesh smplb_mesh;
int32_t smplb_mesh_create(smplb_mesh** out, int32_t device, const int32_t* triangles_host, int32_t T,
                          int32_t V);
int32_t smplb_mesh_destroy(smplb_mesh* m);
size_t smplb_rasterize_workspace_bytes(const smplb_mesh* m, int32_t B, int32_t W, int32_t H);
/* ndc [B,V,3] normalised device coordinates (device) -> bary [B,H,W,3], ids [B,H,W] (int32), zbuf [B,H,W];
 * any output may be NULL.  Uncovered pixels: bary 0, id 0, z 1 (rasterize_triangles_op.cc:121-125).
 * Bit-identical to the reference's C++ for finite depths. */
int32_t smplb_rasterize(const smplb_mesh* m, int32_t B, const float* ndc, int32_t W, int32_t H, float* bary,
                        int32_t* ids, float* zbuf, void* workspace, size_t workspace_bytes, void* stream);
/* g_bary [B,H,W,3] -> g_ndc [B,V,3] (z component 0), from the forward's bary / ids. */
int32_t smplb_rasterize_backward(const smplb_mesh* m, int32_t B, const float* ndc, const float* bary,
                                 const int32_t* ids, const float* g_bary, int32_t W, int32_t H, float* g_ndc,
                                 void* workspace, size_t workspace_bytes, void* stream);

/* normals_from_mesh (deps/pose_3d/utils.py:92-127, with vertex_faces_from_face_verts :130-149 built into the mesh
 * handle): verts [B,V,3] -> unit vertex normals [B,V,3], oriented by the reference's outward test.  One HBM pass
 * (a CTA holds one sample's vertices and normals in shared memory); V <= 9386. */
int32_t smplb_mesh_normals(const smplb_mesh* m, int32_t B, const float* verts, float* normals, void* stream);

/* Fused renderer for the path pose_3d drives (deps/pose_3d/utils.py:20-58 -> mesh_renderer.mesh_renderer without
 * specular terms, deps/tf_mesh_renderer/mesh_renderer.py:197-401): clip transform + perspective divide
 * (rasterize_triangles.py:73-86), rasterisation, perspective-correct interpolation of normals / positions /
 * diffuse colours (:119-164) and diffuse Phong shading (mesh_renderer.py:84-118,158-162) in one pass; only the
 * RGBA image [B,H,W,4] (rows reversed as tf.reverse does, :162) is written.  verts/normals [B,V,3], diffuse
 * [B,V,3] or NULL (= ones), clip_transforms [B,4,4] row-major (perspective x look_at), lights [B,L,3] x2,
 * ambient [B,3] or NULL. */
size_t smplb_render_workspace_bytes(const smplb_mesh* m, int32_t B, int32_t W, int32_t H);
int32_t smplb_render_diffuse(const smplb_mesh* m, int32_t B, const float* verts, const float* normals,
                             const float* diffuse, const float* clip_transforms, const float* light_positions,
                             const float* light_intensities, int32_t L, const float* ambient, int32_t W, int32_t H,
                             float* rgba, void* workspace, size_t workspace_bytes, void* stream);

/* The element-wise graph of deps/tf_mesh_renderer/rasterize_triangles.py around the rasteriser op, forward and adjoint
 * (the gradients TF's autodiff derives for these ops), so that a differentiable render runs on kernels end to end:
 * clip transform (:73-86) -> smplb_rasterize -> interpolation (:119-164) -> smplb_phong_shade.
 * smplb_clip_transform: verts [B,V,3], clip_transforms [B,4,4] row-major -> ndc [B,V,3] (guarded perspective divide) and
 * clip_w [B,V] (the guarded w).  Backward: g_ndc, g_clip_w (nullable) -> g_verts; the matrices are constants.
 * smplb_interpolate: attributes [B,V,C], clip_w, the rasteriser's bary [B,H,W,3] / ids [B,H,W], background [C] ->
 * images [B,H,W,C] (perspective-correct interpolation, alpha from the barycentric sum, background compositing).
 * Backward: g_images -> g_attributes [B,V,C], g_clip_w [B,V] (both zero-filled here, then float atomicAdd: their last bits
 * depend on the order in which pixels arrive) and g_bary [B,H,W,3] (feeds smplb_rasterize_backward). */
int32_t smplb_clip_transform(int32_t B, int32_t V, const float* verts, const float* clip_transforms, float* ndc,
                             float* clip_w, void* stream);
int32_t smplb_clip_transform_backward(int32_t B, int32_t V, const float* verts, const float* clip_transforms,
                                      const float* g_ndc, const float* g_clip_w, float* g_verts, void* stream);
int32_t smplb_interpolate(const smplb_mesh* m, int32_t B, int32_t C, const float* attributes, const float* clip_w,
                          const float* bary, const int32_t* ids, const float* background, int32_t W, int32_t H,
                          float* images, void* stream);
int32_t smplb_interpolate_backward(const smplb_mesh* m, int32_t B, int32_t C, const float* attributes, const float* clip_w,
                                   const float* bary, const int32_t* ids, const float* background, int32_t W, int32_t H,
                                   const float* g_images, float* g_attributes, float* g_clip_w, float* g_bary, void* stream);

/* Phong shading of interpolated attribute images with the specular term, and its adjoint (SURVEY.md 8f N1:
 * deps/tf_mesh_renderer/mesh_renderer.py:23-162 phong_shader, and the part of mesh_renderer :375-400 between the
 * rasterised attribute image and the lit image; TF differentiates this graph, so does the backward entry point).
 * attr [B,H,W,C]: the attribute image exactly as rasterize_triangles returns it, channels addressed by offsets:
 * off_normal (3), off_position (3), off_diffuse (3), off_specular (3, or -1: no specular term), off_shininess (1, or -1:
 * shininess [B] per image).  normalize_normals != 0: the normals are l2-normalised first (mesh_renderer.py:376).
 * alphas [B,H,W] or NULL: alpha = any(diffuse >= 0) (mesh_renderer.py:387).  camera [B,3] is needed with a specular term.
 * The reference normalises reflection . camera over the PIXEL axis (:134-136, tf.nn.l2_normalize(axis=2)); that
 * quirk is reproduced, which makes the specular term a two-pass computation.  rgba [B,H,W,4], rows reversed (:162).
 * Lights are constants (no gradient).  L <= 8.  scratch: smplb_phong_scratch_bytes(B, H, W, L) bytes; the backward call
 * must get the scratch of the forward call of the same inputs untouched (it holds the per-image norms).
 * g_attr [B,H,W,C] receives the gradient of every channel (unused channels: 0); g_alphas [B,H,W] or NULL. */
typedef struct {
  int32_t B, H, W, C, L;
  int32_t off_normal, off_position, off_diffuse, off_specular, off_shininess;
  int32_t normalize_normals;
  const float* attr;
  const float* alphas;
  const float* shininess;        /* [B], used when off_shininess < 0 and off_specular >= 0 */
  const float* light_positions;  /* [B,L,3] */
  const float* light_intensities;/* [B,L,3] */
  const float* camera;           /* [B,3] */
  const float* ambient;          /* [B,3] or NULL */
} smplb_phong_io;
size_t smplb_phong_scratch_bytes(int32_t B, int32_t H, int32_t W, int32_t L);
int32_t smplb_phong_shade(const smplb_phong_io* io, float* rgba, void* scratch, size_t scratch_bytes, void* stream);
int32_t smplb_phong_shade_backward(const smplb_phong_io* io, const float* g_rgba, float* g_attr, float* g_alphas,
                                   void* scratch, size_t scratch_bytes, void* stream);

/* ---- Per-sample operators either side of the path ---------------------------------
 * add_axis_angle_rotations (deps/pose_3d/utils.py:170-199): rv1, rv2 [N,3] -> axis-angle of "rv2 then rv1". */
int32_t smplb_add_axis_angle(int32_t N, const float* rv1, const float* rv2, float* out, void* stream);
/* rotate_global_pose (deps/pose_3d/utils.py:152-167): thetas [B,72], zrot [B] -> thetas with the SURREAL global
 * rotation re-expressed for a y-up body (called on GT poses at pose_model_3d.py:371-372). */
int32_t smplb_rotate_global_pose(int32_t B, const float* thetas, const float* zrot, float* out, void* stream);
/* Camera-angle (ray) loss and camera regularisers (deps/pose_3d/pose_model_3d.py:256-292,
 * utils.get_2d_points_in_3d utils.py:75-89).  gt_joints3d [B,24,3], gt_joints2d [B,24,2] pixels (already
 * flipped, :196-197), cam_pos/cam_rot [B,3], cam_f [B] (focal used for the rays), f_raw [B] nullable (the raw
 * focal output the limit term sees, outputs[:,78]).  loss_out (device, 4 floats) = {cam_angle_loss,
 * cam_neg_loss, cam_rot_loss, their sum}; angle_scale = config.cam_angle_loss_scale (10).  Gradients of the sum
 * (nullable, pos and rot together): g_cam_pos/g_cam_rot [B,3], g_cam_f/g_f_raw [B]. */
size_t smplb_cam_losses_scratch_bytes(int32_t B);
int32_t smplb_cam_losses(int32_t B, const float* gt_joints3d, const float* gt_joints2d, const float* cam_pos,
                         const float* cam_rot, const float* cam_f, const float* f_raw, const smplb_camcfg* cfg,
                         float angle_scale, float* loss_out, float* g_cam_pos, float* g_cam_rot, float* g_cam_f,
                         float* g_f_raw, void* scratch, void* stream);

/* Allocation-free variant of smplb_cam_losses.  No counterpart in the reference: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same inputs, outputs and arithmetic as smplb_cam_losses, but the per-block
 * partial sums live in a fixed device array inside the library (6 160 bytes, held by every GPU that loads the
 * library until exit) instead of a caller-provided scratch buffer; the call allocates nothing.
 * B > SMPLB_CAM_LOSSES_STATIC_MAX_B is rejected with SMPLB_E_UNSUPPORTED.  All calls on a device share the array:
 * calls that may overlap on the GPU (different streams, concurrent graphs) race -- serialise them. */
#define SMPLB_CAM_LOSSES_STATIC_MAX_B 65536
int32_t smplb_cam_losses_static(int32_t B, const float* gt_joints3d, const float* gt_joints2d, const float* cam_pos,
                                const float* cam_rot, const float* cam_f, const float* f_raw, const smplb_camcfg* cfg,
                                float angle_scale, float* loss_out, float* g_cam_pos, float* g_cam_rot, float* g_cam_f,
                                float* g_f_raw, void* stream);

/* 3D losses of deps/pose_3d/pose_model_3d.py:183-192 (tf.losses.mean_squared_error with a scalar weight):
 * loss_out[0] (+)= weight * sum((pred - target)^2) / n_total, and, when g_out is given, its gradient w.r.t. pred,
 * g_out[i] = 2 * weight / n_total * (pred[i] - target[i]).  n_total = element count of the global mean (n, or the
 * count over all ranks when the batch is sharded).  mesh_loss: weight 1.0 on verts [B,6890,3]; joint_loss: weight
 * 2.0 on J_transformed [B,24,3] (config.py:17-18).  accumulate != 0 adds to loss_out[0] (loss3d = mesh + joint).
 * Fixed-order reduction; scratch: smplb_mse_scratch_bytes() of device memory. */
size_t smplb_mse_scratch_bytes(void);
int32_t smplb_mse_grad(int64_t n, const float* pred, const float* target, float weight, int64_t n_total,
                       float* g_out, float* loss_out, int32_t accumulate, void* scratch, void* stream);

/* Measurement aid (no reference counterpart): while enabled, every smplb_forward on this host thread
 * records CUDA events on its stream around pose_prep / blend GEMM / skinning / keypoints.
 * smplb_timing_end synchronises those events and writes ms_out[call][4] = {prep, blend, skin, keypoints}
 * milliseconds. */
int32_t smplb_timing_begin(int32_t max_calls);
int32_t smplb_timing_end(float* ms_out_host, int32_t* ncalls_out_host);
/* The same for smplb_backward calls with vertex gradients made since smplb_timing_begin:
 * ms_out[call][4] = {g_A reduction, g_vposed (skinning adjoint), transposed blend GEMM, joint path}. */
int32_t smplb_timing_bwd_end(float* ms_out_host, int32_t* ncalls_out_host);

#ifdef __cplusplus
}
#endif
#endif /* SMPLB200_H_ */
