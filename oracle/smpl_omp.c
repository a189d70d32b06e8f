/* TEST / BENCH INFRASTRUCTURE -- not part of the product path (the library never links or calls this file).
 *
 * A tight fp32 C + OpenMP version of the hot path, SURVEY.md 8d (iii): "optionally a tight C++/OpenMP version as a
 * fairer CPU number".  Same mathematics as deps/tf_smpl/batch_smpl.py:91-138, batch_lbs.py:38-144 and
 * deps/tf_perspective_projection/project.py:7-81 (+ the pixel rescale of pose_model_3d.py:229-231), but written the way
 * a CPU programmer would rather than as the reference's TF op sequence:
 *   - the blend shapes are ONE [217 x 20670] matrix, stored in tiles of 8 vertex-axis floats, applied to a block of 8
 *     samples at a time with the samples in the lanes of a 256-bit vector and 8 accumulators in registers (the
 *     reference does two matmuls);
 *   - J = Jbar + JS beta (the joint regressor folded through the shape basis, exact in real arithmetic);
 *   - skinning with the <= SM_NW non-zero weights per vertex, vectorised over the 8 samples, instead of the dense
 *     [6890 x 24] product with the tiled W and the [B,6890,4,4] matmul (batch_smpl.py:120-132);
 *   - the key-point regressor in compressed sparse form.
 * It is checked against the float64 oracle (tests/test_cpu_tight.py) and timed by bench.py as cpu_baseline.tight.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NV 6890
#define NV3 20670
#define NJ 24
#define NB 10
#define NP 207
#define NK_MAX 32
#define SM_SB 8      /* samples per block = lanes of one 256-bit vector */
#define SM_NW 4
#define NKX (NB + NP)
#define NT ((NV3 + 7) / 8)   /* tiles of 8 consecutive floats of the vertex axis */
#define CH_T 129             /* tiles per chunk: 1032 floats = 344 vertices (a multiple of 3 and of 8) */
#define CH_F (CH_T * 8)
#define CH_V (CH_F / 3)
#define NCH ((NV3 + CH_F - 1) / CH_F)

typedef float v8 __attribute__((vector_size(32), aligned(4)));

typedef struct {
  float* vt;       /* [NCH * CH_F] zero padded */
  float* D;        /* [NCH * CH_T tiles][NKX][8]: shape basis rows then pose basis rows, tiled along the vertex axis */
  float Jbar[NJ * 3];
  float JS[NJ * 3 * NB];
  int parent[NJ];
  int* widx;       /* [NV][SM_NW] */
  float* wval;     /* [NV][SM_NW] */
  int nk;
  int* kp_ptr;     /* [nk + 1] */
  int* kp_vtx;
  float* kp_val;
} smpl_omp_model;

void smpl_omp_free(smpl_omp_model* m) {
  if (!m) return;
  free(m->vt); free(m->D); free(m->widx); free(m->wval); free(m->kp_ptr); free(m->kp_vtx); free(m->kp_val);
  free(m);
}

/* shapedirs [NB][NV3], posedirs [NP][NV3], J_regressor [NV][NJ], weights [NV][NJ], kp_regressor [NV][nk] (all row-major,
 * the layouts of SMPL.__init__, batch_smpl.py:34-66).  Returns NULL when a vertex has more than SM_NW weights. */
smpl_omp_model* smpl_omp_create(const float* v_template, const float* shapedirs, const float* posedirs,
                                const float* J_regressor, const float* weights, const float* kp_regressor, int nk,
                                const int32_t* parents) {
  if (nk > NK_MAX) return NULL;
  smpl_omp_model* m = (smpl_omp_model*)calloc(1, sizeof(*m));
  m->vt = (float*)calloc((size_t)NCH * CH_F, sizeof(float));
  memcpy(m->vt, v_template, sizeof(float) * NV3);
  m->D = (float*)calloc((size_t)NCH * CH_T * NKX * 8, sizeof(float));
  for (int k = 0; k < NKX; ++k) {
    const float* row = k < NB ? shapedirs + (size_t)k * NV3 : posedirs + (size_t)(k - NB) * NV3;
    for (int i = 0; i < NV3; ++i) m->D[((size_t)(i / 8) * NKX + k) * 8 + (i & 7)] = row[i];
  }
  /* J = J_regressor^T v_shaped = Jbar + JS beta (batch_smpl.py:97-101) */
  for (int j = 0; j < NJ; ++j)
    for (int c = 0; c < 3; ++c) {
      double acc = 0.0;
      for (int v = 0; v < NV; ++v) acc += (double)J_regressor[v * NJ + j] * v_template[v * 3 + c];
      m->Jbar[j * 3 + c] = (float)acc;
      for (int k = 0; k < NB; ++k) {
        double a2 = 0.0;
        for (int v = 0; v < NV; ++v) a2 += (double)J_regressor[v * NJ + j] * shapedirs[(size_t)k * NV3 + v * 3 + c];
        m->JS[(j * 3 + c) * NB + k] = (float)a2;
      }
    }
  m->parent[0] = -1;
  for (int j = 1; j < NJ; ++j) m->parent[j] = parents[j];
  m->widx = (int*)calloc((size_t)NV * SM_NW, sizeof(int));
  m->wval = (float*)calloc((size_t)NV * SM_NW, sizeof(float));
  for (int v = 0; v < NV; ++v) {
    int n = 0;
    for (int j = 0; j < NJ; ++j) {
      const float w = weights[v * NJ + j];
      if (w != 0.f) {
        if (n == SM_NW) { smpl_omp_free(m); return NULL; }
        m->widx[v * SM_NW + n] = j;
        m->wval[v * SM_NW + n] = w;
        ++n;
      }
    }
  }
  m->nk = nk;
  m->kp_ptr = (int*)calloc(nk + 1, sizeof(int));
  int nnz = 0;
  for (int k = 0; k < nk; ++k)
    for (int v = 0; v < NV; ++v) nnz += kp_regressor[v * nk + k] != 0.f;
  m->kp_vtx = (int*)malloc(sizeof(int) * (nnz ? nnz : 1));
  m->kp_val = (float*)malloc(sizeof(float) * (nnz ? nnz : 1));
  int at = 0;
  for (int k = 0; k < nk; ++k) {
    m->kp_ptr[k] = at;
    for (int v = 0; v < NV; ++v)
      if (kp_regressor[v * nk + k] != 0.f) {
        m->kp_vtx[at] = v;
        m->kp_val[at++] = kp_regressor[v * nk + k];
      }
  }
  m->kp_ptr[nk] = at;
  return m;
}

/* batch_rodrigues, batch_lbs.py:38-57 (eps inside the norm only) */
static void rodrigues_smpl(const float* t, float* R) {
  const float ex = t[0] + 1e-8f, ey = t[1] + 1e-8f, ez = t[2] + 1e-8f;
  const float a = sqrtf(ex * ex + ey * ey + ez * ez);
  const float rx = t[0] / a, ry = t[1] / a, rz = t[2] / a;
  const float s = sinf(a), c = cosf(a), k = 1.0f - c;
  R[0] = c + k * rx * rx;      R[1] = k * rx * ry - s * rz; R[2] = k * rx * rz + s * ry;
  R[3] = k * ry * rx + s * rz; R[4] = c + k * ry * ry;      R[5] = k * ry * rz - s * rx;
  R[6] = k * rz * rx - s * ry; R[7] = k * rz * ry + s * rx; R[8] = c + k * rz * rz;
}

/* rodrigues_batch, project.py:51-81 */
static void rodrigues_cam(const float* r, float* R) {
  const float th = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (th == 0.f) {
    R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    return;
  }
  const float ux = r[0] / th, uy = r[1] / th, uz = r[2] / th, s = sinf(th), k = 1.0f - cosf(th);
  R[0] = 1.0f + k * (-uz * uz - uy * uy); R[1] = -s * uz + k * uy * ux;          R[2] = s * uy + k * uz * ux;
  R[3] = s * uz + k * ux * uy;            R[4] = 1.0f + k * (-uz * uz - ux * ux); R[5] = -s * ux + k * uz * uy;
  R[6] = -s * uy + k * ux * uz;           R[7] = s * ux + k * uy * uz;           R[8] = 1.0f + k * (-uy * uy - ux * ux);
}

/* One block of nb <= SM_SB samples; the lanes of the vectors are the samples (idle lanes repeat sample 0).
 * scratch: CH_F vectors. */
static void forward_block(const smpl_omp_model* m, int nb, const float* beta, const float* theta, const float* cam_pos,
                          const float* cam_rot, const float* focal, float ss, float img_w, float img_h, float* verts,
                          float* joints, float* kp, float* px, v8* vp) {
  v8 X[NKX];           /* blend coefficients [k][sample] */
  v8 A[NJ][12];        /* skinning transforms [joint][3x4][sample] */
  for (int l = 0; l < SM_SB; ++l) {
    const int s = l < nb ? l : 0;
    float R[NJ][9], J[NJ][3], GR[NJ][9], Gt[NJ][3];
    for (int k = 0; k < NB; ++k) X[k][l] = beta[s * NB + k];
    for (int j = 0; j < NJ; ++j) {
      rodrigues_smpl(theta + s * NJ * 3 + j * 3, R[j]);
      if (j > 0)   /* pose feature, batch_smpl.py:109-110 */
        for (int e = 0; e < 9; ++e) X[NB + (j - 1) * 9 + e][l] = R[j][e] - ((e == 0 || e == 4 || e == 8) ? 1.0f : 0.0f);
      for (int c = 0; c < 3; ++c) {
        float acc = m->Jbar[j * 3 + c];
        for (int k = 0; k < NB; ++k) acc += m->JS[(j * 3 + c) * NB + k] * beta[s * NB + k];
        J[j][c] = acc;
      }
    }
    /* batch_global_rigid_transformation, batch_lbs.py:84-144 */
    memcpy(GR[0], R[0], sizeof(R[0]));
    memcpy(Gt[0], J[0], sizeof(J[0]));
    for (int j = 1; j < NJ; ++j) {
      const int p = m->parent[j];
      const float d[3] = {J[j][0] - J[p][0], J[j][1] - J[p][1], J[j][2] - J[p][2]};
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c)
          GR[j][r * 3 + c] = GR[p][r * 3] * R[j][c] + GR[p][r * 3 + 1] * R[j][3 + c] + GR[p][r * 3 + 2] * R[j][6 + c];
        Gt[j][r] = GR[p][r * 3] * d[0] + GR[p][r * 3 + 1] * d[1] + GR[p][r * 3 + 2] * d[2] + Gt[p][r];
      }
    }
    float Rc[9];
    rodrigues_cam(cam_rot + s * 3, Rc);
    for (int j = 0; j < NJ; ++j) {
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) A[j][r * 4 + c][l] = GR[j][r * 3 + c];
        A[j][r * 4 + 3][l] = Gt[j][r] - (GR[j][r * 3] * J[j][0] + GR[j][r * 3 + 1] * J[j][1] + GR[j][r * 3 + 2] * J[j][2]);
      }
      if (l >= nb) continue;
      for (int r = 0; r < 3; ++r) joints[(s * NJ + j) * 3 + r] = Gt[j][r];
      /* project(), project.py:19-48, then pose_model_3d.py:229-231 */
      const float dx = Gt[j][0] - cam_pos[s * 3], dy = Gt[j][1] - cam_pos[s * 3 + 1], dz = Gt[j][2] - cam_pos[s * 3 + 2];
      const float qx = Rc[0] * dx + Rc[3] * dy + Rc[6] * dz, qy = Rc[1] * dx + Rc[4] * dy + Rc[7] * dz,
                  qz = Rc[2] * dx + Rc[5] * dy + Rc[8] * dz;
      px[(s * NJ + j) * 2] = (focal[s] * qx / qz) / ss * img_h + img_w * 0.5f;
      px[(s * NJ + j) * 2 + 1] = (focal[s] * qy / qz) / ss * img_h + img_h * 0.5f;
    }
  }
  for (int ch = 0; ch < NCH; ++ch) {
    /* v_posed of the chunk: vp[i][sample] = vt[i] + sum_k D[k][i] X[k][sample]  (batch_smpl.py:94, :112-113) */
    for (int t = 0; t < CH_T; ++t) {
      const float* d = m->D + (size_t)(ch * CH_T + t) * NKX * 8;
      const float* v0 = m->vt + (ch * CH_T + t) * 8;
      v8 a0, a1, a2, a3, a4, a5, a6, a7;
      a0 = a1 = a2 = a3 = a4 = a5 = a6 = a7 = (v8){0, 0, 0, 0, 0, 0, 0, 0};
      for (int k = 0; k < NKX; ++k) {
        const v8 x = X[k];
        a0 += d[k * 8 + 0] * x; a1 += d[k * 8 + 1] * x; a2 += d[k * 8 + 2] * x; a3 += d[k * 8 + 3] * x;
        a4 += d[k * 8 + 4] * x; a5 += d[k * 8 + 5] * x; a6 += d[k * 8 + 6] * x; a7 += d[k * 8 + 7] * x;
      }
      v8* o = vp + t * 8;
      o[0] = a0 + v0[0]; o[1] = a1 + v0[1]; o[2] = a2 + v0[2]; o[3] = a3 + v0[3];
      o[4] = a4 + v0[4]; o[5] = a5 + v0[5]; o[6] = a6 + v0[6]; o[7] = a7 + v0[7];
    }
    /* skinning of the chunk (batch_smpl.py:120-132) */
    const int vbase = ch * CH_V, nv = vbase + CH_V <= NV ? CH_V : NV - vbase;
    for (int v = 0; v < nv; ++v) {
      const int* ix = m->widx + (vbase + v) * SM_NW;
      const float* w = m->wval + (vbase + v) * SM_NW;
      const v8 x = vp[v * 3], y = vp[v * 3 + 1], z = vp[v * 3 + 2];
      v8 o[3];
      for (int r = 0; r < 3; ++r) {
        v8 T[4];
        for (int c = 0; c < 4; ++c)
          T[c] = w[0] * A[ix[0]][r * 4 + c] + w[1] * A[ix[1]][r * 4 + c] + w[2] * A[ix[2]][r * 4 + c] +
                 w[3] * A[ix[3]][r * 4 + c];
        o[r] = T[0] * x + T[1] * y + T[2] * z + T[3];
      }
      for (int l = 0; l < nb; ++l) {
        float* dst = verts + (size_t)l * NV3 + (vbase + v) * 3;
        dst[0] = o[0][l]; dst[1] = o[1][l]; dst[2] = o[2][l];
      }
    }
  }
  if (kp)
    for (int s = 0; s < nb; ++s)
      for (int k = 0; k < m->nk; ++k) {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
        for (int i = m->kp_ptr[k]; i < m->kp_ptr[k + 1]; ++i) {
          const float* v = verts + (size_t)s * NV3 + m->kp_vtx[i] * 3;
          a0 += m->kp_val[i] * v[0]; a1 += m->kp_val[i] * v[1]; a2 += m->kp_val[i] * v[2];
        }
        kp[(s * m->nk + k) * 3] = a0; kp[(s * m->nk + k) * 3 + 1] = a1; kp[(s * m->nk + k) * 3 + 2] = a2;
      }
}

/* verts [B,NV3] (nullable: kept in per-thread scratch), joints [B,NJ,3], kp [B,nk,3] nullable, px [B,NJ,2].
 * nthreads <= 0: all the threads OpenMP offers.  Returns the number of threads used. */
int smpl_omp_forward(const smpl_omp_model* m, int B, const float* beta, const float* theta, const float* cam_pos,
                     const float* cam_rot, const float* focal, float ss, float img_w, float img_h, float* verts,
                     float* joints, float* kp, float* px, int nthreads) {
  int used = 1;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  used = nthreads;
#else
  nthreads = 1;
#endif
  const int nblk = (B + SM_SB - 1) / SM_SB;
#pragma omp parallel num_threads(nthreads)
  {
    v8* vp = (v8*)malloc(sizeof(v8) * CH_F);
    float* vloc = verts ? NULL : (float*)malloc(sizeof(float) * SM_SB * NV3);
#pragma omp for schedule(dynamic, 1)
    for (int blk = 0; blk < nblk; ++blk) {
      const int b0 = blk * SM_SB, nb = b0 + SM_SB <= B ? SM_SB : B - b0;
      forward_block(m, nb, beta + (size_t)b0 * NB, theta + (size_t)b0 * NJ * 3, cam_pos + (size_t)b0 * 3,
                    cam_rot + (size_t)b0 * 3, focal + b0, ss, img_w, img_h, verts ? verts + (size_t)b0 * NV3 : vloc,
                    joints + (size_t)b0 * NJ * 3, kp ? kp + (size_t)b0 * m->nk * 3 : NULL, px + (size_t)b0 * NJ * 2, vp);
    }
    free(vp);
    free(vloc);
  }
  return used;
}
