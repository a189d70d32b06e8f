// Model handle: copies the SMPL constants (SMPL.__init__, deps/tf_smpl/batch_smpl.py:17-69) to the
// device and builds the derived tables the kernels stream: TF32 hi/lo splits of the blend-shape
// basis, the folded joint regression, compressed LBS weights, CSR keypoint regressor.
#include <math.h>
#include <string.h>

#include <vector>

#include <cmath>

#include "common.cuh"

namespace smplb {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
int32_t fail(int32_t code, const std::string& msg) {
  g_err = msg;
  return code;
}
const char* last_error() { return g_err.c_str(); }

int32_t check_current_device(int device, const char* who) {
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return fail(SMPLB_E_CUDA, std::string(who) + ": cudaGetDevice: " + cudaGetErrorString(e));
  if (cur != device)
    return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": the handle lives on device " + std::to_string(device) +
                                         " but the current device is " + std::to_string(cur) +
                                         " (cudaSetDevice to the handle's device before the call)");
  return SMPLB_OK;
}

template <typename T>
static cudaError_t upload(T** dptr, const std::vector<T>& h) {
  size_t bytes = h.size() * sizeof(T);
  cudaError_t e = cudaMalloc((void**)dptr, bytes ? bytes : sizeof(T));
  if (e != cudaSuccess) return e;
  if (bytes) e = cudaMemcpy(*dptr, h.data(), bytes, cudaMemcpyHostToDevice);
  return e;
}

int32_t build_tree(const int32_t* parents, Tree* t, const char* who) {
  t->max_depth = 0;
  t->parent[0] = 0;
  t->depth[0] = 0;
  for (int i = 0; i < NJ; ++i)
    for (int c = 0; c < 3; ++c) t->child[i][c] = -1;
  for (int i = 1; i < NJ; ++i) {
    const int p = parents[i];
    if (p < 0 || p >= i) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": parents[i] must be in [0,i)");
    t->parent[i] = p;
    t->depth[i] = t->depth[p] + 1;
    if (t->depth[i] > t->max_depth) t->max_depth = t->depth[i];
    int c = 0;
    while (c < 3 && t->child[p][c] >= 0) ++c;
    if (c == 3) return fail(SMPLB_E_UNSUPPORTED, std::string(who) + ": more than 3 children per joint");
    t->child[p][c] = i;
  }
  // depth-first pre-order (children in index order)
  {
    int stack[NJ], sp = 0, n = 0;
    stack[sp++] = 0;
    while (sp > 0) {
      const int i = stack[--sp];
      t->rank[i] = n;
      t->order[n++] = i;
      for (int c = 2; c >= 0; --c)
        if (t->child[i][c] >= 0) stack[sp++] = t->child[i][c];
    }
    int size[NJ];
    for (int i = NJ - 1; i >= 0; --i) {
      size[i] = 1;
      for (int c = 0; c < 3; ++c)
        if (t->child[i][c] >= 0) size[i] += size[t->child[i][c]];
    }
    for (int l = 0; l < NJ; ++l) {
      const int i = t->order[l];
      t->sub[l] = size[i];
      t->up[0][l] = i == 0 ? -1 : t->rank[t->parent[i]];
    }
    for (int r = 1; r < 5; ++r)
      for (int l = 0; l < NJ; ++l) {
        const int a = t->up[r - 1][l];
        t->up[r][l] = a < 0 ? -1 : t->up[r - 1][a];
      }
    for (int j = 0; j < NJ; ++j) t->upj[0][j] = j == 0 ? -1 : t->parent[j];
    for (int r = 1; r < 5; ++r)
      for (int j = 0; j < NJ; ++j) {
        const int a = t->upj[r - 1][j];
        t->upj[r][j] = a < 0 ? -1 : t->upj[r - 1][a];
      }
    t->n_rounds = 0;
    while ((1 << t->n_rounds) < t->max_depth + 1) ++t->n_rounds;
  }
  return SMPLB_OK;
}

}  // namespace smplb

extern "C" const char* smplb_last_error(void) { return smplb::last_error(); }
extern "C" const char* smplb_version(void) { return "smplb200 0.1 (sm_100a)"; }

extern "C" int32_t smplb_model_nk(const smplb_model* m) { return m ? m->NK : SMPLB_E_INVALID_ARG; }

extern "C" int32_t smplb_model_create(smplb_model** out, int32_t device, const float* v_template,
                                      const float* shapedirs, const float* posedirs,
                                      const float* J_regressor, const float* weights,
                                      const float* kp_regressor, int32_t NK,
                                      const int32_t* parents) {
  using namespace smplb;
  if (!out || !v_template || !shapedirs || !posedirs || !J_regressor || !weights || !kp_regressor ||
      !parents)
    return fail(SMPLB_E_INVALID_ARG, "smplb_model_create: null argument");
  if (NK < 1 || NK > NKMAX) return fail(SMPLB_E_BAD_SHAPE, "smplb_model_create: NK out of range");
  *out = nullptr;
  // The tensor-core kernels split their operands into an fp16 main part + bf16 residuals (common.cuh): model
  // constants must be finite and inside the fp16 range (SMPL's are metres and unit weights).
  {
    auto in_range = [](const float* p, size_t n) {
      for (size_t i = 0; i < n; ++i)
        if (!(std::fabs(p[i]) <= 65504.0f)) return false;   // also rejects NaN
      return true;
    };
    if (!in_range(shapedirs, (size_t)NB * NV3) || !in_range(posedirs, (size_t)NP * NV3) ||
        !in_range(weights, (size_t)NV * NJ))
      return fail(SMPLB_E_UNSUPPORTED, "smplb_model_create: shapedirs / posedirs / weights must be finite and |x| <= 65504");
  }

  // ---- kinematic tree (kintree_table[0]; parent[i] < i as in batch_lbs.py:122-127) ----
  Tree tree;
  {
    int32_t rc = build_tree(parents, &tree, "smplb_model_create");
    if (rc != SMPLB_OK) return rc;
  }

  // ---- compressed LBS weights (batch_smpl.py:57-60, :120-127) ----
  std::vector<uint32_t> sidx(NV);
  std::vector<float4> sw(NV);
  int has_compressed = 1;
  std::vector<float> Wimg((size_t)SK_NVT * SK_WTILE, 0.f);   // zero: joints 24..31 of the padded K stay 0
  unsigned short* w16 = reinterpret_cast<unsigned short*>(Wimg.data());
  for (int v = 0; v < NV; ++v)
    for (int j = 0; j < NJ; ++j) {
      const float x = weights[(size_t)v * NJ + j];
      const unsigned short h = f16_rne_bits(x);
      w16[wimg16_off(v, j, 0)] = h;
      w16[wimg16_off(v, j, 1)] = f16_rne_bits(x - f16_bits_to_float(h));
    }
  for (int v = 0; v < NV; ++v) {
    int n = 0;
    uint32_t packed = 0;
    float w4[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j = 0; j < NJ; ++j) {
      float w = weights[(size_t)v * NJ + j];
      if (w != 0.0f) {
        if (n == 4) {
          has_compressed = 0;
          break;
        }
        packed |= (uint32_t)j << (8 * n);
        w4[n] = w;
        ++n;
      }
    }
    sidx[v] = packed;
    sw[v] = make_float4(w4[0], w4[1], w4[2], w4[3]);
  }

  // ---- blend-shape basis D = [shapedirs; posedirs], padded; TF32 hi/lo splits ----
  std::vector<float> D32((size_t)KPAD * VPP, 0.f);
  // columns in the internal v_posed order (vp_col): reference column 3v+c -> vp_col(v, c)
  for (int v = 0; v < NV; ++v)
    for (int c = 0; c < 3; ++c) {
      const int col = vp_col(v, c);
      for (int k = 0; k < NB; ++k) D32[(size_t)k * VPP + col] = shapedirs[(size_t)k * NV3 + 3 * v + c];
      for (int k = 0; k < NP; ++k) D32[(size_t)(NB + k) * VPP + col] = posedirs[(size_t)k * NV3 + 3 * v + c];
    }
  // D is stored scaled by 2^e, the power of two that puts max|D| in [2^14, 2^15): both fp16 parts of every entry down to
  // 2^-17 of the largest are then normal numbers (common.cuh); the epilogue of the GEMM multiplies by 2^-e
  float dmax = 0.f;
  for (size_t i = 0; i < D32.size(); ++i) dmax = std::max(dmax, std::fabs(D32[i]));
  int e2 = 0;
  if (dmax > 0.f) {
    int ex;
    std::frexp(dmax, &ex);      // dmax = f * 2^ex, f in [0.5, 1)
    e2 = 15 - ex;               // 2^e2 * dmax in [2^14, 2^15)
  }
  const float d_scale = std::ldexp(1.0f, e2), d_scale_inv = std::ldexp(1.0f, -e2);
  std::vector<float> Dimg((size_t)BL_NT * BL_DTILE, 0.f);
  unsigned short* d16 = reinterpret_cast<unsigned short*>(Dimg.data());
  for (int k = 0; k < KDIM; ++k)
    for (int n = 0; n < VPP; ++n) {
      const float x = D32[(size_t)k * VPP + n] * d_scale;   // exact
      const unsigned short h = f16_rne_bits(x);
      const float hi = f16_bits_to_float(h);
      d16[dimg16_off(n, k, 0)] = h;
      d16[dimg16_off(n, k, 1)] = f16_rne_bits(x - hi);
    }
  std::vector<float> vt(VPP, 0.f);
  for (int v = 0; v < NV; ++v)
    for (int c = 0; c < 3; ++c) vt[vp_col(v, c)] = v_template[3 * v + c];

  // ---- folded joint regression (batch_smpl.py:91-101): J = J_reg^T (v_template + S beta) ----
  std::vector<double> Jbar64(NJ * 3, 0.0), JS64(NJ * 3 * NB, 0.0);
  for (int v = 0; v < NV; ++v)
    for (int j = 0; j < NJ; ++j) {
      double r = J_regressor[(size_t)v * NJ + j];
      if (r == 0.0) continue;
      for (int c = 0; c < 3; ++c) {
        Jbar64[j * 3 + c] += r * (double)v_template[v * 3 + c];
        for (int k = 0; k < NB; ++k)
          JS64[(j * 3 + c) * NB + k] += r * (double)shapedirs[(size_t)k * NV3 + v * 3 + c];
      }
    }
  std::vector<float> Jbar(NJ * 3), JS(NJ * 3 * NB);
  for (size_t i = 0; i < Jbar.size(); ++i) Jbar[i] = (float)Jbar64[i];
  for (size_t i = 0; i < JS.size(); ++i) JS[i] = (float)JS64[i];

  // ---- backward tables ----
  std::vector<int> eptr((size_t)SK_NVT * (NJ + 1), 0);
  std::vector<float2> ents;
  for (int t = 0; t < SK_NVT; ++t) {
    for (int j = 0; j < NJ; ++j) {
      eptr[(size_t)t * (NJ + 1) + j] = (int)ents.size();
      for (int l = 0; l < SK_VT; ++l) {
        const int v = t * SK_VT + l;
        if (v >= NV) break;
        const float w = weights[(size_t)v * NJ + j];
        if (w != 0.0f) {
          float2 e;
          e.x = w;
          memcpy(&e.y, &l, sizeof(int));
          ents.push_back(e);
        }
      }
    }
    eptr[(size_t)t * (NJ + 1) + NJ] = (int)ents.size();
  }
  std::vector<float> DTimg((size_t)BB_NKB * BB_DSTAGE / 2, 0.f);   // rows 217..223 stay zero
  unsigned short* dt16 = reinterpret_cast<unsigned short*>(DTimg.data());
  for (int k = 0; k < KDIM; ++k)
    for (int n = 0; n < VPP; ++n) {
      const float x = D32[(size_t)k * VPP + n];
      const unsigned short h = bf16_rne_bits(x);
      dt16[dtimg16_off(k, n, 0)] = h;
      dt16[dtimg16_off(k, n, 1)] = bf16_rne_bits(x - bf16_bits_to_float(h));
    }

  // ---- keypoint regressor CSR by keypoint (batch_smpl.py:63-69, :134-138) ----
  std::vector<int> kptr(NK + 1, 0), kvtx;
  std::vector<float> kval;
  for (int k = 0; k < NK; ++k) {
    for (int v = 0; v < NV; ++v) {
      const float r = kp_regressor[(size_t)v * NK + k];
      if (r != 0.0f) {
        kvtx.push_back(v);
        kval.push_back(r);
      }
    }
    kptr[k + 1] = (int)kvtx.size();
  }

  // compact keypoint-vertex rows: per vertex tile the touched vertices in ascending order, 3 floats each,
  // the run padded to a multiple of 8 floats (whole 32-byte sectors: no partial-sector writes between CTAs)
  std::vector<char> touched(NV, 0);
  for (int v : kvtx) touched[v] = 1;
  std::vector<int> ctile(SK_NVT + 1, 0), csrc, cslot_of_v(NV, -1), cslot(kvtx.size());
  for (int t = 0; t < SK_NVT; ++t) {
    for (int lv = 0; lv < SK_VT; ++lv) {
      const int v = t * SK_VT + lv;
      if (v >= NV || !touched[v]) continue;
      cslot_of_v[v] = (int)csrc.size();
      for (int c = 0; c < 3; ++c) csrc.push_back(3 * lv + c);
    }
    while (csrc.size() % 8) csrc.push_back(-1);
    ctile[t + 1] = (int)csrc.size();
  }
  if (csrc.empty()) csrc.assign(8, -1);
  for (size_t e = 0; e < kvtx.size(); ++e) cslot[e] = cslot_of_v[kvtx[e]];

  // ---- fused forward images (fused_fwd.cu): D rows and LBS weights per 64-vertex half tile, f16 main + f16 residual ----
  std::vector<float> Dimg2((size_t)FF_NHT * FF_DTILE / 2, 0.f), Wimg2((size_t)FF_NHT * FF_WHT / 2, 0.f);
  {
    unsigned short* d2 = reinterpret_cast<unsigned short*>(Dimg2.data());
    unsigned short* w2 = reinterpret_cast<unsigned short*>(Wimg2.data());
    for (int v = 0; v < NV; ++v) {
      for (int c = 0; c < 3; ++c)
        for (int k = 0; k < KDIM; ++k) {
          const float x = k < NB ? shapedirs[(size_t)k * NV3 + 3 * v + c] : posedirs[(size_t)(k - NB) * NV3 + 3 * v + c];
          const unsigned short h = f16_rne_bits(x);
          d2[ff_dimg_off(v, c, k, 0)] = h;
          d2[ff_dimg_off(v, c, k, 1)] = f16_rne_bits(x - f16_bits_to_float(h));
        }
      for (int j = 0; j < NJ; ++j) {
        const float x = weights[(size_t)v * NJ + j];
        const unsigned short h = f16_rne_bits(x);
        w2[ff_wimg_off(v, j, 0)] = h;
        w2[ff_wimg_off(v, j, 1)] = f16_rne_bits(x - f16_bits_to_float(h));
      }
    }
  }
  std::vector<unsigned short> Wimg3(wimg3_elems(), 0);
  build_wimg3(weights, Wimg3.data());
  std::vector<float> vt_plain(v_template, v_template + NV3);
  std::vector<int> cslot_of(NV3, -1);   // (vertex, component) -> compact keypoint slot
  for (int t = 0; t < SK_NVT; ++t)
    for (int sl = ctile[t]; sl < ctile[t + 1]; ++sl)
      if (csrc[sl] >= 0) cslot_of[(size_t)t * (3 * SK_VT) + csrc[sl]] = sl;

  int prev = 0;
  cudaGetDevice(&prev);
  SMPLB_CUDA_CHECK(cudaSetDevice(device));
  smplb_model* m = new smplb_model();
  memset(m, 0, sizeof(*m));
  m->device = device;
  m->NK = NK;
  m->tree = tree;
  m->d_scale_inv = d_scale_inv;
  m->kp_nnz = (int)kvtx.size();
  m->has_compressed = has_compressed;
  cudaError_t e = cudaSuccess;
  auto chk = [&](cudaError_t x) {
    if (e == cudaSuccess) e = x;
  };
  chk(upload(&m->v_template_pad, vt));
  chk(upload(&m->D32, D32));
  chk(upload(&m->Dimg, Dimg));
  chk(upload(&m->Jbar, Jbar));
  chk(upload(&m->JS, JS));
  {
    std::vector<float> JSt(33 * 32, 0.f);
    for (int j = 0; j < NJ; ++j)
      for (int c = 0; c < 3; ++c) {
        for (int k = 0; k < NB; ++k) JSt[(c * NB + k) * 32 + j] = JS[(j * 3 + c) * NB + k];
        JSt[(3 * NB + c) * 32 + j] = Jbar[j * 3 + c];
      }
    chk(upload(&m->JSt, JSt));
  }
  chk(upload(&m->skin_idx, sidx));
  chk(upload(&m->skin_w, sw));
  chk(upload(&m->kp_ptr, kptr));
  chk(upload(&m->kp_vtx, kvtx));
  chk(upload(&m->Wimg, Wimg));
  chk(upload(&m->ent_ptr, eptr));
  chk(upload(&m->ent, ents));
  chk(upload(&m->DTimg, DTimg));
  chk(upload(&m->kp_val, kval));
  m->kpv_pitch = ctile[SK_NVT];
  chk(upload(&m->kpc_tile_ptr, ctile));
  chk(upload(&m->kpc_src, csrc));
  chk(upload(&m->kpc_slot, cslot));
  {
    unsigned short* d = nullptr;
    chk(upload(&d, Wimg3));
    m->Wimg3 = reinterpret_cast<float*>(d);
  }
  chk(upload(&m->Dimg2, Dimg2));
  chk(upload(&m->Wimg2, Wimg2));
  chk(upload(&m->vt_plain, vt_plain));
  chk(upload(&m->kpc_slot_of, cslot_of));
  {
    std::vector<int4> tab(128);
    for (int l = 0; l < 32; ++l) {
      const bool a = l < NJ;
      tab[l * 2] = make_int4(a ? tree.order[l] : 0, a ? tree.sub[l] : 1, a ? tree.up[0][l] : -1, a ? tree.up[1][l] : -1);
      tab[l * 2 + 1] = make_int4(a ? tree.up[2][l] : -1, a ? tree.up[3][l] : -1, a ? tree.up[4][l] : -1, 0);
      tab[64 + l * 2] = make_int4(a ? l : 0, 1, a ? tree.upj[0][l] : -1, a ? tree.upj[1][l] : -1);
      tab[64 + l * 2 + 1] = make_int4(a ? tree.upj[2][l] : -1, a ? tree.upj[3][l] : -1, a ? tree.upj[4][l] : -1, 0);
    }
    chk(upload(&m->tree_tab, tab));
  }
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    smplb_model_destroy(m);
    return fail(SMPLB_E_CUDA, std::string("smplb_model_create: ") + cudaGetErrorString(e));
  }
  *out = m;
  return SMPLB_OK;
}

extern "C" int32_t smplb_model_destroy(smplb_model* m) {
  if (!m) return SMPLB_OK;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  cudaFree(m->v_template_pad);
  cudaFree(m->D32);
  cudaFree(m->Dimg);
  cudaFree(m->Jbar);
  cudaFree(m->JS);
  cudaFree(m->JSt);
  cudaFree(m->tree_tab);
  cudaFree(m->skin_idx);
  cudaFree(m->skin_w);
  cudaFree(m->kp_ptr);
  cudaFree(m->kp_vtx);
  cudaFree(m->Wimg);
  cudaFree(m->ent_ptr);
  cudaFree(m->ent);
  cudaFree(m->DTimg);
  cudaFree(m->kp_val);
  cudaFree(m->kpc_tile_ptr);
  cudaFree(m->kpc_src);
  cudaFree(m->kpc_slot);
  cudaFree(m->Wimg3);
  cudaFree(m->Dimg2);
  cudaFree(m->Wimg2);
  cudaFree(m->vt_plain);
  cudaFree(m->kpc_slot_of);
  cudaSetDevice(prev);
  delete m;
  return SMPLB_OK;
}
