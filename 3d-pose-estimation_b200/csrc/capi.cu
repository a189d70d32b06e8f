// C ABI orchestration: workspace carving and the forward call (pose_prep -> blend GEMM -> skinning).
#include <vector>

#include "common.cuh"

namespace smplb {

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

Workspace carve(void* base, int B, int flags, int kpv_pitch) {
  Workspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* p = base ? (char*)base + off : nullptr;
    off += align_up(bytes);
    return (float*)p;
  };
  const size_t b = (size_t)(B > 0 ? B : 0);
  w.status = (int*)take(256);   // input-range flags of the last forward call (smplb_forward_status)
  w.Ximg = take(((b + BL_BM - 1) / BL_BM) * BL_XTILE * sizeof(float));
  w.A = take(b * NJ * 12 * sizeof(float));
  w.joints = take(b * NJ * 3 * sizeof(float));
  const bool skin = !(flags & SMPLB_F_NO_SKIN);
  w.Aimg = skin ? take(((b + SK_S - 1) / SK_S) * SK_AGRP * sizeof(float)) : nullptr;
  w.vposed = skin ? take(b * VPP * sizeof(float)) : nullptr;
  w.verts = (skin && (flags & SMPLB_F_SCRATCH_VERTS)) ? take(b * NV3 * sizeof(float)) : nullptr;
  w.bwd = (flags & SMPLB_F_KEEP_VPOSED) ? take(backward_scratch_bytes(B)) : nullptr;
  w.kpv = (skin && kpv_pitch > 0) ? take(((b + SK_S - 1) / SK_S) * SK_S * (size_t)kpv_pitch * sizeof(float)) : nullptr;
  const bool fused = skin && (flags & SMPLB_F_FUSED);
  const size_t n_ch = (b + FF_CH - 1) / FF_CH;
  w.Ximg2 = fused ? take(n_ch * 2 * FF_XHALF * 2) : nullptr;
  w.Aimg2 = fused ? take(n_ch * FF_NG * 2 * FF_AGH * 2) : nullptr;
  w.total = off;
  return w;
}

// ---- optional per-kernel timing (bench.py rooflines): events recorded on the launching stream ----
struct TimingState {
  std::vector<cudaEvent_t> ev;   // 5 events per call: 4 intervals
  int max_calls = 0;
  int ncalls = 0;
  ~TimingState() {
    for (cudaEvent_t e : ev) cudaEventDestroy(e);   // thread exit
  }
};
static thread_local TimingState g_timing;       // forward: prep | blend | skin | keypoints
static thread_local TimingState g_timing_bwd;   // backward (vertex path): g_A reduction | g_vposed | transposed GEMM | joint path

static inline void timing_mark_in(TimingState& t, int slot, cudaStream_t st) {
  if (t.max_calls == 0 || t.ncalls >= t.max_calls) return;
  cudaEventRecord(t.ev[(size_t)t.ncalls * 5 + slot], st);
  if (slot == 4) ++t.ncalls;
}
static inline void timing_mark(int slot, cudaStream_t st) { timing_mark_in(g_timing, slot, st); }
void timing_mark_bwd(int slot, cudaStream_t st) { timing_mark_in(g_timing_bwd, slot, st); }

static int32_t timing_begin_in(TimingState& t, int32_t max_calls) {
  for (cudaEvent_t e : t.ev) cudaEventDestroy(e);
  t.ev.clear();
  t.ncalls = 0;
  t.max_calls = max_calls > 0 ? max_calls : 0;
  t.ev.resize((size_t)t.max_calls * 5);
  for (auto& e : t.ev) SMPLB_CUDA_CHECK(cudaEventCreate(&e));
  return SMPLB_OK;
}

static int32_t timing_end_in(TimingState& t, float* ms_out, int32_t* ncalls_out, const char* who) {
  if (!ncalls_out) return fail(SMPLB_E_INVALID_ARG, std::string(who) + ": null ncalls_out");
  *ncalls_out = t.ncalls;
  for (int i = 0; i < t.ncalls && ms_out; ++i) {
    SMPLB_CUDA_CHECK(cudaEventSynchronize(t.ev[(size_t)i * 5 + 4]));
    for (int k = 0; k < 4; ++k)
      SMPLB_CUDA_CHECK(cudaEventElapsedTime(&ms_out[i * 4 + k], t.ev[(size_t)i * 5 + k], t.ev[(size_t)i * 5 + k + 1]));
  }
  for (cudaEvent_t e : t.ev) cudaEventDestroy(e);
  t.ev.clear();
  t.max_calls = 0;
  t.ncalls = 0;
  return SMPLB_OK;
}

}  // namespace smplb

using namespace smplb;

extern "C" int32_t smplb_timing_begin(int32_t max_calls) {
  int32_t rc = timing_begin_in(g_timing, max_calls);
  return rc != SMPLB_OK ? rc : timing_begin_in(g_timing_bwd, max_calls);
}

extern "C" int32_t smplb_timing_end(float* ms_out, int32_t* ncalls_out) {
  return timing_end_in(g_timing, ms_out, ncalls_out, "smplb_timing_end");
}

extern "C" int32_t smplb_timing_bwd_end(float* ms_out, int32_t* ncalls_out) {
  return timing_end_in(g_timing_bwd, ms_out, ncalls_out, "smplb_timing_bwd_end");
}

extern "C" size_t smplb_workspace_bytes(const smplb_model* m, int32_t B, int32_t flags) {
  return carve(nullptr, B, flags, m ? m->kpv_pitch : 0).total;
}

static int32_t check_forward(const smplb_model* m, const smplb_forward_io* io, void* ws, size_t ws_bytes,
                             int32_t flags) {
  if (!m || !io) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: null model or io");
  if (int32_t rc = check_current_device(m->device, "smplb_forward")) return rc;
  if (io->B < 0) return fail(SMPLB_E_BAD_SHAPE, "smplb_forward: negative batch");
  if (io->B == 0) return SMPLB_OK;
  if (!io->beta || !io->theta) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: beta/theta are required");
  if (io->cam_pos && !io->cam_rot) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: cam_rot required with cam_pos");
  if ((io->uv || io->px) && !io->cam_pos) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: uv/px need a camera");
  if ((flags & SMPLB_F_NO_SKIN) && (io->verts || io->kp))
    return fail(SMPLB_E_INVALID_ARG, "smplb_forward: SMPLB_F_NO_SKIN excludes verts/kp outputs");
  if (io->kp && !io->verts && (flags & SMPLB_F_SKIN_SIMT) && !(flags & SMPLB_F_SCRATCH_VERTS))
    return fail(SMPLB_E_INVALID_ARG,
                "smplb_forward: keypoints without a verts buffer need SMPLB_F_SCRATCH_VERTS on the SIMT skinning path");
  if ((flags & SMPLB_F_FUSED) && (flags & (SMPLB_F_KEEP_VPOSED | SMPLB_F_BLEND_SIMT | SMPLB_F_SKIN_SIMT)))
    return fail(SMPLB_E_INVALID_ARG, "smplb_forward: SMPLB_F_FUSED excludes KEEP_VPOSED / BLEND_SIMT / SKIN_SIMT");
  if (!ws) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: workspace is required");
  if (ws_bytes < smplb_workspace_bytes(m, io->B, flags))
    return fail(SMPLB_E_INVALID_ARG, "smplb_forward: workspace too small (see smplb_workspace_bytes)");
  if (((uintptr_t)ws & 255) != 0) return fail(SMPLB_E_INVALID_ARG, "smplb_forward: workspace must be 256-byte aligned");
  return SMPLB_OK;
}

extern "C" int32_t smplb_forward(const smplb_model* m, const smplb_forward_io* io, void* workspace,
                                 size_t workspace_bytes, int32_t flags, void* stream) {
  int32_t rc = check_forward(m, io, workspace, workspace_bytes, flags);
  if (rc != SMPLB_OK || io->B == 0) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int B = io->B;
  Workspace w = carve(workspace, B, flags, m->kpv_pitch);
  const bool skin = !(flags & SMPLB_F_NO_SKIN) && (io->verts || io->kp || (flags & SMPLB_F_KEEP_VPOSED));

  PrepParams p;
  p.B = B;
  p.beta = io->beta;
  p.theta = io->theta;
  p.Jbar = m->Jbar;
  p.JS = m->JS;
  p.JSt = m->JSt;
  const bool fused = skin && (flags & SMPLB_F_FUSED);
  p.Ximg = (skin && !fused) ? w.Ximg : nullptr;
  p.Ximg2 = fused ? w.Ximg2 : nullptr;
  p.Aimg2 = fused ? w.Aimg2 : nullptr;
  p.joints = io->joints ? io->joints : w.joints;
  p.Rs = io->Rs;
  p.A = io->A ? io->A : w.A;
  p.Aimg = (skin && !fused) ? w.Aimg : nullptr;
  p.cam_pos = io->cam_pos;
  p.cam_rot = io->cam_rot;
  p.focal = io->focal;
  p.uv = io->uv;
  p.px = io->px;
  p.cfg = io->cfg;
  p.tree_tab = m->tree_tab;
  p.n_rounds = m->tree.n_rounds;
  p.status = w.status;
  SMPLB_CUDA_CHECK(cudaMemsetAsync(w.status, 0, sizeof(int), st));
  timing_mark(0, st);
  rc = launch_pose_prep(p, st);
  if (rc != SMPLB_OK) return rc;
  timing_mark(1, st);
  if (!skin) {
    timing_mark(2, st);
    timing_mark(3, st);
    timing_mark(4, st);
    return SMPLB_OK;
  }

  if (fused) {
    // one kernel: v_posed = [beta|phi].D + v_template lives in TMEM and meets T = W.A there (fused_fwd.cu)
    timing_mark(2, st);   // the "blend" interval is empty on this path: everything is counted as skinning
    rc = launch_fused_fwd(m, B, w.Ximg2, w.Aimg2, io->verts, io->kp ? w.kpv : nullptr, st);
    if (rc != SMPLB_OK) return rc;
    timing_mark(3, st);
    if (io->kp) rc = launch_keypoints_compact(m, B, w.kpv, io->kp, st);
    timing_mark(4, st);
    return rc;
  }
  if (flags & SMPLB_F_BLEND_SIMT)
    rc = launch_blend_simt(m, B, w.Ximg, w.vposed, st);
  else
    rc = launch_blend_tc(m, B, w.Ximg, w.vposed, st);
  if (rc != SMPLB_OK) return rc;
  timing_mark(2, st);
  // tensor-core path: the skinning epilogue also emits the compact keypoint-vertex rows, so keypoints need
  // neither a second pass over verts nor (when the caller wants no verts) a verts buffer at all
  const bool compact = io->kp && !(flags & SMPLB_F_SKIN_SIMT);
  float* verts = io->verts ? io->verts : (compact ? nullptr : w.verts);
  if (flags & SMPLB_F_SKIN_SIMT) {
    if (verts) rc = launch_skin_simt(m, B, w.vposed, p.A, verts, st);
  } else {
    rc = launch_skin_tc(m, B, w.vposed, w.Aimg, verts, compact ? w.kpv : nullptr, st);
  }
  if (rc != SMPLB_OK) return rc;
  timing_mark(3, st);
  if (io->kp) rc = compact ? launch_keypoints_compact(m, B, w.kpv, io->kp, st) : launch_keypoints(m, B, verts, io->kp, st);
  timing_mark(4, st);
  return rc;
}

extern "C" int32_t smplb_forward_status(const void* workspace, void* stream) {
  if (!workspace) return fail(SMPLB_E_INVALID_ARG, "smplb_forward_status: null workspace");
  int flags = 0;
  SMPLB_CUDA_CHECK(cudaMemcpyAsync(&flags, workspace, sizeof(int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  SMPLB_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  if (flags & 1)
    return fail(SMPLB_E_NOT_FINITE, "smplb_forward: a beta entry is not finite or exceeds 65504 in magnitude (the fp16 main "
                                    "part of the split-operand GEMM cannot carry it); the outputs of that call are invalid");
  if (flags & 2) return fail(SMPLB_E_NOT_FINITE, "smplb_forward: a theta entry is not finite; the outputs of that call are invalid");
  return SMPLB_OK;
}

extern "C" int32_t smplb_forward_host(const smplb_model* m, const smplb_forward_io* h,
                                      const smplb_forward_io* d, void* workspace, size_t workspace_bytes,
                                      int32_t flags, void* stream) {
  if (!h || !d) return fail(SMPLB_E_INVALID_ARG, "smplb_forward_host: null io");
  if (h->B != d->B) return fail(SMPLB_E_BAD_SHAPE, "smplb_forward_host: host/device batch mismatch");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t B = (size_t)(h->B > 0 ? h->B : 0);
  const int NK = m ? m->NK : 0;
  auto h2d = [&](const float* dst, const float* src, size_t n) -> cudaError_t {
    if (!src) return cudaSuccess;
    if (!dst) return cudaErrorInvalidValue;
    return cudaMemcpyAsync((void*)dst, src, n * sizeof(float), cudaMemcpyHostToDevice, st);
  };
  SMPLB_CUDA_CHECK(h2d(d->beta, h->beta, B * NB));
  SMPLB_CUDA_CHECK(h2d(d->theta, h->theta, B * NJ * 3));
  SMPLB_CUDA_CHECK(h2d(d->cam_pos, h->cam_pos, B * 3));
  SMPLB_CUDA_CHECK(h2d(d->cam_rot, h->cam_rot, B * 3));
  SMPLB_CUDA_CHECK(h2d(d->focal, h->focal, B));
  smplb_forward_io io = *d;
  // outputs the host did not ask for are not produced
  if (!h->verts) io.verts = nullptr;
  if (!h->kp) io.kp = nullptr;
  if (!h->Rs) io.Rs = nullptr;
  if (!h->uv) io.uv = nullptr;
  if (!h->px) io.px = nullptr;
  if (!h->cam_pos) { io.cam_pos = nullptr; io.cam_rot = nullptr; io.focal = nullptr; io.uv = nullptr; io.px = nullptr; }
  if (!h->focal) io.focal = nullptr;
  io.cfg = h->cfg;
  // Chunked pipeline: the kernels of chunk k+1 run on `st` while a second stream drains the outputs of chunk k
  // over PCIe (the D2H of verts is ~20x the kernel time, so everything but the first chunk's kernels is hidden).
  // per host thread and device: created on first use, destroyed when the thread exits (calls that arrive after the CUDA
  // runtime has shut down return an error, which is ignored there)
  struct HostPipe {
    cudaStream_t cs[64] = {};
    cudaEvent_t ev[64] = {}, ev_done[64] = {};
    ~HostPipe() {
      for (int i = 0; i < 64; ++i)
        if (cs[i]) {
          cudaStreamDestroy(cs[i]);
          cudaEventDestroy(ev[i]);
          cudaEventDestroy(ev_done[i]);
        }
    }
  };
  static thread_local HostPipe pipe;
  cudaStream_t(&cs)[64] = pipe.cs;
  cudaEvent_t(&ev)[64] = pipe.ev;
  cudaEvent_t(&ev_done)[64] = pipe.ev_done;
  const int dv = m ? (m->device & 63) : 0;
  if (!cs[dv]) {
    SMPLB_CUDA_CHECK(cudaStreamCreateWithFlags(&cs[dv], cudaStreamNonBlocking));
    SMPLB_CUDA_CHECK(cudaEventCreateWithFlags(&ev[dv], cudaEventDisableTiming));
    SMPLB_CUDA_CHECK(cudaEventCreateWithFlags(&ev_done[dv], cudaEventDisableTiming));
  }
  auto d2h = [&](float* dst, const float* src, size_t n) -> cudaError_t {
    if (!dst) return cudaSuccess;
    if (!src) return cudaErrorInvalidValue;
    return cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToHost, cs[dv]);
  };
  const size_t CH = 8192;
  auto off = [](const float* p, size_t n) { return p ? p + n : nullptr; };
  auto offm = [](float* p, size_t n) { return p ? p + n : nullptr; };
  for (size_t b0 = 0; b0 < B; b0 += CH) {
    const size_t nb = B - b0 < CH ? B - b0 : CH;
    smplb_forward_io c = io;
    c.B = (int32_t)nb;
    c.beta = off(io.beta, b0 * NB); c.theta = off(io.theta, b0 * NJ * 3);
    c.cam_pos = off(io.cam_pos, b0 * 3); c.cam_rot = off(io.cam_rot, b0 * 3); c.focal = off(io.focal, b0);
    c.verts = offm(io.verts, b0 * NV3); c.joints = offm(io.joints, b0 * NJ * 3); c.kp = offm(io.kp, b0 * NK * 3);
    c.Rs = offm(io.Rs, b0 * NJ * 9); c.A = offm(io.A, b0 * NJ * 12); c.uv = offm(io.uv, b0 * NJ * 2);
    c.px = offm(io.px, b0 * NJ * 2);
    int32_t rc = smplb_forward(m, &c, workspace, workspace_bytes, flags, stream);
    if (rc != SMPLB_OK) return rc;
    SMPLB_CUDA_CHECK(cudaEventRecord(ev[dv], st));
    SMPLB_CUDA_CHECK(cudaStreamWaitEvent(cs[dv], ev[dv], 0));
    SMPLB_CUDA_CHECK(d2h(offm(h->verts, b0 * NV3), c.verts, nb * NV3));
    SMPLB_CUDA_CHECK(d2h(offm(h->joints, b0 * NJ * 3), c.joints, nb * NJ * 3));
    SMPLB_CUDA_CHECK(d2h(offm(h->kp, b0 * NK * 3), c.kp, nb * NK * 3));
    SMPLB_CUDA_CHECK(d2h(offm(h->Rs, b0 * NJ * 9), c.Rs, nb * NJ * 9));
    SMPLB_CUDA_CHECK(d2h(offm(h->A, b0 * NJ * 12), c.A, nb * NJ * 12));
    SMPLB_CUDA_CHECK(d2h(offm(h->uv, b0 * NJ * 2), c.uv, nb * NJ * 2));
    SMPLB_CUDA_CHECK(d2h(offm(h->px, b0 * NJ * 2), c.px, nb * NJ * 2));
  }
  SMPLB_CUDA_CHECK(cudaEventRecord(ev_done[dv], cs[dv]));
  SMPLB_CUDA_CHECK(cudaStreamWaitEvent(st, ev_done[dv], 0));
  SMPLB_CUDA_CHECK(cudaStreamSynchronize(st));
  return SMPLB_OK;
}
