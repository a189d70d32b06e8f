#!/usr/bin/env python
"""bench.py — SMPL forward + projection throughput on B200 (one JSON line on rank 0).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (pose_prep -> blend GEMM -> skinning, with the fused
projection of the 24 posed joints) over one batch of synthetic samples, inputs resident in HBM.
Default workload = BASELINE.json configs[3] as written: a GLOBAL batch of 65536 samples sharded over
the N GPUs (rank r owns samples [r*B/N, (r+1)*B/N), smplb200.dist.shard_range; strong scaling, no
collective on the forward path).  At N > 1 the 4-kernel step is replayed from one CUDA graph (the
8192-sample shard of N = 8 takes ~0.5 ms); per-kernel times come from an eager pass with event hooks
right after the timed region.  `--scaling weak` keeps --batch samples PER GPU (the round-1 protocol).
`e2e` times the same step through the C ABI with HOST buffers (smplb_forward_host: H2D of
beta/theta/camera from pinned memory, kernels, D2H of verts/joints/keypoints/2D keypoints).
`--impl reference` times the reference's CPU op sequence (oracle/ref_torch.py port of the TF graph,
all host threads) on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, '3d-pose-estimation_b200')
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = 'smpl_fwd_proj_samples_per_s'
UNIT = 'samples/s'
# Algorithmic bytes per sample of the graded skinning pass (SURVEY.md §8d / BASELINE.md §3):
# read v_posed 82680 + write verts 82680 + read A 1152  (keypoints: separate gather kernel; px: pose_prep)
SKIN_BYTES_PER_SAMPLE = 82680 + 82680 + 1152


def peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            d = json.load(f)
        return float(d['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def tensor_peaks():
    """bf16 dense burst / sustained TFLOP/s (MEASURED_PEAKS.json, else the recipe's fallback) and the TF32 rate
    measured on this pool by tools/measure_tf32_peak.py (profiles/r02_tf32_peak.json)."""
    out = {'bf16_burst': 1590.0, 'bf16_sustained': 1400.0, 'tf32_burst': None, 'source': 'fallback (B200_PROFILING.md)'}
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            d = json.load(f)
        out.update(bf16_burst=float(d['bf16_tflops']), bf16_sustained=float(d['bf16_tflops_sustained']),
                   source='measured (MEASURED_PEAKS.json)')
    except Exception:
        pass
    try:
        with open(os.path.join(ROOT, 'profiles', 'r02_tf32_peak.json')) as f:
            d = json.load(f)
        out['tf32_burst'], out['tf32_sustained'] = float(d['tf32_tflops']), float(d['tf32_tflops_sustained'])
    except Exception:
        pass
    return out


def blob_sha(path):
    data = open(path, 'rb').read()
    return hashlib.sha1(b'blob %d\0' % len(data) + data).hexdigest()


def ncu_record(name, batch):
    """A committed ncu record (tools/ncu_to_profiles.py) is only used while the kernel sources it was captured from
    are unchanged (git blob hashes) and the batch matches; else None."""
    try:
        with open(os.path.join(ROOT, 'profiles', name)) as f:
            rec = json.load(f)
        if int(rec.get('batch', -1)) != batch:
            return None
        for src, sha in rec['sources'].items():
            if blob_sha(os.path.join(PKG, 'csrc', src)) != sha:
                return None
        return rec
    except Exception:
        return None


DTYPE = 'f32 (fp16/bf16 split-operand tensor products, fp32 accumulate)'
# tensor work ISSUED by blend_tc_kernel per sample: 3 kind::f16 products over the padded 224 x 20736 operand images
BLEND_FLOP_ISSUED = 3 * 2 * 224 * 20736
BLEND_FLOP_FP32EQ = 2 * 217 * 20670


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {'hw_slowdown': 0x8, 'sw_power_cap': 0x4, 'hw_thermal_slowdown': 0x40,
                 'sw_thermal_slowdown': 0x20, 'hw_power_brake': 0x80}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        if not self.sm:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons)}
        return {'sm_mhz': float(np.median(self.sm)), 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.sm)}


def cpu_reference_rate(B_chunk, min_seconds, max_chunks, seed=1000):
    """Reference CPU op sequence (tile(W) + [N,6890,4,4] matmul included, batch_smpl.py:120-132)
    + projection + pixel rescale, fp32, all host threads.  Returns (samples/s, cores, n_samples, s)."""
    from oracle import ref_torch as rt          # timed CPU baseline only
    from smplb200 import synthetic as syn
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dd = syn.make_synthetic_smpl(0)
    m = rt.load_model(dd, dtype=torch.float32)
    inp = syn.make_inputs(B_chunk, seed=seed)
    t = {k: torch.from_numpy(v) for k, v in inp.items()}

    def step():
        with torch.no_grad():
            o = rt.smpl_forward(m, t['beta'], t['theta'], tile_W=True)
            rt.project_joints_px(o['J_transformed'], t['cam_pos'], t['cam_rot'], t['focal'])

    step()  # warm-up
    n, t0 = 0, time.perf_counter()
    while True:
        step()
        n += 1
        el = time.perf_counter() - t0
        if (el >= min_seconds and n >= 2) or n >= max_chunks:
            break
    return n * B_chunk / el, cores, n * B_chunk, el, step


def cpu_tight_rate(min_seconds, B_chunk=2048, seed=1000):
    """the tight C + OpenMP version of the same workload (oracle/smpl_omp.c; blend as one tiled matrix
    with 8 samples in the vector lanes, sparse skinning), all host threads.  A fairer CPU number than the TF op sequence;
    reported beside it, never the thing shipped."""
    import numpy as np
    from oracle import ref_numpy as rn          # timed CPU baseline only
    from oracle import smpl_omp
    from smplb200 import synthetic as syn
    tm = smpl_omp.TightModel(rn.load_model(syn.make_synthetic_smpl(0), dtype=np.float32))
    inp = syn.make_inputs(B_chunk, seed=seed)
    a = (inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], rn.CFG.ss, 320.0, 240.0)
    o = tm.forward(*a)
    n, t0 = 0, time.perf_counter()
    while True:
        o = tm.forward(*a, out=o)
        n += 1
        el = time.perf_counter() - t0
        if el >= min_seconds:
            break
    return {'value': n * B_chunk / el, 'unit': UNIT, 'cores': int(o['threads']), 'kind': 'port (tight C + OpenMP, AVX2)',
            'sample': '%d samples (%d-sample chunks, %.1f s) of the same workload: verts, joints, key points, px' % (
                n * B_chunk, B_chunk, el)}


def tensorflow_probe():
    """the reference on real TensorFlow, if the box has it."""
    try:
        import tensorflow as tf                  # noqa: F401
        return 'importable (%s) -- not wired: the reference needs TF1 graph mode' % getattr(tf, '__version__', '?')
    except Exception as e:                       # ImportError on this image
        return 'not importable on this box (%s)' % type(e).__name__


def workload_config(args, world):
    """config block shared by both arms (the driver compares them)."""
    if args.scaling == 'strong':
        wl = ('SMPL forward + 24-joint perspective projection, global batch %d sharded over %d GPU(s) '
              '(BASELINE configs[3]: rank r owns samples [r*B/N, (r+1)*B/N))' % (args.batch, world))
    else:
        wl = ('SMPL forward + 24-joint perspective projection, %d samples per GPU on %d GPU(s) (weak scaling, global '
              'batch %d)' % (args.batch, world, args.batch * world))
    return {'workload': wl,
            'model': 'synthetic SMPL-shaped (6890 verts, 24 joints, 10 betas, <=4 LBS weights/vertex), seed 0',
            'outputs': 'verts[B,6890,3], J_transformed[B,24,3], keypoints[B,19,3], Rs, uv, px[B,24,2]'}


def run_reference(args, rank, world):
    if rank != 0:
        return
    rate, cores, ns, el, step = cpu_reference_rate(args.ref_chunk, 0.0, 1)
    for _ in range(max(args.warmup - 1, 0)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    el = time.perf_counter() - t0
    val = args.steps * args.ref_chunk / el
    sample = '%d steps x %d samples of the same workload, fp32, torch-CPU port of the TF op sequence incl. tile(W)' % (
        args.steps, args.ref_chunk)
    cfg = workload_config(args, world)
    cfg['cpu_sample'] = 'bounded sample: %d samples per step' % args.ref_chunk
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * el / args.steps,
        'higher_is_better': True, 'scaling': args.scaling, 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic', 'config': cfg,
        'cpu_baseline': {'value': val, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': val, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def run_fit(args, rank, local_rank, world, dist):
    """BASELINE config 3: fitting on the 2D re-projection loss, B samples per GPU, one Adam step per bench step;
    the loss scalars are all-reduced over NCCL every step (the only collective of the path)."""
    import smplb200
    from smplb200 import synthetic as syn, dist as sd
    from smplb200.fit import Fitter
    dev = torch.device('cuda', local_rank)
    B = args.batch if args.batch != 65536 else 4096
    model = smplb200.Model(syn.make_synthetic_smpl(0), device=local_rank)
    inp = syn.make_inputs(B, seed=1000 + rank)
    tgt = syn.make_inputs(B, seed=5000 + rank)
    mesh = args.loss == 'mesh'
    if mesh:
        # the 3D-loss variant (pose_model_3d.py:183-192): drives the backward through the vertices
        from smplb200.fit import MeshFitter
        o = model.forward(inp['beta'], tgt['theta'], want_kp=False, want_Rs=False)
        f = MeshFitter(model, inp['beta'], inp['theta'], o['verts'].clone(), o['joints'].clone(), lr=2e-4,
                       n_total_verts=B * world * 6890 * 3, n_total_joints=B * world * 72)
    else:
        gt_px = model.forward(inp['beta'], tgt['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=4)['px'].clone()
        f = Fitter(model, inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], gt_px, focal=inp['focal'], lr=2e-4,
                   n_total=B * world * 48)
    # The loss scalar is only logged: with --allreduce-every k the k per-step values are written into a
    # [k,4] device log by the fit kernels themselves (smplb_fit_io.loss may point anywhere) and summed over the ranks
    # by ONE NCCL all-reduce per k steps, instead of one 25-35 us collective serialised into every 44 us step.
    k = max(1, args.allreduce_every)
    use_graph = bool(args.graph)
    log = torch.zeros((k, 4), device=dev)

    def step_block():
        for i in range(k):
            if mesh:
                l = f.step()
                log[i, :1].copy_(l[:1])
            else:
                f.step(loss_out=log[i])
        sd.allreduce_scalars(log)          # columns 0 (loss share) sum to the global mean; the others are per rank

    n_blocks = (args.steps + k - 1) // k
    steps = n_blocks * k
    for _ in range(max(1, (args.warmup + k - 1) // k)):
        step_block()
    first = float(log[0, 0])
    graph = None
    if use_graph:
        # k steps and their one collective replayed from one CUDA graph; Adam's step counter lives on the device, so
        # every replay advances the bias correction
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            step_block()        # the collective once on the capture stream before capturing it
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=s):
                step_block()
        torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_blocks):
        graph.replay() if graph is not None else step_block()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(t[0])
        print(json.dumps({
            'metric': 'smpl_fit_sample_steps_per_s', 'value': world * B * steps / (ms * 1e-3), 'unit': 'sample-steps/s',
            'n_gpus': world, 'steps': steps, 'warmup': args.warmup, 'ms_per_step': ms / steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32' if not mesh else DTYPE,
            'data': 'synthetic',
            'config': {'workload': ('SMPL fitting, fwd+bwd of the %s + TF1 Adam on theta, B=%d per GPU (BASELINE configs[2])'
                                    % ('3D mesh + joint loss (backward through the vertices)' if mesh else
                                       '2D re-projection loss', B)), 'cuda_graph': graph is not None,
                       'collective': ('NCCL all-reduce of the [%d,4] loss log once every %d step(s)' % (k, k)) if world > 1 else 'none',
                       'adam_step_counter': 'device (advances on graph replays)'},
            'gpu_launches': (15 if mesh else 2) * steps, 'loss_first': first, 'loss_last': float(log[k - 1, 0]),
            'adam_steps_done': int(f.step_dev)}),
            flush=True)
    if graph is not None:
        # a captured NCCL collective keeps the communicator busy: drop the graph before tearing the group down
        # (destroy_process_group otherwise waits forever)
        graph.reset()
        graph = None
        torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_sweep(args, rank, local_rank, world, dist):
    """BASELINE config 5: batch-size sweep 1..65536, forward (+projection) and forward+backward (mesh + joint loss
    gradients through the vertex path), per-kernel times and roofline fractions, at 1/2/4/8 GPUs.  Under --scaling
    strong (default) B is the GLOBAL batch and rank r runs its shard_range slice (ranks without samples idle); under
    weak every rank runs B samples.  No collective; the time is the max over ranks."""
    import smplb200
    from smplb200 import synthetic as syn, _lib, dist as sd
    dev = torch.device('cuda', local_rank)
    model = smplb200.Model(syn.make_synthetic_smpl(0), device=local_rank)
    L = _lib.lib()
    peak, peak_src = peaks()
    tp = tensor_peaks()
    rows = []
    Bs = [1 << i for i in range(0, 17)]
    Bs = [b for b in Bs if args.sweep_min <= b <= args.batch]
    for Bg in Bs:
        if args.scaling == 'strong':
            b0, b1 = sd.shard_range(Bg, rank, world)
            B, B_global = b1 - b0, Bg
        else:
            B, B_global = Bg, Bg * world
        steps = max(3, min(args.steps, int(2e6 / max(B, 1)) + 3))
        res = {'B_global': B_global, 'B_per_gpu_max': (B_global + world - 1) // world if args.scaling == 'strong' else Bg,
               'steps': steps}
        if B > 0:
            inp = syn.make_inputs(B, seed=1000 + rank)
            d = {k: torch.from_numpy(v).to(dev) for k, v in inp.items()}
            out = {}
            g_verts = torch.randn((B, 6890, 3), device=dev) * (1.0 / (B_global * 20670))
            g_joints = torch.randn((B, 24, 3), device=dev) * (1.0 / (B_global * 72))

        def fwd(flags=0):
            model.forward(d['beta'], d['theta'], d['cam_pos'], d['cam_rot'], d['focal'], flags=flags, out=out)

        def fwdbwd():
            fwd(_lib.F_KEEP_VPOSED)
            model.backward(d['beta'], d['theta'], g_verts=g_verts, g_joints=g_joints)

        for name, fn in (('fwd', fwd), ('fwd_bwd', fwdbwd)):
            for _ in range(3 if B > 0 else 0):
                fn()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            kt = None
            ms_local = 0.0
            kb = None
            if B > 0:
                L.smplb_timing_begin(steps)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(steps):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                kms = (C.c_float * (4 * steps))()
                nc = C.c_int32(0)
                L.smplb_timing_end(kms, C.byref(nc))
                if name == 'fwd':
                    kt = np.array(list(kms), dtype=np.float64).reshape(-1, 4)[:nc.value].mean(axis=0)
                L.smplb_timing_bwd_end(kms, C.byref(nc))
                if name == 'fwd_bwd' and nc.value:
                    kb = np.array(list(kms), dtype=np.float64).reshape(-1, 4)[:nc.value].mean(axis=0)
                ms_local = e0.elapsed_time(e1) / steps
            t = torch.tensor([ms_local], dtype=torch.float64, device=dev)
            if dist is not None:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t[0])
            res[name + '_ms'] = ms
            res[name + '_samples_per_s'] = B_global / (ms * 1e-3)
            if kt is not None and rank == 0:
                res['kernel_ms_rank0'] = {'pose_prep': kt[0], 'blend_gemm': kt[1], 'skinning': kt[2], 'keypoints': kt[3]}
                res['skin_hbm_frac'] = SKIN_BYTES_PER_SAMPLE * B / (kt[2] * 1e-3) / 1e9 / peak
                res['blend_frac_of_bf16_burst_issued'] = BLEND_FLOP_ISSUED * B / (kt[1] * 1e-3) / 1e12 / tp['bf16_burst']
                res['blend_tflops_fp32_equivalent'] = BLEND_FLOP_FP32EQ * B / (kt[1] * 1e-3) / 1e12
            if kb is not None and rank == 0:
                # per-kernel algorithmic bytes per sample: g_A reduction reads g_verts + v_posed; the skinning adjoint
                # reads g_verts and writes the bf16 split g_vposed image (4 B per element); the GEMM reads that image
                res['kernel_ms_bwd_rank0'] = {'skin_bwd_gA': kb[0], 'skin_tc<1> g_vposed': kb[1], 'blend_bwd_gemm': kb[2],
                                              'joint_bwd': kb[3]}
                res['roofline_bwd'] = {
                    'skin_bwd_gA': {'bytes_per_sample': 82680 + 82944, 'hbm_frac': (82680 + 82944) * B / (kb[0] * 1e-3) / 1e9 / peak},
                    'skin_tc<1>': {'bytes_per_sample': 82680 + 82944, 'hbm_frac': (82680 + 82944) * B / (kb[1] * 1e-3) / 1e9 / peak},
                    'blend_bwd_gemm': {'bytes_per_sample': 82944, 'hbm_frac': 82944 * B / (kb[2] * 1e-3) / 1e9 / peak,
                                       'tflops_issued': BLEND_FLOP_ISSUED * B / (kb[2] * 1e-3) / 1e12,
                                       'frac_of_bf16_burst': BLEND_FLOP_ISSUED * B / (kb[2] * 1e-3) / 1e12 / tp['bf16_burst']}}
        # backward roofline: SURVEY 8d algorithmic bytes of the vertex path (read g_verts + v_posed, write g_vposed)
        if 'fwd_bwd_ms' in res and 'fwd_ms' in res:
            bwd = max(res['fwd_bwd_ms'] - res['fwd_ms'], 1e-9)
            res['bwd_ms'] = bwd
            res['bwd_hbm_frac_algorithmic'] = 248040.0 * res['B_per_gpu_max'] / (bwd * 1e-3) / 1e9 / peak
        rows.append(res)
        if B > 0:
            del g_verts, out, d
    cpu_rows = []
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the CPU port of the reference op sequence at B in {1, 32, 256, 1024} (BASELINE configs[0] is the
        # B = 1 case), median over the repetitions that fit in ~3 s each, all host threads
        import statistics
        for Bc in (1, 32, 256, 1024):
            _, cores, _, _, step = cpu_reference_rate(Bc, 0.0, 1)
            ts, t_end = [], time.perf_counter() + 3.0
            while len(ts) < 50 and (time.perf_counter() < t_end or len(ts) < 3):
                t0 = time.perf_counter()
                step()
                ts.append(time.perf_counter() - t0)
            med = statistics.median(ts)
            cpu_rows.append({'B': Bc, 'ms': 1e3 * med, 'samples_per_s': Bc / med, 'reps': len(ts), 'cores': cores,
                             'kind': 'port'})
    if rank == 0:
        print(json.dumps({'metric': 'smpl_sweep', 'n_gpus': world, 'scaling': args.scaling, 'dtype': DTYPE, 'data': 'synthetic',
                          'cpu_port_rows': cpu_rows,
                          'peaks': {'hbm_gbs': peak, 'tensor': tp, 'source': peak_src},
                          'config': {'workload': 'BASELINE configs[4]: batch sweep, fwd+projection and fwd+bwd '
                                                 '(g_verts + g_joints -> g_theta, g_beta); %s' % (
                                                     'global batch sharded over the ranks' if args.scaling == 'strong'
                                                     else 'B per GPU')},
                          'rows': rows}), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def uv_sphere_mesh(n_lon=82, n_ring=84):
    from smplb200 import synthetic as syn
    return syn.uv_sphere_mesh(n_lon, n_ring)


def raster_inputs(B, seed):
    """Body-sized ellipsoids (SMPL vertex/triangle counts) under random rotations, as NDC [B,6890,3]."""
    v, tri = uv_sphere_mesh()
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, 2 * np.pi, size=B)
    c, s = np.cos(ang), np.sin(ang)
    R = np.zeros((B, 3, 3), np.float32)
    R[:, 0, 0], R[:, 0, 2], R[:, 1, 1], R[:, 2, 0], R[:, 2, 2] = c, s, 1, -s, c
    body = v * np.array([0.22, 0.75, 0.15], np.float32)
    ndc = np.einsum('bij,vj->bvi', R, body) + rng.uniform(-0.15, 0.15, size=(B, 1, 3)).astype(np.float32)
    ndc[:, :, 2] *= 0.5
    return ndc.astype(np.float32), tri


def run_raster(args, rank, local_rank, world, dist):
    """batched triangle rasterisation of SMPL-sized meshes (6890 vertices, 13776 triangles) at the
    reference's 320x240 (config.py img dims): images/s, with the reference's own C++ rasteriser (oracle/_ref, one
    host thread per image as tf.map_fn would give at best) timed beside it."""
    from smplb200.raster import Mesh
    dev = torch.device('cuda', local_rank)
    W, H = 320, 240
    B = args.batch if args.batch != 65536 else 1024
    ndc_h, tri = raster_inputs(min(B, 64), 1000 + rank)
    ndc = torch.from_numpy(ndc_h).to(dev).repeat((B + ndc_h.shape[0] - 1) // ndc_h.shape[0], 1, 1)[:B].contiguous()
    mesh = Mesh(tri, ndc.shape[1], local_rank)
    g_bary = torch.randn((B, H, W, 3), device=dev)

    def sync():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        for _ in range(args.warmup):
            fn()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    out = {}

    def fwd():
        out['r'] = mesh.rasterize(ndc, W, H)

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_f = timed(fwd, args.steps)
    bary, ids, z = out['r']
    ms_b = timed(lambda: mesh.rasterize_backward(ndc, bary, ids, g_bary, W, H), args.steps)
    # fused renderer (normals -> clip transform -> rasterise -> interpolate -> diffuse Phong), pose_3d's view_f camera
    from tf_mesh_renderer import camera_utils as cu
    wv = ndc * torch.tensor([1.0, 1.0, 2.0], device=dev)                      # world-space bodies around the origin
    eye = torch.tensor([[0.0, -0.35, -4.0]], device=dev).expand(B, 3)
    clip_tf = torch.matmul(cu.perspective(W / H, torch.full((B,), 30.0, device=dev), torch.full((B,), 0.05, device=dev),
                                          torch.full((B,), 100.0, device=dev)),
                           cu.look_at(eye, torch.zeros((B, 3), device=dev), torch.tensor([[0.0, 1.0, 0.0]], device=dev).expand(B, 3)))
    lights = torch.tensor([[[0.0, 1.0, -4.0]]], device=dev).expand(B, 1, 3).contiguous()
    li = torch.ones((B, 1, 3), device=dev)
    nrm = torch.empty_like(wv)

    def render():
        mesh.normals(wv, out=nrm)
        out['img'] = mesh.render_diffuse(wv, nrm, clip_tf, lights, li, W, H)

    ms_r = timed(render, args.steps)
    # e2e: host vertices in, host barycentrics/ids out
    hv = ndc.cpu().pin_memory()
    hb = torch.empty((B, H, W, 3), dtype=torch.float32).pin_memory()
    hi = torch.empty((B, H, W), dtype=torch.int32).pin_memory()

    def e2e():
        d = hv.to(dev, non_blocking=True)
        b_, i_, _ = mesh.rasterize(d, W, H)
        hb.copy_(b_, non_blocking=True)
        hi.copy_(i_, non_blocking=True)
        torch.cuda.synchronize()

    e2e()
    sync()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e()
    el = (time.perf_counter() - t0) / args.e2e_steps
    sampler.stop_flag = True
    sampler.join()
    if rank != 0:
        return
    peak, peak_src = peaks()
    out_bytes = B * W * H * 20 + B * 6890 * 12         # bary + ids + z written, vertices read
    line = {'metric': 'raster_images_per_s', 'value': world * B / (ms_f * 1e-3), 'unit': 'images/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_f, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32 + 16.16 fixed point', 'data': 'synthetic',
            'config': {'workload': 'rasterise B=%d SMPL-sized meshes (6890 vertices, 13776 triangles, UV-ellipsoids) at '
                                   '320x240 per GPU per step -> barycentrics, triangle ids, depth' % B},
            'clocks': sampler.summary(), 'gpu_launches': 4 * args.steps,
            'backward_ms_per_step': ms_b, 'backward_images_per_s': world * B / (ms_b * 1e-3),
            'render_ms_per_step': ms_r, 'render_images_per_s': world * B / (ms_r * 1e-3),
            'render_covered_fraction': float((out['img'][..., 3] > 0.5).float().mean()),
            'covered_fraction': float((bary.sum(-1) > 0.5).float().mean()),
            'roofline': {'bound': 'hbm', 'kernel': 'key fill + raster_tri + raster_big + raster_resolve4 (whole step)',
                         'achieved': out_bytes / (ms_f * 1e-3) / 1e9, 'peak': peak, 'unit': 'GB/s',
                         'frac': out_bytes / (ms_f * 1e-3) / 1e9 / peak, 'traffic': None, 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': out_bytes},
            'e2e': {'value': world * B / el, 'unit': 'images/s', 'h2d_bytes_per_step': hv.numel() * 4,
                    'd2h_bytes_per_step': hb.numel() * 4 + hi.numel() * 4, 'steps': args.e2e_steps}}
    if not args.no_cpu_baseline:
        line['cpu_baseline'] = raster_cpu_rate(ndc_h, tri, W, H, args.cpu_seconds)
    print(json.dumps(line), flush=True)


def raster_cpu_rate(ndc_h, tri, W, H, seconds):
    """The reference's own C++ rasteriser (oracle/_ref) when it was built, else the C restatement; one image per
    host thread (ctypes releases the GIL)."""
    import subprocess
    from concurrent.futures import ThreadPoolExecutor
    subprocess.check_call(['make', '-s', '-C', os.path.join(ROOT, 'oracle'), 'port'])
    from oracle import raster as orc            # timed CPU baseline only
    kind = 'reference' if orc.ref_available() else 'port'
    fn = orc.forward_ref if kind == 'reference' else orc.forward_port
    cores = os.cpu_count() or 1
    n, t0 = 0, time.perf_counter()
    with ThreadPoolExecutor(cores) as ex:
        while time.perf_counter() - t0 < seconds:
            list(ex.map(lambda i: fn(ndc_h[i % len(ndc_h)], tri, W, H), range(4 * cores)))
            n += 4 * cores
    el = time.perf_counter() - t0
    return {'value': n / el, 'unit': 'images/s', 'cores': cores, 'kind': kind,
            'sample': '%d images of the same workload in %.1f s (%s)' % (
                n, el, 'rasterize_triangles_impl.cc compiled from the reference' if kind == 'reference'
                else 'oracle/raster_port.c')}


def run_normals(args, rank, local_rank, world, dist):
    """SURVEY 8f N1 (shading input): normals_from_mesh over verts[B,6890,3] (deps/pose_3d/utils.py:92-127), one HBM
    pass; B = 65536 meshes per GPU (the forward's output batch)."""
    from smplb200 import synthetic as syn
    from smplb200.raster import Mesh
    dev = torch.device('cuda', local_rank)
    B = args.batch
    v, tri = syn.uv_sphere_mesh()
    mesh = Mesh(tri, v.shape[0], local_rank)
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    verts = torch.from_numpy(v).to(dev)[None] * (0.2 + 0.8 * torch.rand((B, 1, 3), device=dev, generator=g)) \
        + 0.002 * torch.randn((B, v.shape[0], 3), device=dev, generator=g)
    out = torch.empty_like(verts)

    def step():
        mesh.normals(verts, out=out)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    sampler.stop_flag = True
    sampler.join()
    t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    if rank != 0:
        return
    peak, peak_src = peaks()
    byt = 2 * 82680 * B
    line = {'metric': 'mesh_normals_samples_per_s', 'value': world * B / (ms * 1e-3), 'unit': 'samples/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'normals_from_mesh of B=%d meshes (6890 vertices, 13776 triangles) per GPU per step; '
                                   'working set %.1f GB exceeds L2' % (B, byt / 1e9)},
            'clocks': sampler.summary(), 'gpu_launches': args.steps,
            'roofline': {'bound': 'hbm', 'kernel': 'normals_kernel', 'achieved': byt / (ms * 1e-3) / 1e9, 'peak': peak,
                         'unit': 'GB/s', 'frac': byt / (ms * 1e-3) / 1e9 / peak, 'traffic': None, 'peak_source': peak_src,
                         'algorithmic_bytes_per_launch': byt}}
    if not args.no_cpu_baseline:
        from oracle import ref_numpy as rn          # timed CPU baseline only
        vf = rn.vertex_faces_from_face_verts(tri)
        hv = verts[:64].cpu().numpy()
        rn.normals_from_mesh(hv, tri, vf)
        n, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < args.cpu_seconds:
            rn.normals_from_mesh(hv, tri, vf)
            n += 64
        el = time.perf_counter() - t0
        line['cpu_baseline'] = {'value': n / el, 'unit': 'samples/s', 'cores': 1, 'kind': 'port',
                                'sample': '%d meshes in %.1f s, numpy restatement of the TF graph (64 per call)' % (n, el)}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--batch', type=int, default=65536,
                    help='samples per step: the GLOBAL batch under --scaling strong (default), per GPU under weak')
    ap.add_argument('--scaling', default='strong', choices=['strong', 'weak'],
                    help='strong: --batch is the global batch, sharded over the ranks (BASELINE configs[3]); '
                         'weak: --batch samples per GPU')
    ap.add_argument('--graph', action=argparse.BooleanOptionalAction, default=None,
                    help='replay the step from one CUDA graph (default: on when more than one GPU; fit mode: off)')
    ap.add_argument('--allreduce-every', type=int, default=1,
                    help='fit mode: all-reduce the logged loss scalars once every k steps (k steps per graph)')
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--blend', default='tc', choices=['tc', 'simt'])
    ap.add_argument('--e2e-steps', type=int, default=3)
    ap.add_argument('--ref-chunk', type=int, default=256)
    ap.add_argument('--cpu-seconds', type=float, default=12.0)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--sweep-min', type=int, default=1, help='sweep mode: smallest batch of the sweep')
    ap.add_argument('--mode', default='forward', choices=['forward', 'fit', 'sweep', 'raster', 'normals'],
                    help="forward: the headline metric; fit: BASELINE config 3 (fitting steps on the 2D re-projection loss)")
    ap.add_argument('--loss', choices=['reproj', 'mesh'], default='reproj', help='fit mode: which loss drives the step')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else max(args.warmup, 1)

    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return

    assert torch.cuda.is_available(), 'bench.py needs a CUDA device (no CPU fallback)'
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        # NCCL prints its version banner to stdout when the first communicator comes up (NCCL_DEBUG=VERSION and WARN both
        # do): create the communicator here, with file descriptor 1 pointing at stderr meanwhile, so that stdout stays
        # the one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    import smplb200
    from smplb200 import synthetic as syn, _lib
    if args.mode == 'fit':
        run_fit(args, rank, local_rank, world, dist)
        return
    if args.mode == 'sweep':
        run_sweep(args, rank, local_rank, world, dist)
        return
    if args.mode in ('raster', 'normals'):
        (run_raster if args.mode == 'raster' else run_normals)(args, rank, local_rank, world, dist)
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return
    run_forward(args, rank, local_rank, world, dist)


def run_forward(args, rank, local_rank, world, dist):
    """The headline metric: SMPL forward + projection samples/s (BASELINE configs[1]/[3])."""
    import smplb200
    from smplb200 import synthetic as syn, _lib, dist as sd
    dev = torch.device('cuda', local_rank)
    dd = syn.make_synthetic_smpl(0)
    model = smplb200.Model(dd, device=local_rank)
    if args.scaling == 'strong':
        b0, b1 = sd.shard_range(args.batch, rank, world)      # contiguous slice of the global batch
        B, B_global = b1 - b0, args.batch
    else:
        B, B_global = args.batch, args.batch * world
    use_graph = args.graph if args.graph is not None else world > 1
    flags = (_lib.F_BLEND_SIMT | _lib.F_SKIN_SIMT) if args.blend == 'simt' else 0
    inp = syn.make_inputs(B, seed=1000 + rank)
    d = {k: torch.from_numpy(v).to(dev) for k, v in inp.items()}
    out = {}

    def step():
        model.forward(d['beta'], d['theta'], d['cam_pos'], d['cam_rot'], d['focal'], flags=flags, out=out)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    graph = None
    if use_graph and B > 0:
        # the whole step (pose_prep, blend GEMM, skinning, keypoints) as one graph launch: at B/N = 8192 the step is
        # ~0.5 ms and four separate launches + ctypes calls would show
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                step()
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(args.warmup):
            graph.replay()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    L = _lib.lib()
    if graph is None:
        L.smplb_timing_begin(args.steps)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        graph.replay() if graph is not None else step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if graph is not None:
        # per-kernel times: the same steps once more, eager, with the library's event hooks (outside the timed region:
        # events cannot be read back from inside a graph)
        L.smplb_timing_begin(args.steps)
        for _ in range(args.steps):
            step()
        torch.cuda.synchronize()
    kms = (C.c_float * (4 * args.steps))()
    ncalls = C.c_int32(0)
    L.smplb_timing_end(kms, C.byref(ncalls))
    kt = np.array(list(kms), dtype=np.float64).reshape(-1, 4)[:ncalls.value]

    # ---- e2e: host buffers through smplb_forward_host --------------------------------------
    e2e = None
    if not args.no_e2e and B > 0:
        NK = model.NK
        hb = {k: torch.from_numpy(v).pin_memory() for k, v in inp.items()}
        ho = {'verts': torch.empty((B, 6890, 3), dtype=torch.float32).pin_memory(),
              'joints': torch.empty((B, 24, 3), dtype=torch.float32).pin_memory(),
              'kp': torch.empty((B, NK, 3), dtype=torch.float32).pin_memory(),
              'px': torch.empty((B, 24, 2), dtype=torch.float32).pin_memory()}
        CamCfg, FIO = _lib.CamCfg, _lib.ForwardIO
        cfg = CamCfg(320.0, 240.0, 0.024, 0.05)
        hio = FIO(B, hb['beta'].data_ptr(), hb['theta'].data_ptr(), ho['verts'].data_ptr(),
                  ho['joints'].data_ptr(), ho['kp'].data_ptr(), None, None, hb['cam_pos'].data_ptr(),
                  hb['cam_rot'].data_ptr(), hb['focal'].data_ptr(), None, ho['px'].data_ptr(), cfg)
        dio = FIO(B, d['beta'].data_ptr(), d['theta'].data_ptr(), out['verts'].data_ptr(),
                  out['joints'].data_ptr(), out['kp'].data_ptr(), None, None, d['cam_pos'].data_ptr(),
                  d['cam_rot'].data_ptr(), d['focal'].data_ptr(), None, out['px'].data_ptr(), cfg)
        ws, need = model.workspace(B, flags)
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

        def e2e_step():
            _lib.check(L.smplb_forward_host(model._h, C.byref(hio), C.byref(dio), C.c_void_p(ws.data_ptr()),
                                            ws.numel(), flags, st))

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()          # synchronises its stream before returning
        barrier()
        el = time.perf_counter() - t0
        h2d = sum(hb[k].numel() * 4 for k in hb)
        d2h = sum(ho[k].numel() * 4 for k in ho)
        e2e = (el, h2d, d2h)
    elif not args.no_e2e:
        barrier()
        barrier()
    sampler.stop_flag = True
    sampler.join()

    t_all = torch.tensor([ms, e2e[0] if e2e else 0.0, float(kt[:, :].sum(axis=1).mean()) if len(kt) else 0.0],
                         dtype=torch.float64, device=dev)
    io_bytes = torch.tensor([e2e[1] if e2e else 0, e2e[2] if e2e else 0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
        dist.all_reduce(io_bytes, op=dist.ReduceOp.SUM)
    ms_max, e2e_max, ksum_max = float(t_all[0]), float(t_all[1]), float(t_all[2])

    if rank == 0:
        peak, peak_src = peaks()
        tp = tensor_peaks()
        skin_ms = float(kt[:, 2].mean()) if len(kt) else float('nan')
        blend_ms = float(kt[:, 1].mean()) if len(kt) else float('nan')
        achieved = SKIN_BYTES_PER_SAMPLE * B / (skin_ms * 1e-3) / 1e9
        tr = ncu_record('skin_traffic.json', B)
        gr = ncu_record('blend_tensor_pipe.json', B)
        issued_tflops = BLEND_FLOP_ISSUED * B / (blend_ms * 1e-3) / 1e12
        cfg = workload_config(args, world)
        cfg.update({
            'samples_per_gpu': B, 'global_batch': B_global, 'cuda_graph': graph is not None,
            'blend_gemm': 'tcgen05 kind::f16: f16 main product + two f16 residual products on a power-of-two scaled basis, fp32 accumulate (fp32-faithful, 2e-7 of the vertex scale; |beta| and model constants <= 65504)' if args.blend == 'tc' else 'fp32 SIMT (verification path)',
            'sharding': 'contiguous batch shards, one rank per GPU, no forward collective',
            'l2': 'working set %.2f GB per GPU per step exceeds the 126 MB L2 (no flush needed)' % (
                B * (82944 + 82680) / 1e9),
        })
        line = {
            'metric': METRIC, 'value': B_global * args.steps / (ms_max * 1e-3), 'unit': UNIT,
            'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms_max / args.steps, 'higher_is_better': True, 'scaling': args.scaling,
            'vs_baseline': None, 'dtype': DTYPE, 'data': 'synthetic', 'config': cfg,
            'clocks': sampler.summary(),
            'gpu_launches': 4 * args.steps,
            'kernel_ms': {'pose_prep': float(kt[:, 0].mean()), 'blend_gemm': blend_ms,
                          'skinning': skin_ms, 'keypoints': float(kt[:, 3].mean()),
                          'sum_max_over_ranks': ksum_max,
                          'source': 'CUDA events around each launch, rank 0, %s' % (
                              'eager pass right after the graph-timed region' if graph is not None
                              else 'inside the timed region')} if len(kt) else None,
            'roofline': {'bound': 'hbm', 'kernel': 'skin_tc_kernel<0> (LBS: tcgen05 weight blend + apply, v_posed -> verts)',
                         'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': tr['dram_bytes_per_launch'] if tr else None,
                         'traffic_source': ('ncu dram__bytes_read+write, %s @ %s' % (tr['report'], tr['captured_at_commit'])) if tr
                         else 'no ncu record for this source revision / batch',
                         'peak_source': peak_src, 'algorithmic_bytes_per_launch': SKIN_BYTES_PER_SAMPLE * B},
            'roofline_gemm': {'bound': 'tensor', 'kernel': 'blend_tc_kernel ([B,217] x [217,20670] as 3 kind::f16 products, M256 N256 K16, cta_group::2)',
                              'flops_issued_per_launch': BLEND_FLOP_ISSUED * B,
                              'achieved_tflops_issued': issued_tflops,
                              'achieved_tflops_fp32_equivalent': BLEND_FLOP_FP32EQ * B / (blend_ms * 1e-3) / 1e12,
                              'frac_of_bf16_burst': issued_tflops / tp['bf16_burst'],
                              'frac_of_bf16_sustained': issued_tflops / tp['bf16_sustained'],
                              'peaks_tflops': tp,
                              'tensor_pipe_active_pct': gr['tensor_pipe_active_pct'] if gr else None,
                              'tensor_pipe_source': ('ncu sm__pipe_tensor_cycles_active, %s @ %s, SM clock %.2f GHz' % (
                                  gr['report'], gr['captured_at_commit'], gr['sm_clock_ghz'])) if gr
                              else 'no ncu record for this source revision / batch'},
        }
        if e2e:
            line['e2e'] = {'value': B_global * args.e2e_steps / e2e_max, 'unit': UNIT,
                           'h2d_bytes_per_step': int(io_bytes[0]), 'd2h_bytes_per_step': int(io_bytes[1]),
                           'steps': args.e2e_steps, 'note': 'smplb_forward_host, pinned host buffers, PCIe-bound; bytes summed over ranks'}
        if not args.no_cpu_baseline and world == 1:     # CPU baseline: rank 0 at N=1 only
            rate, cores, ns, el, _ = cpu_reference_rate(args.ref_chunk, args.cpu_seconds, 1000)
            line['cpu_baseline'] = {
                'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                'sample': '%d samples (%d-sample chunks, %.1f s) of the same workload; torch-CPU fp32 port of '
                          'the reference TF op sequence incl. tile(W)' % (ns, args.ref_chunk, el),
                'tight': cpu_tight_rate(min(args.cpu_seconds, 5.0)),
                'tensorflow': tensorflow_probe()}
        print(json.dumps(line), flush=True)
    if graph is not None:
        graph.reset()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
