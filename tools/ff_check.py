"""Fused forward (SMPLB_F_FUSED) vs the unfused tensor-core path and the float64 oracle; timing at a large batch."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, '3d-pose-estimation_b200')]
import numpy as np, torch
import smplb200
from smplb200 import synthetic as syn, _lib
from oracle import ref_numpy as rn
dd = syn.make_synthetic_smpl(0)
m = smplb200.Model(dd, device=0)
m64 = rn.load_model(dd, dtype=np.float64)
for B in [int(a) for a in (sys.argv[1:] or ['8', '37', '300', '1024'])]:
    inp = syn.make_inputs(B, seed=7 + B)
    o_ref = m.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'])
    vr, kr = o_ref['verts'].clone(), o_ref['kp'].clone()
    o = m.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=_lib.F_FUSED, out={})
    torch.cuda.synchronize()
    dv = (o['verts'] - vr).abs().max().item() / vr.abs().max().item()
    dk = (o['kp'] - kr).abs().max().item() / kr.abs().max().item()
    msg = 'B=%d fused vs unfused: verts %.2e kp %.2e' % (B, dv, dk)
    if B <= 300:
        oo = rn.smpl_forward(m64, inp['beta'].astype(np.float64), inp['theta'].astype(np.float64), tile_W=False)
        msg += ' | vs fp64 oracle: verts %.2e kp %.2e' % (
            np.abs(o['verts'].cpu().numpy() - oo['verts']).max() / np.abs(oo['verts']).max(),
            np.abs(o['kp'].cpu().numpy() - oo['joints']).max() / np.abs(oo['joints']).max())
    print(msg, flush=True)
if os.environ.get('FF_TIME'):
    B = int(os.environ['FF_TIME'])
    inp = syn.make_inputs(B, seed=1)
    d = {k: torch.from_numpy(v).cuda() for k, v in inp.items()}
    for fl, name in ((0, 'unfused'), (_lib.F_FUSED, 'fused')):
        out = {}
        for _ in range(3):
            m.forward(d['beta'], d['theta'], d['cam_pos'], d['cam_rot'], d['focal'], flags=fl, out=out)
        torch.cuda.synchronize()
        import ctypes as C
        L = _lib.lib()
        L.smplb_timing_begin(10)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            m.forward(d['beta'], d['theta'], d['cam_pos'], d['cam_rot'], d['focal'], flags=fl, out=out)
        e1.record(); torch.cuda.synchronize()
        kms = (C.c_float * 40)(); nc = C.c_int32(0)
        L.smplb_timing_end(kms, C.byref(nc))
        kt = np.array(list(kms)).reshape(-1, 4)[:nc.value].mean(axis=0)
        print('B=%d %s forward: %.3f ms  [prep %.3f | blend %.3f | skin/fused %.3f | kp %.3f]' % (
            B, name, e0.elapsed_time(e1) / 10, kt[0], kt[1], kt[2], kt[3]), flush=True)
