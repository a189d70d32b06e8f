"""Small run of every hot-path kernel for compute-sanitizer (tools/sanitize.sh): forward (tcgen05, fused, CUDA-core paths),
backward through the vertices, fitting step, at batches that end inside a tile / group."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, '3d-pose-estimation_b200')]
import numpy as np, torch
import smplb200
from smplb200 import synthetic as syn, _lib
from smplb200.fit import Fitter
m = smplb200.Model(syn.make_synthetic_smpl(0), device=0)
for B in (5, 37, 130):
    inp = syn.make_inputs(B, seed=B)
    for fl in (0, _lib.F_FUSED, _lib.F_BLEND_SIMT | _lib.F_SKIN_SIMT):
        o = m.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=fl, out={})
    m.forward(inp['beta'], inp['theta'], flags=_lib.F_KEEP_VPOSED)
    gv = np.random.default_rng(B).normal(size=(B, 6890, 3)).astype(np.float32) * 1e-6
    gk = np.random.default_rng(B).normal(size=(B, 19, 3)).astype(np.float32) * 1e-3
    gb, gt = m.backward(inp['beta'], inp['theta'], g_verts=gv, g_kp=gk, g_joints=np.ones((B, 24, 3), np.float32))
    m.forward(inp['beta'], inp['theta'], flags=_lib.F_KEEP_VPOSED)
    m.backward(inp['beta'], inp['theta'], g_verts=gv, flags=_lib.F_SKIN_SIMT)
    gt_px = m.forward(inp['beta'], inp['theta'], inp['cam_pos'], inp['cam_rot'], inp['focal'], flags=4)['px'].clone()
    f = Fitter(m, inp['beta'], inp['theta'] + 0.05, inp['cam_pos'], inp['cam_rot'], gt_px, focal=inp['focal'])
    f.step(); f.step()
    torch.cuda.synchronize()
    print('B=%d ok, |g_theta| max %.3e' % (B, float(gt.abs().max())), flush=True)
print('sanitize probe done')
