"""One forward+backward (vertex path) at a given batch, for ncu launch lists: python tools/bwd_probe.py [B] [iters]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, '3d-pose-estimation_b200')]
import torch
import smplb200
from smplb200 import synthetic as syn, _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
it = int(sys.argv[2]) if len(sys.argv) > 2 else 2
m = smplb200.Model(syn.make_synthetic_smpl(0), device=0)
inp = syn.make_inputs(B, seed=1)
d = {k: torch.from_numpy(v).cuda() for k, v in inp.items()}
gv = torch.randn((B, 6890, 3), device='cuda') * 1e-6
gj = torch.randn((B, 24, 3), device='cuda') * 1e-4
out = {}
for _ in range(it):
    m.forward(d['beta'], d['theta'], d['cam_pos'], d['cam_rot'], d['focal'], flags=_lib.F_KEEP_VPOSED, out=out)
    m.backward(d['beta'], d['theta'], g_verts=gv, g_joints=gj)
torch.cuda.synchronize()
print('ok')
