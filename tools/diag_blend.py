"""Diagnostic: tcgen05 kernels vs the CUDA-core verification kernels (same forward otherwise)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, '3d-pose-estimation_b200')):
    sys.path.insert(0, p)
import numpy as np, torch
import smplb200
from smplb200 import synthetic as syn, _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dd = syn.make_synthetic_smpl(0)
m = smplb200.Model(dd, device=0)
inp = syn.make_inputs(B, seed=5)
ref = m.forward(inp['beta'], inp['theta'], flags=_lib.F_BLEND_SIMT | _lib.F_SKIN_SIMT)['verts'].clone()
torch.cuda.synchronize()
for name, fl in (('blend_tc+skin_simt', _lib.F_SKIN_SIMT), ('blend_simt+skin_tc', _lib.F_BLEND_SIMT), ('tc+tc', 0)):
    out = m.forward(inp['beta'], inp['theta'], flags=fl)['verts'].clone()
    torch.cuda.synchronize()
    err = (out - ref).abs()
    print(name, 'B', B, 'max|ref|', float(ref.abs().max()), 'max err', float(err.max()), 'nan', int(torch.isnan(out).sum()))
    e = err.reshape(B, -1)
    rows = e.max(dim=1).values.cpu().numpy()
    cols = e.max(dim=0).values.cpu().numpy()
    print('  bad rows (>1e-5):', int((rows > 1e-5).sum()), 'first', np.nonzero(rows > 1e-5)[0][:16])
    print('  bad cols (>1e-5):', int((cols > 1e-5).sum()), 'first', np.nonzero(cols > 1e-5)[0][:24])
