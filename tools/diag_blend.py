"""Diagnostic: tcgen05 blend GEMM vs the CUDA-core verification GEMM (same forward otherwise)."""
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
ref = m.forward(inp['beta'], inp['theta'], flags=_lib.F_BLEND_SIMT)['verts'].clone()
torch.cuda.synchronize()
out = m.forward(inp['beta'], inp['theta'], flags=0)['verts'].clone()
torch.cuda.synchronize()
err = (out - ref).abs()
print('B', B, 'max|ref|', float(ref.abs().max()), 'max err', float(err.max()), 'nan', int(torch.isnan(out).sum()))
e = err.reshape(B, -1)
rows = e.max(dim=1).values.cpu().numpy()
cols = e.max(dim=0).values.cpu().numpy()
print('bad rows (>1e-5):', int((rows > 1e-5).sum()), 'first', np.nonzero(rows > 1e-5)[0][:16])
print('bad cols (>1e-5):', int((cols > 1e-5).sum()), 'first', np.nonzero(cols > 1e-5)[0][:16])
print('row err sample', rows[:8], 'col err sample', cols[:8])
