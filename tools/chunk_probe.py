"""Does running the forward in sample chunks (so that a chunk's v_posed stays in the 126 MB L2 between the blend GEMM that
writes it and the skinning kernel that reads it, and the scratch is overwritten before it is evicted) beat one pass over
the whole batch?  Whole forward under one CUDA graph per variant; times are CUDA events over replays."""
import sys, json
import torch
sys.path.insert(0, '3d-pose-estimation_b200')
from smplb200 import synthetic as syn
from smplb200.api import Model

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dev = torch.device('cuda', 0)
m = Model(syn.make_synthetic_smpl(0), device=dev)
inp = {k: torch.from_numpy(v).to(dev) for k, v in syn.make_inputs(B, seed=1).items()}
big = dict(verts=torch.empty(B, 6890, 3, device=dev), joints=torch.empty(B, 24, 3, device=dev),
           kp=torch.empty(B, m.NK, 3, device=dev), Rs=torch.empty(B, 24, 3, 3, device=dev),
           uv=torch.empty(B, 24, 2, device=dev), px=torch.empty(B, 24, 2, device=dev))

def run(chunk):
    ws, _ = m.workspace(min(chunk, B), 0)
    for b0 in range(0, B, chunk):
        b1 = min(B, b0 + chunk)
        out = {k: v[b0:b1] for k, v in big.items()}
        w = ws if b1 - b0 == min(chunk, B) else m.workspace(b1 - b0, 0)[0]
        m.forward(inp['beta'][b0:b1], inp['theta'][b0:b1], inp['cam_pos'][b0:b1], inp['cam_rot'][b0:b1],
                  inp['focal'][b0:b1], out=out, ws=w)

res = []
for chunk in [B, 20480, 10240, 5120, 2560, 1280, 640]:
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        run(chunk); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            run(chunk)
        for _ in range(3): g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record(s)
        for _ in range(n): g.replay()
        e1.record(s); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
    ref = big['verts'].clone() if chunk == B else ref
    same = bool(torch.equal(ref, big['verts']))
    res.append(dict(chunk=chunk, ms=ms, samples_per_s=B / ms * 1e3, verts_identical=same))
    print(json.dumps(res[-1]), flush=True)
