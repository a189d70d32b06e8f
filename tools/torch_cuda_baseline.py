""""a naive torch-CUDA composition on the B200 (TF32 off) as the library-kernel GPU comparison, since
the reference ships no GPU kernel of its own".  Not part of the product and not a bench.py arm: it answers "what does
the same op sequence cost on this GPU when every step is a stock library kernel (cuBLAS sgemm / bmm + element-wise)".

Two variants of batch_smpl.py:71-143 + project.py:7-48 on torch CUDA tensors, fp32, TF32 disabled:
  ref_sequence  the reference's op sequence, including tile(W) and the [B,6890,4,4] x [B,6890,4,1] batched matmul
  einsum        the same function without materialising W / homogeneous coordinates
Prints one JSON line per variant (samples/s over `--steps` timed steps after warm-up, CUDA events).

    python tools/torch_cuda_baseline.py --batch 1024 --steps 20
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, '3d-pose-estimation_b200'))
from smplb200 import synthetic as syn  # noqa: E402

PARENTS = None


def load(dev):
    dd = syn.make_synthetic_smpl(0)
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32), device=dev)
    m = dict(v_template=t(dd['v_template']), shapedirs=t(dd['shapedirs'].reshape(-1, 10).T),
             posedirs=t(dd['posedirs'].reshape(-1, 207).T), J_regressor=t(dd['J_regressor'].T.todense()),
             weights=t(dd['weights']), kp=t(dd['cocoplus_regressor'].T.todense()))
    return m, dd['kintree_table'][0].astype(np.int64)


def rodrigues_eps(theta):            # batch_lbs.py:38-57
    angle = torch.linalg.norm(theta + 1e-8, dim=1, keepdim=True)
    r = theta / angle
    c, s = torch.cos(angle)[..., None], torch.sin(angle)[..., None]
    z = torch.zeros_like(r[:, 0])
    K = torch.stack([z, -r[:, 2], r[:, 1], r[:, 2], z, -r[:, 0], -r[:, 1], r[:, 0], z], 1).reshape(-1, 3, 3)
    return c * torch.eye(3, device=theta.device) + (1 - c) * r[:, :, None] * r[:, None, :] + s * K


def rodrigues_k2(rv):                # project.py:51-81
    th = torch.linalg.norm(rv, dim=1, keepdim=True)
    u = rv / th
    z = torch.zeros_like(u[:, 0])
    K = torch.stack([z, -u[:, 2], u[:, 1], u[:, 2], z, -u[:, 0], -u[:, 1], u[:, 0], z], 1).reshape(-1, 3, 3)
    return torch.eye(3, device=rv.device) + torch.sin(th)[..., None] * K + (1 - torch.cos(th))[..., None] * (K @ K)


def forward(m, parents, beta, theta, cam_pos, cam_rot, focal, ref_sequence):
    B = beta.shape[0]
    v_shaped = (beta @ m['shapedirs']).reshape(B, 6890, 3) + m['v_template']
    J = torch.stack([v_shaped[:, :, c] @ m['J_regressor'] for c in range(3)], 2)
    Rs = rodrigues_eps(theta.reshape(-1, 3)).reshape(B, 24, 3, 3)
    phi = (Rs[:, 1:] - torch.eye(3, device=beta.device)).reshape(B, 207)
    v_posed = (phi @ m['posedirs']).reshape(B, 6890, 3) + v_shaped
    G = [None] * 24
    bottom = torch.tensor([0.0, 0.0, 0.0, 1.0], device=beta.device).expand(B, 1, 4)
    for i in range(24):
        t = J[:, i] if i == 0 else J[:, i] - J[:, parents[i]]
        A = torch.cat([torch.cat([Rs[:, i], t[:, :, None]], 2), bottom], 1)
        G[i] = A if i == 0 else G[parents[i]] @ A
    G = torch.stack(G, 1)
    Jt = G[:, :, :3, 3]
    init = G @ torch.cat([J, torch.zeros(B, 24, 1, device=beta.device)], 2)[..., None]
    A = G - torch.nn.functional.pad(init, (3, 0))
    if ref_sequence:
        W = m['weights'].repeat(B, 1).reshape(B, 6890, 24)                                  # batch_smpl.py:122-123
        T = (W @ A.reshape(B, 24, 16)).reshape(B, 6890, 4, 4)                              # :125-127
        vh = torch.cat([v_posed, torch.ones(B, 6890, 1, device=beta.device)], 2)
        verts = (T @ vh[..., None])[:, :, :3, 0]                                           # :128-132
    else:
        T = torch.einsum('vj,bjk->bvk', m['weights'], A[:, :, :3, :].reshape(B, 24, 12)).reshape(B, 6890, 3, 4)
        verts = torch.einsum('bvrc,bvc->bvr', T[..., :3], v_posed) + T[..., 3]
    kp = torch.stack([verts[:, :, c] @ m['kp'] for c in range(3)], 2)
    q = torch.einsum('bji,bpj->bpi', rodrigues_k2(cam_rot), Jt - cam_pos[:, None])
    uv = focal[:, None, None] * q[..., :2] / q[..., 2:3]
    px = uv / 0.024 * 240.0 + torch.tensor([160.0, 120.0], device=beta.device)
    return verts, Jt, kp, Rs, px


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=1024)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    a = ap.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = torch.device('cuda', 0)
    m, parents = load(dev)
    inp = syn.make_inputs(a.batch, seed=1000, dtype=np.float32)
    x = [torch.from_numpy(inp[k]).to(dev) for k in ('beta', 'theta', 'cam_pos', 'cam_rot', 'focal')]
    for name, refseq in (('ref_sequence', True), ('einsum', False)):
        with torch.no_grad():
            for _ in range(a.warmup):
                forward(m, parents, *x, refseq)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                out = forward(m, parents, *x, refseq)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        print(json.dumps({'impl': 'torch-cuda library kernels (fp32, TF32 off)', 'variant': name, 'batch': a.batch,
                          'ms_per_step': ms, 'samples_per_s': a.batch / ms * 1e3,
                          'peak_mem_gb': torch.cuda.max_memory_allocated() / 1e9,
                          'verts_checksum': float(out[0].double().abs().sum())}))


if __name__ == '__main__':
    main()
