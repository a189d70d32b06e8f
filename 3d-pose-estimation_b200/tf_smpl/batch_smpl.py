"""Drop-in for deps/tf_smpl/batch_smpl.py: same class, constructor and call signature
(`SMPL(pkl_path, joint_type='cocoplus', dtype=float32)`, `smpl(beta, theta, get_skin=False)`),
backed by the sm_100a kernels through the C ABI.  Tensors are torch CUDA tensors (carriers);
the call is differentiable (hand-written backward kernels, torch.autograd.Function).
"""
import torch

from smplb200 import _lib
from smplb200.api import Model


class _SMPLFunction(torch.autograd.Function):
    """verts, joints(kp), Rs, J_transformed = f(beta, theta); backward = smplb_backward."""

    @staticmethod
    def forward(ctx, beta, theta, model, flags):
        need_grad = beta.requires_grad or theta.requires_grad
        f = flags | (_lib.F_KEEP_VPOSED if need_grad else 0)
        # every differentiable call keeps its own workspace (v_posed, A image) until its backward has run: the
        # reference's training step calls SMPL twice before it differentiates (pose_model_3d.py:177-192)
        ws = model.new_workspace(beta.shape[0], f) if need_grad else None
        out = model.forward(beta, theta, want_verts=True, want_kp=True, want_Rs=True, flags=f, ws=ws, out={})
        ctx.model, ctx.flags, ctx.ws = model, f, ws
        ctx.save_for_backward(out['_keep'][0], out['_keep'][1])
        ctx.mark_non_differentiable(out['Rs'])
        return out['verts'], out['kp'], out['Rs'], out['joints']

    @staticmethod
    def backward(ctx, g_verts, g_kp, g_Rs, g_joints):
        beta, theta = ctx.saved_tensors
        g_beta, g_theta = ctx.model.backward(beta, theta, g_verts=g_verts, g_joints=g_joints,
                                             g_kp=g_kp, flags=ctx.flags, ws=ctx.ws)
        ctx.ws = None      # release the call's workspace
        return g_beta, g_theta, None, None


class SMPL(object):
    def __init__(self, pkl_path, joint_type='cocoplus', dtype=torch.float32, device=None, flags=0):
        """pkl_path is the path to a SMPL model (deps/tf_smpl/batch_smpl.py:17-69)."""
        if dtype not in (torch.float32, 'float32'):
            raise ValueError('the reference path is fp32 (batch_smpl.py:17); got %r' % (dtype,))
        self._model = Model(pkl_path, joint_type=joint_type, device=device)
        m = self._model
        dev = m.device
        self.v_template = torch.from_numpy(m.v_template).to(dev)
        self.size = m.size
        self.num_betas = m.num_betas
        self.shapedirs = torch.from_numpy(m.shapedirs).to(dev)
        self.J_regressor = torch.from_numpy(m.J_regressor).to(dev)
        self.posedirs = torch.from_numpy(m.posedirs).to(dev)
        self.parents = m.parents
        self.weights = torch.from_numpy(m.weights).to(dev)
        self.joint_regressor = torch.from_numpy(m.joint_regressor).to(dev)
        self.J_transformed = None
        self._flags = flags

    def __call__(self, beta, theta, get_skin=False, name=None):
        """beta: N x 10, theta: N x 72 (axis-angle incl. global rotation).
        Updates self.J_transformed (N x 24 x 3).  Returns joints (N x 19|14 x 3), or
        (verts N x 6890 x 3, joints, Rs N x 24 x 3 x 3) when get_skin (batch_smpl.py:71-143)."""
        dev = self._model.device
        beta = torch.as_tensor(beta, dtype=torch.float32).to(dev)
        theta = torch.as_tensor(theta, dtype=torch.float32).to(dev)
        verts, joints, Rs, J_transformed = _SMPLFunction.apply(beta, theta, self._model, self._flags)
        self.J_transformed = J_transformed
        if get_skin:
            return verts, joints, Rs
        return joints
