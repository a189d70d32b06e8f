"""TEST INFRASTRUCTURE ONLY (oracle): runs the REFERENCE'S OWN PYTHON FILES, imported from /root/reference where they
lie, on top of oracle/tfshim (a torch-CPU stand-in for the `tensorflow` module).  This is how the oracle restatements
are pinned for the SMPL / projection / loss / pose-prep / normals / renderer-graph code, for which the reference has no
tests: `tests/golden/make_reference_golden.py` stores what these files compute on seeded inputs, and
`tests/test_reference_pin.py` compares the restatements (and, on the GPU, the CUDA path) with it.

/root/reference exists only in the build container: everything here is guarded by `available()`, and nothing that
runs on the GPU box imports this module's loaders (the committed fixtures travel instead).

The custom op `rasterize_triangles_kernel.so` (tf.load_op_library, rasterize_triangles.py:22-24) is served by the
reference's C++ rasteriser compiled into oracle/_ref (forward) and oracle/raster_port.c (gradient).
"""
import importlib
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get('SMPLB_REFERENCE_ROOT', '/root/reference')
DEPS = os.path.join(REF, 'deps')
_PKGS = ('tensorflow', 'tf_smpl', 'tf_perspective_projection', 'tf_mesh_renderer', 'pose_3d', 'tf_pose')


def available():
    return os.path.isfile(os.path.join(DEPS, 'tf_smpl', 'batch_smpl.py'))


def _raster_op_module(tf):
    import torch
    from . import raster as orc

    class _Raster(torch.autograd.Function):
        @staticmethod
        def forward(ctx, ndc, tri, W, H):
            fwd = orc.forward_ref if orc.ref_available() else orc.forward_port
            bary, ids, z = fwd(ndc.detach().numpy().astype(np.float32), tri.numpy().astype(np.int32), W, H)
            ctx.save_for_backward(ndc, tri, torch.from_numpy(bary), torch.from_numpy(ids))
            ctx.wh = (W, H)
            ctx.mark_non_differentiable
            return (torch.from_numpy(bary).to(ndc.dtype), torch.from_numpy(ids.astype(np.int32)),
                    torch.from_numpy(z).to(ndc.dtype))

        @staticmethod
        def backward(ctx, g_bary, g_ids, g_z):
            ndc, tri, bary, ids = ctx.saved_tensors
            g = orc.backward_port(ndc.detach().numpy(), tri.numpy(), bary.numpy(), ids.numpy(),
                                  g_bary.numpy().astype(np.float32), *ctx.wh)
            return torch.from_numpy(g).to(ndc.dtype), None, None, None

    def rasterize_triangles(vertices, triangles, image_width, image_height):
        b, i, z = _Raster.apply(tf._raw(vertices), tf._raw(triangles), int(image_width), int(image_height))
        return tf.Tensor(b), tf.Tensor(i), tf.Tensor(z)

    return types.SimpleNamespace(rasterize_triangles=rasterize_triangles)


def load(float64=False, cv2=None):
    """Fresh import of the reference modules under the shim.  Returns a namespace of modules.  `cv2`: a stand-in for
    OpenCV (not installed here); when given, pose_3d/data_helpers.py is imported too (it needs it at import time)."""
    if not available():
        raise RuntimeError('reference tree not present at %s' % REF)
    for k in [k for k in sys.modules if k.split('.')[0] in _PKGS]:
        del sys.modules[k]
    saved = list(sys.path)
    sys.path.insert(0, DEPS)
    sys.path.insert(0, os.path.join(HERE, 'tfshim'))     # shim `tensorflow` and the tf_pose stub win
    try:
        tf = importlib.import_module('tensorflow')
        if not hasattr(tf, 'set_float64'):
            raise RuntimeError('a real tensorflow is installed; the shim is not needed')
        tf.set_float64(float64)
        tf.OP_LIBRARY_LOADER = lambda path: _raster_op_module(tf)
        net = types.ModuleType('pose_3d.network')           # keras CNN: not on the path, never built here
        net.build_model = net.build_discriminator = None
        ns = types.SimpleNamespace(tf=tf)
        ns.batch_smpl = importlib.import_module('tf_smpl.batch_smpl')
        ns.batch_lbs = importlib.import_module('tf_smpl.batch_lbs')
        ns.projection = importlib.import_module('tf_smpl.projection')
        ns.project = importlib.import_module('tf_perspective_projection.project')
        ns.camera_utils = importlib.import_module('tf_mesh_renderer.camera_utils')
        ns.rasterize_triangles = importlib.import_module('tf_mesh_renderer.rasterize_triangles')
        ns.mesh_renderer = importlib.import_module('tf_mesh_renderer.mesh_renderer')
        importlib.import_module('pose_3d')
        sys.modules['pose_3d.network'] = net
        ns.config = importlib.import_module('pose_3d.config')
        ns.utils = importlib.import_module('pose_3d.utils')
        ns.pose_model_3d = importlib.import_module('pose_3d.pose_model_3d')
        if cv2 is not None:
            sys.modules['cv2'] = cv2
            ns.data_helpers = importlib.import_module('pose_3d.data_helpers')
        ns.faces = np.load(os.path.join(DEPS, 'tf_smpl', 'smpl_faces.npy')).astype(np.int32)
        return ns
    finally:
        sys.path[:] = saved
        for k in [k for k in sys.modules if k.split('.')[0] in ('tensorflow', 'tf_pose', 'cv2')]:
            del sys.modules[k]      # never leave the stand-in importable as `tensorflow` for anyone else
