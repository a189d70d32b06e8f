"""ctypes binding of libsmplb200.so (declared in include/smplb200.h).

There is NO CPU fallback: if the shared library is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libsmplb200.so')

OK = 0
E_INVALID_ARG, E_BAD_SHAPE, E_CUDA, E_NOT_FINITE, E_UNSUPPORTED = -1, -2, -3, -4, -5
F_BLEND_SIMT, F_KEEP_VPOSED, F_NO_SKIN, F_SKIN_SIMT, F_SCRATCH_VERTS, F_FUSED = 1, 2, 4, 8, 16, 32

NV, NV3, NJ, NB, NP, KDIM, KPAD, VP_PITCH = 6890, 20670, 24, 10, 207, 217, 224, 20736

c_float_p = C.POINTER(C.c_float)


class CamCfg(C.Structure):
    _fields_ = [('img_w', C.c_float), ('img_h', C.c_float), ('ss', C.c_float),
                ('focal_const', C.c_float)]


class ForwardIO(C.Structure):
    _fields_ = [('B', C.c_int32),
                ('beta', C.c_void_p), ('theta', C.c_void_p),
                ('verts', C.c_void_p), ('joints', C.c_void_p), ('kp', C.c_void_p),
                ('Rs', C.c_void_p), ('A', C.c_void_p),
                ('cam_pos', C.c_void_p), ('cam_rot', C.c_void_p), ('focal', C.c_void_p),
                ('uv', C.c_void_p), ('px', C.c_void_p),
                ('cfg', CamCfg)]


class LossCfg(C.Structure):
    _fields_ = [('cam', CamCfg), ('reproj_scale', C.c_float), ('reproj_cap', C.c_float),
                ('joint_scale', C.c_float), ('mesh_scale', C.c_float)]


class BackwardIO(C.Structure):
    _fields_ = [('B', C.c_int32),
                ('beta', C.c_void_p), ('theta', C.c_void_p),
                ('g_verts', C.c_void_p), ('g_joints', C.c_void_p), ('g_kp', C.c_void_p),
                ('g_px', C.c_void_p),
                ('cam_pos', C.c_void_p), ('cam_rot', C.c_void_p), ('focal', C.c_void_p),
                ('cfg', CamCfg),
                ('g_beta', C.c_void_p), ('g_theta', C.c_void_p)]


class FitIO(C.Structure):
    _fields_ = [('B', C.c_int32),
                ('beta', C.c_void_p), ('theta', C.c_void_p),
                ('cam_pos', C.c_void_p), ('cam_rot', C.c_void_p), ('focal', C.c_void_p),
                ('gt_px', C.c_void_p), ('px', C.c_void_p), ('g_px', C.c_void_p),
                ('g_beta', C.c_void_p), ('g_theta', C.c_void_p),
                ('m_theta', C.c_void_p), ('v_theta', C.c_void_p),
                ('m_beta', C.c_void_p), ('v_beta', C.c_void_p),
                ('loss', C.c_void_p), ('scratch', C.c_void_p),
                ('cfg', LossCfg), ('n_total', C.c_int64), ('step', C.c_int32),
                ('lr', C.c_float), ('beta1', C.c_float), ('beta2', C.c_float), ('eps', C.c_float),
                ('step_dev', C.c_void_p)]


class PhongIO(C.Structure):
    _fields_ = [('B', C.c_int32), ('H', C.c_int32), ('W', C.c_int32), ('C', C.c_int32), ('L', C.c_int32),
                ('off_normal', C.c_int32), ('off_position', C.c_int32), ('off_diffuse', C.c_int32),
                ('off_specular', C.c_int32), ('off_shininess', C.c_int32), ('normalize_normals', C.c_int32),
                ('attr', C.c_void_p), ('alphas', C.c_void_p), ('shininess', C.c_void_p),
                ('light_positions', C.c_void_p), ('light_intensities', C.c_void_p), ('camera', C.c_void_p),
                ('ambient', C.c_void_p)]


# name -> (restype, argtypes); every symbol include/smplb200.h declares
SYMBOLS = {
    'smplb_last_error': (C.c_char_p, []),
    'smplb_version': (C.c_char_p, []),
    'smplb_model_create': (C.c_int32, [C.POINTER(C.c_void_p), C.c_int32, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                       C.c_void_p]),
    'smplb_model_destroy': (C.c_int32, [C.c_void_p]),
    'smplb_model_nk': (C.c_int32, [C.c_void_p]),
    'smplb_workspace_bytes': (C.c_size_t, [C.c_void_p, C.c_int32, C.c_int32]),
    'smplb_forward': (C.c_int32, [C.c_void_p, C.POINTER(ForwardIO), C.c_void_p, C.c_size_t,
                                  C.c_int32, C.c_void_p]),
    'smplb_forward_status': (C.c_int32, [C.c_void_p, C.c_void_p]),
    'smplb_forward_host': (C.c_int32, [C.c_void_p, C.POINTER(ForwardIO), C.POINTER(ForwardIO),
                                       C.c_void_p, C.c_size_t, C.c_int32, C.c_void_p]),
    'smplb_rodrigues': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    'smplb_global_rigid': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_project': (C.c_int32, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    'smplb_orth_proj': (C.c_int32, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    'smplb_reproj_loss': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.POINTER(LossCfg),
                                      C.c_void_p, C.c_void_p]),
    'smplb_backward': (C.c_int32, [C.c_void_p, C.POINTER(BackwardIO), C.c_void_p, C.c_size_t,
                                   C.c_int32, C.c_void_p]),
    'smplb_project_backward': (C.c_int32, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_fit_scratch_bytes': (C.c_size_t, [C.c_int32]),
    'smplb_fit_step': (C.c_int32, [C.c_void_p, C.POINTER(FitIO), C.c_void_p]),
    'smplb_mesh_create': (C.c_int32, [C.POINTER(C.c_void_p), C.c_int32, C.c_void_p, C.c_int32, C.c_int32]),
    'smplb_mesh_destroy': (C.c_int32, [C.c_void_p]),
    'smplb_rasterize_workspace_bytes': (C.c_size_t, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
    'smplb_rasterize': (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    'smplb_rasterize_backward': (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                             C.c_size_t, C.c_void_p]),
    'smplb_mesh_normals': (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_render_workspace_bytes': (C.c_size_t, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
    'smplb_render_diffuse': (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    'smplb_clip_transform': (C.c_int32, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_clip_transform_backward': (C.c_int32, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                  C.c_void_p, C.c_void_p]),
    'smplb_interpolate': (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    'smplb_interpolate_backward': (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_phong_scratch_bytes': (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    'smplb_phong_shade': (C.c_int32, [C.POINTER(PhongIO), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    'smplb_phong_shade_backward': (C.c_int32, [C.POINTER(PhongIO), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_size_t, C.c_void_p]),
    'smplb_add_axis_angle': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_rotate_global_pose': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_cam_losses_scratch_bytes': (C.c_size_t, [C.c_int32]),
    'smplb_cam_losses': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.POINTER(CamCfg), C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_cam_losses_static': (C.c_int32, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.POINTER(CamCfg), C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p]),
    'smplb_mse_scratch_bytes': (C.c_size_t, []),
    'smplb_mse_grad': (C.c_int32, [C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_int64, C.c_void_p, C.c_void_p,
                                   C.c_int32, C.c_void_p, C.c_void_p]),
    'smplb_timing_begin': (C.c_int32, [C.c_int32]),
    'smplb_timing_end': (C.c_int32, [C.c_void_p, C.POINTER(C.c_int32)]),
    'smplb_timing_bwd_end': (C.c_int32, [C.c_void_p, C.POINTER(C.c_int32)]),
    'smplb_adam_tick': (C.c_int32, [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    'smplb_adam_step_dev': (C.c_int32, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_float, C.c_float, C.c_float, C.c_void_p]),
    'smplb_adam_step': (C.c_int32, [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int32, C.c_float, C.c_float, C.c_float, C.c_float,
                                    C.c_void_p]),
}

_lib = None


class SmplbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('smplb200 error %d: %s' % (code, msg))
        self.code = code


def lib():
    """Load the CUDA library (once).  Raises if it has not been built — no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                'libsmplb200.so not found at %s: build it with `python -c "import __graft_entry__ as g; '
                'g.build()"` or `make -C 3d-pose-estimation_b200/csrc` (there is no CPU fallback)' % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != OK:
        raise SmplbError(rc, lib().smplb_last_error().decode('utf-8', 'replace'))
