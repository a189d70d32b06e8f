"""The two scenes of the reference's renderer tests (deps/tf_mesh_renderer/mesh_renderer_test.py:33-166) as NumPy
inputs, shared by the oracle test and the GPU parity test."""
import numpy as np

CUBE_V = np.array([[-1, -1, 1], [-1, -1, -1], [-1, 1, -1], [-1, 1, 1], [1, -1, 1], [1, -1, -1], [1, 1, -1], [1, 1, 1]],
                  np.float32)                                                                    # :36-39
CUBE_N = CUBE_V / np.linalg.norm(CUBE_V, axis=1, keepdims=True)                                  # :40
CUBE_T = np.array([[0, 1, 2], [2, 3, 0], [3, 2, 6], [6, 7, 3], [7, 6, 5], [5, 4, 7], [4, 5, 1], [1, 0, 4], [5, 6, 2],
                   [2, 1, 5], [7, 4, 0], [0, 3, 7]], np.int32)                                   # :41-44
ANGLES = np.array([[-20.0, 0.0, 60.0], [45.0, 60.0, 0.0]], np.float32)                           # :51-52


def world(euler_matrices):
    M = euler_matrices(ANGLES)[:, :3, :3]
    v = np.matmul(np.stack([CUBE_V, CUBE_V]), np.transpose(M, (0, 2, 1)))                        # :53-60
    n = np.matmul(np.stack([CUBE_N, CUBE_N]), np.transpose(M, (0, 2, 1)))
    return v.astype(np.float32), n.astype(np.float32)


def gray_cube(euler_matrices):
    """testRendersSimpleCube (:45-88) -> Gray_Cube_{0,1}.png"""
    v, n = world(euler_matrices)
    return dict(vertices=v, triangles=CUBE_T, normals=n, diffuse_colors=np.ones_like(v),
                camera_position=np.array(2 * [[0.0, 0.0, 6.0]], np.float32),
                camera_lookat=np.zeros((2, 3), np.float32), camera_up=np.array(2 * [[0.0, 1.0, 0.0]], np.float32),
                light_positions=np.array([[[0.0, 0.0, 6.0]], [[0.0, 0.0, 6.0]]], np.float32),
                light_intensities=np.ones((2, 1, 3), np.float32), image_width=640, image_height=480)


def colored_cube(euler_matrices, scalar_shininess=False):
    """testComplexShading (:90-166) -> Colored_Cube_{0,1}.png after tone_mapper(gamma 0.7)"""
    v, n = world(euler_matrices)
    diff = np.array(2 * [[[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 1], [1, 1, 0], [1, 0, 1], [0, 1, 1], [.5, .5, .5]]], np.float32)
    spec = np.array(2 * [[[0, 1, 0], [0, 0, 1], [1, 1, 1], [1, 1, 0], [1, 0, 1], [0, 1, 1], [.5, .5, .5], [1, 0, 0]]], np.float32)
    return dict(vertices=v, triangles=CUBE_T, normals=n, diffuse_colors=diff,
                camera_position=np.array([[0.0, 0.0, 6.0], [0., 0.2, 18.0]], np.float32),
                camera_lookat=np.array([[0.0, 0.0, 0.0], [0.1, -0.1, 0.1]], np.float32),
                camera_up=np.array([[0.0, 1.0, 0.0], [0.1, 1.0, 0.15]], np.float32),
                light_positions=np.array([[[0.0, 0.0, 6.0], [1.0, 2.0, 6.0]], [[0.0, -2.0, 4.0], [1.0, 3.0, 4.0]]], np.float32),
                light_intensities=np.array([[[1, 1, 1], [1, 1, 1]], [[2, 0, 1], [0, 2, 1]]], np.float32),
                image_width=640, image_height=480, specular_colors=spec,
                shininess_coefficients=6.0 if scalar_shininess else 6.0 * np.ones((2, 8), np.float32),
                ambient_color=np.array([[0., 0., 0.], [0.1, 0.1, 0.2]], np.float32),
                fov_y=np.array([40., 13.3], np.float32), near_clip=np.float32(0.1), far_clip=np.float32(25.0))


def images_near(result, baseline_u8, max_outlier_fraction=0.001, pixel_error_threshold=0.01):
    """test_utils.py:80-121: at most 0.1 % of the pixels may differ by more than 1 % of full scale in any channel."""
    result = np.clip(result, 0.0, 1.0)
    base = baseline_u8.astype(float) / 255.0
    outlier = np.any(np.abs(base - result) > pixel_error_threshold, axis=2)
    frac = np.count_nonzero(outlier) / np.prod(base.shape[:2])
    return frac <= max_outlier_fraction, frac
