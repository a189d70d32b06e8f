"""SURREAL `*_info.mat` -> device batches for the SMPL path.

Host-side I/O (scipy.io), mirroring the annotation half of read_maps_poses_images_surreal
(deps/pose_3d/data_helpers.py:43-89): `pose [72,T]`, `shape [10,T]`, `joints2D [2,24,T]`, `zrot [T,1]` -> arrays with
time first, the horizontal flip of the 2D ground truth (:66-67), the validity mask and the every-2nd-frame
subsampling (:80-87).  The heat-map / JPEG half of that function (`read_maps_frames_surreal`) is the CNN encoder's
input: it is read for completeness of the reader but nothing here processes it (the encoder is out of scope).
`gt_batch` then runs what the reference's graph does with these annotations before the losses
(pose_model_3d.py:371-372 rotate_global_pose, :177-181 SMPL) on the kernels."""
import numpy as np
import torch

from . import api


def read_maps_mask_surreal(maps_file):
    """The frame mask of a SURREAL heat-map file as data_helpers.py:44-50 forms it: squeeze(mask) AND (diffs < 250),
    written into the mask array itself, so the dtype is the file's.  (NaN diffs compare False, :48.)  Returns
    (mask, img_size_x = heat-map width, :57).  The heat maps themselves are the CNN's input and are not loaded."""
    import scipy.io
    m = scipy.io.loadmat(maps_file, variable_names=['mask', 'diffs'])
    mask = np.squeeze(m['mask'])
    diffs = np.squeeze(m['diffs'])
    with np.errstate(invalid='ignore'):
        np.logical_and(mask, diffs < 250, out=mask)
    w = None
    for name, shape, _ in scipy.io.whosmat(maps_file):
        if name == 'heat_mat':
            w = int(shape[1])
    return mask, w


# left/right swap of the 19 heat-map channels for the mirrored SURREAL images (data_helpers.py:53)
SURREAL_REORDER = [0, 1, 5, 6, 7, 2, 3, 4, 11, 12, 13, 8, 9, 10, 15, 14, 17, 16, 18]


def read_maps_frames_surreal(maps_file, frames_path, n_joints=14, mask=None, skip=2):
    """The heat-map / frame half of read_maps_poses_images_surreal (data_helpers.py:44-57,70-87): heat maps
    [T,H,W,n_joints] (time first, mirrored horizontally, left/right channels swapped) concatenated with the RGB frames
    `f*.jpg` of `frames_path` (sorted; each min-max normalised to [0,1] like cv2.normalize(NORM_MINMAX); mirrored), then
    the same `x[mask]` / every-`skip`-th-frame selection as the annotations.  This is the CNN encoder's input -- host
    I/O only, decoded with OpenCV when it is installed, else PIL."""
    import glob
    import os
    import scipy.io
    heat = np.transpose(scipy.io.loadmat(maps_file, variable_names=['heat_mat'])['heat_mat'], (3, 0, 1, 2))   # :46
    heat = np.flip(heat, axis=2)[:, :, :, SURREAL_REORDER][:, :, :, :n_joints]                                  # :50-56
    fp = frames_path.decode('utf-8') if isinstance(frames_path, bytes) else frames_path
    files = sorted(glob.glob(os.path.join(fp, 'f*.jpg')))                                                      # :71
    try:
        import cv2
        frames = [cv2.cvtColor(cv2.imread(f), cv2.COLOR_BGR2RGB) for f in files]
    except ImportError:
        from PIL import Image
        frames = [np.array(Image.open(f).convert('RGB')) for f in files]
    frames = np.array(frames, dtype=np.float32)                                                                 # :73
    norm = []
    for fr in frames:                                                                                           # :74-75
        lo, hi = fr.min(), fr.max()
        norm.append((fr - lo) / max(float(hi - lo), 1e-12))
    frames = np.flip(norm, axis=2)                                                                              # :77
    concat = np.concatenate([heat, frames], axis=3)                                                             # :79
    if mask is not None:
        concat = concat[np.asarray(mask)]                                                                       # :81
    return np.ascontiguousarray(concat[::skip], dtype=np.float32)                                               # :85


def read_info_surreal(info_file, img_size_x=320, mask=None, skip=2):
    """Returns dict(poses [N,72], shapes [N,10], joints2d [N,24,2], zrot [N]) float32 (data_helpers.py:58-88).

    `mask` indexes the frames exactly as the reference's `x[mask]` does (:80-82), with NumPy's rules: a bool array
    selects frames; an INTEGER array -- which is what scipy.io.loadmat returns for a MATLAB logical (uint8), hence
    what read_maps_mask_surreal and the reference get from a SURREAL file -- is fancy indexing, i.e. one row per
    frame holding frame 0 or frame 1.  That is the reference's behaviour on such files and it is reproduced, not
    repaired; pass `mask.astype(bool)` for the selection the author intended."""
    import scipy.io
    info = scipy.io.loadmat(info_file)
    poses = np.transpose(info['pose'], (1, 0))                                   # :62
    shapes = np.transpose(info['shape'], (1, 0))                                 # :63
    joints2d = np.transpose(info['joints2D'], (2, 1, 0)).astype(np.float32)      # :65
    joints2d[:, :, 0] = img_size_x - joints2d[:, :, 0]                           # :67 (SURREAL images are mirrored)
    zrot = np.squeeze(np.array(info['zrot']), axis=1) if np.ndim(info['zrot']) == 2 else np.array(info['zrot'])   # :68
    if mask is not None:                                                         # :80-82
        mask = np.asarray(mask)
        poses, shapes, joints2d, zrot = poses[mask], shapes[mask], joints2d[mask], zrot[mask]
    poses, shapes, joints2d, zrot = poses[::skip], shapes[::skip], joints2d[::skip], zrot[::skip]   # :84-87
    return dict(poses=np.ascontiguousarray(poses, dtype=np.float32), shapes=np.ascontiguousarray(shapes, dtype=np.float32),
                joints2d=np.ascontiguousarray(joints2d, dtype=np.float32),
                zrot=np.ascontiguousarray(zrot, dtype=np.float32).reshape(-1))


def gt_batch(model, ann, device=None, want_verts=True):
    """Annotations -> ground-truth tensors on the GPU: pose with the y-up global rotation (utils.py:152-167 on the
    kernel), SMPL meshes / posed joints (batch_smpl.py), 2D joints flipped to the renderer's y-up convention
    (pose_model_3d.py:196-197)."""
    dev = model.device if device is None else device
    theta = api.rotate_global_pose(torch.from_numpy(ann['poses']).to(dev), torch.from_numpy(ann['zrot']).to(dev))
    beta = torch.from_numpy(ann['shapes']).to(dev)
    out = model.forward(beta, theta, want_verts=want_verts, want_kp=False, want_Rs=False)
    j2d = torch.from_numpy(ann['joints2d']).to(dev)
    j2d_flipped = j2d * torch.tensor([1.0, -1.0], device=dev) + torch.tensor([0.0, 240.0], device=dev)
    return dict(theta=theta, beta=beta, joints3d=out['joints'], verts=out.get('verts'), joints2d=j2d,
                joints2d_flipped=j2d_flipped)
