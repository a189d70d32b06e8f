"""Human3.6M TFRecords -> device batches for the SMPL path (second half).

Mirrors deps/pose_3d/data_helpers.py:28-40,92-129 without TensorFlow: the records are the tf.train.Example files
published with HMR (github.com/akanazawa/hmr), which the reference reads with tf.data.TFRecordDataset +
tf.parse_single_example (:92-119).  Here: the TFRecord framing (length, masked CRC-32C, payload, masked CRC-32C) and the
protobuf wire format of tf.train.Example are decoded directly (both formats are public and tiny), so the loader needs
neither TensorFlow nor protobuf.  Host-side I/O like surreal.py; `h36m_batch` puts the annotation fields the hot path
consumes (pose [72], shape [10], 2D joints [14,2]) on the GPU.  The JPEG (`image/encoded`) and the heat maps feed the
CNN encoder, which is out of scope: they are returned (decoded with PIL when it is importable) but not processed."""
import struct

import numpy as np

# ---- CRC-32C (Castagnoli), the checksum of the TFRecord framing ---------------------------------------------------
_CRC_TABLE = None


def _crc_table():
    global _CRC_TABLE
    if _CRC_TABLE is None:
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            t.append(c)
        _CRC_TABLE = t
    return _CRC_TABLE


def crc32c(data):
    t = _crc_table()
    c = 0xFFFFFFFF
    for b in bytes(data):
        c = t[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data):
    """TFRecord's masked CRC: rotate right by 15 and add a constant (tensorflow/core/lib/hash/crc32c.h)."""
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def tfrecord_iter(path, check_crc=True):
    """Yields the payload of every record of a TFRecord file (what tf.data.TFRecordDataset yields, :29)."""
    with open(path, 'rb') as f:
        while True:
            head = f.read(12)
            if not head:
                return
            if len(head) < 12:
                raise ValueError('%s: truncated record header' % path)
            n, = struct.unpack('<Q', head[:8])
            if check_crc and struct.unpack('<I', head[8:])[0] != masked_crc32c(head[:8]):
                raise ValueError('%s: corrupt record length (CRC mismatch)' % path)
            data = f.read(n)
            tail = f.read(4)
            if len(data) < n or len(tail) < 4:
                raise ValueError('%s: truncated record' % path)
            if check_crc and struct.unpack('<I', tail)[0] != masked_crc32c(data):
                raise ValueError('%s: corrupt record payload (CRC mismatch)' % path)
            yield data


def write_tfrecord(path, records):
    """The inverse framing (used by the tests to build fixtures)."""
    with open(path, 'wb') as f:
        for r in records:
            head = struct.pack('<Q', len(r))
            f.write(head + struct.pack('<I', masked_crc32c(head)) + bytes(r) + struct.pack('<I', masked_crc32c(r)))


# ---- protobuf wire format of tf.train.Example ------------------------------------------------------------------------
def _varint(buf, i):
    r, s = 0, 0
    while True:
        b = buf[i]
        i += 1
        r |= (b & 0x7F) << s
        if not b & 0x80:
            return r, i
        s += 7


def _fields(buf):
    """(field number, wire type, value) of one message; length-delimited values come back as memoryviews."""
    i, n = 0, len(buf)
    while i < n:
        key, i = _varint(buf, i)
        num, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(buf, i)
        elif wt == 1:
            v, i = buf[i:i + 8], i + 8
        elif wt == 2:
            ln, i = _varint(buf, i)
            v, i = buf[i:i + ln], i + ln
        elif wt == 5:
            v, i = buf[i:i + 4], i + 4
        else:
            raise ValueError('unsupported protobuf wire type %d' % wt)
        yield num, wt, v


def _feature(buf):
    """Feature { oneof kind { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3 } }"""
    for num, wt, v in _fields(buf):
        if wt != 2:
            continue
        if num == 1:      # BytesList { repeated bytes value = 1 }
            return 'bytes', [bytes(x) for n2, w2, x in _fields(v) if n2 == 1]
        if num == 2:      # FloatList { repeated float value = 1 [packed] }
            vals = []
            for n2, w2, x in _fields(v):
                if n2 == 1:
                    vals.append(np.frombuffer(bytes(x), dtype='<f4'))
            return 'float', (np.concatenate(vals) if vals else np.zeros(0, np.float32))
        if num == 3:      # Int64List { repeated int64 value = 1 [packed] }
            vals = []
            for n2, w2, x in _fields(v):
                if n2 != 1:
                    continue
                if w2 == 0:
                    vals.append(x)
                else:
                    j, m = 0, memoryview(x)
                    while j < len(m):
                        q, j = _varint(m, j)
                        vals.append(q)
            a = np.array(vals, dtype=np.uint64).astype(np.int64)      # two's complement for negative values
            return 'int64', a
    return None, None


def parse_example(record):
    """tf.train.Example bytes -> {key: (kind, values)} (Example.features = 1, Features.feature map = 1, entry key = 1 /
    value = 2)."""
    out = {}
    buf = memoryview(bytes(record))
    for num, wt, feats in _fields(buf):
        if num != 1 or wt != 2:
            continue
        for n2, w2, entry in _fields(feats):
            if n2 != 1 or w2 != 2:
                continue
            key, val = None, None
            for n3, w3, x in _fields(entry):
                if n3 == 1:
                    key = bytes(x).decode('utf-8')
                elif n3 == 2:
                    val = _feature(x)
            if key is not None and val is not None:
                out[key] = val
    return out


# the FixedLenFeature table of read_tfrecord_h36m (data_helpers.py:94-108): key -> (element count, kind)
H36M_KEYS = {'image/center': (2, 'int64'), 'image/encoded': (1, 'bytes'), 'image/filename': (1, 'bytes'),
             'image/format': (1, 'bytes'), 'image/height': (1, 'int64'), 'image/visibility': (14, 'int64'),
             'image/width': (1, 'int64'), 'image/x': (14, 'float'), 'image/y': (14, 'float'),
             'meta/crop_pt': (2, 'int64'), 'meta/has_3d': (1, 'int64'), 'meta/scale_factors': (2, 'float'),
             'mosh/gt3d': (42, 'float'), 'mosh/pose': (72, 'float'), 'mosh/shape': (10, 'float')}


def read_tfrecord_h36m(record, decode_image=True):
    """data_helpers.py:92-119: (img, poses [72], shapes [10], joints2d [14,2]).  Like tf.parse_single_example with
    FixedLenFeature, a missing key or a wrong element count is an error.  img: HxWx3 uint8 when PIL can decode it
    (tf.image.decode_jpeg, :118), else the encoded bytes."""
    f = parse_example(record)
    for key, (n, kind) in H36M_KEYS.items():
        if key not in f:
            raise ValueError('read_tfrecord_h36m: feature %r missing' % key)
        k, v = f[key]
        if k != kind or len(v) != n:
            raise ValueError('read_tfrecord_h36m: feature %r is %s[%d], expected %s[%d]' % (key, k, len(v), kind, n))
    poses = np.asarray(f['mosh/pose'][1], dtype=np.float32)
    shapes = np.asarray(f['mosh/shape'][1], dtype=np.float32)
    joints2d = np.stack([f['image/x'][1], f['image/y'][1]], axis=1).astype(np.float32)      # :115
    img = f['image/encoded'][1][0]
    if decode_image:
        try:
            import io
            from PIL import Image
            img = np.array(Image.open(io.BytesIO(img)))
        except ImportError:
            pass
    return img, poses, shapes, joints2d


def read_maps_h36m(maps_file, n_joints=14):
    """data_helpers.py:122-127: heat maps [T,H,W,n_joints] of a predict_h36m_tfrecords.py file."""
    import scipy.io
    heatmaps = np.transpose(scipy.io.loadmat(maps_file)['heat_mat'], (3, 0, 1, 2))
    return heatmaps[:, :, :, :n_joints]


def concat_img_heatmaps_h36m(heatmaps, img, poses, shapes, joints2d):
    """data_helpers.py:130-131 (zrot is 0 for H36M)."""
    return np.concatenate([heatmaps, img.astype(heatmaps.dtype)], axis=-1), poses, shapes, joints2d, 0.0


def read_annotations_h36m(record_files, limit=None):
    """All records of the files -> dict(poses [N,72], shapes [N,10], joints2d [N,14,2]) float32 (no image decode)."""
    P, S, J = [], [], []
    for path in ([record_files] if isinstance(record_files, (str, bytes)) else record_files):
        for rec in tfrecord_iter(path):
            _, p, s, j = read_tfrecord_h36m(rec, decode_image=False)
            P.append(p); S.append(s); J.append(j)
            if limit is not None and len(P) >= limit:
                break
    return dict(poses=np.stack(P), shapes=np.stack(S), joints2d=np.stack(J))


def h36m_batch(model, ann, device=None, want_verts=True):
    """Annotations -> ground-truth tensors on the GPU (H36M poses need no global-rotation fix: zrot = 0, :131)."""
    import torch
    dev = model.device if device is None else device
    theta, beta = torch.from_numpy(ann['poses']).to(dev), torch.from_numpy(ann['shapes']).to(dev)
    out = model.forward(beta, theta, want_verts=want_verts, want_kp=True, want_Rs=False)
    return dict(theta=theta, beta=beta, joints3d=out['joints'], keypoints=out['kp'], verts=out.get('verts'),
                joints2d=torch.from_numpy(ann['joints2d']).to(dev))
