"""TEST INFRASTRUCTURE ONLY (oracle): a stand-in for the `tensorflow` module, just large enough to EXECUTE THE
REFERENCE'S OWN PYTHON SOURCE (deps/tf_smpl, deps/tf_perspective_projection, deps/pose_3d/utils.py, the loss graph
of deps/pose_3d/pose_model_3d.py, deps/tf_mesh_renderer/*.py) eagerly on torch-CPU tensors.

Why it exists: TensorFlow cannot be installed in this image, and the reference has no test for the SMPL / projection /
loss path, so the oracle restatements (oracle/ref_numpy.py, ref_torch.py) would otherwise be checked only against
themselves.  With this module first on sys.path, `oracle/run_reference.py` imports the reference's files where they
lie under /root/reference and runs them unchanged; what is replaced is TensorFlow's op library, not the algorithm.
Each function below implements the documented TF-1.x semantics of the op of the same name (only the argument
forms the reference uses).  `tf.gradients` is torch autograd, so the reference's graph is also differentiated.

Float width: TF's default float is float32 and the reference hard-codes tf.float32.  `set_float64(True)` makes every
float this module creates -- and the name tf.float32 itself -- double precision, so the reference's statement sequence
can be evaluated at a precision where a restatement error cannot hide in rounding noise.  Off by default.

Nothing here is imported by the product path.
"""
import builtins as _builtins
import contextlib
import types

import numpy as np
import torch

newaxis = None
int32 = torch.int32
int64 = torch.int64
float64 = torch.float64
float32 = torch.float32          # rebound by set_float64
string = 'string'
bool = torch.bool                # noqa: A001  (tf.bool)
_FLOAT = torch.float32


def set_float64(on):
    global float32, _FLOAT
    _FLOAT = torch.float64 if on else torch.float32
    float32 = _FLOAT


class Dimension:
    """tf.Dimension: `.value`, int()-able, comparable."""

    def __init__(self, v):
        self.value = None if v is None else int(v)

    def __int__(self):
        return self.value

    __index__ = __int__

    def __eq__(self, o):
        return self.value == (o.value if isinstance(o, Dimension) else o)

    def __hash__(self):
        return hash(self.value)

    def __repr__(self):
        return 'Dimension(%r)' % self.value


class TensorShape:
    def __init__(self, dims):
        self._d = [int(d) for d in dims]

    def as_list(self):
        return list(self._d)

    @property
    def ndims(self):
        return len(self._d)

    def __len__(self):
        return len(self._d)

    def __getitem__(self, i):
        if isinstance(i, slice):
            return TensorShape(self._d[i])
        return Dimension(self._d[i])

    def __iter__(self):
        return iter(Dimension(d) for d in self._d)

    def __eq__(self, o):
        return self._d == (o._d if isinstance(o, TensorShape) else [int(x) for x in o])

    def __repr__(self):
        return 'TensorShape(%r)' % (self._d,)


def _raw(x, dtype=None):
    """anything -> torch tensor (python / numpy floats become the module float type, ints int32, like tf.constant)"""
    if isinstance(x, Tensor):
        t = x.t
    elif isinstance(x, torch.Tensor):
        t = x
    elif isinstance(x, (list, tuple)) and any(isinstance(e, (Tensor, torch.Tensor)) for e in x):
        t = torch.stack([_raw(e) for e in x])
    else:
        a = np.asarray(x)
        if a.dtype.kind == 'f':
            t = torch.tensor(a.astype(np.float64)).to(_FLOAT)
        elif a.dtype.kind in 'iu':
            t = torch.tensor(a.astype(np.int64)).to(torch.int32)
        elif a.dtype.kind == 'b':
            t = torch.tensor(a)
        else:
            raise TypeError('tf shim: cannot convert %r' % (a.dtype,))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t


def _pair(a, b):
    """operands of a binary op: python scalars stay scalars (weak typing), ints meet floats as floats"""
    ta = a.t if isinstance(a, Tensor) else a
    tb = b.t if isinstance(b, Tensor) else b
    if not isinstance(ta, torch.Tensor) and not isinstance(ta, (int, float)):
        ta = _raw(ta)
    if not isinstance(tb, torch.Tensor) and not isinstance(tb, (int, float)):
        tb = _raw(tb)
    if isinstance(ta, torch.Tensor) and isinstance(tb, torch.Tensor) and ta.dtype != tb.dtype:
        if ta.is_floating_point() and tb.is_floating_point():
            raise TypeError('tf shim: float width mismatch %s vs %s' % (ta.dtype, tb.dtype))
    return ta, tb


def _int(x):
    if isinstance(x, Tensor):
        return int(x.t.item())
    if isinstance(x, torch.Tensor):
        return int(x.item())
    return int(x)


def _shape(s):
    if isinstance(s, Tensor):
        s = s.t
    if isinstance(s, torch.Tensor):
        return [int(v) for v in s.reshape(-1).tolist()]
    if isinstance(s, TensorShape):
        return s.as_list()
    if isinstance(s, (list, tuple)):
        return [_int(e) for e in s]
    return [_int(s)]


class Tensor:
    __array_priority__ = 1000
    __array_ufunc__ = None

    def __init__(self, t):
        self.t = t

    # --- introspection ---
    @property
    def shape(self):
        return TensorShape(self.t.shape)

    def get_shape(self):
        return self.shape

    @property
    def dtype(self):
        return self.t.dtype

    def numpy(self):
        return self.t.detach().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.t.detach().numpy()
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return self.t.shape[0]

    def __int__(self):
        return int(self.t.item())

    __index__ = __int__

    def __float__(self):
        return float(self.t.item())

    def __bool__(self):
        return builtins_bool(self.t.item())

    def __repr__(self):
        return 'tfshim.Tensor(%r)' % (self.t,)

    def __iter__(self):
        return iter(Tensor(self.t[i]) for i in _builtins.range(self.t.shape[0]))

    def __hash__(self):
        return id(self)

    # --- indexing ---
    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        idx = tuple(_index(i) for i in idx)
        return Tensor(self.t[idx])

    # --- arithmetic ---
    def __add__(self, o):
        a, b = _pair(self, o)
        return Tensor(a + b)

    def __radd__(self, o):
        a, b = _pair(o, self)
        return Tensor(a + b)

    def __sub__(self, o):
        a, b = _pair(self, o)
        return Tensor(a - b)

    def __rsub__(self, o):
        a, b = _pair(o, self)
        return Tensor(a - b)

    def __mul__(self, o):
        a, b = _pair(self, o)
        return Tensor(a * b)

    def __rmul__(self, o):
        a, b = _pair(o, self)
        return Tensor(a * b)

    def __truediv__(self, o):
        a, b = _pair(self, o)
        return Tensor(a / b)

    def __rtruediv__(self, o):
        a, b = _pair(o, self)
        return Tensor(a / b)

    def __neg__(self):
        return Tensor(-self.t)

    def __pow__(self, o):
        a, b = _pair(self, o)
        return Tensor(a ** b)

    def __matmul__(self, o):
        return matmul(self, o)

    # --- comparisons (element-wise, like tf.Tensor's overloaded < > <= >=; == is identity in TF1 but the
    #     reference only ever writes tf.equal) ---
    def __lt__(self, o):
        a, b = _pair(self, o)
        return Tensor(a < b)

    def __le__(self, o):
        a, b = _pair(self, o)
        return Tensor(a <= b)

    def __gt__(self, o):
        a, b = _pair(self, o)
        return Tensor(a > b)

    def __ge__(self, o):
        a, b = _pair(self, o)
        return Tensor(a >= b)


builtins_bool = _builtins.bool


def _index(i):
    if isinstance(i, Tensor):
        return i.t.long() if not i.t.dtype == torch.bool else i.t
    if isinstance(i, Dimension):
        return int(i)
    if isinstance(i, slice):
        return slice(*(None if v is None else _int(v) for v in (i.start, i.stop, i.step)))
    return i


# ------------------------------------------------------------------ creation
def constant(value, dtype=None, shape=None, name=None):
    t = _raw(value, dtype)
    if shape is not None:
        t = t.reshape(_shape(shape)) if t.numel() > 1 else t.reshape(-1)[0].expand(_shape(shape)).clone()
    return Tensor(t.clone() if isinstance(value, (Tensor, torch.Tensor)) else t)


convert_to_tensor = constant


def Variable(initial_value, trainable=True, name=None, dtype=None):
    t = _raw(initial_value, dtype).clone().detach()
    if trainable and t.is_floating_point():
        t.requires_grad_(True)
    return Tensor(t)


def zeros(shape, dtype=None, name=None):
    return Tensor(torch.zeros(_shape(shape), dtype=dtype or _FLOAT))


def ones(shape, dtype=None, name=None):
    return Tensor(torch.ones(_shape(shape), dtype=dtype or _FLOAT))


def zeros_like(x, dtype=None, name=None):
    return Tensor(torch.zeros_like(_raw(x), dtype=dtype))


def ones_like(x, dtype=None, name=None):
    return Tensor(torch.ones_like(_raw(x), dtype=dtype))


def eye(num_rows, num_columns=None, batch_shape=None, dtype=None, name=None):
    e = torch.eye(_int(num_rows), _int(num_columns if num_columns is not None else num_rows), dtype=dtype or _FLOAT)
    if batch_shape is not None:
        bs = _shape(batch_shape)
        e = e.expand(*bs, *e.shape).clone()
    return Tensor(e)


def range(start, limit=None, delta=1, dtype=None, name=None):  # noqa: A001
    if limit is None:
        start, limit = 0, start
    return Tensor(torch.arange(_int(start), _int(limit), _int(delta), dtype=dtype or torch.int32))


def shape(x, name=None):
    return Tensor(torch.tensor(list(_raw(x).shape), dtype=torch.int32))


def cast(x, dtype, name=None):
    return Tensor(_raw(x).to(dtype))


def stop_gradient(x, name=None):
    return Tensor(_raw(x).detach())


def identity(x, name=None):
    return Tensor(_raw(x))


# ------------------------------------------------------------------ shape manipulation
def reshape(tensor, shape, name=None):  # noqa: A002
    return Tensor(_raw(tensor).reshape(_shape(shape)))


def expand_dims(x, axis=None, name=None, dim=None):
    return Tensor(_raw(x).unsqueeze(axis if axis is not None else dim))


def squeeze(x, axis=None, name=None):
    t = _raw(x)
    if axis is None:
        return Tensor(t.squeeze())
    for a in sorted(axis if isinstance(axis, (list, tuple)) else [axis], reverse=True):
        t = t.squeeze(a)
    return Tensor(t)


def stack(values, axis=0, name=None):
    ts = [_raw(v) for v in values]
    fl = [t for t in ts if t.is_floating_point()]
    if fl and len(fl) != len(ts):
        ts = [t.to(fl[0].dtype) for t in ts]
    return Tensor(torch.stack(ts, dim=axis))


def concat(values, axis, name=None):
    return Tensor(torch.cat([_raw(v) for v in values], dim=axis))


def tile(x, multiples, name=None):
    return Tensor(_raw(x).repeat(*_shape(multiples)))


def pad(tensor, paddings, mode='CONSTANT', name=None, constant_values=0):
    t = _raw(tensor)
    p = [[_int(a), _int(b)] for a, b in paddings]
    flat = []
    for a, b in reversed(p):          # torch wants last dimension first
        flat += [a, b]
    return Tensor(torch.nn.functional.pad(t, flat, value=constant_values))


def transpose(a, perm=None, name=None):
    t = _raw(a)
    if perm is None:
        perm = list(_builtins.range(t.dim()))[::-1]
    return Tensor(t.permute(*[_int(p) for p in perm]))


def matrix_transpose(a, name=None):
    return Tensor(_raw(a).transpose(-1, -2))


def reverse(tensor, axis, name=None):
    return Tensor(torch.flip(_raw(tensor), dims=[_int(a) for a in axis]))


def gather(params, indices, validate_indices=None, name=None, axis=0):
    """tf.gather: out.shape = params.shape[:axis] + indices.shape + params.shape[axis+1:]; out-of-range indices are
    an error on CPU (InvalidArgumentError), which is what is raised here too."""
    p = _raw(params)
    i = _raw(indices).long()
    axis = _int(axis) % p.dim()
    if i.numel() and (int(i.min()) < 0 or int(i.max()) >= p.shape[axis]):
        raise IndexError('tf.gather: indices out of range [0, %d)' % p.shape[axis])
    out = torch.index_select(p, axis, i.reshape(-1))
    return Tensor(out.reshape(list(p.shape[:axis]) + list(i.shape) + list(p.shape[axis + 1:])))


def scatter_nd(indices, updates, shape, name=None):  # noqa: A002
    """tf.scatter_nd for index depth 1 (the only form used, batch_lbs.py:33): zeros(shape) with updates ADDED at
    indices (duplicates accumulate)."""
    idx = _raw(indices).long()
    upd = _raw(updates)
    shp = _shape(shape)
    if idx.shape[-1] != 1 or len(shp) != 1:
        raise NotImplementedError('tf shim: scatter_nd only for 1-D targets')
    out = torch.zeros(shp, dtype=upd.dtype)
    return Tensor(out.index_add(0, idx.reshape(-1), upd.reshape(-1)))


def meshgrid(*args, **kw):
    indexing = kw.get('indexing', 'xy')
    return [Tensor(t) for t in torch.meshgrid(*[_raw(a) for a in args], indexing=indexing)]


def where(condition, x=None, y=None, name=None):
    """TF1 tf.where: x and y have the same shape; a rank-1 condition picks whole ROWS of higher-rank x / y."""
    c = _raw(condition)
    if x is None:
        return Tensor(torch.nonzero(c))
    tx, ty = _raw(x), _raw(y)
    if tx.shape != ty.shape:
        raise ValueError('tf.where (TF1): x and y must have the same shape, got %s / %s' % (tx.shape, ty.shape))
    if c.shape != tx.shape:
        if c.dim() != 1 or c.shape[0] != tx.shape[0]:
            raise ValueError('tf.where (TF1): condition must match x or be a vector over its first dimension')
        c = c.reshape([-1] + [1] * (tx.dim() - 1))
    return Tensor(torch.where(c, tx, ty))


# ------------------------------------------------------------------ math
def matmul(a, b, transpose_a=False, transpose_b=False, name=None):
    ta, tb = _pair(a, b)
    if transpose_a:
        ta = ta.transpose(-1, -2)
    if transpose_b:
        tb = tb.transpose(-1, -2)
    return Tensor(torch.matmul(ta, tb))


def _unary(fn):
    def op(x, name=None):
        return Tensor(fn(_raw(x)))
    return op


sin, cos, tan, acos, exp, sign, sqrt, reciprocal, square, log = (
    _unary(f) for f in (torch.sin, torch.cos, torch.tan, torch.acos, torch.exp, torch.sign, torch.sqrt,
                        torch.reciprocal, torch.square, torch.log))
abs = _unary(torch.abs)  # noqa: A001


def _binary(fn):
    def op(x, y, name=None):
        a, b = _pair(x, y)
        if not isinstance(a, torch.Tensor):
            a = torch.as_tensor(a, dtype=b.dtype)
        if not isinstance(b, torch.Tensor):
            b = torch.as_tensor(b, dtype=a.dtype)
        return Tensor(fn(a, b))
    return op


add, subtract, multiply, divide, div, minimum, maximum, atan2, equal, less, greater, not_equal = (
    _binary(f) for f in (torch.add, torch.sub, torch.mul, torch.div, torch.div, torch.minimum, torch.maximum,
                         torch.atan2, torch.eq, torch.lt, torch.gt, torch.ne))
pow = _binary(torch.pow)  # noqa: A001
truediv = divide


def cross(a, b, name=None):
    ta, tb = _pair(a, b)
    return Tensor(torch.linalg.cross(ta, tb, dim=-1))


def clip_by_value(t, clip_value_min, clip_value_max, name=None):
    x = _raw(t)
    lo, hi = clip_value_min, clip_value_max
    lo = _raw(lo, x.dtype) if not isinstance(lo, (int, float)) else lo
    hi = _raw(hi, x.dtype) if not isinstance(hi, (int, float)) else hi
    return Tensor(torch.minimum(torch.maximum(x, torch.as_tensor(lo, dtype=x.dtype)), torch.as_tensor(hi, dtype=x.dtype)))


def _axes(axis):
    if axis is None:
        return None
    if isinstance(axis, (list, tuple)):
        return [_int(a) for a in axis]
    return [_int(axis)]


def _reduce(fn):
    def op(x, axis=None, keepdims=False, name=None, keep_dims=None, reduction_indices=None):
        t = _raw(x)
        kd = keepdims if keep_dims is None else keep_dims
        ax = _axes(axis if axis is not None else reduction_indices)
        if ax is None:
            ax = list(_builtins.range(t.dim()))
        return Tensor(fn(t, ax, kd))
    return op


reduce_sum = _reduce(lambda t, ax, kd: t.sum(dim=ax, keepdim=kd))
reduce_mean = _reduce(lambda t, ax, kd: t.mean(dim=ax, keepdim=kd))
reduce_max = _reduce(lambda t, ax, kd: t.amax(dim=ax, keepdim=kd))
reduce_min = _reduce(lambda t, ax, kd: t.amin(dim=ax, keepdim=kd))
reduce_any = _reduce(lambda t, ax, kd: (t.to(torch.int32).sum(dim=ax, keepdim=kd) > 0))
count_nonzero = _reduce(lambda t, ax, kd: (t != 0).to(torch.int64).sum(dim=ax, keepdim=kd))


def norm(tensor, ord='euclidean', axis=None, keepdims=None, name=None, keep_dims=None):  # noqa: A002
    """tf.norm, Euclidean only: sqrt(reduce_sum(x * x, axis))."""
    if ord not in ('euclidean', 2):
        raise NotImplementedError
    t = _raw(tensor)
    kd = builtins_bool(keepdims if keepdims is not None else (keep_dims or False))
    ax = _axes(axis)
    if ax is None:
        ax = list(_builtins.range(t.dim()))
    return Tensor(torch.sqrt((t * t).sum(dim=ax, keepdim=kd)))


# ------------------------------------------------------------------ assertions / control flow (eager: checked at once)
def assert_equal(x, y, message=None, name=None, **kw):
    a, b = _pair(x, y)
    if not builtins_bool(torch.all(torch.as_tensor(a) == torch.as_tensor(b))):
        raise AssertionError(message or 'tf.assert_equal failed')


def assert_greater(x, y, message=None, name=None, **kw):
    a, b = _pair(x, y)
    if not builtins_bool(torch.all(torch.as_tensor(a) > torch.as_tensor(b))):
        raise AssertionError(message or 'tf.assert_greater failed')
    return no_op()


def Assert(condition, data, summarize=None, name=None):
    if not builtins_bool(_raw(condition).all()):
        raise AssertionError('tf.Assert failed')


def no_op(name=None):
    return None


@contextlib.contextmanager
def _null_scope(*a, **k):
    yield None


variable_scope = name_scope = control_dependencies = device = _null_scope


def map_fn(fn, elems, dtype=None, parallel_iterations=None, back_prop=True, swap_memory=False, infer_shape=True,
           name=None):
    """tf.map_fn over the first axis of elems (a tensor or a tuple of tensors); outputs stacked per position."""
    single = not isinstance(elems, (list, tuple))
    es = [elems] if single else list(elems)
    n = _raw(es[0]).shape[0]
    outs = []
    for i in _builtins.range(n):
        arg = Tensor(_raw(es[0])[i]) if single else tuple(Tensor(_raw(e)[i]) for e in es)
        outs.append(fn(arg))
    if isinstance(outs[0], (list, tuple)):
        return tuple(Tensor(torch.stack([_raw(o[k]) for o in outs])) for k in _builtins.range(len(outs[0])))
    return Tensor(torch.stack([_raw(o) for o in outs]))


def gradients(ys, xs, grad_ys=None, name=None, **kw):
    ys = ys if isinstance(ys, (list, tuple)) else [ys]
    single = not isinstance(xs, (list, tuple))
    xs = [xs] if single else list(xs)
    gy = None
    if grad_ys is not None:
        gy = [_raw(g) for g in (grad_ys if isinstance(grad_ys, (list, tuple)) else [grad_ys])]
    else:
        gy = [torch.ones_like(_raw(y)) for y in ys]
    g = torch.autograd.grad([_raw(y) for y in ys], [_raw(x) for x in xs], gy, allow_unused=True, retain_graph=True)
    out = [None if v is None else Tensor(v) for v in g]
    return out


# ------------------------------------------------------------------ tf.nn / tf.losses / tf.summary
def _l2_normalize(x, axis=None, epsilon=1e-12, name=None, dim=None):
    """x * rsqrt(max(sum(x**2, axis), epsilon))"""
    t = _raw(x)
    ax = _axes(axis if axis is not None else dim)
    ss = (t * t).sum(dim=ax, keepdim=True)
    return Tensor(t * torch.rsqrt(torch.clamp(ss, min=epsilon)))


def _softmax(logits, axis=-1, name=None, dim=None):
    return Tensor(torch.softmax(_raw(logits), dim=_int(axis if dim is None else dim)))


nn = types.SimpleNamespace(l2_normalize=_l2_normalize, softmax=_softmax)


def _mean_squared_error(labels, predictions, weights=1.0, scope=None, loss_collection=None, reduction=None):
    """tf.losses.mean_squared_error with the default SUM_BY_NONZERO_WEIGHTS reduction: sum(w * (p - l)^2) divided by
    the number of elements whose (broadcast) weight is non-zero."""
    p, l = _pair(predictions, labels)
    if p.shape != l.shape:
        raise ValueError('mean_squared_error: shapes %s / %s differ' % (p.shape, l.shape))
    sq = (p - l) ** 2
    w = torch.as_tensor(weights.t if isinstance(weights, Tensor) else weights, dtype=sq.dtype)
    wb = torch.broadcast_to(w, sq.shape)
    present = (wb != 0).to(sq.dtype).sum()
    total = (sq * wb).sum()
    return Tensor(torch.where(present > 0, total / torch.clamp(present, min=1.0), torch.zeros_like(total)))


losses = types.SimpleNamespace(mean_squared_error=_mean_squared_error)

SUMMARIES = {}      # name -> last value passed to tf.summary.scalar / image (lets tests read the individual loss terms)


def _summary_scalar(name, tensor, collections=None, family=None):
    SUMMARIES[name] = tensor


def _summary_image(name, tensor, max_outputs=3, collections=None, family=None):
    SUMMARIES[name] = tensor


summary = types.SimpleNamespace(scalar=_summary_scalar, image=_summary_image, merge_all=lambda *a, **k: None,
                                FileWriter=lambda *a, **k: None)


# ------------------------------------------------------------------ graph-mode furniture the reference touches on import
class _Graph:
    def as_default(self):
        return _null_scope()


_GRAPH = _Graph()


def get_default_graph():
    return _GRAPH


def reset_default_graph():
    SUMMARIES.clear()


def RegisterGradient(op_type):
    def deco(fn):
        REGISTERED_GRADIENTS[op_type] = fn
        return fn
    return deco


REGISTERED_GRADIENTS = {}
OP_LIBRARY_LOADER = None     # set by oracle/run_reference.py: path -> module object providing the custom op


def load_op_library(path):
    if OP_LIBRARY_LOADER is None:
        raise RuntimeError('tf shim: no custom-op loader installed for %s' % path)
    return OP_LIBRARY_LOADER(path)


def executing_eagerly():
    return True


def enable_eager_execution(*a, **k):
    return None


keras = types.ModuleType('tensorflow.keras')
test = types.SimpleNamespace(TestCase=object, main=lambda: None)


# ---- tf.train.Example parsing (pose_3d/data_helpers.py:92-119 read_tfrecord_h36m) -----------------------------------
# TensorFlow documents tf.parse_single_example(serialized, {key: FixedLenFeature(shape, dtype)}) as: decode the
# tf.train.Example protobuf, return every requested feature reshaped to `shape`.  The decoding here is done by the
# OFFICIAL protobuf runtime on the published schema (tensorflow/core/example/{example,feature}.proto), built as dynamic
# messages -- deliberately not by the product's own wire-format parser (smplb200/h36m.py), which this pins.
class FixedLenFeature:
    def __init__(self, shape, dtype, default_value=None):
        self.shape, self.dtype, self.default_value = list(shape), dtype, default_value


def example_proto_classes():
    """(Example, Features, Feature) message classes of the public tf.train.Example schema."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fd = descriptor_pb2.FileDescriptorProto(name='smplb_tf_example.proto', package='tensorflow', syntax='proto3')
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name):
        m = fd.message_type.add()
        m.name = name
        return m

    def field(m, name, number, ftype, label=T.LABEL_OPTIONAL, type_name=None, oneof=None):
        f = m.field.add()
        f.name, f.number, f.type, f.label = name, number, ftype, label
        if type_name:
            f.type_name = type_name
        if oneof is not None:
            f.oneof_index = oneof
        return f

    field(msg('BytesList'), 'value', 1, T.TYPE_BYTES, T.LABEL_REPEATED)
    field(msg('FloatList'), 'value', 1, T.TYPE_FLOAT, T.LABEL_REPEATED)      # proto3: packed
    field(msg('Int64List'), 'value', 1, T.TYPE_INT64, T.LABEL_REPEATED)
    ft = msg('Feature')
    ft.oneof_decl.add().name = 'kind'
    field(ft, 'bytes_list', 1, T.TYPE_MESSAGE, type_name='.tensorflow.BytesList', oneof=0)
    field(ft, 'float_list', 2, T.TYPE_MESSAGE, type_name='.tensorflow.FloatList', oneof=0)
    field(ft, 'int64_list', 3, T.TYPE_MESSAGE, type_name='.tensorflow.Int64List', oneof=0)
    fs = msg('Features')
    entry = fs.nested_type.add()
    entry.name = 'FeatureEntry'
    entry.options.map_entry = True
    field(entry, 'key', 1, T.TYPE_STRING)
    field(entry, 'value', 2, T.TYPE_MESSAGE, type_name='.tensorflow.Feature')
    field(fs, 'feature', 1, T.TYPE_MESSAGE, T.LABEL_REPEATED, type_name='.tensorflow.Features.FeatureEntry')
    field(msg('Example'), 'features', 1, T.TYPE_MESSAGE, type_name='.tensorflow.Features')
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName('tensorflow.' + n))
    return get('Example'), get('Features'), get('Feature')


def parse_single_example(serialized, features, name=None, example_names=None):
    Example, _, _ = example_proto_classes()
    ex = Example()
    ex.ParseFromString(bytes(serialized))
    out = {}
    for key, spec in features.items():
        if key not in ex.features.feature:
            raise ValueError('feature %r missing from the Example' % key)
        f = ex.features.feature[key]
        kind = f.WhichOneof('kind')
        vals = list(getattr(f, kind).value)
        n = int(np.prod(spec.shape)) if spec.shape else 1
        if len(vals) != n:
            raise ValueError('feature %r: %d values, FixedLenFeature wants %d' % (key, len(vals), n))
        if spec.dtype == string:
            out[key] = vals[0] if not spec.shape else np.array(vals, dtype=object)
        elif kind == 'int64_list':
            out[key] = Tensor(torch.tensor(vals, dtype=torch.int64).reshape(spec.shape))
        else:
            out[key] = Tensor(torch.tensor(vals, dtype=torch.float32).reshape(spec.shape))
    return out


def _decode_jpeg(contents, channels=0, **kw):
    import io
    from PIL import Image
    return Tensor(torch.from_numpy(np.array(Image.open(io.BytesIO(bytes(contents))))))


image = types.SimpleNamespace(decode_jpeg=_decode_jpeg)
