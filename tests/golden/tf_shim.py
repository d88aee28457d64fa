"""TEST INFRASTRUCTURE ONLY -- an eager NumPy stand-in for the TensorFlow 1.13 symbols the reference's
wavefunction classes touch, so that the UNMODIFIED reference files

    1DTFIM/RNNwavefunction.py, 1DTFIM/RNNwavefunction_paritysym.py, J1J2/ComplexRNNwavefunction.py,
    2DTFIM_1DRNN/RNNwavefunction.py, 2DTFIM_2DRNN/{MDRNNcell,RNNwavefunction}.py,
    2DHeisenberg/{Tensordot2,MDTensorizedcell,MDTensorizedGRUcell,MDComplexRNN}.py,
    RydbergKagome_2DGRU/{MDGRUcell,MDRNNWavefunction}.py

can be imported and *executed* in a container without TensorFlow (``make_wf_golden.py``).  What then runs is the
reference's own code: its sweep order, its ``str(nx)+str(ny)`` neighbour dictionaries, U(1) masks, head
combination, one-hot / stack / transpose bookkeeping, symmetry-group lists and ``rotate60``.  Every "tensor" is a
NumPy array evaluated on the spot (graph mode collapses to eager evaluation; ``tf.Graph`` / ``tf.device`` /
``tf.variable_scope`` are bookkeeping only).

What is NOT the reference's code and therefore stays a restatement from knowledge of the TF 1.13.1 sources
 -- listed here so nobody mistakes it for a pin:
  * ``tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell``  (class ``CudnnCompatibleGRUCell`` below),
  * ``tf.layers.Dense``                               (class ``Dense``),
  * ``tf.multinomial`` / ``tf.random.categorical``     (function ``categorical``: the CPU kernel's inverse-CDF rule,
    driven by uniforms the caller supplies with ``set_uniforms`` instead of TF's Philox stream),
  * the default initialisers (glorot-uniform / VarianceScaling) -- only their *shapes* matter here, the values are
    seeded NumPy draws that the fixtures record.
Everything else is a one-line NumPy equivalent of a documented element-wise / shape op.

Precision: with ``WIDE = True`` (the default) ``tf.float32`` is widened to float64, so the 1D models -- which the
reference runs in float32 with a float64 log-sum -- produce float64 fixtures an oracle can be held to at 1e-12; the
2D models are float64 upstream anyway.
"""
import contextlib
import sys
import types

import numpy as np

WIDE = True

# ----------------------------------------------------------------------------------------------------------------
# tensors
# ----------------------------------------------------------------------------------------------------------------


class _Shape(tuple):
    ndims = property(len)

    def is_fully_defined(self):
        return True

    def as_list(self):
        return [int(v) for v in self]


class Tensor(np.ndarray):
    """ndarray with the two TensorFlow methods the reference calls (``get_shape`` in Tensordot2.py:127)."""

    def get_shape(self):
        return _Shape(self.shape)

    def set_shape(self, shape):
        pass


def _t(x, dtype=None):
    a = np.asarray(x, dtype=dtype)
    return a.view(Tensor)


class _DType:
    def __init__(self, name, np_dtype):
        self.name, self.np = name, np_dtype

    def __repr__(self):
        return "tf." + self.name


float32 = _DType("float32", np.float32)
float64 = _DType("float64", np.float64)
int32 = _DType("int32", np.int32)
int64 = _DType("int64", np.int64)
complex64 = _DType("complex64", np.complex64)
complex128 = _DType("complex128", np.complex128)


def _np_dtype(dt):
    if dt is None:
        return None
    if isinstance(dt, _DType):
        if WIDE and dt is float32:
            return np.float64
        if WIDE and dt is complex64:
            return np.complex128
        return dt.np
    return np.dtype(dt)


def _shape(s):
    if isinstance(s, (int, np.integer)):
        return (int(s),)
    a = np.asarray(s)
    if a.ndim == 0:
        return (int(a),)
    return tuple(int(v) for v in a.reshape(-1))


# ----------------------------------------------------------------------------------------------------------------
# variables and scopes
# ----------------------------------------------------------------------------------------------------------------
AUTO_REUSE = "AUTO_REUSE"
VARIABLES = {}        # full name -> ndarray (creation order preserved)
_scope = []
_rng = np.random.RandomState(0)
_uniforms = None      # [B, T] uniforms consumed column by column by `categorical`
_ucol = 0


def reset(seed=0):
    """Forget every variable and restart the initialiser stream (one call per wavefunction under test)."""
    global _rng, _uniforms, _ucol
    VARIABLES.clear()
    del _scope[:]
    _rng = np.random.RandomState(seed)
    _uniforms, _ucol = None, 0


def set_uniforms(u):
    """Uniforms for the next ``sample`` graph: column t feeds the t-th categorical draw executed."""
    global _uniforms, _ucol
    _uniforms = None if u is None else np.asarray(u, dtype=np.float64)
    _ucol = 0


@contextlib.contextmanager
def variable_scope(name, reuse=None, **_):
    _scope.append(str(name))
    try:
        yield
    finally:
        _scope.pop()


@contextlib.contextmanager
def name_scope(name, default_name=None, values=None):
    yield name or default_name


def _glorot(shape, fan_in, fan_out):
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return _rng.uniform(-lim, lim, size=shape)


def _fans(shape):
    """TF's _compute_fans: rank 1 -> (n, n); rank 2 -> (in, out); else receptive field x the last two dims."""
    shape = tuple(int(s) for s in shape)
    if len(shape) == 1:
        return shape[0], shape[0]
    if len(shape) == 2:
        return shape
    rec = int(np.prod(shape[:-2]))
    return shape[-2] * rec, shape[-1] * rec


def _xavier():
    def init(shape):
        return _glorot(shape, *_fans(shape))
    return init


def _variance_scaling(scale=1.0, mode="fan_avg", distribution="uniform", **_):
    assert mode == "fan_avg" and distribution == "uniform" and scale == 1.0
    return _xavier()


def get_variable(name, shape=None, initializer=None, dtype=None, **_):
    full = "/".join(_scope + [name])
    if full not in VARIABLES:
        init = initializer if initializer is not None else _xavier()
        val = np.asarray(init(tuple(int(s) for s in shape)), dtype=np.float64)
        # the product under test computes from float32 parameters: keep the fixture weights float32-representable
        VARIABLES[full] = val.astype(np.float32).astype(np.float64)
    return _t(VARIABLES[full])


# ----------------------------------------------------------------------------------------------------------------
# ops (alphabetical); keyword names follow TF 1.13 where the reference uses keywords
# ----------------------------------------------------------------------------------------------------------------
def abs_(x):
    return _t(np.abs(x))


def add(a, b):
    return _t(np.add(a, b))


def atan(x):
    return _t(np.arctan(x))


def cast(x, dtype):
    return _t(np.asarray(x).astype(_np_dtype(dtype)))


def complex_(re, im):
    re, im = np.broadcast_arrays(np.asarray(re, dtype=np.float64), np.asarray(im, dtype=np.float64))
    return _t(re + 1j * im)


def concat(values, axis):
    return _t(np.concatenate([np.asarray(v) for v in values], axis=axis))


def constant(value=None, dtype=None, shape=None):
    a = np.asarray(value, dtype=_np_dtype(dtype))
    return _t(a.reshape(_shape(shape)) if shape is not None else a)


def convert_to_tensor(value, dtype=None, name=None):
    return _t(value, _np_dtype(dtype))


def cos(x):
    return _t(np.cos(x))


def einsum(eq, *ops):
    return _t(np.einsum(eq, *[np.asarray(o) for o in ops]))


def exp(x):
    return _t(np.exp(x))


def floordiv(a, b):
    return _t(np.floor_divide(a, b))


def floormod(a, b):
    return _t(np.mod(a, b))


def gather(params, indices):
    return _t(np.asarray(params)[np.asarray(indices)])


def identity(x):
    return _t(np.array(x, copy=True))


def imag(x):
    return _t(np.imag(x))


def log(x):
    with np.errstate(divide="ignore"):
        return _t(np.log(x))


def matmul(a, b, transpose_a=False, transpose_b=False):
    a, b = np.asarray(a), np.asarray(b)
    return _t((a.T if transpose_a else a) @ (b.T if transpose_b else b))


def multiply(a, b):
    return _t(np.multiply(a, b))


def norm(tensor, axis=None, ord=2):
    return _t(np.linalg.norm(np.asarray(tensor), ord=ord, axis=axis))


def one_hot(indices, depth, dtype=None):
    idx = np.asarray(indices).astype(np.int64)
    out = (idx[..., None] == np.arange(int(depth))).astype(_np_dtype(dtype if dtype is not None else float32))
    return _t(out)


def ones(shape, dtype=None):
    return _t(np.ones(_shape(shape), dtype=_np_dtype(dtype if dtype is not None else float32)))


def placeholder(dtype, shape=None, name=None):
    raise RuntimeError("tf_shim is eager: pass the array itself where the reference expects a placeholder")


def range_(*args, dtype=None):
    return _t(np.arange(*[int(a) for a in args], dtype=_np_dtype(dtype)))


def rank(x):
    return np.asarray(x).ndim


def real(x):
    return _t(np.real(x))


def reduce_prod(input_tensor, axis=None):
    return _t(np.prod(np.asarray(input_tensor), axis=axis))


def reduce_sum(input_tensor, axis=None):
    return _t(np.sum(np.asarray(input_tensor), axis=axis))


def reshape(tensor, shape, name=None):
    return _t(np.reshape(np.asarray(tensor), _shape(shape)))


def roll(x, shift, axis):
    return _t(np.roll(np.asarray(x), shift=tuple(shift), axis=tuple(axis)))


def rot90(image, k=1):
    """tf.image.rot90 on [batch, height, width, channels]: counter-clockwise by k*90 degrees in the (height, width) plane."""
    return _t(np.rot90(np.asarray(image), k=k, axes=(1, 2)))


def shape(input):
    return np.asarray(np.asarray(input).shape, dtype=np.int64)


def sign(x):
    return _t(np.sign(x))


def sin(x):
    return _t(np.sin(x))


def slice_(input_, begin, size):
    a = np.asarray(input_)
    idx = tuple(np.s_[int(b):] if int(s) == -1 else np.s_[int(b):int(b) + int(s)] for b, s in zip(begin, size))
    return _t(a[idx])


def sqrt(x):
    return _t(np.sqrt(x))


def stack(values, axis=0):
    return _t(np.stack([np.asarray(v) for v in values], axis=axis))


def stop_gradient(x):
    return _t(x)


def tensordot(a, b, axes):
    ax = ([axes[0]] if np.ndim(axes[0]) == 0 else list(axes[0]), [axes[1]] if np.ndim(axes[1]) == 0 else list(axes[1]))
    return _t(np.tensordot(np.asarray(a), np.asarray(b), axes=ax))


def transpose(a, perm=None):
    return _t(np.transpose(np.asarray(a), axes=perm))


def zeros(shape, dtype=None):
    return _t(np.zeros(_shape(shape), dtype=_np_dtype(dtype if dtype is not None else float32)))


def zeros_like(x):
    return _t(np.zeros_like(np.asarray(x)))


# nn -----------------------------------------------------------------------------------------------------------------
def softmax(x):
    x = np.asarray(x)
    e = np.exp(x - x.max(axis=-1, keepdims=True))
    return _t(e / e.sum(axis=-1, keepdims=True))


def sigmoid(x):
    return _t(1.0 / (1.0 + np.exp(-np.asarray(x))))


def tanh(x):
    return _t(np.tanh(x))


def relu(x):
    return _t(np.maximum(x, 0))


def elu(x):
    x = np.asarray(x)
    return _t(np.where(x > 0, x, np.expm1(np.minimum(x, 0))))


def softsign(x):
    x = np.asarray(x)
    return _t(x / (1.0 + np.abs(x)))


def l2_normalize(x, axis=None, epsilon=1e-12):
    x = np.asarray(x)
    return _t(x / np.sqrt(np.maximum(np.sum(x * x, axis=axis, keepdims=True), epsilon)))


def categorical(logits, num_samples=1, **_):
    """tf.multinomial / tf.random.categorical, CPU kernel rule: per row cdf_j = sum_{i<=j, finite}
    exp(logit_i - max), draw = number of cdf_j <= u * cdf_last.  u comes from `set_uniforms` (column = call index);
    without uniforms (the dead multinomial calls inside 2DTFIM_2DRNN's log_probability) the draw is class 0."""
    global _ucol
    assert num_samples == 1
    lg = np.asarray(logits, dtype=np.float64)
    B = lg.shape[0]
    if _uniforms is None:
        return _t(np.zeros((B, 1), dtype=np.int64))
    u = _uniforms[:, _ucol]
    _ucol += 1
    finite = np.isfinite(lg)
    m = np.where(finite, lg, -np.inf).max(axis=1, keepdims=True)
    with np.errstate(invalid="ignore"):
        e = np.where(finite, np.exp(lg - m), 0.0)
    cdf = np.cumsum(e, axis=1)
    idx = (cdf <= (u * cdf[:, -1])[:, None]).sum(axis=1)
    return _t(np.minimum(idx, lg.shape[1] - 1).astype(np.int64).reshape(B, 1))


# layers / cells -----------------------------------------------------------------------------------------------------
class Dense:
    """tf.layers.Dense: kernel [in, units] glorot-uniform, bias [units] zeros, activation(x @ kernel + bias)."""

    def __init__(self, units, activation=None, name=None, dtype=None, **_):
        self.units, self.activation, self.name = units, activation, name

    def __call__(self, x):
        with variable_scope(self.name):
            k = get_variable("kernel", [x.shape[1], self.units])
            b = get_variable("bias", [self.units], initializer=lambda s: np.zeros(s))
        y = _t(np.asarray(x) @ k + b)
        return self.activation(y) if self.activation is not None else y


class RNNCell:
    """Base of the reference's MD cells (tf.contrib.rnn.RNNCell / tf.compat.v1.nn.rnn_cell.RNNCell)."""

    def __init__(self, _reuse=None, name=None, **_):
        self._name = name

    def zero_state(self, batch_size, dtype):
        return zeros((int(batch_size), int(self.state_size)), dtype)

    def __call__(self, inputs, state):
        return self.call(inputs, state)


class CudnnCompatibleGRUCell(RNNCell):
    """tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell (TF 1.13.1, cudnn_rnn_ops.py) from knowledge of its source:
    [r|u] = sigmoid([x,h] Wg + bg) with bg initialised to 1; c = tanh(x Wci + bci + r * (h Wch + bch));
    h' = (1-u) c + u h.  The reset gate multiplies the hidden projection *after* the matmul (cuDNN convention)."""

    def __init__(self, num_units, **_):
        super().__init__(name="cudnn_compatible_gru_cell")
        self.num_units = self.state_size = num_units

    def call(self, x, h):
        nh, d_in = self.num_units, x.shape[1]
        zeros_, ones_ = (lambda s: np.zeros(s)), (lambda s: np.ones(s))
        with variable_scope("cudnn_compatible_gru_cell"):
            with variable_scope("gates"):
                Wg = get_variable("kernel", [d_in + nh, 2 * nh])
                bg = get_variable("bias", [2 * nh], initializer=ones_)
            with variable_scope("candidate"):
                with variable_scope("input_projection"):
                    Wci = get_variable("kernel", [d_in, nh])
                    bci = get_variable("bias", [nh], initializer=zeros_)
                with variable_scope("hidden_projection"):
                    Wch = get_variable("kernel", [nh, nh])
                    bch = get_variable("bias", [nh], initializer=zeros_)
        x, h = np.asarray(x), np.asarray(h)
        gates = 1.0 / (1.0 + np.exp(-(np.concatenate([x, h], axis=1) @ Wg + bg)))
        r, u = gates[:, :nh], gates[:, nh:]
        c = np.tanh(x @ Wci + bci + r * (h @ Wch + bch))
        new_h = _t((1.0 - u) * c + u * h)
        return new_h, new_h


class MultiRNNCell:
    """tf.nn.rnn_cell.MultiRNNCell: layer l+1 reads layer l's output; state = tuple of per-layer states."""

    def __init__(self, cells):
        self.cells = list(cells)

    def zero_state(self, batch_size, dtype):
        return tuple(c.zero_state(batch_size, dtype) for c in self.cells)

    def __call__(self, inputs, state):
        new = []
        with variable_scope("multi_rnn_cell"):
            for i, c in enumerate(self.cells):
                with variable_scope("cell_%d" % i):
                    inputs, s = c(inputs, state[i])
                new.append(s)
        return inputs, tuple(new)


class Graph:
    @contextlib.contextmanager
    def as_default(self):
        yield self


@contextlib.contextmanager
def device(name):
    yield


class _Device:
    def __init__(self, name, device_type):
        self.name, self.device_type = name, device_type


def _ns(**kw):
    return types.SimpleNamespace(**kw)


def install():
    """Put the stand-in into sys.modules as ``tensorflow`` (and the one sub-module the reference imports by path)."""
    tf = types.ModuleType("tensorflow")
    g = globals()
    for name in ("Tensor Graph device variable_scope get_variable AUTO_REUSE add atan cast concat constant convert_to_tensor cos "
                 "einsum exp floordiv floormod gather identity imag log matmul multiply norm one_hot ones placeholder rank real "
                 "reduce_prod reduce_sum reshape roll shape sign sin sqrt stack stop_gradient tensordot transpose zeros "
                 "zeros_like float32 float64 int32 int64 complex64 complex128").split():
        setattr(tf, name, g[name])
    tf.abs, tf.complex, tf.range, tf.slice = abs_, complex_, range_, slice_
    tf.multinomial = categorical
    tf.set_random_seed = lambda seed: None
    tf.dtypes = _ns(int32=int32, int64=int64, float32=float32, float64=float64)
    tf.math = _ns(log=log, imag=imag, real=real)
    tf.random = _ns(categorical=categorical)
    tf.image = _ns(rot90=rot90)
    tf.nn = _ns(softmax=softmax, sigmoid=sigmoid, tanh=tanh, relu=relu, elu=elu, softsign=softsign, l2_normalize=l2_normalize,
                rnn_cell=_ns(MultiRNNCell=MultiRNNCell, RNNCell=RNNCell))
    tf.layers = _ns(Dense=Dense)
    tf.contrib = _ns(cudnn_rnn=_ns(CudnnCompatibleGRUCell=CudnnCompatibleGRUCell),
                     rnn=_ns(RNNCell=RNNCell, LSTMCell=None),
                     layers=_ns(xavier_initializer=_xavier))
    v1 = _ns(variable_scope=variable_scope, AUTO_REUSE=AUTO_REUSE, get_variable=get_variable, set_random_seed=tf.set_random_seed,
             layers=_ns(Dense=Dense), nn=_ns(rnn_cell=_ns(RNNCell=RNNCell)), name_scope=name_scope,
             keras=_ns(initializers=_ns(VarianceScaling=_variance_scaling)),
             where=lambda c, a, b: _t(np.where(c, a, b)),
             setdiff1d=lambda a, b: (_t(np.setdiff1d(a, b)), None))
    tf.compat = _ns(v1=v1, integral_types=(int, np.integer))
    # `from tensorflow.python.client import device_lib` (2DHeisenberg/MDComplexRNN.py:4): one pretend GPU, so that the
    # reference's "split the symmetry copies over the available GPUs" loop runs once over the whole batch
    python = types.ModuleType("tensorflow.python")
    client = types.ModuleType("tensorflow.python.client")
    device_lib = types.ModuleType("tensorflow.python.client.device_lib")
    device_lib.list_local_devices = lambda: [_Device("/device:CPU:0", "CPU"), _Device("/device:GPU:0", "GPU")]
    client.device_lib, python.client, tf.python = device_lib, client, python
    sys.modules.update({"tensorflow": tf, "tensorflow.python": python, "tensorflow.python.client": client,
                        "tensorflow.python.client.device_lib": device_lib})
    return tf
