"""Parameter initializers.  reference: python/hetu/initializers.py:9-373
(curand kernels src/ops/Initializers.cu; server-side init via the InitTensor RPC).

On a GPU context values are drawn by the counter-based Philox kernel
``hetu_init_random`` (csrc/random.cu) directly into HBM - a 33.7M x 128
embedding table never exists on the host.  A model-parallel shard draws the
*full* tensor's stream and keeps its slice, so sharded and unsharded runs start
from identical weights (needed by the MP-vs-single-GPU equivalence tests).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from .ndarray import NDArray, is_gpu_ctx

K_CONST, K_UNIFORM, K_NORMAL, K_TRUNC = 0, 1, 2, 3


class BaseInit(object):
    kind = K_CONST

    def __init__(self, shape):
        self.shape = tuple(int(s) for s in shape)
        self._shard = None

    def set_shard(self, full_shape, cur_part, parts):
        self._shard = (tuple(full_shape), dict(cur_part), dict(parts))
        shp = list(full_shape)
        for i, p in parts.items():
            shp[i] //= p
        self.shape = tuple(shp)

    def args(self):
        return 0.0, 0.0

    def __call__(self, node, seed, np_rand=None, stream=None):
        self.node = node
        self.seed = (seed or 0) + node.id
        dev = node.ctx.torch_device
        if self._shard is not None:
            full_shape, cur_part, parts = self._shard
            full = self._draw(full_shape, dev, np_rand)
            slc = []
            for i, n in enumerate(full_shape):
                if i in parts:
                    part = n // parts[i]
                    st = part * cur_part.get(i, 0)
                    slc.append(slice(st, st + part))
                else:
                    slc.append(slice(0, n))
            t = full[tuple(slc)].contiguous()
        else:
            t = self._draw(self.shape, dev, np_rand)
        node.tensor_value = NDArray(t, node.ctx)

    def _draw(self, shape, dev, np_rand):
        a, b = self.args()
        if dev.type == 'cuda' and L.cuda_lib() is not None:
            t = torch.empty(shape, dtype=torch.float32, device=dev)
            L.call('init_random', L.p(t), L.c_sz(t.numel()), L.c_int(self.kind), L.c_f32(a), L.c_f32(b),
                   L.c_u64(self.seed & 0xFFFFFFFFFFFFFFFF), L.stream_ptr(dev))
            return t
        return self.init_on_cpu(shape, np_rand).to(dev)

    def init_on_cpu(self, shape, np_rand):
        raise NotImplementedError

    def init_on_ps(self, comm, nid, param_type, seed=0, opt=None):
        """Server-side initialisation (ParamInit PS function)."""
        if param_type == 0:
            length, width = int(np.prod(self.shape)), 1
        else:
            assert len(self.shape) == 2
            length, width = self.shape
        a, b = self.args()
        comm.InitTensor(nid, param_type, length, width, self.kind, float(a), float(b), int(seed), opt)


class ConstantInit(BaseInit):
    kind = K_CONST

    def __init__(self, constant, shape):
        super().__init__(shape)
        self.constant = constant

    def args(self):
        return float(self.constant), 1.0

    def init_on_cpu(self, shape, np_rand):
        return torch.full(shape, float(self.constant), dtype=torch.float32)


class ZerosInit(ConstantInit):
    def __init__(self, shape):
        super().__init__(0.0, shape)


class OnesInit(ConstantInit):
    def __init__(self, shape):
        super().__init__(1.0, shape)


class UniformInit(BaseInit):
    kind = K_UNIFORM

    def __init__(self, low, high, shape):
        super().__init__(shape)
        self.low, self.high = low, high

    def args(self):
        return float(self.low), float(self.high)

    def init_on_cpu(self, shape, np_rand):
        rng = np_rand if np_rand is not None else np.random
        return torch.from_numpy(rng.uniform(self.low, self.high, size=shape).astype(np.float32))


def _fan(shape, mode):
    assert mode in ('fan_in', 'fan_out', 'avg'), 'Mode %s not valid.' % mode
    assert len(shape) >= 2, 'General xavier requires shape to be at least 2D.'
    hw_scale = 1 if len(shape) == 2 else int(np.prod(shape[2:]))
    fan_in, fan_out = hw_scale * shape[1], hw_scale * shape[0]
    return {'fan_in': fan_in, 'fan_out': fan_out, 'avg': (fan_in + fan_out) / 2.0}[mode]


class GeneralXavierUniformInit(UniformInit):
    def __init__(self, gain, mode, shape):
        assert gain > 0, 'Gain value %s not valid.' % str(gain)
        limit = float(np.sqrt(gain / _fan(shape, mode)))
        super().__init__(-limit, limit, shape)


class XavierUniformInit(GeneralXavierUniformInit):
    def __init__(self, shape):
        super().__init__(3.0, 'avg', shape)


class HeUniformInit(GeneralXavierUniformInit):
    def __init__(self, shape):
        super().__init__(6.0, 'fan_in', shape)


class LecunUniformInit(GeneralXavierUniformInit):
    def __init__(self, shape):
        super().__init__(3.0, 'fan_in', shape)


class NormalInit(BaseInit):
    kind = K_NORMAL

    def __init__(self, mean, stddev, shape):
        super().__init__(shape)
        self.mean, self.stddev = mean, stddev

    def args(self):
        return float(self.mean), float(self.stddev)

    def init_on_cpu(self, shape, np_rand):
        rng = np_rand if np_rand is not None else np.random
        return torch.from_numpy(rng.normal(self.mean, self.stddev, size=shape).astype(np.float32))


class GeneralXavierNormalInit(NormalInit):
    def __init__(self, gain, mode, shape):
        assert gain > 0, 'Gain value not allowed.'
        super().__init__(0, float(np.sqrt(gain / _fan(shape, mode))), shape)


class XavierNormalInit(GeneralXavierNormalInit):
    def __init__(self, shape):
        super().__init__(1.0, 'avg', shape)


class HeNormalInit(GeneralXavierNormalInit):
    def __init__(self, shape):
        super().__init__(2.0, 'fan_in', shape)


class LecunNormalInit(GeneralXavierNormalInit):
    def __init__(self, shape):
        super().__init__(1.0, 'fan_in', shape)


class TruncatedNormalInit(NormalInit):
    kind = K_TRUNC

    def init_on_cpu(self, shape, np_rand):
        rng = np_rand if np_rand is not None else np.random
        out = rng.normal(0.0, 1.0, size=shape)
        bad = np.abs(out) > 2.0
        while bad.any():
            out[bad] = rng.normal(0.0, 1.0, size=int(bad.sum()))
            bad = np.abs(out) > 2.0
        return torch.from_numpy((out * self.stddev + self.mean).astype(np.float32))


# ----------------------------------------------------------- user-facing helpers
def _var(init, name, trainable, ctx, default_name):
    from .ops.variable import Variable
    return Variable(name=name or default_name, initializer=init, trainable=trainable, ctx=ctx)


def zeros(shape, name=None, trainable=True, ctx=None):
    return _var(ZerosInit(shape), name, trainable, ctx, 'zeros_initializer')


def ones(shape, name=None, trainable=True, ctx=None):
    return _var(OnesInit(shape), name, trainable, ctx, 'ones_initializer')


def constant(shape, fill_value=0.0, name=None, trainable=True, ctx=None):
    return _var(ConstantInit(fill_value, shape), name, trainable, ctx, 'constant_initializer')


def truncated_normal(shape, mean=0.0, stddev=1.0, name=None, trainable=True, ctx=None):
    return _var(TruncatedNormalInit(mean, stddev, shape), name, trainable, ctx,
                'truncated_normal_initializer')


def random_normal(shape, mean=0.0, stddev=1.0, name=None, trainable=True, ctx=None):
    return _var(NormalInit(mean, stddev, shape), name, trainable, ctx, 'random_normal_initializer')


def random_uniform(shape, minval=-1.0, maxval=1.0, name=None, trainable=True, ctx=None):
    return _var(UniformInit(minval, maxval, shape), name, trainable, ctx, 'random_uniform_initializer')


def general_xavier_normal(shape, gain, mode, name=None, trainable=True, ctx=None):
    return _var(GeneralXavierNormalInit(gain, mode, shape), name, trainable, ctx,
                'general_xavier_normal_initializer')


def general_xavier_uniform(shape, gain, mode, name=None, trainable=True, ctx=None):
    return _var(GeneralXavierUniformInit(gain, mode, shape), name, trainable, ctx,
                'general_xavier_uniform_initializer')


def xavier_normal(shape, name=None, trainable=True, ctx=None):
    return _var(XavierNormalInit(shape), name, trainable, ctx, 'xavier_normal_initializer')


def xavier_uniform(shape, name=None, trainable=True, ctx=None):
    return _var(XavierUniformInit(shape), name, trainable, ctx, 'xavier_uniform_initializer')


def he_normal(shape, name=None, trainable=True, ctx=None):
    return _var(HeNormalInit(shape), name, trainable, ctx, 'he_normal_initializer')


def he_uniform(shape, name=None, trainable=True, ctx=None):
    return _var(HeUniformInit(shape), name, trainable, ctx, 'he_uniform_initializer')


def lecun_normal(shape, name=None, trainable=True, ctx=None):
    return _var(LecunNormalInit(shape), name, trainable, ctx, 'lecun_normal_initializer')


def lecun_uniform(shape, name=None, trainable=True, ctx=None):
    return _var(LecunUniformInit(shape), name, trainable, ctx, 'lecun_uniform_initializer')


def _generate(init_func, **init_kargs):
    def _generator_helper(shape, name=None, trainable=True, ctx=None):
        return init_func(shape=shape, name=name, trainable=trainable, ctx=ctx, **init_kargs)
    return _generator_helper


def GenZeros():
    return _generate(zeros)


def GenOnes():
    return _generate(ones)


def GenConstant(fill_value=0.0):
    return _generate(constant, fill_value=fill_value)


def GenTruncatedNormal(mean=0.0, stddev=1.0):
    return _generate(truncated_normal, mean=mean, stddev=stddev)


def GenNormal(mean=0.0, stddev=1.0):
    return _generate(random_normal, mean=mean, stddev=stddev)


def GenUniform(minval=-1.0, maxval=1.0):
    return _generate(random_uniform, minval=minval, maxval=maxval)


def GenGeneralXavierNormal(gain, mode):
    return _generate(general_xavier_normal, gain=gain, mode=mode)


def GenGeneralXavierUniform(gain, mode):
    return _generate(general_xavier_uniform, gain=gain, mode=mode)


def GenXavierNormal():
    return _generate(xavier_normal)


def GenXavierUniform():
    return _generate(xavier_uniform)


def GenHeNormal():
    return _generate(he_normal)


def GenHeUniform():
    return _generate(he_uniform)


def GenLecunNormal():
    return _generate(lecun_normal)


def GenLecunUniform():
    return _generate(lecun_uniform)
