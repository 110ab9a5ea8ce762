"""CPU-vs-device differential harness (reference tests/tester.py:5-25): build the same op on
``ht.cpu()`` and on a device context, feed both the same random inputs, compare the results.
The CPU path (OpenMP library / torch) is the oracle for the sm_100a kernels."""
from __future__ import annotations

import numpy as np


class HetuTester(object):
    def __init__(self, op, num_inputs, *args, device_ctx=None, **kargs):
        import hetu as ht
        self.device_ctx = device_ctx if device_ctx is not None else ht.gpu(0)
        self.cpu_inputs = [ht.Variable(name='input%d' % i, ctx=ht.cpu()) for i in range(num_inputs)]
        self.cpu_op = op(*self.cpu_inputs, *args, **kargs)
        self.cpu_executor = ht.Executor([self.cpu_op], ctx=ht.cpu())
        self.gpu_inputs = [ht.Variable(name='input%d' % i, ctx=self.device_ctx) for i in range(num_inputs)]
        self.gpu_op = op(*self.gpu_inputs, *args, **kargs)
        self.gpu_executor = ht.Executor([self.gpu_op], ctx=self.device_ctx)

    def test(self, input_shapes, rtol=1e-5, atol=1e-6, gen=None, seed=0):
        rng = np.random.RandomState(seed)
        gen = gen or (lambda shape: rng.normal(size=shape))
        input_vals = [np.asarray(gen(shape), dtype=np.float32) for shape in input_shapes]
        cpu_result = self.cpu_executor.run(feed_dict=dict(zip(self.cpu_inputs, input_vals)), convert_to_numpy_ret_vals=True)
        gpu_result = self.gpu_executor.run(feed_dict=dict(zip(self.gpu_inputs, input_vals)), convert_to_numpy_ret_vals=True)
        np.testing.assert_allclose(cpu_result[0], gpu_result[0], rtol=rtol, atol=atol,
                                   err_msg='%s with shapes %s' % (self.gpu_op.name, input_shapes))
        return cpu_result[0]
