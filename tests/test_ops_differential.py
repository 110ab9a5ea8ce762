"""Op-level differential tests (reference tests/test_ops.py + tester.py): the same graph node on
the CPU executor and on the device executor, same inputs, results must agree.  On the host-only
run the "device" is a second CPU executor, which still exercises graph build, shape inference
and the executor for every op; on a B200 (``-m gpu``) the device side is the sm_100a kernels."""
import numpy as np
import pytest

import hetu as ht
from hetu_61a7_b200.utils.tester import HetuTester


def _cases(dev, mm_tol):
    def T(op, n, *a, **k):
        return HetuTester(op, n, *a, device_ctx=dev, **k)

    T(ht.addbyconst_op, 1, 0.37).test([(2, 3, 4, 5, 6)])
    t = T(ht.add_op, 2)
    t.test([(2, 3, 4, 5), (2, 3, 4, 5)])
    t = T(ht.broadcastto_op, 2)
    t.test([(200, 300), (13, 200, 300)])
    t.test([(3, 1), (2, 3, 5)])

    def cat(*args, axis=0, **kargs):
        return ht.concatenate_op(list(args), axis=axis, **kargs)
    t = T(cat, 3, axis=2)
    t.test([(2, 3, 4, 5), (2, 3, 7, 5), (2, 3, 6, 5)])
    t.test([(2, 3, 4), (2, 3, 7), (2, 3, 6)])
    T(cat, 2, axis=0).test([(31, 4, 5), (23, 4, 5)])

    def total(*args, **kargs):
        return ht.sum_op(list(args), **kargs)
    T(total, 3).test([(5, 7), (5, 7), (5, 7)])
    for op in (ht.relu_op, ht.sigmoid_op, ht.tanh_op, ht.exp_op, ht.abs_op, ht.opposite_op, ht.sin_op, ht.cos_op, ht.floor_op):
        T(op, 1).test([(37, 41)], rtol=1e-4, atol=1e-5)
    T(ht.gelu_op, 1).test([(37, 41)], rtol=2e-3, atol=2e-3)
    pos = lambda shape: np.abs(np.random.RandomState(1).normal(size=shape)) + 0.5     # noqa: E731
    for op in (ht.sqrt_op, ht.rsqrt_op, ht.log_op):
        T(op, 1).test([(19, 23)], gen=pos, rtol=1e-4, atol=1e-5)
    for op in (ht.mul_op, ht.minus_op):
        T(op, 2).test([(8, 9, 10), (8, 9, 10)])
    T(ht.div_op, 2).test([(8, 9), (8, 9)], gen=pos, rtol=1e-4)
    T(ht.mul_byconst_op, 1, -2.5).test([(64, 33)])
    T(ht.matmul_op, 2).test([(64, 48), (48, 32)], rtol=mm_tol, atol=mm_tol)
    T(ht.matmul_op, 2, trans_A=True, trans_B=True).test([(48, 64), (32, 48)], rtol=mm_tol, atol=mm_tol)
    T(ht.batch_matmul_op, 2).test([(4, 16, 24), (4, 24, 8)], rtol=mm_tol, atol=mm_tol)
    T(ht.linear_op, 3).test([(32, 48), (48, 16), (16,)], rtol=mm_tol, atol=mm_tol)
    T(ht.reduce_sum_op, 1, [0]).test([(17, 29)], rtol=1e-4, atol=1e-4)
    T(ht.reduce_mean_op, 1, [1]).test([(17, 29)], rtol=1e-4, atol=1e-5)
    T(ht.reducesumaxiszero_op, 1).test([(33, 12)], rtol=1e-4, atol=1e-4)
    T(ht.transpose_op, 1, [1, 0, 2]).test([(5, 6, 7)])
    T(ht.array_reshape_op, 1, (6, -1)).test([(3, 4, 5)])
    T(ht.slice_op, 1, (1, 2), (2, 3)).test([(4, 6)])
    T(ht.softmax_op, 1).test([(13, 50)], rtol=1e-4, atol=1e-6)
    T(ht.layer_normalization_op, 3, eps=1e-5).test([(12, 64), (64,), (64,)], rtol=1e-3, atol=1e-4)
    T(ht.conv2d_op, 2, padding=1, stride=1).test([(2, 3, 8, 8), (4, 3, 3, 3)], rtol=mm_tol, atol=mm_tol)
    T(ht.max_pool2d_op, 1, 2, 2, 0, 2).test([(2, 3, 8, 8)])
    T(ht.avg_pool2d_op, 1, 2, 2, 0, 2).test([(2, 3, 8, 8)], rtol=1e-4, atol=1e-6)
    onehot = lambda shape: np.eye(shape[1])[np.random.RandomState(2).randint(0, shape[1], shape[0])]     # noqa: E731
    t = T(ht.softmaxcrossentropy_op, 2)
    rs = np.random.RandomState(3)
    vals = [rs.normal(size=(9, 11)).astype(np.float32), onehot((9, 11)).astype(np.float32)]
    a = t.cpu_executor.run(feed_dict=dict(zip(t.cpu_inputs, vals)), convert_to_numpy_ret_vals=True)[0]
    b = t.gpu_executor.run(feed_dict=dict(zip(t.gpu_inputs, vals)), convert_to_numpy_ret_vals=True)[0]
    np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-5)
    t = T(ht.embedding_lookup_op, 2)
    vals = [rs.normal(size=(50, 8)).astype(np.float32), rs.randint(0, 50, (6, 3)).astype(np.float32)]
    a = t.cpu_executor.run(feed_dict=dict(zip(t.cpu_inputs, vals)), convert_to_numpy_ret_vals=True)[0]
    b = t.gpu_executor.run(feed_dict=dict(zip(t.gpu_inputs, vals)), convert_to_numpy_ret_vals=True)[0]
    np.testing.assert_allclose(a, b)
    np.testing.assert_allclose(a, vals[0][vals[1].astype(np.int64)])


def test_ops_cpu_executor_pair():
    _cases(ht.cpu(0), 1e-4)


@pytest.mark.gpu
def test_ops_cpu_vs_b200():
    _cases(ht.gpu(0), 5e-2)
