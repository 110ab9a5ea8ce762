"""Mixed placement (reference tests/test_datatransfer_op.py): a parameter that lives on the host inside a GPU
executor - transfer nodes are inserted around it, the forward value and the updated host parameter are checked
against numpy.  Late-collected file."""
import numpy as np
import pytest

import hetu as ht


def _dense(exec_ctx, rtol):
    rs = np.random.RandomState(0)
    npw = rs.random_sample((5, 10)).astype(np.float32)
    npx = rs.random_sample((7, 5)).astype(np.float32)
    X = ht.Variable(name='x')
    mid = X + 3
    W = ht.Variable(name='w_host', value=npw, ctx=ht.cpu(0))
    y = ht.matmul_op(mid, W)
    train_op = ht.optim.SGDOptimizer(learning_rate=0.1).minimize(y)
    executor = ht.Executor([y, train_op], ctx=exec_ctx)
    pred_y, _ = executor.run(feed_dict={X: npx}, convert_to_numpy_ret_vals=True)
    np.testing.assert_allclose(pred_y, np.matmul(npx + 3, npw), rtol=rtol)
    new_w = npw - 0.1 * np.matmul((npx + 3).T, np.ones((7, 10), dtype=np.float32))
    got = executor.config.placeholder_to_arr_map[W]
    got = got.asnumpy() if hasattr(got, 'asnumpy') else got.detach().cpu().numpy()
    np.testing.assert_allclose(got, new_w, rtol=rtol, atol=1e-5)


def _sparse(exec_ctx, rtol):
    rs = np.random.RandomState(1)
    npemb = rs.random_sample((100, 20)).astype(np.float32)
    npind = rs.randint(100, size=(10,))
    npw = rs.random_sample((20, 30)).astype(np.float32)
    embedding = ht.Variable('embeddingtable_host', value=npemb.copy(), ctx=ht.cpu(0))
    index = ht.Variable(name='index', ctx=ht.cpu(0))
    W = ht.Variable(name='w_dev', value=npw)
    y = ht.matmul_op(ht.embedding_lookup_op(embedding, index), W)
    train_op = ht.optim.SGDOptimizer(0.1).minimize(y)
    executor = ht.Executor([y, train_op], ctx=exec_ctx)
    out, _ = executor.run(feed_dict={index: npind.astype(np.float32)}, convert_to_numpy_ret_vals=True)
    np.testing.assert_allclose(out, np.matmul(npemb[npind], npw), rtol=rtol)
    tmp_grad = np.matmul(np.ones((10, 30), dtype=np.float32), npw.T)
    for i, row in enumerate(npind):
        npemb[row] -= 0.1 * tmp_grad[i]
    got = executor.config.placeholder_to_arr_map[embedding]
    got = got.asnumpy() if hasattr(got, 'asnumpy') else got.detach().cpu().numpy()
    np.testing.assert_allclose(got, npemb, rtol=rtol, atol=1e-5)


def test_host_parameters_in_host_executor():
    _dense(ht.cpu(0), 1e-5)
    _sparse(ht.cpu(0), 1e-5)


@pytest.mark.gpu
def test_host_parameters_in_gpu_executor():
    _dense(ht.gpu(0), 2e-3)                 # tf32 GEMM on the device
    _sparse(ht.gpu(0), 2e-3)
    import torch
    X = ht.Variable(name='x2')
    W = ht.Variable(name='w_host2', value=np.ones((4, 4), np.float32), ctx=ht.cpu(0))
    ex = ht.Executor([ht.matmul_op(X, W)], ctx=ht.gpu(0))
    ex.run(feed_dict={X: np.ones((2, 4), np.float32)})
    t = ex.config.placeholder_to_arr_map[W]
    t = t.tensor if hasattr(t, 'tensor') else t
    assert t.device.type == 'cpu' and torch.is_tensor(t)          # the parameter really stays on the host
