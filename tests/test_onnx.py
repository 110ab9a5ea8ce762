"""ONNX round trip: export a trained graph, load it back, outputs must match.
Pattern: reference tests/onnx/*.py (Hetu -> ONNX -> framework round trips)."""
import numpy as np

import hetu as ht
from hetu_61a7_b200 import onnx as honnx
from hetu_61a7_b200.onnx import proto


def _roundtrip(build, feed_shape, tmp_path, steps=2):
    rs = np.random.RandomState(0)
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    logits = build(x)
    loss = ht.reduce_mean_op(ht.softmaxcrossentropy_op(logits, y_), [0])
    train = ht.optim.SGDOptimizer(0.05).minimize(loss)
    ex = ht.Executor({'train': [loss, train], 'infer': [logits]}, ctx=ht.cpu(0), seed=0)
    X = rs.randn(*feed_shape).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rs.randint(0, 10, feed_shape[0])]
    for _ in range(steps):
        ex.run('train', feed_dict={x: X, y_: Y})
    ref = ex.run('infer', feed_dict={x: X})[0].asnumpy()
    path = str(tmp_path / 'model.onnx')
    honnx.export(ex, [x], [logits], path, name='infer')
    ins, outs = honnx.load_onnx(path)
    ex2 = ht.Executor([outs[0]], ctx=ht.cpu(0))
    got = ex2.run(feed_dict={ins[0]: X})[0].asnumpy()
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-5)
    return path


def test_mlp_roundtrip(tmp_path):
    def build(x):
        h = ht.relu_op(ht.linear_op(x, ht.init.xavier_normal((20, 16), name='w1'), ht.init.zeros((16,), name='b1')))
        h = ht.tanh_op(ht.matmul_op(h, ht.init.xavier_normal((16, 16), name='w2')))
        h = h + ht.broadcastto_op(ht.init.random_normal((16,), name='b2'), h)
        return ht.matmul_op(ht.mul_byconst_op(h, 0.5), ht.init.xavier_normal((16, 10), name='w3'))
    path = _roundtrip(build, (8, 20), tmp_path)
    m = proto.read_model(open(path, 'rb').read())
    assert [n[0] for n in m['nodes']][:2] == ['Gemm', 'Relu'] and m['opset'] == 17


def test_cnn_roundtrip(tmp_path):
    def build(x):
        h = ht.array_reshape_op(x, [-1, 1, 8, 8])
        w = ht.init.he_normal((4, 1, 3, 3), name='cw')
        h = ht.conv2d_op(h, w, padding=1, stride=1)
        h = h + ht.conv2d_broadcastto_op(ht.init.random_normal((4,), name='cb'), h)
        h = ht.batch_normalization_op(h, ht.init.ones((4,), name='bs'), ht.init.zeros((4,), name='bb'), momentum=0.9, eps=1e-5)
        h = ht.max_pool2d_op(ht.relu_op(h), 2, 2, padding=0, stride=2)
        h = ht.array_reshape_op(h, [-1, 4 * 4 * 4])
        return ht.matmul_op(h, ht.init.xavier_normal((64, 10), name='fw'))
    _roundtrip(build, (6, 64), tmp_path)


def test_graphboard_dot(tmp_path):
    x = ht.Variable('x')
    y = ht.relu_op(ht.matmul_op(x, ht.init.xavier_normal((4, 3), name='w')))
    ex = ht.Executor([y], ctx=ht.cpu(0))
    out = ht.graphboard.show(ex, path=str(tmp_path / 'gb'))
    dot = open(str(tmp_path / 'gb' / 'output.dot')).read()
    assert 'MatMulOp' in dot and '->' in dot and out
    import graphboard.graph2fig as g2f
    assert g2f.show is ht.graphboard.show


def test_onnx_where_onehot_const_ops_and_input_order():
    """OneHot / Where / reciprocal-sqrt / const-minus / const-divide handlers (reference
    onnx_opset/{OneHot,Where,Sqrt,Division}.py); graph inputs keep the order given to export()."""
    import os
    import tempfile
    import numpy as np
    import hetu as ht
    from hetu.onnx import hetu2onnx, onnx2hetu
    x, c, ids = ht.Variable('x'), ht.Variable('c'), ht.Variable('ids')
    y = ht.where_op(c, ht.rsqrt_op(x), ht.div_const_op(2.0, x)) + ht.minus_byconst_op(x, 1.0)
    z = ht.one_hot_op(ids, 5)
    ex = ht.Executor([y, z], ctx=ht.cpu(0))
    rs = np.random.RandomState(0)
    xv, cv, iv = rs.rand(4, 5).astype(np.float32) + 0.5, (rs.rand(4, 5) > 0.5).astype(np.float32), np.array([0, 3, 4, 1], np.float32)
    ref = ex.run(feed_dict={x: xv, c: cv, ids: iv}, convert_to_numpy_ret_vals=True)
    path = os.path.join(tempfile.mkdtemp(), 'm.onnx')
    hetu2onnx.export(ex, [x, c, ids], [y, z], path)
    ins, outs = onnx2hetu.load_onnx(path)
    assert [i.name for i in ins] == ['x', 'c', 'ids']
    got = ht.Executor(outs, ctx=ht.cpu(0)).run(feed_dict=dict(zip(ins, [xv, cv, iv])), convert_to_numpy_ret_vals=True)
    np.testing.assert_allclose(got[0], ref[0], rtol=1e-6)
    np.testing.assert_array_equal(got[1], ref[1])
