"""``import hetu as ht`` - alias of :mod:`hetu_61a7_b200` so reference scripts run unchanged."""
import sys as _sys
import hetu_61a7_b200 as _impl
from hetu_61a7_b200 import *            # noqa: F401,F403
from hetu_61a7_b200 import (optim, lr, init, layers, dist, ndarray, stream, data, metrics, Op,
                            Executor, HetuConfig, gradients)

for _name in ('layers', 'optimizer', 'lr_scheduler', 'initializers', 'ndarray', 'stream', 'context',
              'dataloader', 'distributed_strategies', 'ops', 'graph', 'parallel', 'utils', 'models',
              'launcher', 'runner', 'cpu_links', 'gpu_links', 'onnx', 'graphboard', 'memory_pool'):
    try:
        _mod = __import__('hetu_61a7_b200.' + _name, fromlist=['_'])
        _sys.modules['hetu.' + _name] = _mod
    except ImportError:
        pass
_sys.modules['hetu.optim'] = optim
_sys.modules['hetu.init'] = init
_sys.modules['hetu.lr'] = lr
_sys.modules['hetu.dist'] = dist
_sys.modules['hetu.gpu_ops'] = _sys.modules.get('hetu.ops', _impl)

# reference module paths that map onto differently organised modules here
for _alias, _target in (('cstable', 'parallel.cstable'), ('preduce', 'parallel.preduce'), ('profiler', 'utils.profiler'),
                        ('metrics', 'utils.metrics'), ('data', 'utils.data'), ('tokenizers', 'utils.tokenizers'),
                        ('tokenizers.bert_tokenizer', 'utils.tokenizers'), ('_base', '_base'), ('communicator', 'communicator'),
                        ('communicator.mpi_nccl_comm', 'communicator.mpi_nccl_comm'), ('communicator.nccl_comm', 'communicator.nccl_comm'),
                        ('communicator.mpi_comm', 'communicator.mpi_comm'), ('gpu_ops.executor', 'graph.executor'),
                        ('gpu_ops.Node', 'graph.node'), ('gpu_ops.Variable', 'ops.variable'),
                        ('gpu_ops.timer_subexecutor', 'graph.timer'), ('gpu_ops.pipeline_subexecutor', 'parallel.pipeline'),
                        ('gpu_ops.gpipe_subexecutor', 'parallel.pipeline'), ('gpu_ops.pipedream_subexecutor', 'parallel.pipeline')):
    try:
        _sys.modules['hetu.' + _alias] = __import__('hetu_61a7_b200.' + _target, fromlist=['_'])
    except ImportError:
        pass


def _gpu_ops_alias(_file):
    """``hetu.gpu_ops.<File>`` (one module per op file in the reference): a module whose attributes
    resolve in the flat op namespace, so ``from hetu.gpu_ops.MatrixMult import MatMulOp, matmul_op`` works."""
    import types
    m = types.ModuleType('hetu.gpu_ops.' + _file)
    m.__dict__['__getattr__'] = _find_op_name
    return m


def _find_op_name(name):
    """Look a class / constructor up in every op module (classes are not re-exported by ``ops``)."""
    import pkgutil
    pools = [_sys.modules['hetu.gpu_ops']]
    for _pkg in (_impl.ops, _impl.graph):
        for _info in pkgutil.iter_modules(_pkg.__path__):
            try:
                pools.append(__import__(_pkg.__name__ + '.' + _info.name, fromlist=['_']))
            except ImportError:
                pass
    for _extra in ('hetu_61a7_b200.optimizer', 'hetu_61a7_b200.dataloader', 'hetu_61a7_b200.parallel.mp_lowering'):
        pools.append(__import__(_extra, fromlist=['_']))
    for mod in pools:
        if hasattr(mod, name):
            return getattr(mod, name)
    raise AttributeError(name)


for _file in ('Abs AddConst AddElewise Addmm AllGatherCommunicate AllReduceCommunicate AllToAll Arange Argmax Argsort AvgPool Baddbmm '
              'BalanceAssignment BatchMatrixMult BatchNorm BinaryCrossEntropy Bool Broadcast BroadcastCommunicate BroadcastShape Clamp '
              'Concat Concatenate ConstPow Conv2d Conv2dAddBias Conv2dBroadcast Conv2dReduceSum CrossEntropy CrossEntropySparse CuSparse '
              'Cumsum DataTransfer Dispatch DistGCN_15d Division Dropout Dropout2d EmbeddingLookUp Exp Floor Full Gather Gelu GroupTopKIdx '
              'HAllToAll Indexing InstanceNorm2d Interpolate LayerNorm LayoutTransform LeakyRelu Linear LogElewise MaskedFill MatrixDot '
              'MatrixMult Max MaxPool Min MinusByConst MinusElewise MultiplyConst MultiplyElewise NllLoss Norm OneHot OnesLike Opposite Pad '
              'ParameterServerCommunicate PipelineReceive PipelineSend Pow Rand ReduceCommunicate ReduceMean ReduceScatterCommunicate '
              'ReduceSum ReduceSumAxisZero Relu Repeat Reshape ReverseLayoutTransform ReverseLayoutTransformNoGate Roll SamGroupSum SamMax '
              'Scatter Scatter1D Sigmoid Sin Slice SliceAssign SliceByMatrix Softmax SoftmaxCrossEntropy SoftmaxCrossEntropySparse Split '
              'Sqrt Sum Tanh TopKIdx TopKVal Transpose Where ZerosLike').split():
    _sys.modules.setdefault('hetu.gpu_ops.' + _file, _gpu_ops_alias(_file))
__version__ = _impl.__version__
