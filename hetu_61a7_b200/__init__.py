"""hetu_61a7_b200 - a Blackwell-native distributed deep-learning framework with
the Hetu programming model (define-then-run dataflow graph, ``ht.Executor``).

``import hetu as ht`` resolves to this package through the thin ``hetu`` alias
package at the repo root.  Public surface = reference python/hetu/__init__.py:1-13.
"""
from __future__ import absolute_import

from .graph.node import Op
from .ops import *                      # noqa: F401,F403
from .ops import Variable, placeholder_op
from .graph.executor import (Executor, HetuConfig, gradients, wrapped_mpi_nccl_init,
                             new_group_comm, get_mpi_communicate, get_nccl_communicate,
                             get_worker_communicate, worker_init, worker_finish, server_init,
                             server_finish, scheduler_init, scheduler_finish, find_topo_sort,
                             path_to_lib)
from .context import context, get_current_context, DistConfig, DeviceGroup
from .dataloader import dataloader_op, Dataloader, GNNDataLoaderOp, DataloaderOp
from .ndarray import (cpu, gpu, rcpu, rgpu, array, sparse_array, empty, is_gpu_ctx, IndexedSlices,
                      NDArray, ND_Sparse_Array, DLContext)
from . import optimizer as optim
from . import lr_scheduler as lr
from . import initializers as init
from . import layers
from . import distributed_strategies as dist
from . import ndarray
from . import stream
from . import cpu_links
from . import onnx
from . import graphboard
from . import memory_pool
from .memory_pool import HetuMemoryPool
from .utils import data
from .utils import metrics
from .utils.profiler import HetuProfiler, NCCLProfiler
from .utils.tokenizers import BertTokenizer, BasicTokenizer, WordpieceTokenizer
from .parallel.cstable import CacheSparseTable
from .parallel.preduce import PartialReduce

__version__ = '0.1.0'
from . import gpu_links
