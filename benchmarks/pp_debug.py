"""Debug helper: runs tests/pp_worker.py with per-node synchronisation and device tracing."""
import os, sys, runpy, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetu_61a7_b200.parallel import pipeline as pl
from hetu_61a7_b200.ops.variable import PlaceholderOp
from hetu_61a7_b200.ndarray import NDArray

def _run_nodes(self, nodes, vals):
    cfg = self.config
    r = int(os.environ.get('RANK', 0))
    for node in nodes:
        if node in vals:
            continue
        if isinstance(node, PlaceholderOp):
            v = cfg.placeholder_to_arr_map.get(node)
            if v is None and node.tensor_value is not None:
                tv = node.tensor_value
                v = tv.tensor if isinstance(tv, NDArray) else tv
            vals[node] = v
            continue
        ins = [vals.get(i) if i is not None else None for i in node.inputs]
        print('[r%d] cur=%d node=%s ctx=%s ins=%s' % (r, torch.cuda.current_device(), node.name, node.ctx,
              [(str(t.device), tuple(t.shape), t.data_ptr() % 16) if isinstance(t, torch.Tensor) else type(t).__name__ for t in ins]), flush=True)
        vals[node] = node.compute(ins, None, cfg.comp_stream)
        torch.cuda.synchronize()
        print('[r%d]   ok' % r, flush=True)
pl._PipeBase._run_nodes = _run_nodes
sys.argv = ['pp_worker.py', 'gpu', sys.argv[1]]
runpy.run_path(os.path.join(ROOT, 'tests', 'pp_worker.py'), run_name='__main__')
