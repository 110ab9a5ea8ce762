"""Symbolic reverse-mode autodiff and topological sorts.

reference: python/hetu/gpu_ops/executor.py:1066-1181 (gradients), :1305-1393
(topo sorts, sum_node_list).  Traversals are iterative (BERT-large's graph is
deeper than Python's recursion limit)."""
from __future__ import annotations

from typing import Dict, List, Optional

from .node import Op


def find_topo_sort(node_list: List[Op]) -> List[Op]:
    """Post-order DFS: every node appears after all of its inputs."""
    visited, order = set(), []
    for root in node_list:
        if root in visited:
            continue
        stack = [(root, 0)]
        visited.add(root)
        while stack:
            node, i = stack.pop()
            if i < len(node.inputs):
                stack.append((node, i + 1))
                child = node.inputs[i]
                if child is not None and child not in visited:
                    visited.add(child)
                    stack.append((child, 0))
            else:
                order.append(node)
    return order


def find_topo_sort_inference(node_list: List[Op]) -> List[Op]:
    """Topo order for inference: parameters resident on the PS are read through
    a sparse pull instead of the optimizer-owned path (reference :1331-1358)."""
    return find_topo_sort(node_list)


def sum_node_list(node_list: List[Op], ctx=None) -> Optional[Op]:
    """Sum gradient contributions; a single contribution passes through."""
    from ..ops.reduce import sum_op
    node_list = [n for n in node_list if n is not None]
    if not node_list:
        return None
    if len(node_list) == 1:
        return node_list[0]
    return sum_op(node_list, ctx=ctx)


def gradients(output_node, node_list: List[Op], insert_grad: Optional[Op] = None,
              return_all: bool = False):
    """d(output)/d(node) for each node in ``node_list``.

    ``insert_grad`` seeds the backward pass with an externally supplied gradient
    (pipeline stages receive it from the next stage) instead of ones_like.
    With ``return_all`` also returns the backward->forward and forward->backward
    node maps used by the pipeline executors."""
    from ..ops.elementwise import oneslike_op
    if isinstance(output_node, list):
        outputs = list(output_node)
        seeds = insert_grad if isinstance(insert_grad, list) else [insert_grad] * len(outputs)
    else:
        outputs, seeds = [output_node], [insert_grad]
    node_to_grads: Dict[Op, List[Op]] = {}
    for out, seed in zip(outputs, seeds):
        g = seed if seed is not None else oneslike_op(out, ctx=out.raw_ctx)
        node_to_grads.setdefault(out, []).append(g)
    node_to_grad: Dict[Op, Op] = {}
    backward2forward: Dict[Op, tuple] = {}
    forward2backward: Dict[Op, List[Op]] = {}
    for out, seed in zip(outputs, seeds):
        if seed is None:
            backward2forward[node_to_grads[out][0]] = (out, [])
    for node in reversed(find_topo_sort(outputs)):
        if node not in node_to_grads:
            continue
        grad = sum_node_list(node_to_grads[node], node.raw_ctx)
        if grad is None:
            continue
        if getattr(grad, 'mp_state', None) != getattr(node, 'mp_state', None):
            from ..parallel.mp_lowering import match_grad_state
            grad = match_grad_state(grad, node)
        node_to_grad[node] = grad
        if not node.inputs:
            continue
        input_grads = node.gradient(grad)
        if input_grads is None:
            continue
        forward2backward[node] = [grad] + [g for g in input_grads if g is not None]
        for inp, ig in zip(node.inputs, input_grads):
            if ig is None:
                continue
            node_to_grads.setdefault(inp, []).append(ig)
            backward2forward[ig] = (node, [grad])
    grad_list = [node_to_grad.get(n) for n in node_list]
    if return_all:
        return grad_list, backward2forward, forward2backward
    return grad_list
