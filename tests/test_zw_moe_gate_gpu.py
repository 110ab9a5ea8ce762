"""Fused MoE gate kernels (csrc/moe_gate.cu) on a B200 against the composed gate.  Late-collected file."""
import pytest

import hetu as ht
from test_graph_cpu import _fused_gate_vs_composed


@pytest.mark.gpu
def test_fused_moe_gate_kernels_on_b200():
    _fused_gate_vs_composed(ht.gpu(0), 2e-5)
