"""Device GBDT trainer on a B200 (kernels of csrc/gbdt.cu) against the native CPU trainer.  Kept in a
late-collected file: it is the newest GPU path of the suite."""
import numpy as np
import pytest
import scipy.sparse as sp

from hetuml.ensemble import GBDT
from test_ml import _gbdt_level_trainer_matches_node_trainer


@pytest.mark.gpu
def test_gbdt_level_trainer_on_b200():
    _gbdt_level_trainer_matches_node_trainer('cuda')
    rng = np.random.RandomState(12)
    Xd = (rng.rand(2000, 30) < 0.2).astype(np.float32)
    y = (Xd[:, :5].sum(1) > 1).astype(np.float32)
    gb = GBDT(device='gpu', max_depth=3, num_round=5, min_node_ins=10)      # public entry point
    gb.fit((sp.csr_matrix(Xd), y))
    assert gb.evaluate((sp.csr_matrix(Xd), y), 'error')['error'] < 0.1
