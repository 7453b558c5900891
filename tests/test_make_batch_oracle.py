"""CPU: numpy oracle of make_batch against outputs of the reference itself."""
import os
import pickle
import random

import numpy as np
import pytest

from oracle import make_batch_np


@pytest.fixture(scope="module")
def cases(golden_dir):
    with open(os.path.join(golden_dir, "batch_cases.pkl"), "rb") as f:
        return pickle.load(f)


CASES = ['ttt_turn', 'ttt_long', 'ttt_all', 'ttt_solo', 'pttt_solo', 'pttt_turn', 'geister_turn', 'geister_all',
         'cigeister_turn', 'geister_solo']


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(cases, name):
    c = cases[name]
    random.seed(c['seed'])
    out = make_batch_np.make_batch(c['ep_minimum'], c['args'])
    assert set(out) == set(c['out'])
    for k, ref in c['out'].items():
        assert out[k].shape == ref.shape and out[k].dtype == ref.dtype, (k, out[k].shape, ref.shape, out[k].dtype)
        np.testing.assert_array_equal(out[k], ref, err_msg=k)
