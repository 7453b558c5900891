"""CPU: the numpy oracle of losses.py against golden outputs of the reference itself."""
import os

import numpy as np
import pytest

from oracle import targets_np
from tests.synth import TARGET_GOLDEN_CASES, target_inputs

MODES = ['MC', 'TD', 'UPGO', 'VTRACE']


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "targets.npz"))


@pytest.mark.parametrize("ci", range(len(TARGET_GOLDEN_CASES)))
@pytest.mark.parametrize("mode", MODES)
def test_oracle_matches_reference_bitexact(gold, ci, mode):
    B, T, P, head, rpp = TARGET_GOLDEN_CASES[ci]
    d = target_inputs(B, T, P, head, rpp, seed=100 + ci)
    # the stored inputs are what the reference saw
    for k in ('values', 'returns', 'rhos', 'cs'):
        np.testing.assert_array_equal(d[k], gold['c%d_%s' % (ci, k)])
    tg, adv = targets_np.compute_target(mode, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'],
                                        d['rhos'], d['cs'])
    ref_t = gold['c%d_%s_targets' % (ci, mode)]
    ref_a = gold['c%d_%s_advantages' % (ci, mode)]
    assert tg.shape == ref_t.shape and adv.shape == ref_a.shape
    np.testing.assert_array_equal(np.asarray(tg), ref_t)
    np.testing.assert_array_equal(np.asarray(adv), ref_a)


def test_oracle_dispatch_edge_cases(capsys):
    assert targets_np.compute_target('TD', None, None, None, 0.7, 1, None, None) == (None, 0)
    d = target_inputs(2, 3, 2, 'value')
    assert targets_np.compute_target('NOPE', d['values'], d['returns'], None, 0.7, 1, d['rhos'], d['cs']) is None
    assert 'No algorithm named NOPE' in capsys.readouterr().out
