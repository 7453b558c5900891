"""CPU: the PyTorch nets behind Environment.net() are parameter-compatible with the reference's
(same state_dict keys/shapes; identical outputs when the reference tree is available)."""
import os
import sys

import numpy as np
import pytest
import torch

from handyrl_b200 import nets

REF = "/root/reference"


def test_parameter_counts():
    n = lambda m: sum(p.numel() for p in m.parameters())
    assert n(nets.SimpleConv2dModel()) == 29006
    assert n(nets.GeisterNet()) == 233832


def test_forward_shapes():
    torch.manual_seed(0)
    m = nets.SimpleConv2dModel().eval()
    o = m(torch.rand(5, 3, 3, 3))
    assert o['policy'].shape == (5, 9) and o['value'].shape == (5, 1)
    g = nets.GeisterNet().eval()
    hid = g.init_hidden([4])
    o = g({'board': torch.rand(4, 7, 6, 6), 'scalar': torch.rand(4, 18)}, hid)
    assert o['policy'].shape == (4, 214) and o['value'].shape == (4, 1) and o['return'].shape == (4, 1)
    assert len(o['hidden'][0]) == 3 and o['hidden'][0][0].shape == (4, 32, 6, 6)
    h0 = g.init_hidden()
    assert isinstance(h0[0][0], np.ndarray) and h0[0][0].shape == (32, 6, 6)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (GPU box)")
def test_same_outputs_as_reference_nets():
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    try:
        from handyrl.envs.geister import GeisterNet as RefG
        from handyrl.envs.tictactoe import SimpleConv2dModel as RefT
    finally:
        sys.path.remove(REF)
    torch.manual_seed(1)
    for mine, ref, x, hid in (
        (nets.SimpleConv2dModel(), RefT(), torch.rand(6, 3, 3, 3), None),
        (nets.GeisterNet(), RefG(), {'board': torch.rand(6, 7, 6, 6), 'scalar': torch.rand(6, 18)}, 'rnn'),
    ):
        assert list(mine.state_dict().keys()) == list(ref.state_dict().keys())
        mine.load_state_dict(ref.state_dict())
        mine.eval(), ref.eval()
        ha = mine.init_hidden([6]) if hid else None
        hb = ref.init_hidden([6]) if hid else None
        with torch.no_grad():
            a, b = mine(x, ha), ref(x, hb)
        for k in ('policy', 'value', 'return'):
            if k in b:
                assert torch.equal(a[k], b[k]), k
