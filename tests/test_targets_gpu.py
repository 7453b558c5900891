"""GPU parity: handyrl_b200.losses.compute_target (CUDA, through the C-ABI) against
(1) golden outputs of the reference and (2) the numpy oracle on seeded inputs.
Bar: 1e-5 relative (north_star); the kernel follows the reference's rounding
order, so the tests additionally assert bit-exact equality."""
import os

import numpy as np
import pytest
import torch

from oracle import targets_np
from tests.synth import TARGET_GOLDEN_CASES, target_inputs

pytestmark = pytest.mark.gpu
MODES = ['MC', 'TD', 'UPGO', 'VTRACE']
RTOL = 1e-5


def _cuda(d):
    return {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in d.items()}


def _close(x, ref):
    x = x.detach().cpu().numpy()
    assert x.shape == ref.shape
    assert np.all(np.abs(x - ref) <= RTOL * np.maximum(np.abs(ref), 1e-3)), np.abs(x - ref).max()
    np.testing.assert_array_equal(x, ref)


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "targets.npz"))


@pytest.mark.parametrize("ci", range(len(TARGET_GOLDEN_CASES)))
@pytest.mark.parametrize("mode", MODES)
def test_golden(gold, ci, mode):
    from handyrl_b200.losses import compute_target
    B, T, P, head, rpp = TARGET_GOLDEN_CASES[ci]
    d = _cuda(target_inputs(B, T, P, head, rpp, seed=100 + ci))
    tg, adv = compute_target(mode, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    if mode == 'MC':
        assert tg is d['returns']  # losses.py:17 returns the same object
    _close(tg, gold['c%d_%s_targets' % (ci, mode)])
    _close(adv, gold['c%d_%s_advantages' % (ci, mode)])


@pytest.mark.parametrize("B,T,P", [(4096, 32, 2), (4096, 200, 2), (1024, 1024, 2), (300, 16, 2), (1, 1, 1),
                                     (65, 33, 4), (5000, 7, 1)])
@pytest.mark.parametrize("head", ['value', 'return'])
@pytest.mark.parametrize("mode", MODES)
def test_against_oracle(B, T, P, head, mode):
    from handyrl_b200.losses import compute_target
    d = target_inputs(B, T, P, head, seed=B + T)
    ref_t, ref_a = targets_np.compute_target(mode, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'],
                                             d['rhos'], d['cs'])
    g = _cuda(d)
    tg, adv = compute_target(mode, g['values'], g['returns'], g['rewards'], g['lmb'], g['gamma'], g['rhos'], g['cs'])
    _close(tg, np.asarray(ref_t))
    _close(adv, np.asarray(ref_a))


@pytest.mark.parametrize("vt,pt", [('VTRACE', 'UPGO'), ('TD', 'UPGO'), ('VTRACE', 'TD'), ('UPGO', 'VTRACE'),
                                   ('MC', 'UPGO'), ('TD', 'TD')])
def test_fused_pair_equals_two_calls(vt, pt):
    from handyrl_b200.losses import compute_target, compute_target_pair
    d = _cuda(target_inputs(777, 32, 2, 'return', seed=5))
    a = (d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    t1, _ = compute_target(vt, *a)
    _, a2 = compute_target(pt, *a)
    tf, af = compute_target_pair(vt, pt, *a)
    assert torch.equal(tf, t1) and torch.equal(af, a2)


@pytest.mark.parametrize("B,T,P,rpp", [(4096, 32, 2, False), (4096, 32, 2, True), (333, 64, 1, False), (65, 33, 4, False)])
@pytest.mark.parametrize("vt,pt", [('VTRACE', 'UPGO'), ('VTRACE', 'VTRACE'), ('TD', 'VTRACE')])
def test_cs_is_rhos_same_tensor(B, T, P, rpp, vt, pt):
    """compute_loss passes one tensor for rhos and cs (both clipped at 1.0, train.py:222): the row kernel then
    stages it once.  Results must equal the two-tensor call bit for bit, and the oracle."""
    from handyrl_b200.losses import compute_target_pair
    d = target_inputs(B, T, P, 'return', rpp, seed=B + T + 1)
    d['cs'] = d['rhos'].copy()
    rt, _ = targets_np.compute_target(vt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    _, ra = targets_np.compute_target(pt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    g = _cuda(d)
    a = (g['values'], g['returns'], g['rewards'], g['lmb'], g['gamma'])
    t2, a2 = compute_target_pair(vt, pt, *a, g['rhos'], g['cs'])      # two buffers
    t1, a1 = compute_target_pair(vt, pt, *a, g['rhos'], g['rhos'])    # one buffer, aliased
    assert torch.equal(t1, t2) and torch.equal(a1, a2)
    _close(t1, np.asarray(rt))
    _close(a1, np.asarray(ra))


@pytest.mark.parametrize("B,T,P,rpp", [(1000, 200, 2, False), (257, 128, 2, True), (333, 128, 1, False),
                                       (4096, 1024, 2, False), (17, 68, 2, False), (19, 76, 2, False),
                                       (33, 84, 2, True)])
@pytest.mark.parametrize("vt,pt", [('TD', 'TD'), ('UPGO', 'UPGO'), ('VTRACE', 'VTRACE'), ('VTRACE', 'UPGO'),
                                   ('UPGO', 'VTRACE'), ('TD', 'UPGO')])
@pytest.mark.parametrize("head", ['value', 'return'])
def test_long_rows_ws_kernel(monkeypatch, B, T, P, rpp, vt, pt, head):
    """Rows longer than 64 steps run in the warp-specialised kernel (loader / compute / storer warps, mbarrier
    ring).  Both chunk widths and several ring depths must give the oracle's bits, with ragged last tiles, a
    partial first chunk (T % chunk != 0), per-player rhos, P = 1 and cs aliasing rhos.  The pinned-order scans work on
    groups of four timesteps, two per loop trip: first chunks of 1 (T = 68), 2 (T = 200), 3 (T = 76) and 5 (T = 84)
    groups cover their single-group, even and odd paths, full chunks the steady loop."""
    from handyrl_b200 import _lib
    from handyrl_b200.losses import compute_target_pair
    d = target_inputs(B, T, P, head, rpp, seed=B + T + 7)
    rt, _ = targets_np.compute_target(vt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    _, ra = targets_np.compute_target(pt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['cs'])
    g = _cuda(d)
    a = (g['values'], g['returns'], g['rewards'], g['lmb'], g['gamma'])
    for rowlane in (0, 1):          # lane = (row, player) with 32/P-row tiles / lane = row with 32-row tiles
        for row in ('256', '512'):
            for stages in ('2', '5'):
                _lib.set_tuning('target_ws_rowlane', rowlane)
                _lib.set_tuning('target_ws_row', int(row))
                _lib.set_tuning('target_ws_stages', int(stages))
                try:
                    t1, a1 = compute_target_pair(vt, pt, *a, g['rhos'], g['cs'])
                finally:
                    _lib.set_tuning('target_ws_rowlane', None)
                    _lib.set_tuning('target_ws_row', None)
                    _lib.set_tuning('target_ws_stages', None)
                _close(t1, np.asarray(rt))
                _close(a1, np.asarray(ra))
    if 'VTRACE' in (vt, pt):
        rt2, _ = targets_np.compute_target(vt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['rhos'])
        _, ra2 = targets_np.compute_target(pt, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'], d['rhos'], d['rhos'])
        t1, a1 = compute_target_pair(vt, pt, *a, g['rhos'], g['rhos'])
        _close(t1, np.asarray(rt2))
        _close(a1, np.asarray(ra2))


@pytest.mark.parametrize("B,T,P,rpp", [(4096, 32, 2, False), (777, 16, 2, True), (333, 64, 1, False), (100, 200, 2, False),
                                       (65, 33, 4, False), (50, 7, 2, False)])
@pytest.mark.parametrize("vt,pt", [('VTRACE', 'UPGO'), ('UPGO', 'VTRACE'), ('TD', 'TD'), ('MC', 'UPGO')])
def test_heads_equal_per_head_calls(B, T, P, rpp, vt, pt):
    """compute_target_heads (value head: outcome bootstrap, rewards None, gamma 1; return head: rewards, gamma
    0.8) runs both heads in one launch for short contiguous rows and falls back to one call per head otherwise;
    either way the bits are those of compute_target_pair per head, and of the oracle."""
    from handyrl_b200.losses import compute_target_heads, compute_target_pair
    dv = target_inputs(B, T, P, 'value', rpp, seed=B + T)
    dr = target_inputs(B, T, P, 'return', rpp, seed=B + T + 1)
    gv, gr = _cuda(dv), _cuda(dr)
    rho, c = gr['rhos'], gr['cs']
    heads = [(gv['values'], gv['returns'], None, 1), (gr['values'], gr['returns'], gr['rewards'], 0.8)]
    for cs_t in (c, rho):     # two tensors / one aliased tensor
        res = compute_target_heads(vt, pt, heads, 0.7, rho, cs_t)
        for (tg, adv), h in zip(res, heads):
            t1, a1 = compute_target_pair(vt, pt, h[0], h[1], h[2], 0.7, h[3], rho, cs_t)
            assert torch.equal(tg, t1) and torch.equal(adv, a1)
    (tg, adv), _ = compute_target_heads(vt, pt, heads, 0.7, rho, c)
    rt, _ = targets_np.compute_target(vt, dv['values'], dv['returns'], None, 0.7, 1, dr['rhos'], dr['cs'])
    _, ra = targets_np.compute_target(pt, dv['values'], dv['returns'], None, 0.7, 1, dr['rhos'], dr['cs'])
    _close(tg, np.asarray(rt))
    _close(adv, np.asarray(ra))
    # a missing head behaves like compute_target(None, ...)
    res = compute_target_heads(vt, pt, [(None, None, None, 1), heads[1]], 0.7, rho, c)
    assert res[0] == (None, 0) and torch.equal(res[1][0], compute_target_pair(vt, pt, *heads[1][:3], 0.7, 0.8, rho, c)[0])


def test_noncontiguous_and_broadcast_views():
    from handyrl_b200.losses import compute_target
    d = target_inputs(50, 20, 2, 'return', seed=9)
    ref = {m: targets_np.compute_target(m, d['values'], d['returns'], d['rewards'], d['lmb'], d['gamma'],
                                        d['rhos'], d['cs']) for m in MODES}
    g = _cuda(d)
    # values as a strided slice of a wider tensor, rewards transposed storage
    wide = torch.zeros(50, 20, 4, 1, device='cuda')
    wide[:, :, 1:3] = g['values']
    v = wide[:, :, 1:3]
    rew = g['rewards'].permute(1, 0, 2, 3).contiguous().permute(1, 0, 2, 3)
    assert not v.is_contiguous() and not rew.is_contiguous()
    for m in MODES:
        tg, adv = compute_target(m, v, g['returns'], rew, g['lmb'], g['gamma'], g['rhos'], g['cs'])
        _close(tg, np.asarray(ref[m][0]))
        _close(adv, np.asarray(ref[m][1]))


def test_dispatch_edge_cases(capsys):
    from handyrl_b200.losses import compute_target
    assert compute_target('TD', None, None, None, 0.7, 1, None, None) == (None, 0)
    d = _cuda(target_inputs(2, 3, 2, 'value'))
    assert compute_target('NOPE', d['values'], d['returns'], None, 0.7, 1, d['rhos'], d['cs']) is None
    assert 'No algorithm named NOPE' in capsys.readouterr().out


def test_full_size_properties():
    """BASELINE sizes (B=65536): linear-recursion identities that need no oracle.
    TD with lambda=1, gamma=1, no rewards reproduces the bootstrap everywhere; V-Trace with
    rho=c=1, lambda=1 equals TD(1); UPGO >= TD elementwise."""
    from handyrl_b200.losses import compute_target
    B, T, P = 65536, 32, 2
    gen = torch.Generator(device='cuda').manual_seed(0)
    v = torch.tanh(torch.randn(B, T, P, 1, device='cuda', generator=gen))
    oc = torch.randint(-1, 2, (B, 1, P, 1), device='cuda', generator=gen).float()
    one = torch.ones(B, T, 1, 1, device='cuda')
    tg, adv = compute_target('TD', v, oc, None, 1.0, 1.0, one, one)
    assert torch.equal(tg, oc.expand_as(tg)) and torch.equal(adv, oc - v)
    tv, _ = compute_target('VTRACE', v, oc, None, 1.0, 1.0, one, one)
    assert (tv - oc).abs().max().item() < 1e-5
    td, _ = compute_target('TD', v, oc, None, 0.7, 1.0, one, one)
    up, _ = compute_target('UPGO', v, oc, None, 0.7, 1.0, one, one)
    assert bool((up >= td).all())
    # slice parity against the oracle on 4096 rows of the big tensor
    sl = slice(30000, 34096)
    ref_t, ref_a = targets_np.compute_target('UPGO', v[sl].cpu().numpy(), oc[sl].cpu().numpy(), None, 0.7, 1.0,
                                             None, None)
    _close(up[sl], np.asarray(ref_t))
