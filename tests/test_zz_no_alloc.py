"""b200rl_env_rollout_static: the allocation-free variant of the persistent rollout (its ticket scratch is a file-scope
`static __device__` array of the library, include/b200rl.h) against b200rl_env_rollout with a caller-owned scratch.

it has no counterpart in the reference.  Collected last on purpose:
a failure here must not hide another test.  Both entry points are called through the
C-ABI on identical starting states, TWO calls in a row (the static array keeps the previous call's contents: the second
call proves that what must start at zero is cleared), and every output is compared bit for bit - the tolerance of this
repository's byte / integer tests.  Works alone in a fresh process and after the other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_GAMES = 262144          # B200RL_ROLLOUT_STATIC_MAX_GAMES (include/b200rl.h)


def _p(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _buffers(torch, lib, game, N, W, D):
    """torch-owned device buffers of one rollout (state reset by the library); no allocation happens after this"""
    dims = (ctypes.c_int64 * 8)()
    lib.call("b200rl_env_dims", game, ctypes.cast(dims, ctypes.c_void_p))
    A, obits, sbits = int(dims[0]), int(dims[1]), int(dims[2])
    sb = int(lib.lib().b200rl_env_state_bytes(game))
    dev = "cuda"
    b = dict(
        state=torch.zeros(N * sb // 4 + 4, dtype=torch.int32, device=dev),
        main=torch.full((W, N, obits), -7.0, device=dev),
        scalar=torch.full((W, N, sbits), -7.0, device=dev) if sbits else None,
        amask=torch.full((W, N, A), -7.0, device=dev),
        action=torch.full((W, N), -9, dtype=torch.int32, device=dev),
        traj=torch.zeros(N, D, 8, dtype=torch.int32, device=dev),
        ept=torch.zeros(1 << 16, 4, dtype=torch.int32, device=dev),
        ctr=torch.zeros(8, dtype=torch.int64, device=dev),
        scratch=torch.full((int(lib.lib().b200rl_env_rollout_scratch_bytes(N)) // 4 + 4,), 0x5A5A5A5A,
                           dtype=torch.int32, device=dev),     # garbage: the launcher must clear what it uses
    )
    lib.call("b200rl_env_reset", game, _p(b["state"]), N, None, lib.stream_ptr())
    return b


def _rollout(lib, static, game, b, N, plies, it0, W, D):
    head = (game, _p(b["state"]), N, plies, 77, 1000, it0, _p(b["main"]), _p(b["scalar"]), _p(b["amask"]), _p(b["action"]),
            W, _p(b["traj"]), D, _p(b["ept"]), b["ept"].shape[0], _p(b["ctr"]))
    if static:
        lib.call("b200rl_env_rollout_static", *head, lib.stream_ptr())
    else:
        lib.call("b200rl_env_rollout", *head, _p(b["scratch"]), b["scratch"].numel() * 4, lib.stream_ptr())


def _episode_rows(b):
    n = int(b["ctr"][5].item())
    rows = b["ept"][:n].cpu().numpy()
    return rows[np.lexsort((rows[:, 0], rows[:, 2], rows[:, 3]))]      # append order is not deterministic; content is


@pytest.mark.parametrize("game,N,plies", [(2, 16384, 24), (2, 1000, 9), (0, 4096, 12)])   # Geister x2, TicTacToe
def test_static_scratch_rollout_equals_caller_scratch_rollout(game, N, plies):
    import torch
    from handyrl_b200 import _lib
    W, D = 2, 64
    ref = _buffers(torch, _lib, game, N, W, D)
    new = _buffers(torch, _lib, game, N, W, D)
    it0 = 0
    for call in range(2):                                  # two calls in a row on the same stream
        _rollout(_lib, False, game, ref, N, plies, it0, W, D)
        _rollout(_lib, True, game, new, N, plies, it0, W, D)
        it0 += plies
        torch.cuda.synchronize()
        for k in ("state", "main", "scalar", "amask", "action", "traj"):
            if ref[k] is not None:
                assert torch.equal(ref[k], new[k]), (call, k)
        assert torch.equal(ref["ctr"][:6], new["ctr"][:6]), call
        assert int(new["ctr"][1].item()) == N * plies * (call + 1)          # every game stepped once per ply
        np.testing.assert_array_equal(_episode_rows(ref), _episode_rows(new))
        assert float(new["main"].min()) >= 0 and float(new["amask"].min()) >= 0   # every wave byte was written


def test_static_scratch_rollout_rejects_sizes_beyond_its_maximum():
    import torch
    from handyrl_b200 import _lib
    state = torch.zeros(64, dtype=torch.int32, device="cuda")
    ctr = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(_lib.B200RLError) as ei:
        _lib.call("b200rl_env_rollout_static", 0, _p(state), MAX_GAMES + 32, 4, 1, 0, 0, None, None, None, None, 0,
                  None, 0, None, 0, _p(ctr), _lib.stream_ptr())
    assert "B200RL_ROLLOUT_STATIC_MAX_GAMES" in str(ei.value) and str(MAX_GAMES) in str(ei.value)
    torch.cuda.synchronize()
    assert int(ctr.sum().item()) == 0                      # nothing was launched


def test_static_scratch_rollout_at_its_maximum():
    """N = the stated maximum (TicTacToe: 4 bytes of state per game) runs and steps every game once per ply."""
    import torch
    from handyrl_b200 import _lib
    N = MAX_GAMES
    state = torch.zeros(N + 4, dtype=torch.int32, device="cuda")
    ctr = torch.zeros(8, dtype=torch.int64, device="cuda")
    _lib.call("b200rl_env_reset", 0, _p(state), N, None, _lib.stream_ptr())
    for _ in range(2):
        _lib.call("b200rl_env_rollout_static", 0, _p(state), N, 8, 3, 0, 0, None, None, None, None, 0, None, 0, None, 0,
                  _p(ctr), _lib.stream_ptr())
    torch.cuda.synchronize()
    c = ctr.cpu().numpy()
    assert c[1] == 2 * 8 * N and c[0] == c[2] + c[3] + c[4] and c[0] > 0
