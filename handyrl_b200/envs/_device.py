"""Shared plumbing of the single-game facades: one slot of the batched CUDA engine.

Rules, legal moves and observation encoding always run in the CUDA kernels; the host side
only keeps what the reference keeps in Python lists/dicts for its string protocol
(`record`, `layouts`) and reads back small results.

Process model.  The reference builds the environment in the parent (`Learner.__init__`,
handyrl/train.py:428; `eval_main`, handyrl/evaluation.py:294) and then *forks* its workers, each of
which builds the environment again (handyrl/worker.py:32,189; handyrl/evaluation.py:151,210).  A CUDA
context does not survive `fork`, so the facade never touches the device in `__init__`, `players()`,
`action_length()` or `net()`: the 1-slot `BatchedEnv` is created on the first call that needs the game
state, in the process that makes that call.  A facade object inherited through `fork` drops the parent's
engine handle and starts from a fresh game in the child (what the reference's workers do anyway: they
`reset()` before every episode).
"""
import ctypes
import os

import numpy as np
import torch

from .. import _lib
from ..engine import GAMES, BatchedEnv, VIEW_NONE
from ..environment import BaseEnvironment


def _default_device():
    return torch.device("cuda", torch.cuda.current_device())


def _static_dims(game):
    """(A, ...) of a game from the library's host-side table: no CUDA call."""
    dims = (ctypes.c_int64 * 8)()
    _lib.call("b200rl_env_dims", GAMES[game], ctypes.cast(dims, ctypes.c_void_p))
    return [int(x) for x in dims]


class DeviceEnvironment(BaseEnvironment):
    GAME = None

    def __init__(self, args=None):
        super().__init__()
        self._device_arg = (args or {}).get('device') if isinstance(args, dict) else None
        self._engine = None
        self._engine_pid = None
        self._info = None
        self._A = _static_dims(self.GAME)[0]
        self._host_reset({})

    # -- engine access -----------------------------------------------------
    @property
    def _env(self):
        """The 1-slot engine of THIS process, created on first use (never in __init__)."""
        pid = os.getpid()
        if self._engine is None or self._engine_pid != pid:
            if self._engine is not None:
                # inherited through fork: the parent's context is unusable here
                self._engine = None
                self._host_reset({})
            if torch.cuda._is_in_bad_fork():
                raise _lib.B200RLError(
                    "the parent process initialised CUDA before forking this process; a forked child cannot "
                    "create a CUDA context.  Construct the environment before the first CUDA call of the parent "
                    "(the facades defer theirs) or use the 'spawn' start method")
            self._engine = BatchedEnv(self.GAME, 1, device=self._device_arg or _default_device())
            self._engine_pid = pid
            self._info = None
        return self._engine

    def _host_reset(self, args):
        """Host-side part of reset(): the fields of the reference's string protocol."""
        self.record = []

    def _dirty(self):
        self._info = None

    def _get_info(self):
        if self._info is None:
            self._info = [int(x) for x in self._env.info()[0].tolist()]   # turn, terminal, plies, winner code
        return self._info

    def _apply(self, action, selected=0):
        self._env.step([int(action)], [int(selected)])
        self._dirty()

    def _raw_state(self):
        """uint32 words of the slot's state (layout in DESIGN.md)"""
        env = self._env
        w = env.state_bytes // 4
        return env.state[:w].cpu().numpy().view(np.uint32).copy()

    def _set_raw_state(self, words):
        env = self._env
        w = env.state_bytes // 4
        env.state[:w].copy_(torch.from_numpy(np.asarray(words, dtype=np.uint32).view(np.int32)))
        self._dirty()

    # -- reference interface -------------------------------------------------
    def reset(self, args={}):
        self._env.reset()
        self._dirty()
        self._host_reset(args)

    def terminal(self):
        return bool(self._get_info()[1])

    def outcome(self):
        o = self._env.outcome()[0].tolist()
        # the reference returns Python ints (tictactoe.py:142-147, geister.py:440-445)
        return {p: int(o[idx]) for idx, p in enumerate(self.players())}

    def legal_actions(self, player=None):
        ids, cnt = self._env.legal_list()
        n = int(cnt[0].item())
        return [int(a) for a in ids[0, :n].tolist()]

    def action_length(self):
        return self._A

    def players(self):
        return [0, 1]

    def _view(self, player):
        return VIEW_NONE if player is None else int(player)
