"""Shared plumbing of the single-game facades: one slot of the batched CUDA engine.

Rules, legal moves and observation encoding always run in the CUDA kernels; the host side
only keeps what the reference keeps in Python lists/dicts for its string protocol
(`record`, `layouts`) and reads back small results.
"""
import numpy as np
import torch

from ..engine import BatchedEnv, VIEW_NONE
from ..environment import BaseEnvironment


def _default_device():
    return torch.device("cuda", torch.cuda.current_device())


class DeviceEnvironment(BaseEnvironment):
    GAME = None

    def __init__(self, args=None):
        super().__init__()
        dev = (args or {}).get('device') if isinstance(args, dict) else None
        self._env = BatchedEnv(self.GAME, 1, device=dev or _default_device())
        self._info = None
        self.reset()

    # -- engine access -----------------------------------------------------
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
        w = self._env.state_bytes // 4
        return self._env.state[:w].cpu().numpy().view(np.uint32).copy()

    def _set_raw_state(self, words):
        w = self._env.state_bytes // 4
        self._env.state[:w].copy_(torch.from_numpy(np.asarray(words, dtype=np.uint32).view(np.int32)))
        self._dirty()

    # -- reference interface -------------------------------------------------
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
        return self._env.A

    def players(self):
        return [0, 1]

    def _view(self, player):
        return VIEW_NONE if player is None else int(player)
