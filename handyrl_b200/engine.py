"""Batched game engine: N concurrent games of one kind resident on one GPU.

`BatchedEnv` is the device-side counterpart of N reference `Environment`
objects (handyrl/environment.py:43-146).  Every method is the batched form of
the reference method of the same name and runs as one CUDA launch through the
C-ABI (include/b200rl.h); the single-game facades in `handyrl_b200.envs` are
1-slot views of it.  State is a torch-owned device buffer, so there is no
hidden allocation behind the ABI.
"""
import ctypes

import numpy as np
import torch

from . import _lib

GAMES = {
    "TicTacToe": 0,
    "ParallelTicTacToe": 1,
    "Geister": 2,
    "CIGeister": 3,
}
OBS_SHAPES = {0: ((3, 3, 3), None), 1: ((3, 3, 3), None), 2: ((7, 6, 6), (18,)), 3: ((7, 6, 6), (18,))}

VIEW_NONE, VIEW_TURN = -1, 2


def _ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


class BatchedEnv:
    def __init__(self, game, n_games, device="cuda"):
        self.game_name = game
        self.game = GAMES[game] if isinstance(game, str) else int(game)
        self.N = int(n_games)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.B200RLError("BatchedEnv needs a CUDA device (no CPU fallback)")
        lib = _lib.lib()
        sb = lib.b200rl_env_state_bytes(self.game)
        if sb < 0:
            raise _lib.B200RLError("unknown game %r" % (game,))
        self.state_bytes = int(sb)
        dims = (ctypes.c_int64 * 8)()
        _lib.call("b200rl_env_dims", self.game, ctypes.cast(dims, ctypes.c_void_p))
        (self.A, self.obs_floats, self.scalar_floats, self.LMAX, self.LW, self.P, self.max_steps,
         self.has_reward) = [int(x) for x in dims]
        self.obs_shape, self.scalar_shape = OBS_SHAPES[self.game]
        # SoA planes: [state_bytes/16][N] uint4 for Geister, [N] uint32 for TicTacToe
        self.state = torch.zeros(self.N * self.state_bytes // 4 + 4, dtype=torch.int32, device=self.device)
        # ticket counters of rollout(): caller-owned, so the library allocates nothing behind the ABI
        sb = int(lib.b200rl_env_rollout_scratch_bytes(self.N))
        self._rollout_scratch = torch.zeros((sb + 3) // 4 + 4, dtype=torch.int32, device=self.device)
        self.it = 0  # global ply counter of rollout()
        self.reset()

    # -- helpers ---------------------------------------------------------
    def _st(self):
        return _ptr(self.state)

    def _stream(self):
        return _lib.stream_ptr()

    def _i32(self, x):
        if x is None:
            return None
        if not torch.is_tensor(x):
            x = torch.as_tensor(np.asarray(x, dtype=np.int32))
        return x.to(device=self.device, dtype=torch.int32).contiguous()

    # -- reference interface, batched -------------------------------------
    def reset(self, slot_mask=None):
        m = None
        if slot_mask is not None:
            m = torch.as_tensor(slot_mask).to(device=self.device, dtype=torch.uint8).contiguous()
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_reset", self.game, self._st(), self.N, _ptr(m), self._stream())

    def step(self, actions, selected=None, want_reward=False, want_terminal=False):
        """actions: (N,) int32, negative = None.  Returns (reward (N,2) | None, terminal (N,) | None)."""
        a = self._i32(actions)
        s = self._i32(selected)
        rew = torch.empty(self.N, 2, dtype=torch.float32, device=self.device) if want_reward else None
        term = torch.empty(self.N, dtype=torch.uint8, device=self.device) if want_terminal else None
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_step", self.game, self._st(), self.N, _ptr(a), _ptr(s), _ptr(rew), _ptr(term),
                      self._stream())
        return rew, term

    def legal_list(self):
        ids = torch.empty(self.N, self.LMAX, dtype=torch.int32, device=self.device)
        cnt = torch.empty(self.N, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_legal_list", self.game, self._st(), self.N, _ptr(ids), _ptr(cnt), self._stream())
        return ids, cnt

    def alloc_obs(self, n=None, waves=None):
        n = self.N if n is None else n
        lead = (n,) if waves is None else (waves, n)
        main = torch.empty(lead + self.obs_shape, dtype=torch.float32, device=self.device)
        scalar = None
        if self.scalar_shape is not None:
            scalar = torch.empty(lead + self.scalar_shape, dtype=torch.float32, device=self.device)
        return main, scalar

    def observe(self, view=VIEW_TURN, views=None, obs=True, amask=False, bits=False, out=None):
        """observation(player) for all games.  view: -1 None, 0/1 player, 2 turn player.
        Returns dict with 'board'/'scalar' (Geister) or 'obs' (TicTacToe), 'action_mask', 'legal_bits'."""
        out = out or {}
        main = scalar = am = lb = None
        if obs:
            main, scalar = out.get("main"), out.get("scalar")
            if main is None:
                main, scalar = self.alloc_obs()
        if amask:
            am = out.get("action_mask")
            if am is None:
                am = torch.empty(self.N, self.A, dtype=torch.float32, device=self.device)
        if bits:
            lb = torch.empty(self.N, self.LW, dtype=torch.int32, device=self.device)
        v = self._i32(views)
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_observe", self.game, self._st(), self.N, int(view), _ptr(v), _ptr(main),
                      _ptr(scalar), _ptr(am), _ptr(lb), self._stream())
        return {"main": main, "scalar": scalar, "action_mask": am, "legal_bits": lb}

    def outcome(self):
        o = torch.empty(self.N, 2, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_outcome", self.game, self._st(), self.N, _ptr(o), self._stream())
        return o

    def info(self):
        """(N,4) int32: turn, terminal, plies, winner code (0 none, 1 player0, 2 player1, 3 draw)."""
        o = torch.empty(self.N, 4, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_info", self.game, self._st(), self.N, _ptr(o), self._stream())
        return o

    def step_observe(self, actions, selected=None, auto_reset=True, out=None):
        """Fused ply (generation.py:31-68).  `out` may carry preallocated tensors."""
        out = out if out is not None else {}
        a = self._i32(actions)
        s = self._i32(selected)
        if "main" not in out:
            out["main"], out["scalar"] = self.alloc_obs()
        if "action_mask" not in out:
            out["action_mask"] = torch.empty(self.N, self.A, dtype=torch.float32, device=self.device)
        if "reward" not in out:
            out["reward"] = torch.empty(self.N, 2, dtype=torch.float32, device=self.device)
            out["terminal"] = torch.empty(self.N, dtype=torch.uint8, device=self.device)
            out["outcome"] = torch.empty(self.N, 2, dtype=torch.float32, device=self.device)
            out["turn"] = torch.empty(self.N, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_step_observe", self.game, self._st(), self.N, _ptr(a), _ptr(s),
                      1 if auto_reset else 0, _ptr(out["main"]), _ptr(out.get("scalar")), _ptr(out["action_mask"]),
                      _ptr(out["reward"]), _ptr(out["terminal"]), _ptr(out["outcome"]), _ptr(out["turn"]),
                      self._stream())
        return out

    def rollout(self, n_plies, seed=0, game_id_base=0, waves=None, traj=None, ep_table=None, counters=None,
                out_action=None):
        """Persistent uniform-policy self-play for n_plies plies per game (auto-reset).
        waves: dict(main, scalar, action_mask) of [W][N][...] buffers or None (compact mode)."""
        main = scalar = am = None
        W = 0
        if waves is not None:
            main, scalar, am = waves.get("main"), waves.get("scalar"), waves.get("action_mask")
            ref = main if main is not None else am
            W = ref.shape[0]
        D = 0 if traj is None else traj.shape[1]     # (N, D, 8) slot-major ring
        E = 0 if ep_table is None else ep_table.shape[0]
        with torch.cuda.device(self.device):
            _lib.call("b200rl_env_rollout", self.game, self._st(), self.N, int(n_plies), int(seed),
                      int(game_id_base), int(self.it), _ptr(main), _ptr(scalar), _ptr(am), _ptr(out_action),
                      max(W, 1 if out_action is not None else 0), _ptr(traj), D, _ptr(ep_table), E, _ptr(counters),
                      _ptr(self._rollout_scratch), self._rollout_scratch.numel() * 4, self._stream())
        self.it += int(n_plies)

    # -- state access for the single-game facades and tests ------------------
    def get_state(self):
        return self.state.clone()

    def set_state(self, st):
        self.state.copy_(st)
