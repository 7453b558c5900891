"""ctypes wrapper of oracle/libgames_oracle.so (TEST INFRASTRUCTURE, see oracle/__init__.py)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libgames_oracle.so")
GAMES = {"tictactoe": 0, "parallel_tictactoe": 1, "geister": 2, "ci_geister": 3,
         "TicTacToe": 0, "ParallelTicTacToe": 1, "Geister": 2, "CIGeister": 3}
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE])
        h = ctypes.CDLL(_PATH)
        h.orc_selfplay.restype = ctypes.c_int64
        h.orc_selfplay.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_int, ctypes.c_void_p]
        h.orc_random_tapes.restype = ctypes.c_int64
        h.orc_reward.restype = ctypes.c_float
        _lib = h
    return _lib


class Env:
    """Single game, mirrors the reference Environment methods used on the hot path."""

    def __init__(self, game):
        self.h = lib()
        self.game = GAMES[game] if isinstance(game, str) else game
        self.buf = ctypes.create_string_buffer(self.h.orc_sizeof())
        self.geister = self.game >= 2
        self.reset()

    def reset(self):
        self.h.orc_reset(self.buf, self.game)

    def step(self, action, sel=0):
        self.h.orc_step(self.buf, int(action), int(sel))

    def turn(self):
        return self.h.orc_turn(self.buf)

    def terminal(self):
        return bool(self.h.orc_terminal(self.buf))

    def reward(self):
        return float(self.h.orc_reward(self.buf))

    def outcome(self):
        o = (ctypes.c_float * 2)()
        self.h.orc_outcome(self.buf, o)
        return [float(o[0]), float(o[1])]

    def legal_actions(self):
        out = (ctypes.c_int * 80)()
        n = self.h.orc_legal_actions(self.buf, out)
        return [int(out[i]) for i in range(n)]

    def observation(self, player):
        """player None -> -1.  Returns (main float32[252|27], scalar float32[18|0])."""
        main = np.zeros(252 if self.geister else 27, dtype=np.float32)
        scalar = np.zeros(18, dtype=np.float32)
        self.h.orc_observation(self.buf, -1 if player is None else int(player),
                               main.ctypes.data_as(ctypes.c_void_p), scalar.ctypes.data_as(ctypes.c_void_p))
        return main, (scalar if self.geister else scalar[:0])


def random_tapes(game, n_games, seed=0):
    """Random games at C speed; same columns as tests/golden/tape_*.npz (dict of arrays)."""
    g = GAMES[game] if isinstance(game, str) else game
    geister = g >= 2
    lmax = 70 if geister else 9
    cap = n_games * (204 if geister else 11)
    mb = 32 if geister else 4
    d = dict(game=np.zeros(cap, np.int32), action=np.zeros(cap, np.int32), sel=np.zeros(cap, np.int32),
             turn=np.zeros(cap, np.int32), terminal=np.zeros(cap, np.int32), reward=np.zeros(cap, np.float32),
             legal_count=np.zeros(cap, np.int32), legal_ids=np.zeros((cap, lmax), np.int32),
             obs0=np.zeros((cap, mb), np.uint8), obs1=np.zeros((cap, mb), np.uint8), obsN=np.zeros((cap, mb), np.uint8),
             sc0=np.zeros((cap, 3), np.uint8), sc1=np.zeros((cap, 3), np.uint8), scN=np.zeros((cap, 3), np.uint8),
             outcome=np.zeros((n_games, 2), np.float32), steps=np.zeros(n_games, np.int32))
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    n = lib().orc_random_tapes(g, n_games, ctypes.c_uint64(seed), ctypes.c_int64(cap), P(d['game']), P(d['action']),
                               P(d['sel']), P(d['turn']), P(d['terminal']), P(d['reward']), P(d['legal_count']),
                               P(d['legal_ids']), lmax, P(d['obs0']), P(d['obs1']), P(d['obsN']), P(d['sc0']),
                               P(d['sc1']), P(d['scN']), P(d['outcome']), P(d['steps']))
    for k in list(d):
        if k not in ('outcome', 'steps'):
            d[k] = d[k][:n]
    return d


def selfplay(game, n_steps_per_thread, seed=0, n_threads=1):
    g = GAMES[game] if isinstance(game, str) else game
    cs = ctypes.c_double(0)
    return int(lib().orc_selfplay(g, ctypes.c_int64(n_steps_per_thread), ctypes.c_uint64(seed), n_threads,
                                  ctypes.byref(cs)))


# ---- whole hot path on the CPU (oracle/hotpath_oracle.c): timed baseline of bench.py --impl reference ----
_hp = None


def _hotpath_lib():
    global _hp
    if _hp is None:
        path = os.path.join(_HERE, "libhotpath_oracle.so")
        if not os.path.exists(path):
            import subprocess
            subprocess.check_call(["make", "-C", _HERE])
        h = ctypes.CDLL(path)
        h.orc_hotpath_create.restype = ctypes.c_void_p
        h.orc_hotpath_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_uint64,
                                         ctypes.c_int, ctypes.c_double, ctypes.c_double]
        h.orc_hotpath_step.restype = ctypes.c_int64
        h.orc_hotpath_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        _hp = h
    return _hp


class HotPath:
    """G Geister games per thread: generation -> select_episode + make_batch -> UPGO/V-Trace targets, per step."""

    def __init__(self, games_per_thread, plies, T, n_threads, replay_cap=256, seed=0, gamma=0.8, lmb=0.7):
        self.h = _hotpath_lib()
        self.G, self.plies, self.T, self.n_threads = games_per_thread, plies, T, n_threads
        self.ctx = self.h.orc_hotpath_create(games_per_thread, plies, T, replay_cap, seed, n_threads, float(gamma),
                                             float(lmb))

    def step(self):
        out = (ctypes.c_int64 * 2)()
        n = self.h.orc_hotpath_step(self.ctx, out)
        return int(n), int(out[0]), int(out[1])

    def phase_seconds(self):
        """cumulative thread-seconds spent in (generation, batching, targets)"""
        out = (ctypes.c_double * 3)()
        self.h.orc_hotpath_phase_seconds(ctypes.c_void_p(self.ctx), out)
        return [float(x) for x in out]

    def peek(self):
        """thread 0's arrays: dict of numpy views (value-head / return-head targets and their inputs)"""
        fp = ctypes.POINTER(ctypes.c_float)
        ptrs = [fp() for _ in range(7)]
        B = ctypes.c_int(0)
        self.h.orc_hotpath_peek(ctypes.c_void_p(self.ctx), *[ctypes.byref(p) for p in ptrs], ctypes.byref(B))
        B, T = B.value, self.T
        n = B * T
        arr = lambda p, k: np.ctypeslib.as_array(p, shape=(k,)).copy()
        tv, adv, v = arr(ptrs[0], n * 4), arr(ptrs[1], n * 4), arr(ptrs[2], n * 4)
        return dict(B=B, T=T, tv=tv.reshape(2, B, T, 2, 1), adv=adv.reshape(2, B, T, 2, 1), v=v.reshape(2, B, T, 2, 1),
                    rho=arr(ptrs[3], n).reshape(B, T, 1, 1), outcome=arr(ptrs[4], B * 2).reshape(B, 1, 2, 1),
                    ret=arr(ptrs[5], n * 2).reshape(B, T, 2, 1), reward=arr(ptrs[6], n * 2).reshape(B, T, 2, 1))
