"""Helpers to replay recorded game tapes (tests/golden/tape_*.npz or oracle.games_c.random_tapes)
through the CUDA engine in lockstep: game g occupies slot g, ply k of every game is one launch."""
import numpy as np


def unpack(a, n):
    return np.unpackbits(a, axis=-1, bitorder='little')[..., :n].astype(np.float32)


def tape_index(t):
    """-> idx [L, G] of state rows (-1 where game g has no k-th state), L = longest game + 1."""
    game = t['game']
    G = int(game.max()) + 1
    starts = np.flatnonzero(np.r_[True, game[1:] != game[:-1]])
    lens = np.diff(np.r_[starts, len(game)])
    L = int(lens.max())
    idx = -np.ones((L, G), dtype=np.int64)
    for g, (s, n) in enumerate(zip(starts, lens)):
        idx[:n, g] = np.arange(s, s + n)
    return idx


def legal_lists(t, rows):
    """list of python lists for the given state rows"""
    if 'legal_flat' in t:
        off = t['legal_off']
        return [list(t['legal_flat'][off[r]:off[r + 1]]) for r in rows]
    return [list(t['legal_ids'][r][:t['legal_count'][r]]) for r in rows]
