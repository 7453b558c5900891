"""Batched self-play on the GPU — the device counterpart of `handyrl/generation.py`.

The reference `Generator.generate` (handyrl/generation.py:20-88) plays ONE game per call
in a worker process and returns an episode dict whose moments are bz2-pickled.  Here N
games advance together on one GPU and their moments stay in HBM as a ring of 32-byte
records (`TrajectoryStore`); an episode is the same dict (`args, steps, outcome, moment`)
whose `moment` is a `DeviceMoments` handle instead of a list of compressed blobs, so the
reference's `select_episode` slicing (handyrl/train.py:294-301) works on it unchanged and
`handyrl_b200.train.make_batch` gathers windows straight from the ring.
"""
import bz2
import ctypes
import pickle

import numpy as np
import torch

from . import _lib
from .engine import BatchedEnv, _ptr


class TrajectoryStore:
    """Ring of per-ply records for N game slots: [N][D] x 32 B, slot-major (the plies of a window are contiguous),
    plus the finished-episode table and counters filled by the kernels."""

    def __init__(self, env, depth=None, max_episodes=1 << 16, record_values=False, record_logits=False):
        self.env = env
        self.N = env.N
        self.D = int(depth or max(2 * env.max_steps + 64, 64))
        dev = env.device
        self.traj = torch.zeros(self.N, self.D, 8, dtype=torch.int32, device=dev)
        self.E = int(max_episodes)
        self.ep_table = torch.zeros(self.E, 4, dtype=torch.int32, device=dev)
        self.counters = torch.zeros(8, dtype=torch.int64, device=dev)
        # net outputs beside the records, slot-major too: values [N][D][2]; logits [N][D][PL][LS] = the logits of
        # the legal actions only (ascending action id), LS = longest legal list rounded up to 4 floats
        # (288 B per Geister ply instead of the 856 B of a full row; see include/b200rl.h)
        self.PL = 2 if env.game == 1 else 1
        self.LS = int(_lib.lib().b200rl_env_logit_slots(env.game))
        self.values = torch.zeros(self.N, self.D, 2, dtype=torch.float32, device=dev) if record_values else None
        self.logits = (torch.zeros(self.N, self.D, self.PL, self.LS, dtype=torch.float32, device=dev)
                       if record_logits else None)
        self.it = 0            # next ring iteration to be written
        self._ep_read = 0      # episode-table cursor already handed out


class DeviceMoments:
    """`episode['moment']` of a device-resident episode: plies [0, steps) of one game slot,
    ply k stored at ring iteration first_it + k.  Slicing returns the handle itself, which is
    what `select_episode` (handyrl/train.py:296-299) needs."""

    def __init__(self, store, slot, first_it, steps):
        self.store, self.slot, self.first_it, self.steps = store, int(slot), int(first_it), int(steps)

    def __getitem__(self, item):
        return self

    def __len__(self):
        return self.steps

    def alive(self):
        """True while the ring has not wrapped over this episode."""
        return self.store.it - self.first_it <= self.store.D


class HostReplayIndex:
    """Replay index on the host, like the reference Learner's `episodes` deque (train.py:480-503), for a
    device-resident trajectory store: a pinned mirror of the finished-episode table.  Only these 16-byte rows
    cross PCIe; the moments stay in the trajectory ring.

    `fetch_async` enqueues the read-back of the counters and of a fixed slab of `max_new_rows` table rows starting
    at the committed cursor, on the current stream, and returns a ticket; `commit(ticket)` waits for that copy and
    advances the cursor; `select` draws windows with the SAME rule and RNG as the device kernel
    (`b200rl_select_windows_host`) into a pinned buffer ready for upload.

    Threading: `fetch_async` belongs to the thread that enqueues the rollouts, `commit` + `select` may run on a
    background thread (both block without the GIL: event wait / ctypes call).  Several fetches may be in flight
    (`slots`): each has its own pinned counter copy and event, and every slab starts at the cursor committed when it
    was enqueued, so a lagging commit only makes slabs overlap (the same bytes are copied twice).  Rows beyond a
    slab are picked up by the next fetch: the cursor never passes what has landed.  `select` only reads rows below
    the committed cursor; keep `maximum_episodes + max_new_rows <= E` so that incoming rows do not wrap onto the
    ones being drawn from."""

    def __init__(self, gen, max_new_rows=1 << 15, n_threads=4, slots=4):
        st = gen.store
        self.gen, self.st = gen, st
        self.table = torch.zeros(st.E, 4, dtype=torch.int32).pin_memory()
        self.ctrs = [torch.zeros(8, dtype=torch.int64).pin_memory() for _ in range(slots)]
        self.events = [torch.cuda.Event() for _ in range(slots)]
        self._next_slot = 0
        self._last_ticket = None
        self.max_new, self.n_threads = int(min(max_new_rows, st.E)), int(n_threads)
        self.table.copy_(st.ep_table)                          # one full read-back; later only the new rows
        self.cursor = self.read = int(st.counters[5].item())
        self.draws = 0
        self.d2h_bytes = 64 + self.max_new * 16

    def fetch_async(self):
        slot = self._next_slot
        self._next_slot = (slot + 1) % len(self.ctrs)
        start = self.read                                      # committed cursor right now (a stale value is safe)
        E, n = self.st.E, self.max_new
        self.ctrs[slot].copy_(self.st.counters, non_blocking=True)
        a0 = start % E
        n1 = min(n, E - a0)
        self.table[a0:a0 + n1].copy_(self.st.ep_table[a0:a0 + n1], non_blocking=True)
        if n1 < n:
            self.table[:n - n1].copy_(self.st.ep_table[:n - n1], non_blocking=True)
        self.events[slot].record()
        self._last_ticket = (slot, start)
        return self._last_ticket

    def commit(self, ticket=None):
        slot, start = ticket if ticket is not None else self._last_ticket
        self.events[slot].synchronize()
        cur = min(int(self.ctrs[slot][5]), start + self.max_new)    # rows past the slab wait for the next fetch
        if cur > self.cursor:
            self.read = self.cursor = cur
        return self.cursor

    def select(self, batch_size, args, out_pinned, seed=0, lookahead=0, it_now=None):
        """`lookahead` = plies that will be generated between `it_now` (default: the store's current iteration) and
        the use of the windows (a pipelined caller draws the windows of a later step while rollouts run): episodes
        whose first plies the ring will have overwritten by then are skipped like rejected draws.  A caller that
        selects on a background thread passes `it_now` explicitly (the store's counter belongs to the thread that
        enqueues the rollouts).  The C routine releases the GIL."""
        if self.cursor <= 0:
            raise _lib.B200RLError("HostReplayIndex.select: no finished episodes indexed yet")
        a = int(args['maximum_episodes'])
        _lib.call("b200rl_select_windows_host", ctypes.c_void_p(self.table.data_ptr()), self.st.E, self.cursor, a,
                  int(args['forward_steps']), int(seed) ^ 0x5E1EC7, self.draws,
                  int(self.st.it if it_now is None else it_now) + int(lookahead), self.st.D,
                  ctypes.c_void_p(out_pinned.data_ptr()), int(batch_size), self.n_threads)
        self.draws += 256
        return out_pinned


WIN_OUTCOME = {0: (0, 0), 1: (1, -1), 2: (-1, 1), 3: (0, 0)}


def _map(x, fn):
    """map over nested tuples / lists / dicts of tensors (hidden states)"""
    if isinstance(x, (list, tuple)):
        return type(x)(_map(y, fn) for y in x)
    if isinstance(x, dict):
        return {k: _map(v, fn) for k, v in x.items()}
    return fn(x) if x is not None else None


def _zip(a, b, fn):
    """zip-map over two nested containers of the same structure"""
    if isinstance(a, (list, tuple)):
        return type(a)(_zip(x, y, fn) for x, y in zip(a, b))
    if isinstance(a, dict):
        return {k: _zip(a[k], b[k], fn) for k in a}
    return fn(a, b) if a is not None else None


def _set_player(h, rows, player, new):
    """h (N, 2, ...) with h[n, player[n]] <- new[n]"""
    h[rows, player] = new.to(h.dtype)      # in place (play() runs under no_grad on its own copy)
    return h


class BatchGenerator:
    """N concurrent self-play games on one GPU (one process per GPU; shard by `game_id_base`)."""

    def __init__(self, game, n_games, device="cuda", seed=0, game_id_base=0, depth=None, max_episodes=1 << 16,
                 record_values=False, record_logits=False, args=None):
        self.env = BatchedEnv(game, n_games, device=device)
        self.store = TrajectoryStore(self.env, depth, max_episodes, record_values, record_logits)
        self.seed, self.game_id_base = int(seed), int(game_id_base)
        self.args = args or {}

    # -- epoch-0 policy: everything on the device ---------------------------------
    def rollout(self, n_plies, waves=None, out_action=None):
        """n_plies plies per game with the uniform (RandomModel) policy, auto-reset."""
        st = self.store
        assert self.env.it == st.it
        self.env.rollout(n_plies, seed=self.seed, game_id_base=self.game_id_base, waves=waves, traj=st.traj,
                         ep_table=st.ep_table, counters=st.counters, out_action=out_action)
        st.it = self.env.it

    # -- caller-chosen actions (policy network in PyTorch, tape replay) -------------
    def _net_outputs(self, values, logits):
        """(values (N,2), logits (N,PL,A)) as contiguous fp32 CUDA tensors when the store records them, else None."""
        env, st = self.env, self.store
        v = l = None
        if st.values is not None and values is not None:
            v = torch.as_tensor(values, dtype=torch.float32, device=env.device).reshape(env.N, 2).contiguous()
        if st.logits is not None and logits is not None:
            l = torch.as_tensor(logits, dtype=torch.float32, device=env.device)
            if l.numel() != env.N * st.PL * env.A:
                raise _lib.B200RLError("logits must have shape (N, %s%d)" % ("2, " if st.PL == 2 else "", env.A))
            l = l.contiguous()
        return v, l

    def step(self, actions, selected=None, both=None, values=None, logits=None, auto_reset=True):
        """One recorded ply.  actions (N,) int32, <0 = slot idle (only allowed between episodes: an idle ply
        inside an episode would leave a hole in the ring).  `values` (N,2) / `logits` (N,A) [(N,2,A) for
        ParallelTicTacToe] are recorded beside the ring by the kernel when the store keeps them.
        Returns terminal (N,) uint8."""
        env, st = self.env, self.store
        a, s, b = env._i32(actions), env._i32(selected), env._i32(both)
        v, l = self._net_outputs(values, logits)
        term = torch.empty(env.N, dtype=torch.uint8, device=env.device)
        with torch.cuda.device(env.device):
            _lib.call("b200rl_env_step_record", env.game, env._st(), env.N, _ptr(a), _ptr(s), _ptr(b),
                      1 if auto_reset else 0, _ptr(st.traj), st.D, st.it, _ptr(st.ep_table), st.E,
                      _ptr(st.counters), _ptr(term), _ptr(l), _ptr(v), _ptr(st.logits), _ptr(st.values),
                      env._stream())
        st.it += 1
        env.it = st.it
        return term

    # -- policy network in the loop (generation.py:31-68 with a real model) -----------------
    def observe_turn(self, out=None):
        """Restart finished games, then the turn player's observation + action mask of every game (the inputs of
        `model.inference`, generation.py:36-41).  Returns the dict of `BatchedEnv.step_observe`."""
        return self.env.step_observe(None, auto_reset=True, out=out)

    def act(self, logits, values=None, u=None, out_action=None):
        """Sample one action per game from `logits` (N, A) [(N, 2, A) for ParallelTicTacToe: one row per player]
        like generation.py:48-53, apply and record it.  The legal logits and `values` (N, 2) are recorded beside
        the ring by the same kernel when the store keeps them.  Returns (actions (N, PL) int32, terminal (N,) uint8)."""
        env, st = self.env, self.store
        N, PL = env.N, st.PL
        if not (logits.is_cuda and logits.dtype == torch.float32 and logits.numel() == N * PL * env.A):
            raise _lib.B200RLError("act: logits must be CUDA float32 of shape (N, %s%d)" % ("2, " if PL == 2 else "", env.A))
        logits = logits.contiguous()
        v, _ = self._net_outputs(values, None)
        if u is not None:
            u = u.to(device=env.device, dtype=torch.float64).contiguous()
        act = out_action if out_action is not None else torch.empty(N, PL, dtype=torch.int32, device=env.device)
        term = torch.empty(N, dtype=torch.uint8, device=env.device)
        with torch.cuda.device(env.device):
            _lib.call("b200rl_env_sample_record", env.game, env._st(), N, _ptr(logits), _ptr(u), self.seed,
                      self.game_id_base, _ptr(st.traj), st.D, st.it, _ptr(st.ep_table), st.E, _ptr(st.counters),
                      _ptr(act), _ptr(term), _ptr(v), _ptr(st.logits), _ptr(st.values), env._stream())
        st.it += 1
        env.it = st.it
        return act, term

    def init_hidden(self, model):
        """Per-PLAYER recurrent state of all games, leaves shaped (N, 2, ...) - the reference keeps
        `hidden[player]` for every player of a game (generation.py:24-26).  None for feed-forward nets."""
        if not hasattr(model, 'init_hidden'):
            return None
        h = model.init_hidden([self.env.N, 2])
        return _map(h, lambda t: t.to(self.env.device)) if h is not None else None

    @torch.no_grad()
    def play(self, model, n_plies, hidden=None):
        """n_plies plies of self-play for all games with `model(obs, hidden) -> {'policy','value'[,'hidden']}`
        evaluated on the device (one forward per ply for all N games; `observation: False` semantics: only the
        acting players observe, generation.py:36-41).

        `hidden` is the per-player state of `init_hidden` (leaves (N, 2, ...)).  Like the reference's
        `hidden[player]` (generation.py:24-26, 39-41), a player's state advances only on the plies that player
        observes: each ply the turn player's slice is gathered, fed to the net, and the new state is scattered
        back into that slice alone; both slices are zeroed when a game restarts.  This is the recurrent context
        `forward_prediction` rebuilds from the batch in training (train.py:160-175), so the recorded behaviour
        logits and values are what the learner's network would produce at the same weights.
        Returns the final hidden state."""
        env, st = self.env, self.store
        N = env.N
        simultaneous = st.PL == 2
        out = None
        rows = torch.arange(N, device=env.device)
        for _ in range(int(n_plies)):
            out = self.observe_turn(out)
            if env.scalar_shape is not None:
                obs = {'scalar': out['scalar'], 'board': out['main']}
            else:
                obs = out['main']
            if hidden is not None:
                fresh = (out['terminal'] != 0)   # slots restarted by observe_turn
                hidden = _map(hidden, lambda h: h * (~fresh).view(-1, *([1] * (h.dim() - 1))).to(h.dtype))
            if simultaneous:
                # both players act every ply (parallel_tictactoe.py:20-29) on their own view and state
                pol, val = [], []
                for q in (0, 1):
                    o = model(obs, _map(hidden, lambda h: h[:, q]) if hidden is not None else None)
                    if hidden is not None and o.get('hidden') is not None:
                        hidden = _zip(hidden, o['hidden'], lambda h, nh: _set_player(h, rows, q, nh))
                    pol.append(o['policy'].float())
                    val.append(o['value'].float().view(N) if o.get('value') is not None else None)
                logits = torch.stack(pol, dim=1)
                values = torch.stack(val, dim=1) if val[0] is not None and st.values is not None else None
            else:
                turn = out['turn'].long()
                o = model(obs, _map(hidden, lambda h: h[rows, turn]) if hidden is not None else None)
                if hidden is not None and o.get('hidden') is not None:
                    hidden = _zip(hidden, o['hidden'], lambda h, nh: _set_player(h, rows, turn, nh))
                logits = o['policy'].float()
                v = o.get('value')
                values = None
                if v is not None and st.values is not None:
                    # the net evaluates the turn player's view: store it in that player's slot (generation.py:40-44)
                    values = torch.zeros(N, 2, device=env.device).index_put_((rows, turn), v.view(N).float())
            self.act(logits, values=values)
        return hidden

    # -- window selection on the device (Batcher.select_episode, train.py:284-302) -------
    def select_windows(self, batch_size, args, seed=None, out=None, counters=None, it_now=None):
        """(batch_size, 8) int32 windows for `train.make_batch_windows`, drawn on the device from the
        finished-episode table without a host round trip.  `counters` may be a SNAPSHOT of the store's counters (a
        caller that trains beside a running generation wave must not read the live cursor: a row is written after
        its cursor increment) and `it_now` the ring iteration the windows must survive until (default: now)."""
        st, env = self.store, self.env
        if out is None:
            out = torch.empty(batch_size, 8, dtype=torch.int32, device=env.device)
        self._draws = getattr(self, "_draws", 0)
        ctr = st.counters if counters is None else counters
        with torch.cuda.device(env.device):
            _lib.call("b200rl_select_windows", _ptr(st.ep_table), st.E, _ptr(ctr),
                      int(args['maximum_episodes']), int(args['forward_steps']),
                      int(self.seed if seed is None else seed) ^ 0x5E1EC7, self._draws,
                      int(st.it if it_now is None else it_now), st.D, _ptr(out),
                      int(batch_size), env._stream())
        self._draws += 256
        return out

    # -- finished episodes ------------------------------------------------------------
    def stats(self):
        c = self.store.counters.cpu().numpy()
        return {"episodes": int(c[0]), "env_steps": int(c[1]), "wins0": int(c[2]), "wins1": int(c[3]),
                "draws": int(c[4])}

    def finished_rows(self, out_host=None):
        """Rows of the finished-episode table added since the last call (device -> host), oldest first:
        (n,4) int32 {slot, steps | win << 16, end_it lo, end_it hi}.  This is what a host-side replay index
        needs (the moments themselves stay in the trajectory ring)."""
        st = self.store
        cur = int(st.counters[5].item())
        lo = max(st._ep_read, cur - st.E)
        st._ep_read = cur
        if cur <= lo:
            return np.zeros((0, 4), dtype=np.int32)
        a, b = lo % st.E, cur % st.E
        if a < b:
            rows = st.ep_table[a:b]
        else:
            rows = torch.cat([st.ep_table[a:], st.ep_table[:b]])
        if out_host is not None and out_host.shape[0] >= rows.shape[0]:
            h = out_host[:rows.shape[0]]
            h.copy_(rows, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return h.numpy()
        return rows.cpu().numpy()

    def episodes(self):
        """Episodes finished since the last call, in reference form
        `{'args', 'steps', 'outcome', 'moment'}` (generation.py:79-86) with a DeviceMoments handle."""
        st = self.store
        cur = int(st.counters[5].item())
        lo = max(self._read_cursor(), cur - st.E)
        out = []
        if cur > lo:
            idx = torch.arange(lo, cur, device=st.ep_table.device) % st.E
            rows = st.ep_table[idx].cpu().numpy()
            for slot, sw, it_lo, it_hi in rows:
                steps, win = int(sw) & 0xFFFF, (int(sw) >> 16) & 3
                end_it = (int(np.uint32(it_lo)) | (int(np.uint32(it_hi)) << 32))
                o = WIN_OUTCOME[win]
                out.append({"args": self.args, "steps": steps, "outcome": {0: o[0], 1: o[1]},
                            "moment": DeviceMoments(st, slot, end_it - steps + 1, steps), "win": win})
        st._ep_read = cur
        return out

    def _read_cursor(self):
        return self.store._ep_read


def export_episode(ep, args):
    """Device episode -> the reference wire format (generation.py:79-86): moments re-encoded on the GPU,
    pickled and bz2-compressed in blocks of args['compress_steps'].  RandomModel semantics: zero logits and
    zero values unless the store recorded them."""
    from .train import make_batch  # local import: train imports this module
    steps = ep["steps"]
    win = {"args": ep["args"], "outcome": ep["outcome"], "moment": ep["moment"], "base": 0, "start": 0,
           "end": steps, "total": steps, "win": ep.get("win")}
    a = dict(args)
    a.update(turn_based_training=True, observation=False, forward_steps=steps)
    b = make_batch([win], a)
    cpu = lambda t: t[0].cpu().numpy()
    obs = b["observation"]
    pol, am, act = cpu(b["policy"]), cpu(b["action_mask"]), cpu(b["action"])
    val, rew, ret, tm = cpu(b["value"]), cpu(b["reward"]), cpu(b["return"]), cpu(b["turn_mask"])
    env = ep["moment"].store.env
    moments = []
    keys = ["observation", "policy", "action_mask", "action", "value", "reward", "return"]
    for t in range(steps):
        m = {k: {0: None, 1: None} for k in keys}
        turn = [p for p in (0, 1) if tm[t, p, 0] > 0]
        for p in turn[:1]:
            if isinstance(obs, dict):
                m["observation"][p] = {k: v[0, t, 0].cpu().numpy() for k, v in obs.items()}
            else:
                m["observation"][p] = obs[0, t, 0].cpu().numpy()
            m["policy"][p], m["action_mask"][p], m["action"][p] = pol[t, 0], am[t, 0], int(act[t, 0, 0])
            m["value"][p] = val[t, p]
        for p in (0, 1):
            m["reward"][p] = float(rew[t, p, 0]) if env.has_reward else None
            m["return"][p] = float(ret[t, p, 0])
        m["turn"] = turn
        moments.append(m)
    cs = args["compress_steps"]
    return {"args": ep["args"], "steps": steps, "outcome": ep["outcome"],
            "moment": [bz2.compress(pickle.dumps(moments[i:i + cs])) for i in range(0, steps, cs)]}
