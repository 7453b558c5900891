"""Geister facade: the reference interface (handyrl/envs/geister.py:170-538) over one slot of the
batched CUDA engine.  Rules, legal moves and the hidden-piece observation run in the kernels; the
host keeps only what the reference's string protocol needs (move record, layouts) and the action
id <-> board coordinate codec (pure arithmetic on ids)."""
import random

import numpy as np

from ._device import DeviceEnvironment

_DIRS = ((-1, 0), (0, -1), (0, 1), (1, 0))          # geister.py:188


class Environment(DeviceEnvironment):
    GAME = 'Geister'
    X, Y = 'ABCDEF', '123456'
    BLACK, WHITE = 0, 1
    BLUE, RED = 0, 1
    C = 'BW'
    T = 'BR'
    P = {-1: '_', 0: 'B', 1: 'R', 2: 'b', 3: 'r', 4: '*'}
    GOALS = (((-1, 5), (6, 5)), ((-1, 0), (6, 0)))   # geister.py:183-186

    def _host_reset(self, args):
        self.args = args
        self.record = []
        self.layouts = {}

    # -- decoded device state ------------------------------------------------
    def _decoded(self):
        w = self._raw_state()
        bb = [int(w[2 * k]) | (int(w[2 * k + 1]) << 32) for k in range(4)]
        return bb, int(w[12]), int(w[13]), int(w[14])   # bitboards, turn_count + 2, winner, aux

    @property
    def turn_count(self):
        return self._get_info()[2] - 2

    @property
    def color(self):
        return self._get_info()[0]

    @property
    def captured_type(self):
        c = self._decoded()[3] & 3
        return None if c == 0 else c - 1

    def opponent(self, color):
        return self.BLACK + self.WHITE - color

    def onboard(self, pos):
        return 0 <= pos[0] < 6 and 0 <= pos[1] < 6

    # -- action id codec (geister.py:275-295): WHITE ids are the 180-degree rotated BLACK ids ----
    def fromdirection2action(self, pos_from, d, c):
        x, y = int(pos_from[0]), int(pos_from[1])
        if c == self.WHITE:
            x, y, d = 5 - x, 5 - y, 3 - d
        return d * 36 + x * 6 + y

    def action2from(self, a, c):
        cell = a % 36
        x, y = cell // 6, cell % 6
        if c == self.WHITE:
            x, y = 5 - x, 5 - y
        return np.array((x, y), dtype=np.int32)

    def action2direction(self, a, c):
        d = a // 36
        return 3 - d if c == self.WHITE else d

    def action2to(self, a, c):
        return self.action2from(a, c) + np.array(_DIRS[self.action2direction(a, c)], dtype=np.int32)

    def position2str(self, pos):
        return self.X[pos[0]] + self.Y[pos[1]] if self.onboard(pos) else '**'

    def str2position(self, s):
        if s == '**':
            return None
        return np.array((self.X.find(s[0]), self.Y.find(s[1])), dtype=np.int32)

    def action2str(self, a, player):
        if a >= 144:
            return 's' + str(a - 144)
        return self.position2str(self.action2from(a, player)) + self.position2str(self.action2to(a, player))

    def str2action(self, s, player):
        if s[0] == 's':
            return 144 + int(s[1:])
        c = player
        pos_from, pos_to = self.str2position(s[:2]), self.str2position(s[2:])
        if pos_to is None:
            # leaving the board: the adjacent goal square of this colour fixes the direction
            target = next(g for g in self.GOALS[c] if abs(g[0] - pos_from[0]) + abs(g[1] - pos_from[1]) == 1)
            diff = (target[0] - int(pos_from[0]), target[1] - int(pos_from[1]))
        else:
            diff = (int(pos_to[0] - pos_from[0]), int(pos_to[1] - pos_from[1]))
        return self.fromdirection2action(pos_from, _DIRS.index(diff), c)

    def record_string(self):
        return ' '.join(self.action2str(a, i % 2) for i, a in enumerate(self.record))

    def __str__(self):
        bb, _, _, _ = self._decoded()

        def cell(i):
            for code in range(4):
                if (bb[code] >> i) & 1:
                    return code if self.layouts.get(code // 2, -1) >= 0 else 4
            return -1

        s = '  ' + ' '.join(self.Y) + '\n'
        for x in range(6):
            s += self.X[x] + ' ' + ' '.join(self.P[cell(x * 6 + y)] for y in range(6)) + '\n'
        s += 'color = ' + self.C[self.color] + '\n'
        s += 'record = ' + self.record_string()
        return s

    # -- transitions ------------------------------------------------------------
    def _set(self, layout):
        self.layouts[self.color] = layout
        if layout < 0:
            layout = random.randrange(70)       # hidden layout of a network-match mirror (geister.py:353-354)
        self._apply(144 + layout)

    def play(self, action, _=None):
        if self.turn_count < 0:
            return self._set(action - 144)
        self._apply(action)
        self.record.append(action)

    def diff_info(self, player):                  # geister.py:398-410
        color = player
        played_color = (self.turn_count - 1) % 2
        info = {}
        if len(self.record) == 0:
            if self.turn_count > -2:
                info['set'] = self.layouts[played_color] if color == played_color else -1
        else:
            info['move'] = self.action2str(self.record[-1], played_color)
            if color == played_color and self.captured_type is not None:
                info['captured'] = self.T[self.captured_type]
        return info

    def update(self, info, reset):                # geister.py:412-425
        if reset:
            self.args = {**self.args, **info}
            self.reset(info)
        elif 'set' in info:
            self._set(info['set'])
        elif 'move' in info:
            action = self.str2action(info['move'], self.color)
            if 'captured' in info:
                self._reveal(self.action2to(action, self.color), self.T.index(info['captured']))
            self.play(action)

    def _reveal(self, pos, t):
        """A mirror environment learns the true type of the opponent piece it is about to capture:
        the board cell changes type, piece_cnt does not (geister.py:418-423)."""
        w = self._raw_state()
        opp = self.opponent(self.color)
        cell = int(pos[0]) * 6 + int(pos[1])
        new, old = opp * 2 + t, opp * 2 + (1 - t)
        bb = [int(w[2 * k]) | (int(w[2 * k + 1]) << 32) for k in range(4)]
        if (bb[old] >> cell) & 1:
            bb[old] &= ~(1 << cell)
            bb[new] |= 1 << cell
            corr = [int(np.int8(np.uint8((int(w[15]) >> (8 * k)) & 0xFF))) for k in range(4)]
            corr[old] += 1
            corr[new] -= 1
            w[15] = sum((c & 0xFF) << (8 * k) for k, c in enumerate(corr))
            for k in (old, new):
                w[2 * k], w[2 * k + 1] = bb[k] & 0xFFFFFFFF, bb[k] >> 32
            self._set_raw_state(w)

    # -- queries --------------------------------------------------------------------
    def turn(self):
        return self.players()[self._get_info()[0]]

    def reward(self):
        return {p: -0.01 for p in self.players()}      # geister.py:434-436

    def net(self):
        from ..nets import GeisterNet
        return GeisterNet

    def observation(self, player=None):
        o = self._env.observe(view=self._view(player))
        return {'scalar': o['scalar'][0].cpu().numpy(), 'board': o['main'][0].cpu().numpy()}
