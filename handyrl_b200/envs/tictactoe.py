"""TicTacToe facade: the reference interface (handyrl/envs/tictactoe.py:72-172) over one slot of
the batched CUDA engine.  Rules / legal moves / observation run in the kernels; the host keeps
only the move record used by the string protocol."""
import numpy as np

from ._device import DeviceEnvironment


class Environment(DeviceEnvironment):
    GAME = 'TicTacToe'
    X, Y = 'ABC', '123'
    BLACK, WHITE = 1, -1
    C = {0: '_', BLACK: 'O', WHITE: 'X'}

    # -- string protocol (tictactoe.py:86-101, 121-131) ---------------------
    def action2str(self, a, _=None):
        return self.X[a // 3] + self.Y[a % 3]

    def str2action(self, s, _=None):
        return self.X.find(s[0]) * 3 + self.Y.find(s[1])

    def record_string(self):
        return ' '.join(self.action2str(a) for a in self.record)

    def _cells(self):
        w = int(self._raw_state()[0])
        o, x = w & 0x1ff, (w >> 9) & 0x1ff
        return [self.BLACK if (o >> a) & 1 else (self.WHITE if (x >> a) & 1 else 0) for a in range(9)]

    def _board_string(self):
        cells = self._cells()
        s = '  ' + ' '.join(self.Y) + '\n'
        for i in range(3):
            s += self.X[i] + ' ' + ' '.join(self.C[cells[i * 3 + j]] for j in range(3)) + '\n'
        return s

    def __str__(self):
        return self._board_string() + 'record = ' + self.record_string()

    def play(self, action, _=None):
        self._apply(action)
        self.record.append(action)

    def diff_info(self, _):
        if len(self.record) == 0:
            return ""
        return self.action2str(self.record[-1])

    def update(self, info, reset):
        if reset:
            self.reset()
        else:
            self.play(self.str2action(info))

    def turn(self):
        return self.players()[len(self.record) % 2]

    def net(self):
        from ..nets import SimpleConv2dModel
        return SimpleConv2dModel

    def observation(self, player=None):
        o = self._env.observe(view=self._view(player))
        return o['main'][0].cpu().numpy()
