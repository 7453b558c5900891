"""ParallelTicTacToe facade (handyrl/envs/parallel_tictactoe.py:13-58): both players submit a
move every ply and one of them, chosen at random, lands."""
import random

from .tictactoe import Environment as TicTacToe


class Environment(TicTacToe):
    GAME = 'ParallelTicTacToe'

    def __str__(self):
        return self._board_string()

    def step(self, actions):
        selected_player = random.choice(list(actions.keys()))     # parallel_tictactoe.py:22
        self._step(actions[selected_player], selected_player)

    def _step(self, action, selected_player):
        self._apply(action, selected_player)
        self.record.append(([self.BLACK, self.WHITE][selected_player], action))

    def diff_info(self, _):
        if len(self.record) == 0:
            return ""
        color, action = self.record[-1]
        return self.action2str(action) + ":" + self.C[color]

    def update(self, info, reset):
        if reset:
            self.reset()
        else:
            saction, scolor = info.split(":")
            self._step(self.str2action(saction), 'OX'.index(scolor))

    def turn(self):
        # the reference returns (does not raise) an exception object here (parallel_tictactoe.py:54-55)
        return NotImplementedError()

    def turns(self):
        return self.players()
