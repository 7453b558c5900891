"""Environment plugin API — mirror of `handyrl/environment.py` (interface unchanged).

`make_env(env_args)` / `prepare_env(env_args)` resolve `env_args['env']` through `ENVS` or as a
dotted module path exposing `Environment` (handyrl/environment.py:18-37); the built-in names map
to the GPU-backed facades of `handyrl_b200.envs`, each a 1-slot view of the batched CUDA engine.
The reference's own `make_env` also accepts these modules by dotted path
(`env: 'handyrl_b200.envs.geister'`), which is how `main.py --train/--eval` pick them up unmodified.
"""
import importlib

ENVS = {
    'TicTacToe': 'handyrl_b200.envs.tictactoe',
    'Geister': 'handyrl_b200.envs.geister',
    'CIGeister': 'handyrl_b200.envs.ci_geister',
    'ParallelTicTacToe': 'handyrl_b200.envs.parallel_tictactoe',
}


def _module(env_args):
    name = env_args['env']
    return importlib.import_module(ENVS.get(name, name)), name


def prepare_env(env_args):
    module, name = _module(env_args)
    if module is None:
        print("No environment %s" % name)
    elif hasattr(module, 'prepare'):
        module.prepare()


def make_env(env_args):
    module, name = _module(env_args)
    if module is None:
        print("No environment %s" % name)
    else:
        return module.Environment(env_args)


class BaseEnvironment:
    """Method set and defaults of handyrl/environment.py:43-146."""

    def __init__(self, args={}):
        pass

    def __str__(self):
        return ''

    def reset(self, args={}):
        raise NotImplementedError()

    def play(self, action, player):
        raise NotImplementedError()

    def step(self, actions):
        # sequential games: apply every submitted action in turn (environment.py:65-68)
        for player, action in actions.items():
            if action is not None:
                self.play(action, player)

    def turn(self):
        return 0

    def turns(self):
        return [self.turn()]

    def terminal(self):
        raise NotImplementedError()

    def reward(self):
        return {}

    def outcome(self):
        raise NotImplementedError()

    def legal_actions(self, player):
        raise NotImplementedError()

    def action_length(self):
        raise NotImplementedError()

    def players(self):
        return [0]

    def observation(self, player=None):
        raise NotImplementedError()

    def action2str(self, a, player=None):
        return str(a)

    def str2action(self, s, player=None):
        return int(s)

    def diff_info(self, player=None):
        return ''

    def update(self, info, reset):
        raise NotImplementedError()
