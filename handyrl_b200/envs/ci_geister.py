"""CIGeister ("complete information" Geister, handyrl/envs/ci_geister.py): same rules as Geister;
the observation always exposes the opponent's blue/red planes (ci_geister.py:562-566)."""
from .geister import Environment as Geister


class Environment(Geister):
    GAME = 'CIGeister'
