"""B200-native (sm_100a) implementation of HandyRL's self-play hot path.

Parts: batched game stepping + observation encoding
(`handyrl_b200.engine`, `handyrl_b200.envs`), episode-to-batch assembly
(`handyrl_b200.train.make_batch`) and the off-policy target recursion
(`handyrl_b200.losses.compute_target`).  All compute goes through the C-ABI
library `libb200rl.so` (include/b200rl.h); there is no CPU fallback.
"""

__version__ = "0.1.0"
