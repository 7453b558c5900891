"""GPU: tools/sanitize_smoke.py - every entry point of the library at small RAGGED sizes in one process, with blocking
launches so that a fault is reported at the call that caused it.  (It found the missing wave-base alignment check of
b200rl_env_rollout: N = 130 TicTacToe games with 2 waves faulted with 'misaligned address' inside the kernel.)"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_every_entry_point_runs_at_ragged_sizes():
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_smoke.py")], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "all entry points ran" in r.stdout, r.stdout[-3000:]
