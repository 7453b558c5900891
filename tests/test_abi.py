"""CPU: libb200rl.so loads and exports every symbol include/b200rl.h declares."""
import ctypes
import os
import re

from handyrl_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200rl.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rl_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    syms = _declared_symbols()
    assert len(syms) >= 4
    h = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(h, s), "libb200rl.so does not export %s" % s


def test_python_binding_covers_header():
    syms = set(_declared_symbols()) - {"b200rl_last_error"}
    assert syms == set(_lib.SIGNATURES), (syms ^ set(_lib.SIGNATURES))


def test_version_and_error_string():
    h = _lib.lib()
    assert h.b200rl_version() >= 1
    assert isinstance(h.b200rl_last_error(), bytes)


def test_argument_errors_do_not_need_a_gpu():
    h = _lib.lib()
    z = ctypes.c_void_p(0)
    st = h.b200rl_target_mc(z, 0, 0, 0, z, 0, 0, 0, z, 1, 1, 1, z)
    assert st != 0 and b"null" in h.b200rl_last_error()


def test_rollout_rejects_misaligned_wave_bases_before_launching():
    """Several waves are written with 128-bit stores: N = 130 TicTacToe games would put wave 1 of the observation at
    130 * 27 * 4 bytes (not a multiple of 16) - found by tools/sanitize_smoke.py as a 'misaligned address' fault in the
    kernel when the check only asked for an even N.  The pointers are never dereferenced: validation comes first."""
    h = _lib.lib()
    fake = ctypes.c_void_p(0x10000)
    for game, n, ok in ((0, 130, False), (0, 132, True), (2, 130, True), (2, 131, False)):
        rc = h.b200rl_env_rollout(game, fake, n, 0 if ok else 8, 1, 0, 0, fake, None if game == 0 else fake, fake, None, 2,
                                  None, 0, None, 0, None, fake, 1 << 20, None)
        if ok:
            assert rc == 0, h.b200rl_last_error()       # n_plies = 0: validated, nothing launched
        else:
            assert rc != 0 and b"multiple of 4" in h.b200rl_last_error()


def test_no_cpu_fallback():
    import pytest
    import torch
    from handyrl_b200.losses import compute_target
    v = torch.zeros(2, 3, 2, 1)
    with pytest.raises(_lib.B200RLError):
        compute_target('TD', v, v, None, 0.7, 1.0, None, None)


def test_facades_construct_without_cuda():
    """B1: the reference builds the environment in the parent before forking its workers (train.py:428,
    worker.py:189): construction, players(), action_length() and net() must not touch the device."""
    import torch
    from handyrl_b200.environment import make_env
    was_init = torch.cuda.is_initialized()
    for name, A in (("TicTacToe", 9), ("ParallelTicTacToe", 9), ("Geister", 214), ("CIGeister", 214)):
        e = make_env({"env": name, "id": 3})
        assert e.players() == [0, 1] and e.action_length() == A
        assert isinstance(e.net(), type)
        assert e._engine is None
    assert torch.cuda.is_initialized() == was_init
