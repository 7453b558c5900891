"""B1 / BASELINE configs[0]: the reference's own `main.py --train` and `--eval`, UNMODIFIED, on the CUDA facades.

The reference tree is the git-ignored copy `baseline/_ref/` made by `tools/install_ref.sh` (run by
`__graft_entry__.build()` where /root/reference exists; it ships to the GPU box with the gpurun snapshot).
Nothing of this repository is imported by the launched process except through the plugin boundary:
`env_args.env` is the dotted path of a facade module (handyrl/environment.py:18-37).

The reference's learner forks its workers after building the environment in the parent
(handyrl/train.py:428 -> handyrl/worker.py:189,32), so these runs also prove that the facades defer their CUDA
initialisation to the process that plays (handyrl_b200/envs/_device.py).

`main.py --train` never returns in the reference (its worker processes are not joined after "finished server",
observed with the reference's own TicTacToe env too), so the test waits for the checkpoint + the
"finished server" line and then ends the process group it started.
"""
import os
import signal
import subprocess
import sys
import time

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(os.path.join(REF, "main.py")),
                                 reason="baseline/_ref not installed (tools/install_ref.sh needs /root/reference)")]

TRAIN_ARGS = dict(
    turn_based_training=True, observation=False, gamma=0.8, forward_steps=16, compress_steps=4,
    entropy_regularization=1.0e-1, entropy_regularization_decay=0.1, update_episodes=40, batch_size=64,
    minimum_episodes=80, maximum_episodes=100000, epochs=1, num_batchers=2, eval_rate=0.1,
    worker=dict(num_parallel=4), policy_target='MC', value_target='MC', seed=0, restart_epoch=0)
TRAIN_ARGS['lambda'] = 0.7


def _write_config(path, env, **over):
    cfg = dict(env_args=dict(env=env), train_args={**TRAIN_ARGS, **over},
               worker_args=dict(server_address='', num_parallel=4))
    with open(os.path.join(path, "config.yaml"), "w") as f:
        yaml.safe_dump(cfg, f)


def _env():
    e = dict(os.environ)
    e["PYTHONPATH"] = ROOT + os.pathsep + e.get("PYTHONPATH", "")
    e["PYTHONDONTWRITEBYTECODE"] = "1"
    e["PYTHONUNBUFFERED"] = "1"
    return e


def _run_train(cwd, timeout):
    """Launch main.py --train in its own process group; return (log text, finished flag)."""
    log = os.path.join(cwd, "train.log")
    with open(log, "w") as f:
        p = subprocess.Popen([sys.executable, os.path.join(REF, "main.py"), "--train"], cwd=cwd, env=_env(),
                             stdout=f, stderr=subprocess.STDOUT, start_new_session=True)
    t0 = time.time()
    done = False
    try:
        while time.time() - t0 < timeout:
            time.sleep(1.0)
            txt = open(log).read()
            if "finished server" in txt and os.path.exists(os.path.join(cwd, "models", "1.pth")):
                done = True
                break
            if p.poll() is not None:
                break
    finally:
        try:
            os.killpg(p.pid, signal.SIGKILL)      # exactly the group started above
        except ProcessLookupError:
            pass
        p.wait()
    return open(log).read(), done


def _run_eval(cwd, n_games, n_proc, timeout):
    # Python 3.12 removed random.randrange(1e8) (handyrl/evaluation.py:305): the launcher restores it and then
    # runs the unmodified main.py as __main__ (tools/ref_main_py312.py) - needed with the reference's own envs too
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ref_main_py312.py"),
                        os.path.join(REF, "main.py"), "--eval", "models/1.pth", str(n_games), str(n_proc)],
                       cwd=cwd, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=timeout, text=True)
    return r.returncode, r.stdout


def _check_train_log(txt):
    assert "Cannot re-initialize CUDA" not in txt, txt[-3000:]
    assert "Traceback" not in txt, txt[-3000:]
    assert "updated model(" in txt, txt[-3000:]
    assert "loss = " in txt, txt[-3000:]            # the trainer really consumed batches made from facade episodes
    assert "generation stats = " in txt


@pytest.mark.parametrize("env,over,ngames", [
    ("handyrl_b200.envs.tictactoe", {}, 20),                                        # BASELINE configs[0]
    ("handyrl_b200.envs.parallel_tictactoe", dict(turn_based_training=False, policy_target='TD',
                                                  value_target='TD'), 12),
    ("handyrl_b200.envs.geister", dict(update_episodes=12, minimum_episodes=16, batch_size=16,
                                       policy_target='UPGO', value_target='VTRACE'), 4),
])
def test_reference_main_train_and_eval_unmodified(tmp_path, env, over, ngames):
    cwd = str(tmp_path)
    _write_config(cwd, env, **over)
    txt, done = _run_train(cwd, timeout=900)
    assert done, "main.py --train did not reach its checkpoint:\n" + txt[-3000:]
    _check_train_log(txt)
    rc, out = _run_eval(cwd, ngames, 2, timeout=600)
    assert rc == 0, out[-3000:]
    assert "Traceback" not in out and "Cannot re-initialize CUDA" not in out, out[-3000:]
    assert "---agent 0---" in out and "total" in out, out[-3000:]
    assert out.count("*** Game") == ngames
