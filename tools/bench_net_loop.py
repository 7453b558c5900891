"""Net-in-the-loop self-play rate ("a net-in-the-loop run (GeisterNet forward in PyTorch)
reported separately").  N games advance one ply per network forward: observe_turn (k_step_observe) ->
model(obs, hidden) in PyTorch -> act (k_sample_record).  The network is library PyTorch/cuDNN and NOT part of
the hot path this repo accelerates; the point of the number is the split between the engine kernels and the
forward pass.  Prints one JSON object.

  python tools/bench_net_loop.py [--env Geister] [--games 65536] [--plies 32] [--amp]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200.generation import BatchGenerator, _map  # noqa: E402
from handyrl_b200.loop import make_net  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--env', default='Geister')
    ap.add_argument('--games', type=int, default=65536)
    ap.add_argument('--plies', type=int, default=32)
    ap.add_argument('--warmup', type=int, default=4)
    ap.add_argument('--amp', action='store_true', help='bf16 autocast + channels_last for the forward (labelled)')
    a = ap.parse_args()
    dev = torch.device('cuda')
    torch.manual_seed(0)
    net = make_net(a.env).to(dev).eval()
    if a.amp:
        net = net.to(memory_format=torch.channels_last)
    gen = BatchGenerator(a.env, a.games, device=dev, depth=64, max_episodes=1 << 20, record_values=True,
                         record_logits=True)
    env = gen.env
    hidden = None
    if hasattr(net, 'init_hidden'):
        hidden = tuple([t.to(dev) for t in part] for part in net.init_hidden([a.games]))

    ev = lambda: torch.cuda.Event(enable_timing=True)
    t_obs = t_net = t_act = 0.0
    out = None
    with torch.no_grad():
        for ply in range(a.warmup + a.plies):
            e = [ev() for _ in range(4)]
            e[0].record()
            out = gen.observe_turn(out)
            e[1].record()
            obs = {'scalar': out['scalar'], 'board': out['main']} if env.scalar_shape is not None else out['main']
            if hidden is not None:
                fresh = (out['terminal'] != 0)
                hidden = _map(hidden, lambda h: h * (~fresh).view(-1, *([1] * (h.dim() - 1))).to(h.dtype))
            with torch.autocast('cuda', dtype=torch.bfloat16, enabled=a.amp):
                o = net(obs, hidden)
            hidden = o.get('hidden', None)
            if hidden is not None and a.amp:
                hidden = _map(hidden, lambda h: h.float())
            turn = out['turn'].long().view(-1, 1)
            values = torch.zeros(a.games, 2, device=dev).scatter_(1, turn, o['value'].view(a.games, 1).float())
            e[2].record()
            gen.act(o['policy'].float(), values=values)
            e[3].record()
            torch.cuda.synchronize()
            if ply >= a.warmup:
                t_obs += e[0].elapsed_time(e[1])
                t_net += e[1].elapsed_time(e[2])
                t_act += e[2].elapsed_time(e[3])
    total = t_obs + t_net + t_act
    steps = a.games * a.plies
    print(json.dumps({
        "env": a.env, "games": a.games, "plies": a.plies, "forward": "bf16 autocast" if a.amp else "fp32",
        "env_steps_per_s": steps / (total * 1e-3),
        "ms_per_ply": {"observe_turn": t_obs / a.plies, "network_forward": t_net / a.plies,
                       "sample_record": t_act / a.plies},
        "engine_share": (t_obs + t_act) / total,
        "engine_only_env_steps_per_s": steps / ((t_obs + t_act) * 1e-3),
    }))


if __name__ == '__main__':
    main()
