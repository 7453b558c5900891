"""Loss and data-parallel learner around the CUDA target recursion.

* `forward_prediction`, `compose_losses`, `compute_loss` keep the signatures and semantics of
  handyrl/train.py:136-258; the four `compute_target` calls of train.py:248-253 collapse into one
  `compute_target_heads` call (per head: targets of `value_target`, advantages of `policy_target`; both heads in
  one kernel launch for rows of <= 64 steps).
* `DataParallelLearner` replaces `nn.DataParallel` (train.py:365-366) and the pickled weight hand-off to the
  workers (train.py:613, worker.py:52) with one process per GPU: gradients are averaged with ONE all-reduce over
  a flat fp32 bucket, parameters + buffers are broadcast from rank 0 as one flat tensor.
  Everything else of the reference's Learner/Trainer (threads, queues, sockets) is out of scope.
"""
import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import _lib
from .losses import compute_target_heads


def _map(x, fn):
    if isinstance(x, dict):
        return {k: _map(v, fn) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_map(v, fn) for v in x)
    return fn(x) if x is not None else None


def _zip_map(fn, *xs):
    x0 = xs[0]
    if isinstance(x0, dict):
        return {k: _zip_map(fn, *[x[k] for x in xs]) for k in x0}
    if isinstance(x0, (list, tuple)):
        return type(x0)(_zip_map(fn, *vs) for vs in zip(*xs))
    return fn(*xs) if x0 is not None else None


def forward_prediction(model, hidden, batch, args):
    """Network outputs for a (B, T, P, ...) batch, masked like handyrl/train.py:136-184."""
    obs = batch['observation']
    tmask, omask = batch['turn_mask'], batch['observation_mask']
    B, T = tmask.shape[0], tmask.shape[1]
    if hidden is None:
        outputs = model(_map(obs, lambda o: o.reshape(-1, *o.shape[3:])), None)   # one pass over B*T*P rows
        outputs = {k: v for k, v in outputs.items() if v is not None and k != 'hidden'}
    else:
        seq = {}
        squeeze_players = args['turn_based_training'] and not args['observation']
        for t in range(T):
            o_t = _map(obs, lambda o: o[:, t].reshape(-1, *o.shape[3:]))
            m_t = omask[:, t]                                                      # (B, P, 1)
            gate = _map(hidden, lambda h: m_t.reshape(*h.shape[:2], *([1] * (h.dim() - 2))))
            h_in = _zip_map(lambda h, m: h * m, hidden, gate)
            if squeeze_players:
                h_in = _map(h_in, lambda h: h.sum(1))          # only the turn player's state is non-zero
            else:
                h_in = _map(h_in, lambda h: h.reshape(-1, *h.shape[2:]))
            out_t = model(o_t, h_in)
            nxt = out_t.pop('hidden')
            for k, v in out_t.items():
                if v is not None:
                    seq.setdefault(k, []).append(v)
            nxt = _zip_map(lambda nh, h: nh.reshape(h.shape[0], -1, *h.shape[2:]), nxt, hidden)
            hidden = _zip_map(lambda h, nh, m: h * (1 - m) + nh * m, hidden, nxt, gate)
        outputs = {k: torch.stack(v, dim=1) for k, v in seq.items()}
    for k in list(outputs):
        o = outputs[k].reshape(B, T, -1, outputs[k].shape[-1])
        if k == 'policy':
            outputs[k] = (o * tmask).sum(2, keepdim=True) - batch['action_mask']   # turn player's logits, masked
        else:
            outputs[k] = o * omask
    return outputs


def compose_losses(outputs, log_selected_policies, total_advantages, targets, batch, args):
    """handyrl/train.py:188-215"""
    tmask, omask = batch['turn_mask'], batch['observation_mask']
    losses = {}
    dcnt = tmask.sum().item()
    turn_adv = (total_advantages * tmask).sum(2, keepdim=True)
    losses['p'] = (-log_selected_policies * turn_adv).sum()
    if 'value' in outputs:
        losses['v'] = (((outputs['value'] - targets['value']) ** 2) * omask).sum() / 2
    if 'return' in outputs:
        losses['r'] = (F.smooth_l1_loss(outputs['return'], targets['return'], reduction='none') * omask).sum()
    entropy = torch.distributions.Categorical(logits=outputs['policy']).entropy() * tmask.sum(-1)
    losses['ent'] = entropy.sum()
    base = losses['p'] + losses.get('v', 0) + losses.get('r', 0)
    decay = 1 - batch['progress'] * (1 - args['entropy_regularization_decay'])
    losses['total'] = base + (entropy * decay).sum() * -args['entropy_regularization']
    return losses, dcnt


def _p(t):
    return t.data_ptr()


class _PolicyTerms(torch.autograd.Function):
    """log pi_t(a) * emask, clipped rho and c in one pass over the two logit rows (b200rl_policy_terms)."""

    @staticmethod
    def forward(ctx, pol_t, pol_b, actions, emask):
        lead, A = pol_t.shape[:-1], pol_t.shape[-1]
        rows = pol_t.numel() // A
        pt, pb = pol_t.detach().contiguous(), pol_b.contiguous()
        act = actions.contiguous()
        em = emask.expand(*lead, 1).contiguous()
        shape = tuple(lead) + (1,)
        logp, rho, c = (torch.empty(shape, dtype=torch.float32, device=pol_t.device) for _ in range(3))
        with torch.cuda.device(pol_t.device):
            _lib.call("b200rl_policy_terms", _p(pb), _p(pt), _p(act), _p(em), rows, A, 1.0, 1.0, _p(logp), _p(rho),
                      _p(c), _lib.stream_ptr())
        ctx.save_for_backward(pt, act, em)
        ctx.mark_non_differentiable(rho, c)
        return logp, rho, c

    @staticmethod
    def backward(ctx, g_logp, _g_rho, _g_c):
        pt, act, em = ctx.saved_tensors
        A = pt.shape[-1]
        grad = torch.empty_like(pt)
        with torch.cuda.device(pt.device):
            _lib.call("b200rl_policy_terms_backward", _p(pt), _p(act), _p(em), _p(g_logp.contiguous()),
                      pt.numel() // A, A, _p(grad), _lib.stream_ptr())
        return grad, None, None, None


def value_blend(values, omask, emask, outcome, symmetric):
    """train.py:236-239 in one kernel (bit-identical to the torch expression)."""
    B, T, P = values.shape[0], values.shape[1], values.shape[2]
    v = values.contiguous()
    out = torch.empty_like(v)
    with torch.cuda.device(v.device):
        _lib.call("b200rl_value_blend", _p(v), _p(omask.contiguous()), _p(emask.contiguous()),
                  _p(outcome.contiguous()), B, T, P, 1 if symmetric else 0, _p(out), _lib.stream_ptr())
    return out


def compute_loss(batch, model, hidden, args, fused=True):
    """handyrl/train.py:218-258 with the target recursion (and, with `fused`, the policy-ratio and value-blend
    pre-processing) on the CUDA kernels.  `fused=False` keeps the reference's torch expressions around the scan."""
    outputs = forward_prediction(model, hidden, batch, args)
    actions, emask = batch['action'], batch['episode_mask']
    fused = fused and outputs['policy'].is_cuda and outputs['policy'].dtype == torch.float32
    if fused:
        logp_t, clipped_rhos, cs = _PolicyTerms.apply(outputs['policy'], batch['policy'], actions, emask)
        cs = clipped_rhos   # both thresholds are 1.0 (train.py:222): one tensor, the scan reads it once
    else:
        logp_b = F.log_softmax(batch['policy'], dim=-1).gather(-1, actions) * emask
        logp_t = F.log_softmax(outputs['policy'], dim=-1).gather(-1, actions) * emask
        rhos = torch.exp(logp_t.detach() - logp_b)
        clipped_rhos = torch.clamp(rhos, 0, 1.0)      # clip_rho_threshold
        cs = torch.clamp(rhos, 0, 1.0)                # clip_c_threshold
    nograd = {k: v.detach() for k, v in outputs.items()}
    if 'value' in nograd:
        v = nograd['value']
        sym = bool(args['turn_based_training'] and v.shape[2] == 2)   # two-player zero-sum: fill in -opponent's value
        if fused:
            nograd['value'] = value_blend(v, batch['observation_mask'], emask, batch['outcome'], sym)
        else:
            if sym:
                v = (v - v.flip(2)) / (batch['observation_mask'].sum(dim=2, keepdim=True) + 1e-8)
            nograd['value'] = v * emask + batch['outcome'] * (1 - emask)
    targets, advantages = {}, {}
    # train.py:248-253: value head (bootstrap = outcome, no rewards, gamma 1) and return head, one kernel launch
    (targets['value'], advantages['value']), (targets['return'], advantages['return']) = compute_target_heads(
        args['value_target'], args['policy_target'],
        [(nograd.get('value'), batch['outcome'], None, 1),
         (nograd.get('return'), batch['return'], batch['reward'], args['gamma'])],
        args['lambda'], clipped_rhos, cs)
    total_advantages = clipped_rhos * sum(advantages.values())
    return compose_losses(outputs, logp_t, total_advantages, targets, batch, args)


# ------------------------------------------------------------------------------------------------
# data-parallel learner: one process per GPU
# ------------------------------------------------------------------------------------------------
def flat_grads(params):
    return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in params])


def allreduce_gradients(params, group=None, extra=None):
    """Average gradients over the ranks with ONE all-reduce of a flat fp32 bucket (GeisterNet: 0.94 MB).
    `extra` (a list of floats) rides in the same bucket and comes back as the mean over the ranks - the learner
    uses it for the data count of the step, so that every replica applies the same learning rate."""
    distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if not distributed:
        return list(extra) if extra is not None else None
    params = [p for p in params if p.requires_grad]
    bucket = flat_grads(params)
    if extra is not None:
        bucket = torch.cat([bucket, torch.tensor(list(extra), dtype=bucket.dtype, device=bucket.device)])
    dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=group)
    bucket /= dist.get_world_size(group)
    off = 0
    for p in params:
        n = p.numel()
        g = bucket[off:off + n].view_as(p)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        off += n
    return bucket[off:].tolist() if extra is not None else None


def all_ranks_have(count, minimum, group=None):
    """True iff EVERY rank's `count` (a 1-element tensor or a number) reaches `minimum` - the collective form of the
    trainer's `len(episodes) < minimum_episodes` gate (train.py:330-331).  Every rank gets the same answer, so all of
    them enter (or skip) the same sequence of NCCL calls afterwards."""
    n = count.detach().clone().reshape(1) if torch.is_tensor(count) else torch.tensor([int(count)], dtype=torch.int64)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(n, op=dist.ReduceOp.MIN, group=group)
    return int(n.item()) >= int(minimum)


def broadcast_weights(model, src=0, group=None):
    """Parameters and buffers (BatchNorm statistics) from rank `src` to every rank as one flat broadcast."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    ts = [t for t in list(model.parameters()) + list(model.buffers()) if t.is_floating_point()]
    flat = torch.cat([t.detach().reshape(-1).float() for t in ts])
    dist.broadcast(flat, src=src, group=group)
    off = 0
    with torch.no_grad():
        for t in ts:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t))
            off += n
    for t in model.buffers():   # integer buffers (num_batches_tracked)
        if not t.is_floating_point():
            dist.broadcast(t, src=src, group=group)


class DataParallelLearner:
    """Optimiser step of handyrl/train.py:371-400 (Adam, lr schedule by data count, gradient clip 4.0) with
    gradient all-reduce; every rank trains on windows of its OWN trajectory store."""

    def __init__(self, model, args, group=None):
        self.model, self.args, self.group = model, args, group
        self.default_lr = 3e-8
        self.data_cnt_ema = args['batch_size'] * args['forward_steps']
        self.params = [p for p in model.parameters() if p.requires_grad]
        lr = self.default_lr * self.data_cnt_ema
        self.optimizer = torch.optim.Adam(self.params, lr=lr, weight_decay=1e-5) if self.params else None
        self.steps = 0
        self.batch_cnt, self.data_cnt = 0, 0.0        # of the running epoch (train.py:370, 396)
        broadcast_weights(model, 0, group)

    def train_step(self, batch, hidden=None):
        self.model.train()
        losses, dcnt = compute_loss(batch, self.model, hidden, self.args)
        self.optimizer.zero_grad()
        losses['total'].backward()
        # the step's data count travels in the gradient bucket: all replicas see the same (mean) count and therefore
        # apply the same learning rate below - rank-local counts would let the replicas drift apart between the
        # per-epoch weight broadcasts
        dcnt_mean = allreduce_gradients(self.params, self.group, extra=[dcnt])[0]
        torch.nn.utils.clip_grad_norm_(self.params, 4.0)
        self.optimizer.step()
        self.steps += 1
        self.batch_cnt += 1
        self.data_cnt += dcnt_mean
        return {k: float(v.detach()) for k, v in losses.items()}, dcnt

    def end_epoch(self):
        """Learning-rate schedule of handyrl/train.py:396-398, applied once per epoch like the reference:
        data_cnt_ema <- 0.8 * ema + 0.2 * (mean data count per batch of the epoch); lr = 3e-8 * ema / (1 + steps * 1e-5)."""
        if self.batch_cnt > 0:
            self.data_cnt_ema = self.data_cnt_ema * 0.8 + self.data_cnt / (1e-2 + self.batch_cnt) * 0.2
            for g in self.optimizer.param_groups:
                g['lr'] = self.default_lr * self.data_cnt_ema / (1 + self.steps * 1e-5)
        self.batch_cnt, self.data_cnt = 0, 0.0
