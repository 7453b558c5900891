"""Off-policy target recursion on the GPU — drop-in for `handyrl/losses.py`.

`compute_target(algorithm, values, returns, rewards, lmb, gamma, rhos, cs)`
keeps the reference signature, argument meaning, return shapes and error
behaviour (handyrl/losses.py:61-74):

* `values is None` -> `(None, 0)`                      (losses.py:62-63)
* 'MC'     -> `(returns, returns - values)`; the target IS `returns` (losses.py:16-17)
* 'TD'     -> TD(lambda)                               (losses.py:20-28)
* 'UPGO'   -> UPGO                                     (losses.py:31-40)
* 'VTRACE' -> V-Trace                                  (losses.py:43-58)
* unknown  -> prints `No algorithm named ...`, returns None (losses.py:73-74)

Tensors must be CUDA fp32 in the reference's batch-first layout (B, T, P, 1);
any of them may be a broadcast view (stride 0), `returns` may be (B, 1, P, 1).
The work happens in `b200rl_target_scan` (handyrl_b200/csrc/target_scan.cu)
on the current torch stream.  There is no CPU path.
"""
import ctypes

import torch

from . import _lib

MODES = {"MC": 0, "TD": 1, "UPGO": 2, "VTRACE": 3}


def _as_f32_cuda(x, name):
    _lib.require_cuda(x, name)
    if x.dtype != torch.float32:
        raise _lib.B200RLError("%s must be float32 (got %s)" % (name, x.dtype))
    return x


def _bcast_shape(*tensors):
    shape = torch.broadcast_shapes(*[t.shape for t in tensors if t is not None])
    if len(shape) < 2:
        raise _lib.B200RLError("compute_target expects (B, T, ...) tensors, got shape %s" % (tuple(shape),))
    return shape


def _view3(x, shape3, full_shape):
    """(data_ptr, stride_b, stride_t, stride_p) of `x` broadcast to full_shape and
    flattened to (B, T, P).  Trailing dims are collapsed; a copy is made only
    when the broadcast view cannot be described by three strides."""
    if x is None:
        return [ctypes.c_void_p(0), 0, 0, 0], None
    xe = x.expand(full_shape)
    B, T, P = shape3
    tail = full_shape[2:]
    if len(tail) == 0:
        sp = 0
    else:
        # collapse trailing dims into one stride if they are mutually contiguous or broadcast
        strides = xe.stride()[2:]
        sizes = tail
        live = [(s, n) for s, n in zip(strides, sizes) if n > 1]
        if len(live) == 0:
            sp = 0
        elif all(s == 0 for s, _ in live):
            sp = 0
        else:
            ok = True
            expect = live[-1][0]
            for s, n in reversed(live):
                if s != expect:
                    ok = False
                    break
                expect = s * n
            if ok:
                sp = live[-1][0]
            else:
                xe = xe.contiguous()
                sp = 1
    sb = xe.stride(0) if B > 1 else 0
    st = xe.stride(1) if T > 1 else 0
    if xe.is_contiguous() and sp == 1:
        sb, st = T * P, P
    return [ctypes.c_void_p(xe.data_ptr()), sb, st, sp], xe


def _call_on(dev, fn, *args):
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    stream = torch._C._cuda_getCurrentRawStream(idx)
    if torch.cuda.current_device() != idx:
        with torch.cuda.device(dev):
            return fn(*args, stream)
    return fn(*args, stream)


def _v4(x, B, T, P):
    """(ptr, stride_b, stride_t, stride_p) of a 4-D (b,t,p,1) tensor whose dims are full or 1 (broadcast)."""
    s, st = x.shape, x.stride()
    return (x.data_ptr(), st[0] if (s[0] == B and B > 1) else 0, st[1] if (s[1] == T and T > 1) else 0,
            st[2] if (s[2] == P and P > 1) else 0)


def _is4(x, B, T, P):
    s = x.shape
    return (len(s) == 4 and s[3] == 1 and (s[0] == B or s[0] == 1) and (s[1] == T or s[1] == 1)
            and (s[2] == P or s[2] == 1) and x.dtype == torch.float32 and x.is_cuda)


def _scan_4d(mode_t, mode_a, values, returns, rewards, lmb, gamma, rhos, cs):
    """Lean argument marshalling for the shapes compute_loss produces (handyrl/train.py:245-253): every tensor
    is 4-D (b,t,p,1) with dims full or 1.  Returns None when the inputs do not fit (general path below)."""
    vs = values.shape
    if len(vs) != 4 or vs[3] != 1 or values.dtype != torch.float32 or not values.is_cuda:
        return None
    B, T, P = vs[0], vs[1], vs[2]
    if B == 0 or not _is4(returns, B, returns.shape[1] if returns.dim() == 4 else -1, P):
        return None
    need_rho = mode_t == 3 or mode_a == 3
    if rewards is not None and not _is4(rewards, B, T, P):
        return None
    if need_rho and not (_is4(rhos, B, T, P) and _is4(cs, B, T, P)):
        return None
    v = _v4(values, B, T, P)
    rs, rst = returns.shape, returns.stride()
    boot_ptr = returns.data_ptr() + 4 * rst[1] * (rs[1] - 1)          # returns[:, -1]
    boot_sb = rst[0] if (rs[0] == B and B > 1) else 0
    boot_sp = rst[2] if (rs[2] == P and P > 1) else 0
    r = _v4(rewards, B, T, P) if rewards is not None else (0, 0, 0, 0)
    rho = _v4(rhos, B, T, P) if need_rho else (0, 0, 0, 0)
    c = _v4(cs, B, T, P) if need_rho else (0, 0, 0, 0)
    dev = values.device
    if (B * T * P) % 4 == 0:   # one allocation, both halves 16-byte aligned
        out = torch.empty((2, B, T, P, 1), dtype=torch.float32, device=dev)
        out_t, out_a = out[0], out[1]
    else:
        out_t = torch.empty((B, T, P, 1), dtype=torch.float32, device=dev)
        out_a = torch.empty((B, T, P, 1), dtype=torch.float32, device=dev)
    rc = _call_on(dev, _lib.lib().b200rl_target_scan_fused, mode_t, mode_a, *v, boot_ptr, boot_sb, boot_sp, *r, *rho,
                  *c, out_t.data_ptr(), out_a.data_ptr(), B, T, P, float(lmb), float(gamma))
    _lib.check(rc, "b200rl_target_scan_fused")
    return out_t, out_a


def _scan(mode_t, mode_a, values, returns, rewards, lmb, gamma, rhos, cs):
    res = _scan_4d(mode_t, mode_a, values, returns, rewards, lmb, gamma, rhos, cs)
    if res is not None:
        return res
    values = _as_f32_cuda(values, "values")
    returns = _as_f32_cuda(returns, "returns")
    need_rho = MODES["VTRACE"] in (mode_t, mode_a)
    ins = [values, returns[:, -1:]]
    if rewards is not None:
        rewards = _as_f32_cuda(rewards, "rewards")
        ins.append(rewards)
    if need_rho:
        rhos = _as_f32_cuda(rhos, "rhos")
        cs = _as_f32_cuda(cs, "cs")
        ins += [rhos, cs]
    else:
        rhos = cs = None
    full = _bcast_shape(*ins)
    B, T = full[0], full[1]
    P = 1
    for n in full[2:]:
        P *= n
    shape3 = (B, T, P)
    keep = []
    v_args, k = _view3(values, shape3, full)
    keep.append(k)
    # bootstrap = returns[:, -1]  (losses.py:21,32,45,55)
    boot = returns[:, -1:].expand((B, 1) + tuple(full[2:]))
    b_args, k = _view3(boot, (B, 1, P), (B, 1) + tuple(full[2:]))
    keep.append(k)
    r_args, k = _view3(rewards, shape3, full)
    keep.append(k)
    rho_args, k = _view3(rhos, shape3, full)
    keep.append(k)
    c_args, k = _view3(cs, shape3, full)
    keep.append(k)
    out_t = torch.empty(full, dtype=torch.float32, device=values.device)
    out_a = torch.empty(full, dtype=torch.float32, device=values.device)
    with torch.cuda.device(values.device):
        _lib.call(
            "b200rl_target_scan_fused", mode_t, mode_a,
            *v_args, b_args[0], b_args[1], b_args[3], *r_args, *rho_args, *c_args,
            ctypes.c_void_p(out_t.data_ptr()), ctypes.c_void_p(out_a.data_ptr()),
            B, T, P, float(lmb), float(gamma), _lib.stream_ptr(),
        )
    del keep
    return out_t, out_a


def _monte_carlo_4d(values, returns):
    vs = values.shape
    if len(vs) != 4 or vs[3] != 1 or values.dtype != torch.float32 or not values.is_cuda or vs[0] == 0:
        return None
    B, T, P = vs[0], vs[1], vs[2]
    if not _is4(returns, B, T, P):
        return None
    adv = torch.empty((B, T, P, 1), dtype=torch.float32, device=values.device)
    rc = _call_on(values.device, _lib.lib().b200rl_target_mc, *_v4(values, B, T, P), *_v4(returns, B, T, P),
                  adv.data_ptr(), B, T, P)
    _lib.check(rc, "b200rl_target_mc")
    return returns, adv


def monte_carlo(values, returns):
    res = _monte_carlo_4d(values, returns)
    if res is not None:
        return res
    values = _as_f32_cuda(values, "values")
    returns = _as_f32_cuda(returns, "returns")
    full = _bcast_shape(values, returns)
    B, T = full[0], full[1]
    P = 1
    for n in full[2:]:
        P *= n
    v_args, k1 = _view3(values, (B, T, P), full)
    r_args, k2 = _view3(returns, (B, T, P), full)
    adv = torch.empty(full, dtype=torch.float32, device=values.device)
    with torch.cuda.device(values.device):
        _lib.call("b200rl_target_mc", *v_args, *r_args, ctypes.c_void_p(adv.data_ptr()), B, T, P,
                  _lib.stream_ptr())
    del k1, k2
    return returns, adv


def temporal_difference(values, returns, rewards, lmb, gamma):
    return _scan(MODES["TD"], MODES["TD"], values, returns, rewards, lmb, gamma, None, None)


def upgo(values, returns, rewards, lmb, gamma):
    return _scan(MODES["UPGO"], MODES["UPGO"], values, returns, rewards, lmb, gamma, None, None)


def vtrace(values, returns, rewards, lmb, gamma, rhos, cs):
    return _scan(MODES["VTRACE"], MODES["VTRACE"], values, returns, rewards, lmb, gamma, rhos, cs)


def compute_target(algorithm, values, returns, rewards, lmb, gamma, rhos, cs):
    if values is None:
        return None, 0

    if algorithm == 'MC':
        return monte_carlo(values, returns)
    elif algorithm == 'TD':
        return temporal_difference(values, returns, rewards, lmb, gamma)
    elif algorithm == 'UPGO':
        return upgo(values, returns, rewards, lmb, gamma)
    elif algorithm == 'VTRACE':
        return vtrace(values, returns, rewards, lmb, gamma, rhos, cs)
    else:
        print('No algorithm named %s' % algorithm)


def compute_target_pair(value_target, policy_target, values, returns, rewards, lmb, gamma, rhos, cs):
    """Additive API: what `compute_loss` (handyrl/train.py:248-253) obtains from two
    `compute_target` calls per head — targets of `value_target` and advantages of
    `policy_target` — in ONE pass over the inputs (20 B instead of 36 B per element
    for V-Trace + UPGO).  Returns `(targets, advantages)`."""
    if values is None:
        return None, 0
    if value_target not in MODES or policy_target not in MODES:
        print('No algorithm named %s' % (value_target if value_target not in MODES else policy_target))
        return None
    if value_target == policy_target:
        return compute_target(value_target, values, returns, rewards, lmb, gamma, rhos, cs)
    if 'MC' in (value_target, policy_target):
        t, _ = compute_target(value_target, values, returns, rewards, lmb, gamma, rhos, cs)
        _, a = compute_target(policy_target, values, returns, rewards, lmb, gamma, rhos, cs)
        return t, a
    return _scan(MODES[value_target], MODES[policy_target], values, returns, rewards, lmb, gamma, rhos, cs)


def compute_target_heads(value_target, policy_target, heads, lmb, rhos, cs):
    """Additive API: `compute_target_pair` for all heads of one loss in one call.

    `heads` is a list of `(values, returns, rewards, gamma)` — what handyrl/train.py:248-253 passes for the value
    head `(values, outcome, None, 1)` and the return head `(returns_pred, returns, rewards, gamma)`; a head whose
    `values` is None yields `(None, 0)` like `compute_target`.  With two live heads of the same 4-D shape both
    recursions run in one kernel launch (`b200rl_target_scan_heads`); every other case falls back to one
    `compute_target_pair` per head.  Results are bit-identical to the per-head calls."""
    live = [h for h in heads if h[0] is not None]
    fused = None
    if (len(live) == 2 and value_target in MODES and policy_target in MODES
            and 'MC' not in (value_target, policy_target)):
        fused = _scan_heads_4d(MODES[value_target], MODES[policy_target], live, lmb, rhos, cs)
    out = []
    for h in heads:
        if h[0] is None:
            out.append((None, 0))
        elif fused is not None:
            out.append(fused.pop(0))
        else:
            out.append(compute_target_pair(value_target, policy_target, h[0], h[1], h[2], lmb, h[3], rhos, cs))
    return out


def _scan_heads_4d(mode_t, mode_a, heads, lmb, rhos, cs):
    """Both heads through b200rl_target_scan_heads; None when the inputs are not the 4-D compute_loss shapes."""
    vs = heads[0][0].shape
    if len(vs) != 4 or vs[3] != 1:
        return None
    B, T, P = vs[0], vs[1], vs[2]
    if B == 0 or (B * T * P) % 4 != 0:
        return None
    need_rho = mode_t == 3 or mode_a == 3
    if need_rho and not (rhos is not None and cs is not None and _is4(rhos, B, T, P) and _is4(cs, B, T, P)):
        return None
    dev = heads[0][0].device
    arr = (_lib.TargetHead * 2)()
    out = torch.empty((2, 2, B, T, P, 1), dtype=torch.float32, device=dev)
    rho = _v4(rhos, B, T, P) if need_rho else (0, 0, 0, 0)
    c = _v4(cs, B, T, P) if need_rho else (0, 0, 0, 0)
    for i, (values, returns, rewards, gamma) in enumerate(heads):
        if (tuple(values.shape) != tuple(vs) or values.dtype != torch.float32 or not values.is_cuda
                or values.device != dev or returns.dim() != 4 or not _is4(returns, B, returns.shape[1], P)):
            return None
        if rewards is not None and not _is4(rewards, B, T, P):
            return None
        rs, rst = returns.shape, returns.stride()
        h = arr[i]
        h.values, h.v_sb, h.v_st, h.v_sp = _v4(values, B, T, P)
        h.boot = returns.data_ptr() + 4 * rst[1] * (rs[1] - 1)          # returns[:, -1]
        h.boot_sb = rst[0] if (rs[0] == B and B > 1) else 0
        h.boot_sp = rst[2] if (rs[2] == P and P > 1) else 0
        h.rewards, h.r_sb, h.r_st, h.r_sp = _v4(rewards, B, T, P) if rewards is not None else (None, 0, 0, 0)
        h.rhos, h.rho_sb, h.rho_st, h.rho_sp = rho if need_rho else (None, 0, 0, 0)
        h.cs, h.c_sb, h.c_st, h.c_sp = c if need_rho else (None, 0, 0, 0)
        h.out_targets, h.out_adv = out[i, 0].data_ptr(), out[i, 1].data_ptr()
        h.gamma = float(gamma)
    rc = _call_on(dev, _lib.lib().b200rl_target_scan_heads, mode_t, mode_a, ctypes.addressof(arr), 2, B, T, P,
                  float(lmb))
    _lib.check(rc, "b200rl_target_scan_heads")
    return [(out[0, 0], out[0, 1]), (out[1, 0], out[1, 1])]
