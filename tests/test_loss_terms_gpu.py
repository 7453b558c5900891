"""GPU: fused loss pre-processing (handyrl/train.py:224-239) against the torch expressions of the reference."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _case(B, T, Po, A, seed):
    g = torch.Generator(device='cuda').manual_seed(seed)
    legal = torch.rand(B, T, Po, A, device='cuda', generator=g) < 0.2
    legal[..., 0] = True
    amask = torch.where(legal, 0.0, 1e32).float()
    pol_b = torch.randn(B, T, Po, A, device='cuda', generator=g) - amask
    pol_t = (torch.randn(B, T, Po, A, device='cuda', generator=g) * 2 - amask).requires_grad_(True)
    # actions: a legal one per row
    act = torch.multinomial(legal.reshape(-1, A).float(), 1, generator=g).reshape(B, T, Po, 1)
    emask = (torch.rand(B, T, 1, 1, device='cuda', generator=g) < 0.8).float()
    return pol_b, pol_t, act, emask


@pytest.mark.parametrize("B,T,Po,A", [(7, 5, 1, 214), (4, 16, 2, 214), (33, 3, 1, 9), (1, 1, 1, 214), (64, 32, 1, 214)])
def test_policy_terms_forward_backward(B, T, Po, A):
    from handyrl_b200.learner import _PolicyTerms
    pol_b, pol_t, act, emask = _case(B, T, Po, A, 3)
    logp, rho, c = _PolicyTerms.apply(pol_t, pol_b, act, emask)
    ref_b = F.log_softmax(pol_b, dim=-1).gather(-1, act) * emask
    pol_r = pol_t.detach().clone().requires_grad_(True)
    ref_t = F.log_softmax(pol_r, dim=-1).gather(-1, act) * emask
    ref_rho = torch.clamp(torch.exp(ref_t.detach() - ref_b), 0, 1.0)
    assert logp.shape == ref_t.shape and rho.shape == ref_rho.shape
    torch.testing.assert_close(logp, ref_t, rtol=1e-5, atol=2e-6)     # tolerance: fp32 reduction order
    torch.testing.assert_close(rho, ref_rho, rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(c, ref_rho, rtol=1e-5, atol=2e-6)
    w = torch.randn_like(logp)
    (logp * w).sum().backward()
    (ref_t * w).sum().backward()
    torch.testing.assert_close(pol_t.grad, pol_r.grad, rtol=1e-5, atol=2e-6)
    assert not rho.requires_grad and not c.requires_grad


@pytest.mark.parametrize("P,sym", [(2, True), (2, False), (1, False)])
def test_value_blend_bit_exact(P, sym):
    from handyrl_b200.learner import value_blend
    g = torch.Generator(device='cuda').manual_seed(1)
    B, T = 37, 9
    v = torch.tanh(torch.randn(B, T, P, 1, device='cuda', generator=g))
    omask = (torch.rand(B, T, P, 1, device='cuda', generator=g) < 0.5).float()
    emask = (torch.rand(B, T, 1, 1, device='cuda', generator=g) < 0.7).float()
    outcome = torch.randint(-1, 2, (B, 1, P, 1), device='cuda', generator=g).float()
    v = v * omask
    out = value_blend(v, omask, emask, outcome, sym)
    ref = v
    if sym:
        opp = -torch.stack([v[:, :, 1], v[:, :, 0]], dim=2)
        ref = (v + opp) / (omask.sum(dim=2, keepdim=True) + 1e-8)
    ref = ref * emask + outcome * (1 - emask)
    assert torch.equal(out, ref)
