"""Fused AdamW (csrc/optim.cu) against torch.optim.AdamW -- the optimizer of the reference's caller
(experiments/lorentz/main.py:164,292) -- on the model's own parameter shapes, several steps, and under CUDA-graph replay."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _params(seed):
    import aether_b200 as ab
    torch.manual_seed(seed)
    m = ab.Aether(6, 64, 0.0, 3, device="cuda")
    return [p.detach().clone().requires_grad_(True) for p in m.parameters()]


@pytest.mark.parametrize("wd", [1e-12, 1e-2])
def test_fused_adamw_matches_torch(wd):
    from aether_b200.optim import FusedAdamW
    a, b = _params(3), _params(3)
    oa = FusedAdamW(a, lr=5e-4, weight_decay=wd)
    ob = torch.optim.AdamW(b, lr=5e-4, weight_decay=wd)
    g = torch.Generator(device="cuda").manual_seed(5)
    for step in range(6):
        for pa, pb in zip(a, b):
            gr = torch.randn(pa.shape, device="cuda", generator=g) * (0.1 + step)
            pa.grad, pb.grad = gr.clone(), gr.clone()
        oa.step()
        ob.step()
    assert oa.step_count() == 6
    for pa, pb in zip(a, b):
        assert torch.allclose(pa, pb, rtol=2e-6, atol=1e-8), float((pa - pb).abs().max())
        assert torch.allclose(oa.state[pa]["exp_avg"], ob.state[pb]["exp_avg"], rtol=1e-5, atol=1e-7)
        assert torch.allclose(oa.state[pa]["exp_avg_sq"], ob.state[pb]["exp_avg_sq"], rtol=1e-5, atol=1e-9)


def test_fused_adamw_is_capturable():
    from aether_b200.optim import FusedAdamW
    a, b = _params(4), _params(4)
    oa, ob = FusedAdamW(a, lr=1e-3), FusedAdamW(b, lr=1e-3)
    grads = [torch.randn_like(p) for p in a]
    for pa, pb, gr in zip(a, b, grads):
        pa.grad, pb.grad = gr.clone(), gr.clone()
    oa.step()                                  # state created outside capture
    ob.step()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            oa.step()
    for _ in range(3):
        g.replay()
        ob.step()
    torch.cuda.synchronize()
    assert oa.step_count() == ob.step_count() == 4
    for pa, pb in zip(a, b):
        assert torch.equal(pa, pb)
