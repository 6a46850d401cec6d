"""torch.library registration of the inference step: schema, fake (meta) kernel, and no CPU kernel."""
import pytest
import torch
from torch._subclasses.fake_tensor import FakeTensorMode

import aether_b200.torch_ops  # noqa: F401  (registers torch.ops.aether_b200.step)
from aether_b200 import _lib


def _args(dev, Nn=12, E=40, D=3, n_params=_lib.NUM_PARAMS):
    x = torch.empty(Nn, D, device=dev)
    i32 = lambda n: torch.empty(n, dtype=torch.int32, device=dev)
    params = [torch.empty(4, 4, device=dev) for _ in range(n_params)]
    return (x, x, torch.empty(Nn, 1, device=dev), torch.empty(E, 2, device=dev), i32(E), i32(Nn + 1), i32(E), i32(E),
            i32(E), i32(Nn + 1), params, E, 64, 0)


def test_schema():
    s = str(torch.ops.aether_b200.step.default._schema)
    assert s.startswith("aether_b200::step(Tensor x, Tensor vel, Tensor charges, Tensor edge_attr_orig")
    assert "Tensor[] params" in s and s.endswith("-> Tensor")


@pytest.mark.parametrize("D", [2, 3])
def test_fake_kernel_propagates_shape(D):
    with FakeTensorMode():
        out = torch.ops.aether_b200.step(*_args("meta", D=D))
        assert tuple(out.shape) == (12, D) and out.dtype == torch.float32


def test_fake_kernel_checks_arguments():
    with FakeTensorMode():
        with pytest.raises(RuntimeError):
            torch.ops.aether_b200.step(*_args("meta", n_params=3))
        a = list(_args("meta"))
        a[5] = torch.empty(5, dtype=torch.int32, device="meta")        # rowptr of the wrong length
        with pytest.raises(RuntimeError):
            torch.ops.aether_b200.step(*a)


def test_no_cpu_kernel():
    with pytest.raises(NotImplementedError):
        torch.ops.aether_b200.step(*_args("cpu"))
