"""aether_graph_prepare_static: the no-allocation variant of aether_graph_prepare (scratch in a static __device__ array
inside the library; no counterpart in the reference).

Called through the C ABI and compared with the existing entry point on the same inputs, two calls in a row (the second on
a smaller graph, so stale contents of the shared array would show).  Integer path: the bar is bit-exact.  This file is
named so that pytest collects it last.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import aether_oracle as O

pytestmark = pytest.mark.gpu

MAX_NODES, MAX_EDGES = 65536, 524288


class Arrays:
    """Caller-owned inputs / outputs of one graph_prepare call (allocated by the test, not by the entry point)."""

    def __init__(self, _lib, send, recv, n_nodes):
        dev = "cuda"
        self.n_nodes, self.E = int(n_nodes), int(send.numel())
        self.send64 = send.to(dev, torch.int64).contiguous()
        self.recv64 = recv.to(dev, torch.int64).contiguous()
        i32 = dict(dtype=torch.int32, device=dev)
        # poisoned outputs: every word the call must write is checked against the existing entry point's
        self.out = {k: torch.full((max(self.E, 1),), -7, **i32) for k in ("perm", "send", "recv", "sperm")}
        self.out["rowptr"] = torch.full((self.n_nodes + 1,), -7, **i32)
        self.out["colptr"] = torch.full((self.n_nodes + 1,), -7, **i32)
        self.status = torch.full((1,), -7, **i32)
        self.c = _lib.AetherGraphC(*(self.out[k].data_ptr() for k in ("perm", "rowptr", "send", "recv", "sperm", "colptr")))


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def run_existing(lib, _lib, a):
    nbytes = lib.aether_graph_scratch_bytes(a.n_nodes, a.E)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _lib.check(lib.aether_graph_prepare(_lib.ptr(a.send64), _lib.ptr(a.recv64), a.E, a.n_nodes, ctypes.byref(a.c),
                                        _lib.ptr(a.status), _lib.ptr(scratch), nbytes, stream()), "aether_graph_prepare")
    torch.cuda.synchronize()


def run_static(lib, _lib, a):
    rc = lib.aether_graph_prepare_static(_lib.ptr(a.send64), _lib.ptr(a.recv64), a.E, a.n_nodes, ctypes.byref(a.c),
                                         _lib.ptr(a.status), stream())
    _lib.check(rc, "aether_graph_prepare_static")
    torch.cuda.synchronize()


def graphs():
    """Two inputs: ragged random multigraph with isolated nodes and an unsorted edge order, then a smaller batched
    fully connected list (the lorentz shape)."""
    g = torch.Generator().manual_seed(3)
    n1, e1 = 5000, 60000
    s1 = torch.randint(0, n1 - 100, (e1,), generator=g)          # the last 100 nodes are isolated
    r1 = torch.randint(0, n1 - 100, (e1,), generator=g)
    s2, r2 = O.fc_edges(5, 16)
    return [(s1, r1, n1), (s2, r2, 80)]


def test_static_graph_prepare_matches_existing_entry_point_twice():
    from aether_b200 import _lib
    lib = _lib.load()
    for k, (send, recv, n) in enumerate(graphs()):
        want = Arrays(_lib, send, recv, n)
        got = Arrays(_lib, send, recv, n)
        run_existing(lib, _lib, want)
        if k == 1:
            # second call: nothing is allocated on the device by the call (the first one may load the module lazily)
            free0 = torch.cuda.mem_get_info()[0]
            alloc0 = torch.cuda.memory_allocated()
        run_static(lib, _lib, got)
        if k == 1:
            assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_allocated() == alloc0
        assert int(got.status.item()) == 0 and int(want.status.item()) == 0
        for name in want.out:
            assert torch.equal(got.out[name], want.out[name]), (k, name)
        # and against the oracle's receiver CSR (the existing entry point's own bar)
        perm, rowptr, ss, rs = O.receiver_csr(send, recv, n)
        assert np.array_equal(got.out["rowptr"].cpu().numpy(), rowptr.astype(np.int32))
        assert np.array_equal(got.out["perm"].cpu().numpy()[:got.E], perm.astype(np.int32))


def test_static_graph_prepare_limits_and_errors():
    from aether_b200 import _lib
    lib = _lib.load()
    g = _lib.AetherGraphC()
    one = ctypes.c_void_p(256)
    n0 = lib.aether_launch_count()
    assert lib.aether_graph_prepare_static(one, one, MAX_EDGES + 1, 10, ctypes.byref(g), one, None) == -4
    assert lib.aether_graph_prepare_static(one, one, 10, MAX_NODES + 1, ctypes.byref(g), one, None) == -4
    assert lib.aether_graph_prepare_static(one, one, -1, 10, ctypes.byref(g), one, None) == -1
    assert lib.aether_launch_count() == n0                       # rejected before anything is launched
    msg = lib.aether_error_string(-4)
    assert b"65536" in msg and b"524288" in msg
    # the largest accepted size runs (edges of a ring over MAX_NODES nodes, repeated to MAX_EDGES)
    idx = torch.arange(MAX_EDGES, dtype=torch.int64)
    send, recv = idx % MAX_NODES, (idx * 7 + 1) % MAX_NODES
    want, got = Arrays(_lib, send, recv, MAX_NODES), Arrays(_lib, send, recv, MAX_NODES)
    run_existing(lib, _lib, want)
    run_static(lib, _lib, got)
    for name in want.out:
        assert torch.equal(got.out[name], want.out[name]), name
