"""Host-side multi-GPU logic on CPU: world_size-2 `gloo` processes.

* data parallel over graphs: per-rank edge lists with local numbering equal the slices of the global list;
  the mean of per-rank gradients (ONE all-reduce of the flat buffer) equals the full-batch oracle gradient;
* receiver partition of one scene: the local slices tile the receiver-sorted edge array, and an oracle
  forward that only ever computes the owned rows + an all-gather of node states between layers equals
  the unpartitioned oracle.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from aether_b200 import dist as ad
from oracle import aether_oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world=2, *args):
    port = _free_port()
    mp.spawn(_entry, args=(world, port, fn, args), nprocs=world, join=True)


def _entry(rank, world, port, fn, args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def test_shard_range_tiles():
    for n in (0, 1, 7, 16, 1024, 1025):
        for w in (1, 2, 3, 8):
            prev = 0
            for r in range(w):
                b, e = ad.shard_range(n, w, r)
                assert b == prev and e >= b
                prev = e
            assert prev == n
            sizes = [ad.shard_range(n, w, r)[1] - ad.shard_range(n, w, r)[0] for r in range(w)]
            assert max(sizes) - min(sizes) <= 1


def test_shard_batch_edges_equals_local_builder():
    N, B = 7, 10
    send, recv = O.fc_edges(N, B)
    for w in (1, 2, 4):
        for r in range(w):
            s, rr, g0, g1 = ad.shard_batch_edges(send, recv, N, B, w, r)
            so, ro = O.fc_edges(N, g1 - g0)
            assert torch.equal(s, so) and torch.equal(rr, ro)


def test_flat_grad_detection():
    ps = [torch.nn.Parameter(torch.zeros(3, 2)), torch.nn.Parameter(torch.zeros(5)), torch.nn.Parameter(torch.zeros(4, 4))]
    offs, tot = [], 0
    for p in ps:
        offs.append(tot)
        tot += (p.numel() + 3) // 4 * 4
    flat = torch.zeros(tot)
    for o, p in zip(offs, ps):
        p.grad = flat[o:o + p.numel()].view_as(p)
    assert ad.flat_grad(ps) is flat
    ps[1].grad = torch.zeros(5)
    assert ad.flat_grad(ps) is None


def _dp_worker(rank, world, D, N, B, H):
    p_full = {k: v.clone().requires_grad_(True) for k, v in O.init_params(D, H, seed=1).items()}
    x, v, q = O.synth_state(B, N, D, seed=3)
    tgt = x + 0.05
    send, recv = O.fc_edges(N, B)
    # full batch on every rank (the thing data parallelism must reproduce)
    _, ea = O.caller_prep(x, v, q, send, recv)
    out = O.aether_step(p_full, x, (send, recv), v, ea, q)
    torch.nn.functional.mse_loss(out, tgt).backward()
    # this rank's shard, local numbering
    s, r, g0, g1 = ad.shard_batch_edges(send, recv, N, B, world, rank)
    n0, n1 = g0 * N, g1 * N
    p = {k: t.detach().clone().requires_grad_(True) for k, t in p_full.items()}
    xs, vs, qs = x[n0:n1], v[n0:n1], q[n0:n1]
    _, eas = O.caller_prep(xs, vs, qs, s, r)
    outs = O.aether_step(p, xs, (s, r), vs, eas, qs)
    assert torch.equal(outs, out[n0:n1].detach()) or float((outs - out[n0:n1]).abs().max()) < 1e-6
    torch.nn.functional.mse_loss(outs, tgt[n0:n1]).backward()
    params = [t for t in p.values() if t.grad is not None]
    # lay the gradients out as the CUDA backward does (views of one buffer), then ONE collective
    offs, tot = [], 0
    for t in params:
        offs.append(tot)
        tot += (t.numel() + 3) // 4 * 4
    flat = torch.zeros(tot)
    for o, t in zip(offs, params):
        g = flat[o:o + t.numel()].view_as(t)
        g.copy_(t.grad)
        t.grad = g
    assert ad.flat_grad(params) is flat
    assert ad.allreduce_grads(params, world) == 1
    for k, t in p.items():
        if t.grad is None:
            continue
        want = p_full[k].grad
        scale = float(want.abs().max().clamp(min=1e-12))
        assert float((t.grad - want).abs().max()) / scale < 2e-4, k
    # fallback layout (separate tensors) gives the same result
    for t in params:
        t.grad = t.grad.clone()
    before = [t.grad.clone() for t in params]
    ad.allreduce_grads(params, world)
    for t, b in zip(params, before):
        assert torch.allclose(t.grad, b, rtol=0, atol=1e-6)   # already averaged and equal on all ranks


@pytest.mark.timeout(300)
def test_data_parallel_gradients_gloo():
    _spawn(_dp_worker, 2, 2, 5, 4, 32)


def _partitioned_forward(p, x, vel, q, send, recv, world, rank, group=None):
    """Oracle forward that computes only the rows this rank owns; all other rows come from all-gathers."""
    Nn = x.shape[0]
    perm, rowptr, ss, rs = O.receiver_csr(send, recv, Nn)
    part = ad.ReceiverPartition(torch.from_numpy(rowptr), Nn, world, rank)
    n0, n1, e0, e1 = part.node_begin, part.node_end, part.edge_begin, part.edge_end
    sl = torch.from_numpy(ss[e0:e1])
    rl = torch.from_numpy(rs[e0:e1])
    assert bool(((rl >= n0) & (rl < n1)).all())
    lrp = part.local_rowptr(torch.from_numpy(rowptr))
    assert int(lrp[-1]) == e1 - e0 and lrp.numel() == n1 - n0 + 1
    silu = torch.nn.functional.silu
    f = O.field_forward(p, x, vel, q)                      # per node, cheap: every rank computes all rows
    _, ea_l = O.caller_prep(x, vel, q, sl, rl)             # caller prep on the local sorted edges
    rel_feat, R, attr = O.localize(x, vel, f, sl, rl)
    e = torch.cat([attr, ea_l], -1)
    H = p["gnn.layer_1.message_fn.2.weight"].shape[0]
    xbuf = torch.zeros(part.padded_nodes, H)
    h_own = None
    for l in range(1, 5):
        name = "gnn.layer_%d" % l
        if l == 1:
            m = e
        else:
            part.all_gather_rows(xbuf, group)              # the exchange step
            m = torch.cat([xbuf[sl], xbuf[rl], e], -1)
        e = silu(O._lin(p, name + ".message_fn.2", silu(O._lin(p, name + ".message_fn.0", m))))
        agg = torch.zeros(n1 - n0, H).index_add_(0, rl - n0, e)
        deg = (lrp[1:] - lrp[:-1]).clamp(min=1).to(agg.dtype).unsqueeze(1)
        base = O._lin(p, name + ".res", rel_feat[n0:n1]) if l == 1 else h_own
        y = base + agg / deg
        h_own = y + O._lin(p, name + ".update_fn.2", silu(O._lin(p, name + ".update_fn.0", y)))
        xbuf[n0:n1] = h_own
    h = silu(O._lin(p, "gnn.out_mlp.0", h_own))
    h = silu(O._lin(p, "gnn.out_mlp.3", h))
    pred = O._lin(p, "gnn.out_mlp.6", h)
    out_own = x[n0:n1] + O.apply_rot(R[n0:n1], pred)
    D = x.shape[1]
    obuf = torch.zeros(part.padded_nodes, D)
    obuf[n0:n1] = out_own
    part.all_gather_rows(obuf, group)
    return obuf[:Nn], part


def _part_worker(rank, world, D, N, k):
    p = O.init_params(D, 32, seed=2)
    x, v, q = O.synth_state(1, N, D, seed=5)
    send, recv = O.knn_edges(x, k)                         # irregular in-degrees, some receivers isolated
    _, ea = O.caller_prep(x, v, q, send, recv)
    with torch.no_grad():
        want = O.aether_step(p, x, (send, recv), v, ea, q)
        got, part = _partitioned_forward(p, x, v, q, send, recv, world, rank)
    assert float((got - want).abs().max()) <= 2e-5 * float(want.abs().max())
    # the slices of all ranks tile the edge array
    cuts = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(cuts, torch.tensor([part.edge_begin, part.edge_end]))
    assert int(cuts[0][0]) == 0 and int(cuts[-1][1]) == send.numel()
    for a, b in zip(cuts[:-1], cuts[1:]):
        assert int(a[1]) == int(b[0])


@pytest.mark.timeout(300)
@pytest.mark.parametrize("D,N,k", [(2, 37, 6), (3, 64, 9)])
def test_receiver_partition_gloo(D, N, k):
    _spawn(_part_worker, 2, D, N, k)


def test_receiver_partition_plan_uneven():
    rowptr = torch.tensor(np.concatenate([[0], np.cumsum([3, 0, 2, 5, 1, 0, 4])]))
    tot = 0
    for r in range(3):
        part = ad.ReceiverPartition(rowptr, 7, 3, r)
        assert part.block == 3 and part.padded_nodes == 9
        tot += part.n_local_edges
        assert part.n_own == (3 if r < 2 else 1)
    assert tot == 15
