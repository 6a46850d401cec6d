"""torchrun --nproc-per-node N tools/peer_check.py [H] [n]: the receiver-partitioned rollout with the library's own
peer-memory exchange (csrc/peer.cu) against the same rollout with NCCL all-gathers: bitwise equality, then timing of both
as captured graphs (CUDA events, max over ranks)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import aether_b200 as ab  # noqa: E402


def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1)
    model = ab.Aether(6, H, 0.0, 3, device=dev, precision="bf16").eval()
    g = torch.Generator().manual_seed(1)
    side = (n / 4.0) ** (1 / 3)                       # density ~4 per unit volume: ~32 neighbours within radius 1.25
    x = (torch.rand(n, 3, generator=g) * side).to(dev)
    v = (0.1 * torch.randn(n, 3, generator=g)).to(dev)
    q = (torch.randint(0, 2, (n,), generator=g).float() * 2 - 1).to(dev)
    send, recv = ab.radius_graph(x, 1.25, 32)
    T = 5
    out = {}
    rolls = {}
    for ex in ("nccl", "peer"):
        for graph in (False, True):
            r = ab.PartitionedRollout(model, send, recv, n, T, world=world, rank=rank, cuda_graph=graph, exchange=ex)
            traj = r.run(x, v, q).clone()
            torch.cuda.synchronize()
            assert not r.peer_error(), "peer exchange timed out"
            rolls[(ex, graph)] = (r, traj)
    ref = rolls[("nccl", False)][1]
    for k, (r, traj) in rolls.items():
        same = bool(torch.equal(traj, ref))
        out["equal_%s_%s" % (k[0], "graph" if k[1] else "eager")] = same
        assert same, k
    for ex in ("nccl", "peer"):
        r = rolls[(ex, True)][0]
        for _ in range(3):
            r.run(x, v, q)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); r.run(x, v, q); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        t = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["ms_per_rollout_" + ex] = float(t.item())
        out["launches_" + ex] = r.launches_per_replay
        assert not r.peer_error()
    if rank == 0:
        out.update(world=world, H=H, particles=n, edges=int(send.numel()), steps=T)
        print(json.dumps(out), flush=True)
    for r, _ in rolls.values():
        r.release()
    del rolls
    torch.cuda.synchronize(); dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
