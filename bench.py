#!/usr/bin/env python
"""Benchmark of the Aether state-to-state step (BASELINE.json metric: edge-messages/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2b|c2|c1] [--hidden H]
                    [--precision fp32|bf16] [--impl ours|reference]

Prints ONE JSON line (rank 0).  A step is one pass of the hot path over one batch of synthetic
particles: training workloads run caller prep + forward + MSE + backward + AdamW
(experiments/lorentz/main.py:243-247,289-292); rollout workloads run T rollout steps.
  value : whole-job edge-messages/s (4 per edge per step) with inputs resident in HBM
  e2e   : the same through the public module API with pinned HOST inputs copied every step,
          the edge list rebuilt on the device every step and the loss read back
  roofline : edge-MLP kernels, algorithmic FLOPs / CUDA-event time of those launches
  cpu_baseline : the CPU oracle (torch CPU restatement of the reference) on a bounded sample
  secondary (default c4 run only): config 4 STRONG scaling (1024 graphs split over the ranks), config 5 (one
          100k-particle scene, receiver-partitioned) at H = 256 and 64 -- the other scaling cases of SURVEY.md 8(d/e) -- and,
          at N = 1, `gpu_eager_baseline` (the oracle port as eager PyTorch on the same GPU) and `parity` (errors against
          the oracle at the config-4 shape, 20-step rollout MSE at the config-1 shape)
`--impl reference` times that CPU path alone.  Nothing here reads /root/reference.
"""
import argparse
import gc
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (D, N, graphs per GPU, H default, mode, rollout steps)
    "c4": dict(D=2, N=100, B=1024, H=256, mode="train", T=1,
               name="config4: synthetic 2D charged, 100 particles fully connected (9900 edges/graph), batch 1024 per GPU, train step"),
    "c2b": dict(D=3, N=20, B=128, H=64, mode="train", T=1, name="lorentz dynamic_20body shape, batch 128, train step"),
    "c2": dict(D=3, N=5, B=128, H=64, mode="train", T=1, name="config2: gravitational 3D, 5 particles, batch 128, train step"),
    "c1": dict(D=2, N=5, B=16, H=64, mode="rollout", T=20, name="config1: electrostatic 2D, 5 particles, batch 16, 20-step rollout"),
    "c3": dict(D=2, N=0, B=256, H=64, mode="rollout", T=12, knn=10,
               name="config3: inD-style scenes, 10-40 agents (seed 1), kNN k=10 rebuilt every step, 256 scenes per GPU, 12-step rollout"),
    "c5": dict(D=3, N=100000, B=1, H=64, mode="partitioned-rollout", T=5, k=32, radius=2.25,
               name="config5: synthetic 3D gravitational scene, 100k particles (unit density), radius graph r=2.25 capped at the "
                    "32 nearest (~3.2M edges), ONE scene receiver-partitioned over the GPUs, 5-step rollout"),
}


class Ctx:
    """One process per GPU: rank / world / device and the NCCL process group (initialised once per process)."""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist

    def close(self):
        if self.dist is not None:
            torch.cuda.synchronize()
            self.dist.barrier()
            self.dist.destroy_process_group()
            self.dist = None


def release_memory():
    gc.collect()
    torch.cuda.empty_cache()


def scene_sizes(cfg, rank=0):
    """Agents per scene: N for the fixed-size workloads, U{10..40} (seed 1 + rank) for the inD-style one."""
    if cfg["N"]:
        return [cfg["N"]] * cfg["B"]
    g = torch.Generator().manual_seed(1 + rank)
    return [int(v) for v in torch.randint(10, 41, (cfg["B"],), generator=g)]


def synth_inputs(cfg, sizes, seed):
    """(x, v, q): the simulators' distributions for the particle workloads; inD-style scenes use positions in the
    unit square and speeds U(0,1)*0.05 (speed-normalised tracks)."""
    from aether_b200 import synth
    D = cfg["D"]
    if cfg["N"]:
        return synth.synth_state(len(sizes), cfg["N"], D, seed=seed)
    return synth.ind_style_scenes(sizes, D, seed=seed)


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def algorithmic_flops_edge_mlp(D, H, E):
    """Forward FLOPs of the four message_fn MLPs in the reference's formulation."""
    fe = 7 * D + D * (D - 1) // 2 + 2
    return 2.0 * E * (fe * H + 13 * H * H)


def hbm_rooflines(prof, D, H, E, Nn, bf16, pk):
    """HBM-bound kernels of the path (north_star subsystem 2): the receiver-sorted segmented mean and the
    gather + frame-transform feature kernel.  achieved = algorithmic bytes per launch / mean CUDA-event time of
    the launches (library-side event pairs); peak = measured copy bandwidth."""
    out = {}
    peak = pk.get("hbm_gbs")
    esz = 2 if bf16 else 4
    fe = 7 * D + D * (D - 1) // 2 + 2
    specs = {"segmented_mean (receiver-sorted, no atomics)": (6, E * H * esz + Nn * H * 4 * 2 + (Nn + 1) * 4),
             "edge_features (gather + frame transform)": (4, E * (12 + 8 + (128 if bf16 else 4 * ((fe + 3) // 4 * 4))))}
    for name, (cat, nbytes) in specs.items():
        n = prof["spans"][cat] if cat < len(prof["spans"]) else 0
        if not n or prof["ms"][cat] <= 0:
            continue
        ms = prof["ms"][cat] / n
        ach = nbytes / (ms * 1e-3) / 1e9
        out[name] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak if peak else None,
                     "algorithmic_bytes_per_launch": nbytes, "avg_ms_per_launch": ms, "launches": n}
    return out


def executed_flops_edge_mlp(D, H, E):
    fe = 7 * D + D * (D - 1) // 2 + 2
    return 2.0 * E * (fe * H + H * H + 3 * 2 * H * H)     # node-factorised: 2 H^2 per edge in layers 2-4


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_step_sample(cfg, H, graphs, seed=1):
    """Returns (step_fn, edges_per_step): one CPU-oracle pass of the workload on `graphs` graphs."""
    from oracle import aether_oracle as O
    D, N = cfg["D"], cfg["N"]
    p = {k: v.requires_grad_(True) for k, v in O.init_params(D, H, seed).items()}
    sizes = scene_sizes(cfg)[:graphs]
    x, v, q = synth_inputs(cfg, sizes, seed)
    if cfg.get("knn"):
        k = cfg["knn"]
        E = sum(n * min(k, max(n - 1, 0)) for n in sizes)
        T = cfg["T"]

        def kstep():
            with torch.no_grad():
                O.rollout(p, x, v, q, lambda xx: O.knn_edges_batched(xx, sizes, k), T)
        return kstep, E * T
    send, recv = O.fc_edges(N, graphs)
    E = send.numel()
    if cfg["mode"] == "train":
        opt = torch.optim.AdamW(list(p.values()), lr=5e-4, weight_decay=1e-12)
        target = x + 0.05

        def step():
            opt.zero_grad()
            _, ea = O.caller_prep(x, v, q, send, recv)
            out = O.aether_step(p, x, (send, recv), v, ea, q)
            loss = torch.nn.functional.mse_loss(out, target)
            loss.backward()
            opt.step()
            return float(loss.detach())
        return step, E
    T = cfg["T"]

    def step():
        with torch.no_grad():
            O.rollout(p, x, v, q, lambda xx: (send, recv), T)
    return step, E * T


def cpu_sample_graphs(cfg, H):
    E = cfg["N"] * (cfg["N"] - 1) if cfg["N"] else 25 * cfg.get("knn", 10)
    budget = 8.0e5 * (64.0 / H) ** 2 if cfg["mode"] == "train" else 2.0e6   # edges per sample step (c4, H=256: 4 graphs)
    return int(max(1, min(cfg["B"], budget // max(E, 1))))


def c5_scene(cfg, n, seed=1):
    """Uniform unit-density particles in a cube, speeds 0.05, charges +-1 (config 5)."""
    from aether_b200 import synth
    return synth.unit_density_scene(n, seed)


def c5_cpu_sample(cfg, H, n_sub=6250):
    """CPU oracle on a same-density sub-scene (graph built once outside the timed region, like the GPU path)."""
    from oracle import aether_oracle as O
    p = O.init_params(3, H, 1)
    x, v, q = c5_scene(cfg, n_sub)
    send, recv = O.radius_knn_edges(x, cfg["radius"], cfg["k"])
    T = cfg["T"]

    def step():
        with torch.no_grad():
            O.rollout(p, x, v, q, lambda xx: (send, recv), T)
    return step, int(send.numel()) * T, "%d of %d particles at the same density (%d edges), %d-step rollout" % (
        n_sub, cfg["N"], int(send.numel()), T)


def run_c5(args, cfg, H, ctx, steps, warmup, do_e2e=True, do_cpu=True):
    """Config 5: one 100k-particle scene, receiver-partitioned over the ranks (strong scaling: fixed total work).
    Returns the JSON line as a dict on rank 0, None elsewhere."""
    import aether_b200 as ab
    from aether_b200 import _lib
    lib = _lib.load(build_if_missing=False)
    rank, world, local, dev, dist = ctx.rank, ctx.world, ctx.local, ctx.dev, ctx.dist
    n = cfg["N"] if not args.batch else args.batch          # --batch overrides the particle count for c5
    T = cfg["T"]
    torch.manual_seed(1)
    model = ab.Aether(6, H, 0.0, 3, device=dev, precision=args.precision)
    model.eval()
    x, v, q = c5_scene(cfg, n)
    host = [t.pin_memory() for t in (x, v, q)]
    xd, vd, qd = [t.to(dev) for t in host]
    send, recv = ab.radius_graph(xd, cfg["radius"], cfg["k"])          # replicated on every rank, built once
    E = int(send.numel())
    # The T-step loop -- on several GPUs including its NCCL all-gathers -- is replayed as ONE CUDA graph (8 GPUs, H = 256:
    # 9.5 vs 10.4 ms per 5-step rollout eager; H = 64: 3.95 vs 4.85 ms; profiles/r02_c5_8gpu.md).  The captured graphs are
    # released before the communicator is torn down (roll.release() below); --no-graph launches eagerly.
    use_graph = not args.no_graph
    # The exchange between phases is the library's own (csrc/peer.cu: stores into the peers' symmetric-memory copies,
    # flags, no NCCL call on the data path); AETHER_B200_EXCHANGE=nccl selects torch.distributed's all-gather.
    exchange = os.environ.get("AETHER_B200_EXCHANGE", "peer") if world > 1 else "nccl"
    try:
        roll = ab.PartitionedRollout(model, send, recv, n, T, world=world, rank=rank, cuda_graph=use_graph, exchange=exchange)
    except Exception as e:                       # no symmetric memory on this box: say so and use NCCL
        if exchange != "peer":
            raise
        print("bench: peer-memory exchange unavailable (%r), using NCCL all-gathers" % (e,), file=sys.stderr, flush=True)
        exchange = "nccl"
        roll = ab.PartitionedRollout(model, send, recv, n, T, world=world, rank=rank, cuda_graph=use_graph)
    eager = ab.PartitionedRollout(model, send, recv, n, T, world=world, rank=rank, exchange=exchange) if use_graph else roll

    def step_resident():
        roll.run(xd, vd, qd)

    def step_e2e():
        xs, vs, qs = [t.to(dev, non_blocking=True) for t in host]
        traj = roll.run(xs, vs, qs)
        return float(traj[-1].abs().mean().item())

    def timed(fn, steps, warmup, profile=False):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        if profile:
            lib.aether_profile_enable(1)
        n0 = lib.aether_launch_count()
        evs = []
        for _ in range(steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        launches = lib.aether_launch_count() - n0
        total = sum(a.elapsed_time(b) for a, b in evs) * 1e-3
        prof = None
        if profile:
            lib.aether_profile_enable(0)
            import ctypes
            ms = (ctypes.c_double * 7)()
            cnt = (ctypes.c_int64 * 7)()
            lib.aether_profile_collect(ms, cnt, 7)
            prof = {"ms": list(ms), "spans": list(cnt)}
        if dist is not None:
            t = torch.tensor([total], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total = float(t.item())
        return total, launches, prof

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    total, launches, prof = timed(step_resident, steps, warmup, profile=True)
    launches += steps * roll.launches_per_replay          # kernels launched by graph replays
    clocks = sampler.stop() if rank == 0 else None
    if eager is not roll:       # graph replays carry no library-side event pairs: time the same kernels launched eagerly
        _, _, prof = timed(lambda: eager.run(xd, vd, qd), max(1, min(steps, 3)), 1, profile=True)
    e2e = None
    if do_e2e:
        e2e_total, _, _ = timed(step_e2e, steps, max(1, warmup - 1))
        e2e = 4.0 * E * T * steps / e2e_total
    # captured graphs (and the communicator work they hold) are dropped before anything else uses the communicator
    roll.release()
    if roll.peer_error() or eager.peer_error():
        raise RuntimeError("bench: a wait of the peer-memory exchange gave up (a rank did not arrive)")
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    if rank != 0:
        return None
    pk, pk_src = peaks()
    e_local = roll.local[0].E
    spans = max(prof["spans"][0], 1)
    ms_per_layer = prof["ms"][0] / spans
    alg = algorithmic_flops_edge_mlp(3, H, e_local) / 4.0
    achieved = alg / (ms_per_layer * 1e-3) / 1e12 if ms_per_layer > 0 else 0.0
    peak = pk.get("bf16_tflops_sustained", pk.get("bf16_tflops"))
    line = {"metric": "edge_messages_per_sec", "value": 4.0 * E * T * steps / total, "unit": "edge-messages/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": total / steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": {"workload": cfg["name"], "num_dims": 3, "particles": n, "edges": E, "edges_rank0": e_local,
                       "hidden": H, "mode": cfg["mode"], "rollout_steps": T, "precision": args.precision,
                       "parallelism": "receiver-partition x%d, all-gather of [Nn,H] sender terms before layers 2-4 "
                                      "and of [Nn,3] positions per step" % world,
                       "l2": "working set exceeds L2" if E * H * 2 * 2 > 126e6 * world else "flushed: no (L2-resident edge state)"},
            "particle_steps_per_sec": n * T * steps / total, "gpu_launches": int(launches), "clocks": clocks,
            "collectives_per_step": 4 * T if world > 1 else 0,
            "exchange": ("peer-memory stores + flags (csrc/peer.cu), no NCCL on the data path" if exchange == "peer"
                         else "NCCL all-gather") if world > 1 else None,
            "roofline": {"bound": "tensor", "kernel": "edge-MLP forward (message_fn, per layer, rank 0's edges)",
                         "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                         "traffic": None, "peak_source": pk_src + " bf16 sustained",
                         "algorithmic_flops_per_launch_group": alg, "avg_ms_per_launch_group": ms_per_layer},
            "roofline_hbm": hbm_rooflines(prof, 3, H, e_local, roll.local[0].part.n_own,
                                          args.precision == "bf16" and H in (64, 128, 256), pk)}
    if e2e is not None:
        h2d = sum(t.numel() * t.element_size() for t in host)
        line["e2e"] = {"value": e2e, "unit": "edge-messages/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4}
    if do_cpu:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        cstep, e_step, sample = c5_cpu_sample(cfg, H)
        cstep()
        t0, it = time.perf_counter(), 0
        while it < 2 or (time.perf_counter() - t0 < args.cpu_seconds and it < 50):
            cstep()
            it += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": 4.0 * e_step * it / dt, "unit": "edge-messages/s", "cores": cores, "kind": "port",
                                "sample": sample + ", %d repeats, %d host threads" % (it, cores)}
    return line


def run_reference(args, cfg, H):
    """--impl reference: the CPU oracle port on the host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if cfg["mode"] == "partitioned-rollout":
        step, e_step, sample0 = c5_cpu_sample(cfg, H)
        graphs = 1
    else:
        graphs = cpu_sample_graphs(cfg, H)
        step, e_step = cpu_step_sample(cfg, H, graphs)
        sample0 = None
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = 4.0 * e_step * args.steps / dt
    sample = "%d of %d graphs per step (%d edges), all %d host threads" % (graphs, cfg["B"], e_step, cores)
    if sample0:
        sample = sample0 + ", all %d host threads" % cores
    line = {"impl": "reference", "metric": "edge_messages_per_sec", "value": val, "unit": "edge-messages/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong" if sample0 else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": cfg["name"], "hidden": H, "mode": cfg["mode"], "sample": sample},
            "cpu_baseline": {"value": val, "unit": "edge-messages/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "edge-messages/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_dp(args, cfg, H, ctx, steps, warmup, do_e2e=True, do_cpu=True, scaling="weak"):
    """Data-parallel workloads (configs 1-4): cfg["B"] graphs PER GPU, one process per GPU, no data-path collective,
    one gradient all-reduce per training step.  Returns the JSON line as a dict on rank 0, None elsewhere."""
    import aether_b200 as ab
    from aether_b200 import _lib
    lib = _lib.load(build_if_missing=False)

    rank, world, local, dev, dist = ctx.rank, ctx.world, ctx.local, ctx.dev, ctx.dist

    D, N, B, T = cfg["D"], cfg["N"], cfg["B"], cfg["T"]
    torch.manual_seed(1)
    model = ab.Aether(2 * D, H, 0.0, D, device=dev, precision=args.precision)
    sizes = scene_sizes(cfg, rank)
    x, v, q = synth_inputs(cfg, sizes, 1 + rank)
    target = x + 0.05 * torch.randn(x.shape, generator=torch.Generator().manual_seed(7 + rank))
    Nn = x.shape[0]
    host = [t.pin_memory() for t in (x, v, q, target)]
    xd, vd, qd, td = [t.to(dev) for t in host]
    qd1 = qd.reshape(-1)
    plan = None
    if cfg.get("knn"):
        plan = ab.KnnPlan(sizes, cfg["knn"], dev)
        E, send, recv, graph = plan.n_edges, None, None, None
    else:
        send, recv = ab.fc_edges(N, B, dev)
        E = int(send.numel())
        graph = ab.PreparedGraph(send, recv, Nn)
    train = cfg["mode"] == "train"
    params = model.param_list()
    flat_n = sum(p.numel() for p in params)
    if train:
        opt = ab.FusedAdamW(model.parameters(), lr=5e-4, weight_decay=1e-12)       # one launch for all 47 tensors
        model.train()
    else:
        model.eval()
        # scenes that fit one SM take the single-launch kernel (csrc/small.cu); --multi-launch keeps the graph of
        # per-step launches for comparison
        roll = ab.Rollout(model, Nn, T, edges=None if plan else (send, recv), knn_plan=plan, cuda_graph=True, sizes=sizes,
                          persistent=False if args.multi_launch else None)

    from aether_b200 import dist as adist

    def allreduce_grads():
        adist.allreduce_grads(params, world)     # ONE in-place all-reduce of the flat gradient buffer

    # small training shapes are launch-latency bound: replay the whole step as one CUDA graph
    graphed = None
    # (up to ~1.5 M edges at H = 256 -- config 4's strong-scaling share at 8 GPUs -- replay still beats eager launches:
    # 10.4 vs 10.8 ms per step at 128 graphs)
    if train and (E * H <= 400e6 or args.graph) and not args.no_graph:
        opt = ab.FusedAdamW(model.parameters(), lr=5e-4, weight_decay=1e-12)       # capturable: the step count is on the device
        graphed = ab.GraphedTrainStep(model, opt, (send, recv), Nn, allreduce=(lambda: adist.allreduce_grads(params, world))
                                      if world > 1 else None)

    def step_resident():
        if graphed is not None:
            return graphed.step(xd, vd, qd, td)
        if not train:
            roll.run(xd, vd, qd1)
            return None
        opt.zero_grad(set_to_none=True)
        h, ea = ab.caller_prep(xd, vd, qd1, send, recv)
        out = model(h, xd, graph, vd, ea, qd)
        loss = torch.nn.functional.mse_loss(out, td)
        loss.backward()
        allreduce_grads()
        opt.step()
        return loss

    def step_e2e():
        xs, vs, qs, ts = [t.to(dev, non_blocking=True) for t in host]
        if graphed is not None:
            return float(graphed.step(xs, vs, qs, ts).item())
        if not train:
            traj = roll.run(xs, vs, qs.reshape(-1))
            return float(traj[-1].abs().mean().item())
        s, r = ab.fc_edges(N, B, dev)                       # the reference rebuilds the edge list per batch
        opt.zero_grad(set_to_none=True)
        h, ea = ab.caller_prep(xs, vs, qs.reshape(-1), s, r)
        out = model(h, xs, [s, r], vs, ea, qs)
        loss = torch.nn.functional.mse_loss(out, ts)
        loss.backward()
        allreduce_grads()
        opt.step()
        return float(loss.item())

    small = E * H * 4 * 2 < 256e6            # working set may sit in the 126 MB L2: flush between steps
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev) if small else None

    def timed(fn, steps, warmup, profile=False):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        if profile:
            lib.aether_profile_enable(1)
        n0 = lib.aether_launch_count()
        total = 0.0
        evs = []
        for _ in range(steps):
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        launches = lib.aether_launch_count() - n0
        # kernels launched by graph replays are not seen by the host-side counter: add what each replay launches
        replayed = graphed if graphed is not None else (roll if not train else None)
        if replayed is not None and fn is not eager_fn:
            launches += steps * replayed.launches_per_replay
        total = sum(a.elapsed_time(b) for a, b in evs) * 1e-3
        prof = None
        if profile:
            lib.aether_profile_enable(0)
            import ctypes
            ms = (ctypes.c_double * 7)()
            cnt = (ctypes.c_int64 * 7)()
            lib.aether_profile_collect(ms, cnt, 7)
            prof = {"ms": list(ms), "spans": list(cnt)}
        if dist is not None:
            t = torch.tensor([total], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total = float(t.item())
        return total, launches, prof

    eager_fn = None
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    total, launches, prof = timed(step_resident, steps, warmup, profile=True)
    clocks = sampler.stop() if rank == 0 else None
    if graphed is not None:
        # graph replays carry no library-side event pairs: one eager pass of the same step for the kernel timings
        g_keep, graphed = graphed, None
        eager_fn = step_resident
        _, _, prof = timed(step_resident, max(1, min(steps, 3)), 1, profile=True)
        graphed, eager_fn = g_keep, None
    if not train:
        # graph replays carry no library-side event pairs: time the same kernels once more, launched eagerly
        eager = ab.Rollout(model, Nn, T, edges=None if plan else (send, recv), knn_plan=plan, cuda_graph=False, persistent=False)
        eager_fn = lambda: eager.run(xd, vd, qd1)
        _, _, prof = timed(eager_fn, max(1, min(steps, 3)), 1, profile=True)
        multi_ms = None
        if roll.persistent:
            # the path the single launch replaces, measured in the same run: the T steps' launches replayed as one graph
            alt = ab.Rollout(model, Nn, T, edges=None if plan else (send, recv), knn_plan=plan, cuda_graph=True, persistent=False)
            eager_fn = lambda: alt.run(xd, vd, qd1)
            t_alt, _, _ = timed(eager_fn, steps, warmup)
            multi_ms = t_alt / steps * 1e3
        eager_fn = None
    e2e = None
    if do_e2e:
        e2e_total, _, _ = timed(step_e2e, steps, max(1, warmup - 1))
        e2e = 4.0 * E * T * world * steps / e2e_total

    if rank != 0:
        return None
    pk, pk_src = peaks()
    value = 4.0 * E * T * world * steps / total
    # roofline of the dominant kernels: the edge-MLP launches of the forward pass (category 0)
    spans = max(prof["spans"][0], 1)
    ms_per_layer = prof["ms"][0] / spans
    alg = algorithmic_flops_edge_mlp(D, H, E) / 4.0          # per span = one layer, averaged over the four
    achieved = alg / (ms_per_layer * 1e-3) / 1e12 if ms_per_layer > 0 else 0.0
    peak = pk.get("bf16_tflops_sustained", pk.get("bf16_tflops"))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("%s_%s_h%d" % (args.workload, args.precision, H))
    except Exception:
        pass
    esz = 2 if args.precision == "bf16" and H in (64, 128, 256) else 4
    roofline = {"bound": "tensor", "kernel": "edge-MLP forward (message_fn, per layer: ONE chained launch)", "achieved": achieved,
                "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                "peak_source": pk_src + " bf16 sustained", "algorithmic_flops_per_launch_group": alg,
                "executed_flops_per_launch_group": executed_flops_edge_mlp(D, H, E) / 4.0,
                "avg_ms_per_launch_group": ms_per_layer,
                # the same launch against the HBM roofline: the chained kernel reads the previous layer's [E,H] edge state
                # and writes this layer's (bf16), plus the first Linear's output when the backward needs it -- 3 passes
                # in training, 2 in inference (layer 1 reads the [E,64] feature operand instead of an [E,H] tensor)
                "hbm_view": (lambda nb: {"algorithmic_bytes_per_launch_group": nb,
                                         "achieved_gbs": nb / (ms_per_layer * 1e-3) / 1e9 if ms_per_layer > 0 else 0.0,
                                         "peak_gbs": pk.get("hbm_gbs"),
                                         "frac": (nb / (ms_per_layer * 1e-3) / 1e9) / pk["hbm_gbs"]
                                         if ms_per_layer > 0 and pk.get("hbm_gbs") else None})(
                    ((3.0 if train else 2.0) * E * H - (E * (H - 64)) / 4.0) * esz)}
    # backward of the edge MLPs: data gradients = 3 chained layer pairs (5 [E,H] passes each: g1, pre2, pre1 in; g2, g1 out),
    # the top layer (4 passes) and the layer-1 feature gradient (reads g1, writes [E,32] fp32); weight gradients 2 passes each
    extra = {}
    if train and prof["ms"][1] > 0:
        nb = E * H * esz * (3 * 5.0 + 4.0) + E * (H * esz + 32 * 4)
        ms = prof["ms"][1] / steps
        extra["roofline_bwd"] = {"bound": "hbm", "kernel": "edge-MLP backward, data gradients (per step)", "unit": "GB/s",
                                 "achieved": nb / (ms * 1e-3) / 1e9, "peak": pk.get("hbm_gbs"),
                                 "frac": nb / (ms * 1e-3) / 1e9 / pk["hbm_gbs"] if pk.get("hbm_gbs") else None,
                                 "algorithmic_bytes_per_step": nb, "ms_per_step": ms, "traffic": None}
    if train and prof["ms"][2] > 0:
        nb = E * H * esz * 2.0 * 7 + E * (H + 64) * esz
        ms = prof["ms"][2] / steps
        extra["roofline_wgrad"] = {"bound": "hbm", "kernel": "edge-MLP backward, weight gradients (per step)", "unit": "GB/s",
                                   "achieved": nb / (ms * 1e-3) / 1e9, "peak": pk.get("hbm_gbs"),
                                   "frac": nb / (ms * 1e-3) / 1e9 / pk["hbm_gbs"] if pk.get("hbm_gbs") else None,
                                   "algorithmic_bytes_per_step": nb, "ms_per_step": ms, "traffic": None}
    if not train and roll.persistent:
        # single-launch rollout: nothing but inputs, weights and the trajectory crosses HBM, so the roofline SURVEY 8(d)
        # asks for is the compulsory traffic; what bounds the kernel is fp32 FMA issue and per-scene latency (fp32_view)
        n_par = sum(p.numel() for p in params)
        nb = 4.0 * (Nn * (2 * D + 1) + T * Nn * D + n_par)
        ms = total / steps * 1e3
        fl = 2.0 * T * (E * ((7 * D + D * (D - 1) // 2 + 2) * H + 7 * H * H) + Nn * (4 * 4 * H * H + 6 * H * H + 2 * H * H))
        cc_peak = 148 * 128 * 2 * (clocks["sm_mhz"] if clocks and clocks.get("sm_mhz") else 1800.0) * 1e6 / 1e12
        roofline = {"bound": "hbm", "kernel": "small_rollout_kernel (all %d steps of every scene in one launch)" % T,
                    "achieved": nb / (ms * 1e-3) / 1e9, "peak": pk.get("hbm_gbs"), "unit": "GB/s",
                    "frac": nb / (ms * 1e-3) / 1e9 / pk["hbm_gbs"] if pk.get("hbm_gbs") else None, "traffic": None,
                    "algorithmic_bytes_per_launch": nb, "peak_source": pk_src,
                    "note": "compulsory traffic only (inputs + weights + trajectory); the kernel is bound by per-scene latency "
                            "(barriers between its phases) and, in fp32 mode, by CUDA-core issue",
                    "flops_view": {"executed_tflops": fl / (ms * 1e-3) / 1e12, "ctas": len(sizes), "sms": 148,
                                   "mode": "CUDA cores, fp32" if args.precision == "fp32" else "mma.sync m16n8k16 bf16, fp32 accumulate",
                                   "cuda_core_peak_tflops": cc_peak if args.precision == "fp32" else None,
                                   "frac_of_cuda_core_peak": fl / (ms * 1e-3) / 1e12 / cc_peak if args.precision == "fp32" else None},
                    "multi_launch_graph_ms_per_step": multi_ms}
    step_ms = total / steps * 1e3
    share = {k: (prof["ms"][i] / steps) / step_ms for i, k in
             enumerate(["edge_mlp_fwd", "edge_bwd_data", "edge_wgrad", "segment_bwd", "edge_feat", "node", "segment_fwd"])} if train else None
    line = {"metric": "edge_messages_per_sec", "value": value, "unit": "edge-messages/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None,
            "dtype": "f32" if args.precision == "fp32" else "bf16",
            "data": "synthetic",
            "config": {"workload": cfg["name"], "num_dims": D, "particles_per_graph": N, "graphs_per_gpu": B,
                       "hidden": H, "mode": cfg["mode"] + (" (whole step replayed as one CUDA graph, static edge list)" if graphed is not None else "")
                       + (" (single launch: one CTA per scene for all steps)" if not train and roll.persistent else ""),
                       "rollout_steps": T, "edges_per_gpu": E, "nodes_per_gpu": Nn,
                       "precision": args.precision, "parallelism": "dp%d" % world,
                       "l2": "flushed between steps" if small else "working set exceeds L2"},
            "particle_steps_per_sec": Nn * T * world * steps / total,
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "roofline_hbm": hbm_rooflines(prof, D, H, E, Nn, args.precision == "bf16" and H in (64, 128, 256), pk),
            "time_share": share}
    line.update(extra)
    if share is not None:
        share["other (loss, optimizer, caller prep, graph CSR, launch gaps)"] = max(0.0, 1.0 - sum(share.values()))
    if e2e is not None:
        h2d = sum(t.numel() * t.element_size() for t in host)
        line["e2e"] = {"value": e2e, "unit": "edge-messages/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4}
    if do_cpu:
        # all host threads, 3 warm-up + >= 10 timed iterations (bounded by --cpu-seconds), median and
        # min; then the same sample on ONE thread (2 iterations)
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        graphs = cpu_sample_graphs(cfg, H)
        cstep, e_step = cpu_step_sample(cfg, H, graphs)
        for _ in range(3):
            cstep()
        ts, t_begin = [], time.perf_counter()
        while len(ts) < 3 or (len(ts) < 10 and time.perf_counter() - t_begin < 2.5 * args.cpu_seconds) or \
                (time.perf_counter() - t_begin < args.cpu_seconds and len(ts) < 50):
            t0 = time.perf_counter()
            cstep()
            ts.append(time.perf_counter() - t0)
        med = statistics.median(ts)
        torch.set_num_threads(1)
        t1 = []
        for _ in range(2):
            t0 = time.perf_counter()
            cstep()
            t1.append(time.perf_counter() - t0)
        torch.set_num_threads(cores)
        line["cpu_baseline"] = {"value": 4.0 * e_step / med, "unit": "edge-messages/s", "cores": cores, "kind": "port",
                                "value_best": 4.0 * e_step / min(ts), "value_1_thread": 4.0 * e_step / min(t1),
                                "sample": "%d of %d graphs per step (%d edges), median of %d steps after 3 warm-up, %d host "
                                          "threads; 1-thread figure: best of 2 steps" % (graphs, B, e_step, len(ts), cores)}
    return line

def brief(line):
    """The fields of a secondary measurement that go into the primary JSON line."""
    if line is None:
        return None
    keep = ("value", "unit", "ms_per_step", "n_gpus", "steps", "scaling", "gpu_launches", "collectives_per_step",
            "particle_steps_per_sec", "exchange")
    out = {k: line[k] for k in keep if k in line}
    c = line["config"]
    out["config"] = {k: c[k] for k in ("workload", "hidden", "graphs_per_gpu", "edges_per_gpu", "particles", "edges",
                                       "edges_rank0", "parallelism", "mode", "rollout_steps") if k in c}
    out["roofline_frac"] = line["roofline"]["frac"]
    return out


def gpu_eager_baseline(cfg, H, dev, graphs=32, iters=10):
    """BASELINE.md 4, second baseline: the oracle port (plain eager PyTorch, index_add_ scatter, fp32) with its
    parameters and inputs on the same GPU, same step definition (forward + MSE + backward + AdamW), on a sub-batch.
    This is bench.py's baseline leg: the only place besides cpu_baseline that executes oracle/."""
    from oracle import aether_oracle as O
    D, N = cfg["D"], cfg["N"]
    p = {k: v.to(dev).requires_grad_(True) for k, v in O.init_params(D, H, 1).items()}
    x, v, q = [t.to(dev) for t in synth_inputs(cfg, [N] * graphs, 1)]
    send, recv = [t.to(dev) for t in O.fc_edges(N, graphs)]
    E = int(send.numel())
    opt = torch.optim.AdamW(list(p.values()), lr=5e-4, weight_decay=1e-12)
    target = x + 0.05

    def step():
        opt.zero_grad()
        _, ea = O.caller_prep(x, v, q, send, recv)
        out = O.aether_step(p, x, (send, recv), v, ea, q)
        loss = torch.nn.functional.mse_loss(out, target)
        loss.backward()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    med = statistics.median(ms)
    return {"value": 4.0 * E / (med * 1e-3), "unit": "edge-messages/s", "ms_per_step_median": med, "ms_per_step_min": min(ms),
            "kind": "port, eager PyTorch fp32 on the same GPU (device-timed, inputs resident)", "iterations": iters,
            "sample": "%d of %d graphs per step (%d edges); the eager step keeps [E,3H] concatenations and every "
                      "activation for autograd, which is what bounds the sub-batch" % (graphs, cfg["B"], E)}


def parity_report(dev, precision):
    """Errors against the oracle, reported next to the throughput (north_star: 'with rollout MSE reported'):
    one train step at config 4's per-graph shape (2 graphs) and a 20-step rollout at config 1's shape."""
    import aether_b200 as ab
    from oracle import aether_oracle as O
    out = {}
    # --- config-4 shape, forward + all gradients
    D, N, B, H = 2, 100, 2, 256
    torch.manual_seed(1)
    model = ab.Aether(2 * D, H, 0.0, D, device=dev, precision=precision)
    p = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    x, v, q = O.synth_state(B, N, D, seed=1)
    send, recv = ab.fc_edges(N, B, dev)
    xc = x.to(dev).requires_grad_(True)
    h, ea = ab.caller_prep(xc.detach(), v.to(dev), q.to(dev).reshape(-1), send, recv)
    got = model(h, xc, [send, recv], v.to(dev), ea, q.to(dev))
    torch.nn.functional.mse_loss(got, (x + 0.1).to(dev)).backward()
    so, ro = O.fc_edges(N, B)
    po = {k: t.clone().requires_grad_(True) for k, t in p.items()}
    xo = x.clone().requires_grad_(True)
    _, eo = O.caller_prep(x, v, q, so, ro)
    want = O.aether_step(po, xo, (so, ro), v, eo, q)
    torch.nn.functional.mse_loss(want, x + 0.1).backward()

    def rl2(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp(min=1e-30))

    def rmax(a, b):
        return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-30))

    grads = [(k, prm.grad.cpu(), po[k].grad) for k, prm in model.named_parameters()
             if po[k].grad is not None and float(po[k].grad.abs().max()) > 0]
    out["config4_shape_step"] = {
        "shape": "D=2 N=100 B=2 H=256 %s, train step vs fp32 CPU oracle" % precision,
        "out_rel_max": rmax(got.detach().cpu(), want.detach()), "displacement_rel_l2": rl2((got.detach().cpu() - x), (want.detach() - x)),
        "grad_x_rel_l2": rl2(xc.grad.cpu(), xo.grad), "grad_w_rel_max_worst": max(rmax(a, b) for _, a, b in grads),
        "grad_w_rel_l2_worst": max(rl2(a, b) for _, a, b in grads), "tolerance": 1e-5 if precision == "fp32" else 2e-2}
    # --- config-1 shape, 20-step rollout
    D, N, B, H, T = 2, 5, 16, 64, 20
    torch.manual_seed(1)
    model = ab.Aether(2 * D, H, 0.0, D, device=dev, precision=precision)
    model.eval()
    p = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    x, v, q = O.synth_state(B, N, D, seed=1)
    send, recv = ab.fc_edges(N, B, dev)
    # both rollout paths: the single launch (csrc/small.cu, what `--workload c1` times) and the per-step launches
    traj = ab.rollout(model, x.to(dev), v.to(dev), q.to(dev).reshape(-1), T, edges=(send, recv), sizes=[N] * B, persistent=True).cpu()
    traj_ml = ab.rollout(model, x.to(dev), v.to(dev), q.to(dev).reshape(-1), T, edges=(send, recv), persistent=False).cpu()
    so, ro = O.fc_edges(N, B)
    with torch.no_grad():
        ref = O.rollout(p, x, v, q, lambda xx: (so, ro), T)
    ref = torch.stack(list(ref)) if not torch.is_tensor(ref) else ref
    n = min(traj.shape[0], ref.shape[0])
    mse = [float(((traj[-n:][t] - ref[-n:][t]) ** 2).mean()) for t in range(n)]
    out["config1_shape_rollout"] = {"shape": "D=2 N=5 B=16 H=64 %s, %d-step rollout vs fp32 CPU oracle" % (precision, T),
                                    "path": "single launch (small_rollout_kernel)",
                                    "mse_step_1": mse[0], "mse_step_last": mse[-1], "mse_mean": sum(mse) / len(mse),
                                    "rel_max": float((traj - ref).abs().max() / ref.abs().max()),
                                    "multi_launch_mse_mean": float(((traj_ml - ref) ** 2).mean()),
                                    "multi_launch_rel_max": float((traj_ml - ref).abs().max() / ref.abs().max()),
                                    "position_scale_rms": float((ref ** 2).mean().sqrt())}
    return out


def seq2seq_report(dev, steps=20, reps=10):
    """seq2seq decoder path at the shape of scripts/electrostatic_field_aether.sh: 5 particles, batch 64,
    decoder_hidden 512, 2 edge types, 20 prediction steps as ONE CUDA graph of field predictor + decoder steps; error of
    the first 3 steps against the CPU oracle.  A reported secondary figure, not the headline metric."""
    from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
    from oracle import seq2seq_oracle as so
    D, N, B, Hd, K = 2, 5, 64, 512, 2
    torch.manual_seed(1)
    m = Seq2SeqAether({"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": Hd, "num_edge_types": K,
                       "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": Hd, "gumbel_temp": 0.5}).to(dev).eval()
    E = N * (N - 1)
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1)
    edges = torch.nn.functional.one_hot(torch.randint(0, K, (steps, B, E)), K).float()
    r = DecoderRollout(m, B, steps, cuda_graph=True)
    xd, ed = x.to(dev), edges.to(dev)
    for _ in range(3):
        traj = r.run(xd, ed)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        r.run(xd, ed)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    ms = ts[len(ts) // 2]
    w = {k[len("decoder."):]: v.detach().cpu() for k, v in m.state_dict().items() if k.startswith("decoder.")}
    w.update({k: v.detach().cpu() for k, v in m.state_dict().items() if k.startswith("field_net.")})
    w["rff_B"] = m.coordinate_embedding.B.detach().cpu()
    t0 = time.perf_counter()
    ref = so.predict_future(w, x.unsqueeze(1), edges[:3].transpose(0, 1), 3, D, K, False, return_everything=False)
    cpu_ms_per_step = (time.perf_counter() - t0) * 1e3 / 3
    err = float((traj[:3].transpose(0, 1).cpu() - ref).abs().max() / ref.abs().max())
    # the same rollout in tensor-core mode (precision="bf16": mma.sync, 2e-2 tolerance), error against the same oracle steps
    mb = Seq2SeqAether({"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": Hd, "num_edge_types": K,
                        "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": Hd, "gumbel_temp": 0.5},
                       precision="bf16").to(dev).eval()
    mb.load_state_dict(m.state_dict())
    rb = DecoderRollout(mb, B, steps, cuda_graph=True)
    for _ in range(3):
        trajb = rb.run(xd, ed)
    tb = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rb.run(xd, ed)
        b.record()
        torch.cuda.synchronize()
        tb.append(a.elapsed_time(b))
    tb.sort()
    errb = float((trajb[:3].transpose(0, 1).cpu() - ref).abs().max() / ref.abs().max())
    bf16 = {"us_per_step": 1e3 * tb[len(tb) // 2] / steps, "rel_err_vs_oracle_3_steps": errb, "tolerance": 2e-2}
    return {"tensor_core_mode": bf16,
            "workload": "seq2seq RecurrentDecoder rollout: 2D, 5 particles, batch 64, decoder_hidden 512, 2 edge types, "
                        "%d prediction steps, one CUDA graph" % steps, "dtype": "f32", "ms_per_rollout": ms,
            "us_per_step": 1e3 * ms / steps, "particle_steps_per_s": B * N * steps / (ms * 1e-3),
            "launches_per_rollout": r.launches_per_replay, "rel_err_vs_oracle_3_steps": err,
            "cpu_oracle_ms_per_step": cpu_ms_per_step, "cpu_threads": torch.get_num_threads()}


_REAL_STDOUT = sys.stdout


def emit(line):
    """The ONE JSON line of the contract goes to the real stdout; everything else a run prints (the modules mirror the
    reference's constructor, which prints "Network Size N", nn/state2state/aether.py:163-167) goes to stderr."""
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def main():
    sys.stdout = sys.stderr
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))   # c5: see --workload c5 below
    ap.add_argument("--hidden", type=int, default=0)
    ap.add_argument("--batch", type=int, default=0, help="graphs per GPU (default: the workload's)")
    ap.add_argument("--precision", default=os.environ.get("AETHER_B200_PRECISION", "bf16"), choices=["fp32", "bf16"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="default c4 run: skip the strong-scaling / config-5 / baseline legs")
    ap.add_argument("--multi-launch", action="store_true", help="rollout workloads: per-step launches instead of the single-launch kernel")
    ap.add_argument("--strong", action="store_true", help="c4: split the workload's batch over the ranks (strong scaling)")
    ap.add_argument("--graph", action="store_true", help="capture the train step in a CUDA graph at any size")
    ap.add_argument("--no-graph", action="store_true", help="small training shapes: do not capture the step in a CUDA graph")
    args = ap.parse_args()
    cfg = dict(WORKLOADS[args.workload])
    H = args.hidden or cfg["H"]
    if args.batch:
        cfg["B"] = args.batch
    if args.impl == "reference":
        return run_reference(args, cfg, H)
    ctx = Ctx()
    default_run = args.workload == "c4" and not args.hidden and not args.batch and not args.strong
    try:
        if args.workload == "c5":
            line = run_c5(args, cfg, H, ctx, args.steps, args.warmup, not args.no_e2e, not args.no_cpu_baseline)
        elif args.strong:
            cfg["B"] = max(1, cfg["B"] // ctx.world)
            line = run_dp(args, cfg, H, ctx, args.steps, args.warmup, not args.no_e2e, not args.no_cpu_baseline, "strong")
        else:
            line = run_dp(args, cfg, H, ctx, args.steps, args.warmup, not args.no_e2e, not args.no_cpu_baseline, "weak")
        if default_run and not args.no_secondary:
            release_memory()
            sec = {}
            ssteps, swarm = max(3, min(args.steps, 10)), 3
            if ctx.world > 1:
                c4s = dict(WORKLOADS["c4"])
                c4s["B"] = c4s["B"] // ctx.world
                c4s["name"] = "config4 STRONG scaling: 1024 graphs in total, %d per GPU, train step" % c4s["B"]
                sec["c4_strong"] = brief(run_dp(args, c4s, H, ctx, ssteps, swarm, False, False, "strong"))
                release_memory()
            for hh in (256, 64):
                sec["c5_h%d" % hh] = brief(run_c5(args, dict(WORKLOADS["c5"]), hh, ctx, ssteps, swarm, False, False))
                release_memory()
            if ctx.rank == 0:
                line["secondary"] = sec
                if ctx.world == 1:
                    try:
                        line["gpu_eager_baseline"] = gpu_eager_baseline(WORKLOADS["c4"], H, ctx.dev)
                    except Exception as e:          # an out-of-memory eager baseline must not lose the bench line
                        line["gpu_eager_baseline"] = {"unavailable": repr(e)[:200]}
                    release_memory()
                    line["parity"] = parity_report(ctx.dev, args.precision)
                    try:
                        sec["seq2seq_rollout"] = seq2seq_report(ctx.dev)
                    except Exception as e:
                        sec["seq2seq_rollout"] = {"unavailable": repr(e)[:200]}
        if ctx.rank == 0:
            emit(line)
    finally:
        ctx.close()


if __name__ == "__main__":
    main()
