"""bench.py --workload c5: data-parallel training on synthetic trajectories (BASELINE.json configs[4], SURVEY.md row D
"C5"): 8 graphs per GPU x 2 000 particles (16 000 nodes / ~320 k edges), random-walk-noised one-step targets
(noise_std 3e-4 -- the reference's default is None, examples/train_dyn.py:235, so the value is stated here),
forward + L1Loss(sum)/N + backward + NCCL gradient all-reduce + Adam(1e-4) per step.

One bench "step" = one training step of every rank on its own batch (weak scaling).  value = graphs/s over all ranks.
The e2e figure rebuilds the batch from pinned HOST windows every step (H2D of the observation windows and next
positions, device-side noise / features / radius graph) and reads the loss back.
"""
import json
import os
import time

import numpy as np
import torch

import scenes

GRAPHS_PER_GPU, PARTICLES, NOISE_STD = 8, 2000, 3e-4


def make_samples(rank, count=GRAPHS_PER_GPU, n=PARTICLES):
    out = []
    for i in range(count):
        obs, nxt = scenes.make_state(n, seed=1000 * rank + i, n_rigid=scenes.N_RIGID)
        g = torch.Generator().manual_seed(7 + 1000 * rank + i)
        nxt = nxt + torch.randn((n, 3), generator=g) * 5e-5
        out.append((obs, nxt))
    return out


def cpu_reference_step(samples, repeats=1):
    """The reference's training step on the host cores (oracle port: torch autograd over the same modules,
    sklearn KDTree graphs per sample, torch.optim.Adam): seconds per step."""
    import subprocess
    root = os.path.dirname(os.path.abspath(__file__))
    subprocess.check_call(["make", "-s", "-C", os.path.join(root, "oracle")])
    from oracle import restate
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    model = restate.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    feats = restate.GraphFeatures(scenes.CONN_R, scenes.make_stats(), scenes.make_bounds(), scenes.CART_IDX,
                                  scenes.MAT_IDX, scenes.CTRL_IDX, scenes.MAX_NBRS,
                                  connectivity=restate.get_connectivity_sklearn)
    opt = torch.optim.Adam(model.parameters(), lr=1e-4)
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        nodes, edge_attr, ei, tgt = feats.process_collate(samples)
        pred = model(nodes, edge_attr, ei)
        loss = torch.nn.L1Loss(reduction="sum")(pred, tgt) / pred.shape[0]
        opt.zero_grad()
        loss.backward()
        opt.step()
        times.append(time.perf_counter() - t0)
    return times, torch.get_num_threads()


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sample = 2                                        # graphs per timed CPU step (bounded: ~10 s each)
    times, cores = cpu_reference_step(make_samples(0, sample), repeats=max(1, args.steps))
    value = sample * len(times) / sum(times)
    line = {"impl": "reference", "metric": "training graphs/s", "value": value, "unit": "graphs/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C5 training step: {sample} graphs x {PARTICLES} particles per CPU step "
                                   f"(sample of the {GRAPHS_PER_GPU}-graph batch)", "mp_steps": 10, "hidden": 128},
            "cpu_baseline": {"value": value, "unit": "graphs/s", "cores": cores, "kind": "port",
                             "sample": f"{sample} graphs per step, {len(times)} steps (oracle/restate.py + torch autograd + Adam)"},
            "e2e": {"value": value, "unit": "graphs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run(args):
    import bench
    if args.impl == "reference":
        return run_reference(args)
    import gnn_manip_b200 as gm
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- gnn_manip_b200 has no CPU path (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps if args.steps != 3 else 20, max(args.warmup, 3)
    torch.manual_seed(0)
    model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    ga = gm.GraphBoundedMultimaterialControl(scenes.CONN_R, scenes.make_stats(), scenes.CART_IDX, [scenes.MAT_IDX],
                                             scenes.CTRL_IDX, scenes.make_bounds(), noise=NOISE_STD,
                                             max_neighbours=scenes.MAX_NBRS)
    samples = make_samples(rank)
    n_nodes = GRAPHS_PER_GPU * PARTICLES
    results = {}
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for mode in ("bucketed", "single"):
        tr = gm.Trainer(model, n_nodes, n_nodes * scenes.MAX_NBRS, lr=1e-4, device=dev, bucketed_allreduce=(mode == "bucketed"))
        batch = gm.make_batch(ga, samples, device=dev, seed=1)
        for _ in range(W):
            tr.loss_batch(batch)
        barrier()
        l0 = gm._lib.launch_count()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with bench.ClockSampler(local) as clocks:
            a.record(stream)
            for _ in range(K):
                tr.loss_batch(batch)
            b.record(stream)
            barrier()
        results[mode] = {"ms": a.elapsed_time(b), "launches": gm._lib.launch_count() - l0, "loss": float(tr.loss),
                         "clocks": clocks.summary(), "edges": int(batch.graph.n_edges)}
        # forward / backward / optimiser split (single process timing, rank-local)
        if mode == "single":
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record(stream)
            tr.forward_backward(batch, backward=False)
            ev[1].record(stream)
            tr.forward_backward(batch, backward=True)
            ev[2].record(stream)
            tr.optimizer_step()
            ev[3].record(stream)
            torch.cuda.synchronize(dev)
            results["split_ms"] = {"forward": ev[0].elapsed_time(ev[1]),
                                   "forward+backward+allreduce": ev[1].elapsed_time(ev[2]),
                                   "adam+repack": ev[2].elapsed_time(ev[3])}
            if dist is not None:
                g = tr.grads.clone()
                ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                torch.cuda.synchronize(dev)
                ev2[0].record(stream)
                for _ in range(10):
                    dist.all_reduce(g)
                ev2[1].record(stream)
                torch.cuda.synchronize(dev)
                results["allreduce_ms"] = ev2[0].elapsed_time(ev2[1]) / 10
        # end to end: batch rebuilt from pinned host windows every step, loss read back
        if mode == "bucketed":
            host = [(o.pin_memory(), t.pin_memory()) for o, t in samples]
            h2d = sum(o.numel() * 4 + t.numel() * 4 for o, t in host)
            barrier()
            a2, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a2.record(stream)
            losses = []
            for i in range(K):
                bt = gm.make_batch(ga, [(o.to(dev, non_blocking=True), t.to(dev, non_blocking=True)) for o, t in host],
                                   device=dev, seed=100 + i)
                losses.append(float(tr.loss_batch(bt)))           # D2H of the loss: 4 bytes
            b2.record(stream)
            barrier()
            results["e2e_ms"], results["h2d"], results["loss_e2e"] = a2.elapsed_time(b2), h2d, losses[-1]
        tr.close()
        del tr
    best = min(("bucketed", "single"), key=lambda m: results[m]["ms"])
    t = torch.tensor([results["bucketed"]["ms"], results["single"]["ms"], results["e2e_ms"]], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_b, ms_s, ms_e2e = (float(v) for v in t)
    ms = min(ms_b, ms_s)
    if rank == 0:
        r = results[best]
        line = {"metric": "training graphs/s", "value": world * GRAPHS_PER_GPU * K / (ms * 1e-3), "unit": "graphs/s",
                "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16x3 (fp32-equivalent operands, fp32 accumulate)",
                "data": "synthetic",
                "config": {"workload": f"C5 data-parallel training: {GRAPHS_PER_GPU} graphs x {PARTICLES} particles per GPU, "
                                       f"random-walk noise {NOISE_STD}, L1Loss(sum)/N, Adam(1e-4)",
                           "edges_per_gpu_batch": r["edges"], "mp_steps": 10, "hidden": 128,
                           "l2": "activations of one step (about 8 GB) exceed L2", "all_reduce": best},
                "ms_per_step_bucketed_allreduce": ms_b / K, "ms_per_step_single_allreduce": ms_s / K,
                "split_ms": results.get("split_ms"), "allreduce_ms_6p4MB": results.get("allreduce_ms"),
                "loss": r["loss"],
                "e2e": {"value": world * GRAPHS_PER_GPU * K / (ms_e2e * 1e-3), "unit": "graphs/s",
                        "h2d_bytes_per_step": results["h2d"], "d2h_bytes_per_step": 4},
                "gpu_launches": int(r["launches"]), "clocks": r["clocks"]}
        if world == 1 and not args.no_cpu_baseline:
            times, cores = cpu_reference_step(make_samples(0, 2))
            line["cpu_baseline"] = {"value": 2 / times[0], "unit": "graphs/s", "cores": cores, "kind": "port",
                                    "sample": "one training step on 2 graphs x 2 000 particles (oracle/restate.py, torch "
                                              "autograd, sklearn KDTree, torch.optim.Adam)"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
