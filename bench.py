#!/usr/bin/env python
"""bench.py -- rollout particle-steps/s of the gnn-manip GraphNet hot path on B200 (contract: see DESIGN.md section 6).

  python bench.py [--gpus N --steps K --warmup W]          this repo's CUDA path (C ABI, CUDA graph per rollout step)
  python bench.py --impl reference [...]                   the CPU restatement of the reference path on the host cores

Workload (BASELINE.json configs[1], SURVEY.md row D "C2"): synthetic granular pour, 2 000 particles (828 rigid),
300 rollout steps, k=6, conn_r=0.015, max 20 neighbours, 10 MP steps, hidden 128, seed-0 random-init weights
(dataset/trained_model.pth is absent from the reference checkout).  One bench "step" = one 300-step rollout.
With N GPUs every rank rolls out its own CMA-ES candidate (perturbed control curve): weak scaling, no collective
on the data path.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import scenes  # noqa: E402

METRIC = "rollout particle-steps/s"
UNIT = "particle-steps/s"
FLOP_PER_EDGE_MP = 163_840            # reference formulation, SURVEY.md section 8 row D
BYTES_PER_EDGE_MP = 1_032             # read e 512 + write e' 512 + two int32 indices


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=2000)
    ap.add_argument("--rollout-steps", type=int, default=300)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c3", action="store_true", help="skip the 100k-particle roofline leg")
    ap.add_argument("--no-c4", action="store_true", help="skip the CMA-ES population line (64 candidates x 150 steps)")
    ap.add_argument("--only-c3", action="store_true", help="profiling aid: only the 100k-particle forward profile")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sust=float(d["bf16_tflops_sustained"]), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback")


def control_curve(steps, rank=0):
    """(rot [rad], ty [m]) per step: dataset/sample_traj.npy when the fixture copy exists, else the stand-in curve;
    ranks > 0 get a CMA-ES-style perturbation (numpy seed 1234 + rank)."""
    p = os.path.join(ROOT, "tests", "golden", "sample_traj.npy")
    st = np.load(p) if os.path.exists(p) else scenes.sample_traj()
    reps = -(-(steps + 1) // st.shape[0])
    st = np.concatenate([st] * reps)[: steps + 1]
    rot, ty = np.deg2rad(st[:steps, 0]), st[:steps, 1] - st[0, 1]
    if rank > 0:
        rng = np.random.RandomState(1234 + rank)
        rot = rot + np.cumsum(rng.randn(steps) * 1e-3)
        ty = ty + np.cumsum(rng.randn(steps) * 1e-5)
    return rot, ty


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# CPU legs (the only place bench.py touches oracle/): the restated reference path on the host cores
# ---------------------------------------------------------------------------------------------------------
def cpu_rollout_sample(n, sample_steps, repeats=1, warm=1):
    """Run the oracle's rollout loop (graph = the reference's own sklearn KDTree call sequence, features, fp32 torch
    model, Euler) for `sample_steps` steps of the C2 workload on all host threads.  Returns seconds per repeat."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    from oracle import restate
    torch.set_num_threads(os.cpu_count() or 1)
    obs, _ = scenes.make_state(n, seed=0, n_rigid=min(scenes.N_RIGID, n // 2))
    torch.manual_seed(0)
    model = restate.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    feats = restate.GraphFeatures(scenes.CONN_R, scenes.make_stats(), scenes.make_bounds(), scenes.CART_IDX,
                                  scenes.MAT_IDX, scenes.CTRL_IDX, scenes.MAX_NBRS,
                                  connectivity=restate.get_connectivity_sklearn)
    n_rigid = int((obs[-1, :, scenes.MAT_IDX] == 1).sum())
    rot, ty = control_curve(sample_steps)
    traj = restate.rigid_body_trajectory(rot, ty, sample_steps, scenes.TY_INIT, obs[-1, :n_rigid][:, scenes.CART_IDX])
    if warm:
        restate.rollout(model, feats, (obs, None), traj[:1], feats.stats, 1)
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        restate.rollout(model, feats, (obs, None), traj, feats.stats, sample_steps)
        times.append(time.perf_counter() - t0)
    return times, torch.get_num_threads()


def run_reference(args):
    """--impl reference: the reference's CPU path (restated in oracle/, kind "port": the reference is pure Python
    and does not travel to the GPU box) on the host cores; each bench step = a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, sample = args.particles, 8
    times, cores = cpu_rollout_sample(n, sample, repeats=args.steps, warm=max(1, min(args.warmup, 1)))
    total = sum(times)
    value = n * sample * len(times) / total
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2 rollout: {n} particles x {args.rollout_steps} steps "
                                   f"(each bench step = a {sample}-rollout-step sample)", "k_steps": 6,
                       "conn_r": scenes.CONN_R, "max_neighbours": scenes.MAX_NBRS, "mp_steps": 10, "hidden": 128},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{sample} rollout steps of the C2 scene per bench step, {len(times)} steps"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# GPU leg
# ---------------------------------------------------------------------------------------------------------
def profile_forward(gm, model, obs, reps, warm, flush):
    """Per-kernel-class device time of one model forward on the graph of `obs` (CUDA events on the launch stream,
    gm_model_forward_profile).  Returns (avg ms per launch for [enc_edge, enc_node, mp_edge, mp_node], E, N)."""
    import ctypes
    L, lib = gm._lib.lib(), gm._lib
    dev = torch.device("cuda", torch.cuda.current_device())
    ga = make_graph_attr(gm)
    o = obs.to(dev)
    nodes = ga.compute_nodes(o)
    pos = o[-1][:, scenes.CART_IDX].contiguous()
    g = gm.radius_graph(pos, scenes.CONN_R, scenes.MAX_NBRS)
    E, N = g.n_edges, pos.shape[0]
    pm = model.packed(dev)
    pred = torch.empty((N, 3), device=dev)
    ws_bytes = L.gm_forward_workspace_bytes(pm.handle, N, E, lib.GM_IMPL_TC)
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=dev)
    ms, cnt = (ctypes.c_float * 4)(), (ctypes.c_int32 * 4)()
    tot, launches = np.zeros(4), np.zeros(4)
    for i in range(warm + reps):
        if flush is not None:
            flush.add_(1)
        lib.check(L.gm_model_forward_profile(pm.handle, lib.ptr(nodes), None, lib.ptr(pos), 3, scenes.CONN_R, N, E,
                                             lib.ptr(g.n_edges_dev), lib.ptr(g.recv_ptr), lib.ptr(g.esend),
                                             lib.ptr(g.erecv), lib.ptr(g.eref), lib.ptr(pred), lib.ptr(ws), ws_bytes,
                                             lib.stream_ptr(dev), ms, cnt), "gm_model_forward_profile")
        if i >= warm:
            tot += np.array(ms[:])
            launches += np.array(cnt[:])
    return tot / np.maximum(launches, 1), E, N


def make_graph_attr(gm):
    return gm.GraphBoundedMultimaterialControl(scenes.CONN_R, scenes.make_stats(), scenes.CART_IDX, [scenes.MAT_IDX],
                                               scenes.CTRL_IDX, scenes.make_bounds(), max_neighbours=scenes.MAX_NBRS)


def roofline_entry(ms_edge, E, pk, sustained):
    peak_tf = pk["tf_sust"] if sustained else pk["tf_burst"]
    tf = FLOP_PER_EDGE_MP * E / (ms_edge * 1e-3) / 1e12
    gbs = BYTES_PER_EDGE_MP * E / (ms_edge * 1e-3) / 1e9
    return {"kernel": "mp_edge_kernel", "bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s",
            "frac": tf / peak_tf, "traffic": None, "peak_source": pk["source"] + (" sustained" if sustained else " burst"),
            "executed_frac": 1.8 * tf / peak_tf,   # bf16x3: 3 x 98 304 MAC-flops executed per 163 840 algorithmic
            "hbm_achieved_gbs": gbs, "hbm_frac": gbs / pk["hbm"], "edges_per_launch": int(E),
            "us_per_launch": 1e3 * ms_edge, "edges_per_s": E / (ms_edge * 1e-3)}


def run_ours(args):
    import gnn_manip_b200 as gm
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- gnn_manip_b200 has no CPU path (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    N, T, K, W = args.particles, args.rollout_steps, args.steps, max(args.warmup, 3)
    if args.only_c3:
        torch.manual_seed(0)
        model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
        obs3, _ = scenes.make_state(100_000, seed=0, n_rigid=0)
        per3, E3, _ = profile_forward(gm, model, obs3, reps=2, warm=1, flush=None)
        print(json.dumps({"only_c3": True, "edges": E3, "kernel_ms": list(map(float, per3))}), flush=True)
        return

    obs, _ = scenes.make_state(N, seed=0, n_rigid=min(scenes.N_RIGID, N // 2))
    torch.manual_seed(0)
    model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    ga = make_graph_attr(gm)
    rot, ty = control_curve(T, rank)
    plan = gm.Rollout(model, ga, obs, rot=rot, ty=ty, ty_init=tuple(scenes.TY_INIT), use_cuda_graph=True,
                      impl=gm._lib.GM_IMPL_TC, device=dev)
    flush = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)     # 256 MB > 126 MB L2
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident rollouts: `value` ----
    def one_rollout(timed):
        flush.add_(1.0)                      # evict L2 between bench steps (untimed)
        plan.reset()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        plan.run(T)
        b.record(stream)
        return (a, b) if timed else None

    for _ in range(W):
        one_rollout(False)
    barrier()
    l0 = gm._lib.launch_count()
    with ClockSampler(local) as clocks:
        evs = [one_rollout(True) for _ in range(K)]
        barrier()
    ms = sum(a.elapsed_time(b) for a, b in evs)
    final_dev = plan.obs[-1].clone()
    launches = plan.launches_per_step * T * K + (gm._lib.launch_count() - l0)

    # ---- end to end through the public API with HOST buffers: `e2e` ----
    obs_host = obs.clone().pin_memory()
    rot_host = torch.from_numpy(np.ascontiguousarray(rot)).pin_memory()
    ty_host = torch.from_numpy(np.ascontiguousarray(ty)).pin_memory()
    out_host = torch.empty((N, scenes.DATA_DIM), dtype=torch.float32).pin_memory()
    rot_dev, ty_dev = plan._keep[0], plan._keep[1]
    h2d = obs_host.numel() * 4 + rot_host.numel() * 8 + ty_host.numel() * 8
    d2h = out_host.numel() * 4

    def one_e2e():
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        rot_dev.view(-1).copy_(rot_host, non_blocking=True)        # this candidate's control curve
        ty_dev.view(-1).copy_(ty_host, non_blocking=True)
        plan.reset(obs_host)                                       # H2D of the initial window + step counter
        plan.run(T)
        out_host.copy_(plan.obs[-1], non_blocking=True)            # D2H of the end state
        b.record(stream)
        return a, b

    one_e2e()
    barrier()
    evs2 = [one_e2e() for _ in range(K)]
    barrier()
    ms_e2e = sum(a.elapsed_time(b) for a, b in evs2)
    assert torch.equal(out_host, final_dev.cpu()), "e2e rollout must reproduce the device-resident rollout"

    # ---- C4: CMA-ES population, 64 candidates x 150-step rollouts, population sharded over the ranks:
    #      each rank batches its candidates as disjoint graphs in ONE plan; no collective on the data path ----
    ms_c4, c4 = 0.0, None
    if not args.no_c4:
        from gnn_manip_b200.sharding import partition_population
        pop, T4 = 64, 150
        lo, hi = partition_population(pop, world, rank)
        curves = [control_curve(T4, i) for i in range(lo, hi)]
        rot4 = np.stack([c[0] for c in curves]); ty4 = np.stack([c[1] for c in curves])
        del plan
        plan4 = gm.Rollout(model, ga, obs, rot=rot4, ty=ty4, ty_init=tuple(scenes.TY_INIT), n_graphs=hi - lo,
                           use_cuda_graph=True, impl=gm._lib.GM_IMPL_TC, device=dev)
        plan4.run(3)                                               # warm-up (graph capture, caches)
        barrier()
        flush.add_(1.0)
        plan4.reset()
        a4, b4 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a4.record(stream)
        plan4.run(T4)
        b4.record(stream)
        barrier()
        ms_c4 = a4.elapsed_time(b4)
        c4 = {"population": pop, "rollout_steps": T4, "candidates_per_rank": hi - lo, "particles": N}
        del plan4

    t = torch.tensor([ms, ms_e2e, ms_c4], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_c4 = float(t[0]), float(t[1]), float(t[2])
    value = world * N * T * K / (ms * 1e-3)
    e2e = world * N * T * K / (ms_e2e * 1e-3)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16x3 (fp32-equivalent operands, fp32 accumulate)", "data": "synthetic",
            "config": {"workload": f"C2 rollout: {N} particles x {T} steps per GPU (one CMA-ES candidate per rank)",
                       "k_steps": 6, "conn_r": scenes.CONN_R, "max_neighbours": scenes.MAX_NBRS, "mp_steps": 10,
                       "hidden": 128, "weights": "random-init seed 0", "l2": "flushed (256 MB write) between bench steps",
                       "cuda_graph": True},
            "us_per_rollout_step": 1e3 * ms / (K * T),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches), "clocks": clocks.summary()}

    if c4 is not None:
        c4.update({"ms": ms_c4, "cma_candidates_per_s": c4["population"] / (ms_c4 * 1e-3),
                   "particle_steps_per_s": c4["population"] * N * c4["rollout_steps"] / (ms_c4 * 1e-3),
                   "us_per_rollout_step": 1e3 * ms_c4 / c4["rollout_steps"]})
        line["c4_cma_population"] = c4
    if rank == 0:
        per, E, _ = profile_forward(gm, model, obs, reps=20, warm=3, flush=None)
        line["roofline"] = roofline_entry(per[2], E, pk, sustained=True)
        line["roofline"]["note"] = ("C2 graph (L2-resident latents, launch/latency-bound regime); "
                                    "see roofline_c3 for the 100k-particle cloud")
        try:   # DRAM bytes of the same launch from the committed ncu capture (never measured under this run)
            with open(os.path.join(ROOT, "profiles", "mp_edge_traffic.json")) as f:
                tr = json.load(f)["c2"]
            if int(tr["edges_per_launch"]) == int(E):
                line["roofline"]["traffic"] = tr["dram_bytes_per_launch"]
                line["roofline"]["traffic_source"] = tr["source"]
        except (OSError, KeyError, ValueError):
            pass
        line["kernel_us_c2"] = {"enc_edge": 1e3 * per[0], "enc_node": 1e3 * per[1], "mp_edge": 1e3 * per[2],
                                "mp_node": 1e3 * per[3]}
        if not args.no_c3:
            obs3, _ = scenes.make_state(100_000, seed=0, n_rigid=0)
            per3, E3, _ = profile_forward(gm, model, obs3, reps=5, warm=2, flush=flush)
            line["roofline_c3"] = roofline_entry(per3[2], E3, pk, sustained=True)
            try:   # DRAM bytes of the same launch from the committed ncu capture (never measured under this run)
                with open(os.path.join(ROOT, "profiles", "mp_edge_traffic.json")) as f:
                    tr = json.load(f)["c3"]
                if int(tr["edges_per_launch"]) == int(E3):
                    line["roofline_c3"]["traffic"] = tr["dram_bytes_per_launch"]
                    line["roofline_c3"]["traffic_source"] = tr["source"]
            except (OSError, KeyError, ValueError):
                pass
            line["kernel_us_c3"] = {"enc_edge": 1e3 * per3[0], "enc_node": 1e3 * per3[1], "mp_edge": 1e3 * per3[2],
                                    "mp_node": 1e3 * per3[3]}
            fwd_ms = per3[0] + per3[1] + 10 * (per3[2] + per3[3])
            line["c3_model_particle_steps_per_s"] = 100_000 / (fwd_ms * 1e-3)
        if world == 1 and not args.no_cpu_baseline:
            sample = 24
            times, cores = cpu_rollout_sample(N, sample)
            line["cpu_baseline"] = {"value": N * sample / times[0], "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{sample} rollout steps of the same C2 scene (oracle/restate.py, "
                                              "sklearn KDTree + torch fp32), 1 warm-up step"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    # stdout carries exactly ONE JSON line: anything native libraries write to fd 1 meanwhile (NCCL prints its version
    # banner there on some boxes, make prints when it rebuilds the oracle) is sent to stderr instead
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(_real_stdout, "w")
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
    sys.stdout.flush()
