#!/usr/bin/env python
"""bench.py -- rollout particle-steps/s of the gnn-manip GraphNet hot path on B200 (contract: DESIGN.md section 5).

  python bench.py [--gpus N --steps K --warmup W]          this repo's CUDA path (C ABI, CUDA graph per rollout step)
  python bench.py --impl reference [...]                   the CPU restatement of the reference path on the host cores
  python bench.py --workload c5 [...]                      data-parallel training step (BASELINE.json configs[4])

Workload of the headline line (BASELINE.json configs[1], SURVEY.md row D "C2"): synthetic granular pour, 2 000
particles (828 rigid), 300 rollout steps, k=6, conn_r=0.015, max 20 neighbours, 10 MP steps, hidden 128, seed-0
random-init weights (dataset/trained_model.pth is absent from the reference checkout).  One bench "step" = one 300-step
rollout.  With N GPUs every rank rolls out its own CMA-ES candidate (perturbed control curve): weak scaling, no
collective on the data path.  Extra keys of the same JSON line: c4_cma_population (configs[3]: 64 candidates x 150
steps, evaluated -- rollout + Sinkhorn objective -- and sharded over the ranks), roofline (the dominant kernel,
mp_edge, on the 100 000-particle cloud of configs[2]), roofline_c2, rooflines_c3 (graph build / features / integrator
as GB/s), c3_rollout, cpu_baseline.
"""
import argparse
import ctypes
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
CPU_STARTS = (0, 100, 200, 294)       # start steps of the CPU-baseline samples (states: tests/golden/rollout_c2.npz)
CPU_SEG = 6                           # CPU steps timed from each start state


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"])
    ap.add_argument("--particles", type=int, default=2000)
    ap.add_argument("--rollout-steps", type=int, default=300)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c3", action="store_true", help="skip the 100k-particle legs")
    ap.add_argument("--no-c4", action="store_true", help="skip the CMA-ES population line (64 candidates x 150 steps)")
    ap.add_argument("--only-c3", action="store_true", help="profiling aid: only the 100k-particle forward profile")
    ap.add_argument("--only-c3-step", action="store_true", help="profiling aid: only one 100k-particle rollout step")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sust=float(d["bf16_tflops_sustained"]), source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback")


def sample_traj_file():
    p = os.path.join(ROOT, "tests", "golden", "sample_traj.npy")
    return np.load(p) if os.path.exists(p) else scenes.sample_traj()


def control_curve(steps, rank=0):
    """(rot [rad], ty [m]) per step: dataset/sample_traj.npy when the fixture copy exists, else the stand-in curve;
    ranks > 0 get a CMA-ES-style perturbation (numpy seed 1234 + rank)."""
    st = sample_traj_file()
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
def cpu_rollout_sample(n, repeats=1, warm=1):
    """Time the oracle's rollout loop (graph = the reference's own sklearn KDTree call sequence, features, fp32 torch
    model, Euler) on all host threads, on a bounded sample of the C2 workload that COVERS the horizon the GPU rolls
    out: CPU_SEG free-running steps from each of the states entering steps CPU_STARTS of the reference's own 300-step
    rollout (tests/golden/rollout_c2.npz; the cloud disperses, so edges per step fall from 40 k to 19 k and a sample of
    the first steps only would time the CPU on the densest graphs).  Returns (seconds per repeat, threads, steps per
    repeat, mean edges of the sampled steps, description)."""
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
    rigid0 = obs[-1, :n_rigid][:, scenes.CART_IDX]
    gpath = os.path.join(ROOT, "tests", "golden", "rollout_c2.npz")
    segments = []
    if n == 2000 and os.path.exists(gpath) and "obs_0" in np.load(gpath):
        g = np.load(gpath)
        rot, ty = g["rot"], g["ty"]
        for t0 in CPU_STARTS:
            seg = min(CPU_SEG, len(rot) - t0)
            traj = restate.rigid_body_trajectory(rot[t0: t0 + seg], ty[t0: t0 + seg], seg, scenes.TY_INIT, rigid0)
            segments.append((torch.from_numpy(g[f"obs_{t0}"]), traj, seg))
        desc = (f"{CPU_SEG} free-running steps from each of the states entering steps {list(CPU_STARTS)} of the "
                "reference's 300-step C2 rollout")
    else:
        seg = 4 * CPU_SEG
        rot, ty = control_curve(seg)
        segments.append((obs, restate.rigid_body_trajectory(rot, ty, seg, scenes.TY_INIT, rigid0), seg))
        desc = f"the first {seg} rollout steps of the scene"
    edges = []
    if warm:
        restate.rollout(model, feats, (segments[0][0], None), segments[0][1][:1], feats.stats, 1)
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        for state, traj, seg in segments:
            restate.rollout(model, feats, (state, None), traj, feats.stats, seg,
                            record=lambda i, d: edges.append(int(d["senders"].numel())))
        times.append(time.perf_counter() - t0)
    steps = sum(s[2] for s in segments)
    return times, torch.get_num_threads(), steps, float(np.mean(edges)), desc


def run_reference(args):
    """--impl reference: the reference's CPU path (restated in oracle/, kind "port": the reference is pure Python
    and does not travel to the GPU box) on the host cores; each bench step = a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.particles
    times, cores, sample, mean_edges, desc = cpu_rollout_sample(n, repeats=args.steps, warm=max(1, min(args.warmup, 1)))
    total = sum(times)
    value = n * sample * len(times) / total
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2 rollout: {n} particles x {args.rollout_steps} steps "
                                   f"(each bench step = {sample} rollout steps: {desc})", "k_steps": 6,
                       "conn_r": scenes.CONN_R, "max_neighbours": scenes.MAX_NBRS, "mp_steps": 10, "hidden": 128,
                       "mean_edges_per_sampled_step": mean_edges},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{sample} rollout steps per bench step ({desc}), {len(times)} bench steps; "
                                       "oracle/restate.py: sklearn KDTree + torch fp32 (measured 1.3-1.7x faster than "
                                       "the unmodified reference loop on the same cores, BASELINE.md section 3)"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# GPU legs
# ---------------------------------------------------------------------------------------------------------
PHASES = ["apply_control", "node_features", "cell_grid", "neighbour_kernel", "recv_csr", "enc_edge", "enc_node",
          "mp_edge", "mp_node", "integrate"]


def make_graph_attr(gm):
    return gm.GraphBoundedMultimaterialControl(scenes.CONN_R, scenes.make_stats(), scenes.CART_IDX, [scenes.MAT_IDX],
                                               scenes.CTRL_IDX, scenes.make_bounds(), max_neighbours=scenes.MAX_NBRS)


def profile_forward(gm, model, obs, reps, warm, flush):
    """Per-kernel-class device time of one model forward on the graph of `obs` (CUDA events on the launch stream,
    gm_model_forward_profile).  Returns (avg ms per launch for [enc_edge, enc_node, mp_edge, mp_node], E, N, pred)."""
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
    return tot / np.maximum(launches, 1), E, N, pred


def profile_rollout_step(gm, plan, reps, warm, flush):
    """Per-phase device time of one un-captured, serial rollout step (gm_rollout_step_profile): avg ms per launch."""
    L, lib = gm._lib.lib(), gm._lib
    ms, cnt = (ctypes.c_float * 10)(), (ctypes.c_int32 * 10)()
    tot, launches = np.zeros(10), np.zeros(10)
    for i in range(warm + reps):
        if flush is not None:
            flush.add_(1)
        lib.check(L.gm_rollout_step_profile(plan.handle, lib.stream_ptr(plan.device), ms, cnt),
                  "gm_rollout_step_profile")
        if i >= warm:
            tot += np.array(ms[:])
            launches += np.array(cnt[:])
    return tot / np.maximum(launches, 1)


def check_c3_forward(pred):
    """The tensor-core forward the C3 roofline is quoted on must reproduce the unmodified reference's forward on the
    same scene (tests/golden/c3_forward.npz): a kernel that computes garbage cannot report a roofline number."""
    path = os.path.join(ROOT, "tests", "golden", "c3_forward.npz")
    if not os.path.exists(path):
        return None
    g = np.load(path)
    got = pred.cpu()[torch.from_numpy(g["idx"].astype(np.int64))].double()
    want = torch.from_numpy(g["pred"]).double()
    rel = float((got - want).norm() / want.norm())
    if not rel < 1e-3:
        raise SystemExit(f"bench.py: C3 forward differs from the reference golden (rel_l2 {rel:.3e}); no number reported")
    return rel


def tensor_roofline(ms_edge, E, pk):
    """mp_edge_kernel timed ALONE with CUDA events -> burst bf16 peak is the denominator; the sustained figure beside."""
    tf = FLOP_PER_EDGE_MP * E / (ms_edge * 1e-3) / 1e12
    gbs = BYTES_PER_EDGE_MP * E / (ms_edge * 1e-3) / 1e9
    return {"kernel": "mp_edge16_kernel", "bound": "tensor", "achieved": tf, "peak": pk["tf_burst"], "unit": "TFLOP/s",
            "frac": tf / pk["tf_burst"], "traffic": None, "peak_source": pk["source"] + " burst (kernel timed alone)",
            "frac_of_sustained_peak": tf / pk["tf_sust"],
            "executed_frac": 1.8 * tf / pk["tf_burst"],   # bf16x3: 3 x 98 304 MAC-flops executed per 163 840 algorithmic
            "hbm_achieved_gbs": gbs, "hbm_frac": gbs / pk["hbm"], "edges_per_launch": int(E),
            "algorithmic_flop_per_edge": FLOP_PER_EDGE_MP, "algorithmic_bytes_per_edge": BYTES_PER_EDGE_MP,
            "us_per_launch": 1e3 * ms_edge, "edges_per_s": E / (ms_edge * 1e-3)}


def attach_traffic(entry, key, E):
    """DRAM bytes of the same launch from the committed ncu capture (never measured under this run); dropped when the
    capture is of another launch shape."""
    try:
        with open(os.path.join(ROOT, "profiles", "mp_edge_traffic.json")) as f:
            tr = json.load(f)[key]
        if int(tr["edges_per_launch"]) == int(E):
            entry["traffic"] = tr["dram_bytes_per_launch"]
            entry["traffic_source"] = tr["source"]
    except (OSError, KeyError, ValueError):
        pass


def hbm_rooflines(per_ms, N, E, K, pk):
    """Graph build / feature build / integrator at C3 as achieved GB/s on their algorithmic bytes."""
    def entry(kernel, ms, nbytes, what):
        gbs = nbytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        return {"kernel": kernel, "bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                "frac": gbs / pk["hbm"], "traffic": None, "us": 1e3 * ms, "algorithmic_bytes": int(nbytes),
                "bytes": what}
    p = dict(zip(PHASES, per_ms))
    a1 = p["cell_grid"] + p["neighbour_kernel"] + p["recv_csr"]
    return {
        "radius_graph_all": entry("bbox..cell_fill + neighbour_kernel + scans + receiver CSR sort (row A1)", a1,
                                  12 * N + 4 * K * N + 4 * N + 4 * (N + 1) + 3 * 4 * K * N,
                                  "read 12 N; write nbr 4 K N + cnt 4 N + recv_ptr 4 (N+1) + esend/erecv/eref 12 K N"),
        "neighbour_kernel": entry("neighbour_kernel", p["neighbour_kernel"], 16 * N + 4 * K * N + 4 * N,
                                  "read cell-sorted float4 positions 16 N; write nbr 4 K N + cnt 4 N"),
        "recv_csr": entry("edge_ptr scan + recv_fill + recv_sort_emit_warp", p["recv_csr"],
                          4 * K * N + 4 * N + 4 * (N + 1) + 3 * 4 * K * N,
                          "read nbr 4 K N + cnt 4 N; write recv_ptr + esend/erecv/eref"),
        "node_features": entry("node_features_kernel (row A2)", p["node_features"], 4 * 6 * 8 * N + 4 * 25 * N,
                               "read obs window 4*6*8 N; write nodes 4*25 N"),
        "integrate": entry("integrate_kernel (row A5)", p["integrate"], 4 * (3 * N + 2 * 8 * N) + 4 * 6 * 8 * N,
                           "read pred 12 N + two frames 64 N; write the shifted window 4*6*8 N (no ring buffer)"),
    }


def c4_population(gm, model, ga, obs, N, world, rank, dev, flush, barrier):
    """BASELINE.json configs[3] as SURVEY.md section 8 row D states it: InterpolatedCMAsolver(total_steps=150,
    traj_points=50) on sample_traj[:151] (6-dimensional search vector), population 64 = x0 + 1.5 N(0, I), numpy seed
    1234; the population is split contiguously over the ranks, each rank batching its candidates as disjoint graphs of
    ONE plan; a candidate is EVALUATED: parametrisation kernel -> 150-step rollout -> batched Sinkhorn objective ->
    loss assembly.  No collective on the data path (the popsize scalars are all-gathered outside the timed region)."""
    from gnn_manip_b200.sharding import partition_population
    pop, T4 = 64, 150
    nxt = obs[-1, :, 2:5].clone()
    solver = gm.InterpolatedCMAsolver(model, ga, (obs, nxt), scenes.RX_INIT_DEG, scenes.TY_INIT, 0.05, 1.0, 1e-3, 50.0,
                                      2e-3, 1.0, 0.5, dev, cma_var=1.5, cma_popsize=pop, cma_rand=1234, max_rot=3.0,
                                      max_ty=0.002, total_steps=T4, traj_points=50, impl=gm._lib.GM_IMPL_TC)
    solver.set_sample_traj(sample_traj_file()[: T4 + 1])
    x0 = solver._initial_params()
    X = x0[None] + 1.5 * np.random.RandomState(1234).randn(pop, x0.size)
    lo, hi = partition_population(pop, world, rank)
    n_rigid = int((obs[-1, :, scenes.MAT_IDX] == 1).sum())
    solver.desired_pos = (obs[-1, n_rigid:, 2:5] + torch.tensor([0.05, 0.0, -0.05])).to(dev).contiguous()
    stream = torch.cuda.current_stream(dev)

    def evaluate():
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record(stream)
        plan, terms = solver.rollout_population(X[lo:hi])
        ev[1].record(stream)
        loss = solver.combine_loss(solver.wasserstein_population(plan).double(), terms)
        ev[2].record(stream)
        return ev, loss, plan

    # parity guard: the first steps of this rank's first candidate on the tensor path against the fp32 CUDA-core path
    plan = solver._plan(hi - lo)
    solver.interpolate_population(X[lo:hi], plan.rot_dev, plan.ty_dev)
    plan.reset()
    single = gm.Rollout(model, ga, obs, rot=plan.rot_dev[0].cpu().numpy(), ty=plan.ty_dev[0].cpu().numpy(),
                        ty_init=tuple(scenes.TY_INIT), impl=gm._lib.GM_IMPL_SIMT, use_cuda_graph=False, device=dev)
    worst = 0.0
    for _ in range(3):
        plan.run(1)
        single.run(1)
        a, b = plan.pred[:N].double(), single.pred.double()
        worst = max(worst, float((a - b).norm() / b.norm()))
    single.close()
    if not worst < 1e-3:
        raise SystemExit(f"bench.py: C4 batched tensor-path rollout differs from the fp32 path (rel_l2 {worst:.3e})")
    evaluate()                                                     # warm-up of the whole objective
    barrier()
    flush.add_(1.0)
    ev, loss, plan = evaluate()
    barrier()
    ms_eval, ms_roll = ev[0].elapsed_time(ev[2]), ev[0].elapsed_time(ev[1])
    values = loss.cpu().numpy()
    info = {"population": pop, "rollout_steps": T4, "candidates_per_rank": hi - lo, "particles": N,
            "parametrisation": "InterpolatedCMAsolver(total_steps=150, traj_points=50) on sample_traj[:151], "
                               "x0 + 1.5 N(0, I), numpy seed 1234 (6-dim search vector)",
            "objective": "50 * Sinkhorn(p=2, blur=.05) + 1e-3 vel + 2e-3 acc + 0.5 interp (synthetic target cloud)",
            "tensor_vs_fp32_rel_l2_first_steps": worst, "objective_first_candidates": [float(v) for v in values[:3]]}
    solver.close()
    return ms_eval, ms_roll, info


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
    torch.manual_seed(0)
    model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    ga = make_graph_attr(gm)
    if args.only_c3 or args.only_c3_step:
        obs3, _ = scenes.make_state(100_000, seed=0, n_rigid=0)
        if args.only_c3:
            per3, E3, _, pred3 = profile_forward(gm, model, obs3, reps=2, warm=1, flush=None)
            print(json.dumps({"only_c3": True, "edges": E3, "kernel_ms": list(map(float, per3)),
                              "rel_l2_vs_reference": check_c3_forward(pred3)}), flush=True)
        else:
            plan3 = gm.Rollout(model, ga, obs3, use_cuda_graph=False, impl=gm._lib.GM_IMPL_TC, device=dev)
            per = profile_rollout_step(gm, plan3, reps=2, warm=1, flush=None)
            print(json.dumps({"only_c3_step": True, "phase_ms": dict(zip(PHASES, map(float, per)))}), flush=True)
        return

    obs, _ = scenes.make_state(N, seed=0, n_rigid=min(scenes.N_RIGID, N // 2))
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
    h2d = obs_host.numel() * 4 + rot_host.numel() * 8 + ty_host.numel() * 8
    d2h = out_host.numel() * 4

    def one_e2e():
        flush.add_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        plan.rot_dev.view(-1).copy_(rot_host, non_blocking=True)   # this candidate's control curve
        plan.ty_dev.view(-1).copy_(ty_host, non_blocking=True)
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

    # edges per rollout step over the horizon (untimed): the cloud disperses, the graphs thin out
    edge_trace = None
    if rank == 0:
        plan.reset()
        n_edges_dev = plan.graph_handles()[3]
        trace = torch.zeros((T,), dtype=torch.int32, device=dev)
        for i in range(T):
            plan.run(1)
            trace[i].copy_(n_edges_dev[0], non_blocking=True)
        edge_trace = trace.cpu().numpy()

    # ---- C4: evaluated CMA-ES population ----
    ms_c4, ms_c4_roll, c4 = 0.0, 0.0, None
    if not args.no_c4:
        del plan
        ms_c4, ms_c4_roll, c4 = c4_population(gm, model, ga, obs, N, world, rank, dev, flush, barrier)
        plan = None

    t = torch.tensor([ms, ms_e2e, ms_c4, ms_c4_roll], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_c4, ms_c4_roll = (float(v) for v in t)
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
    if edge_trace is not None:
        line["config"]["edges_per_rollout_step"] = {"first": int(edge_trace[0]), "mean": float(edge_trace.mean()),
                                                    "last": int(edge_trace[-1])}
    if c4 is not None:
        c4.update({"ms": ms_c4, "ms_rollouts_only": ms_c4_roll,
                   "cma_candidates_per_s": c4["population"] / (ms_c4 * 1e-3),
                   "rollout_candidates_per_s": c4["population"] / (ms_c4_roll * 1e-3),
                   "particle_steps_per_s": c4["population"] * N * c4["rollout_steps"] / (ms_c4 * 1e-3),
                   "us_per_rollout_step": 1e3 * ms_c4_roll / c4["rollout_steps"]})
        line["c4_cma_population"] = c4
    if rank == 0:
        per, E, _, _ = profile_forward(gm, model, obs, reps=20, warm=3, flush=None)
        line["roofline_c2"] = tensor_roofline(per[2], E, pk)
        line["roofline_c2"]["note"] = ("first step of the C2 rollout (39 999 edges, L2-resident latents): launch / "
                                       "latency-bound regime")
        attach_traffic(line["roofline_c2"], "c2", E)
        line["kernel_us_c2"] = {"enc_edge": 1e3 * per[0], "enc_node": 1e3 * per[1], "mp_edge": 1e3 * per[2],
                                "mp_node": 1e3 * per[3]}
        if not args.no_c3:
            obs3, _ = scenes.make_state(100_000, seed=0, n_rigid=0)
            per3, E3, N3, pred3 = profile_forward(gm, model, obs3, reps=5, warm=2, flush=flush)
            line["roofline"] = tensor_roofline(per3[2], E3, pk)
            line["roofline"]["workload"] = "C3: 100 000-particle cloud, 2 M edges, one mp_edge launch (L2 flushed)"
            line["roofline"]["forward_rel_l2_vs_reference"] = check_c3_forward(pred3)
            attach_traffic(line["roofline"], "c3", E3)
            line["kernel_us_c3"] = {"enc_edge": 1e3 * per3[0], "enc_node": 1e3 * per3[1], "mp_edge": 1e3 * per3[2],
                                    "mp_node": 1e3 * per3[3]}
            fwd_ms = per3[0] + per3[1] + 10 * (per3[2] + per3[3])
            line["c3_model_particle_steps_per_s"] = N3 / (fwd_ms * 1e-3)
            line["mp_step_edges_per_s_c3"] = E3 / ((per3[2] + per3[3]) * 1e-3)
            # one whole rollout step at C3, phase by phase, and a 20-step rollout
            plan3 = gm.Rollout(model, ga, obs3, use_cuda_graph=True, impl=gm._lib.GM_IMPL_TC, device=dev)
            ph = profile_rollout_step(gm, plan3, reps=3, warm=1, flush=flush)
            line["phase_us_c3"] = {k: 1e3 * float(v) for k, v in zip(PHASES, ph)}
            line["rooflines_c3"] = hbm_rooflines(ph, N3, E3, scenes.MAX_NBRS, pk)
            plan3.reset()
            plan3.run(3)
            flush.add_(1.0)
            plan3.reset()
            a3, b3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a3.record(stream)
            plan3.run(20)
            b3.record(stream)
            torch.cuda.synchronize(dev)
            ms3 = a3.elapsed_time(b3)
            line["c3_rollout"] = {"particles": N3, "steps": 20, "ms": ms3, "us_per_rollout_step": 1e3 * ms3 / 20,
                                  "particle_steps_per_s": N3 * 20 / (ms3 * 1e-3),
                                  "frac_of_hbm_bound_28M": N3 * 20 / (ms3 * 1e-3) / 28.0e6}
            plan3.close()
        else:
            line["roofline"] = line["roofline_c2"]
        if world == 1 and not args.no_cpu_baseline:
            times, cores, sample, mean_edges, desc = cpu_rollout_sample(N)
            line["cpu_baseline"] = {"value": N * sample / times[0], "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{sample} rollout steps ({desc}; oracle/restate.py, sklearn KDTree + "
                                              "torch fp32), 1 warm-up step",
                                    "mean_edges_per_sampled_step": mean_edges}
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
    if a.workload == "c5":
        import bench_train
        bench_train.run(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
    sys.stdout.flush()
