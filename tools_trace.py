"""Debug aid (-DGM_TRACE build): timeline of one CUDA-graph rollout step.  Every kernel stamps %globaltimer right
after its dependency wait (= when its predecessor completed); the deltas are the effective critical-path durations."""
import ctypes, sys
import numpy as np, torch
import gnn_manip_b200 as gm, scenes, bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
ng = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.manual_seed(0)
model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
obs, _ = scenes.make_state(n, seed=0, n_rigid=min(scenes.N_RIGID, n // 2))
curves = [bench.control_curve(20, i) for i in range(ng)]
rot = np.stack([c[0] for c in curves]) if ng > 1 else curves[0][0]
ty = np.stack([c[1] for c in curves]) if ng > 1 else curves[0][1]
plan = gm.Rollout(model, bench.make_graph_attr(gm), obs, rot=rot, ty=ty, ty_init=tuple(scenes.TY_INIT),
                  n_graphs=ng, use_cuda_graph=True, impl=gm._lib.GM_IMPL_TC)
buf = (ctypes.c_ulonglong * (5 * 161))()
dump = lambda: gm._lib.check(gm._lib.lib().gm_debug_stamps(buf, 5 * 161), "gm_debug_stamps")
plan.run(12); torch.cuda.synchronize(); dump()
plan.run(2); torch.cuda.synchronize(); dump()
s = np.array(buf[:], dtype=np.int64)
names = ["physics", "radius", "enc_node", "edge", "node"]
ev = []
for i, nm in enumerate(names):
    c = int(s[i * 161])
    ev += [(int(t), nm) for t in s[i * 161 + 1: i * 161 + 1 + c]]
ev.sort()
t0 = ev[0][0]
half = len(ev) // 2
print("events", len(ev))
for k, (t, nm) in enumerate(ev[half:] if len(ev) % 2 == 0 else ev):
    base = ev[half][0] if len(ev) % 2 == 0 else t0
    nxt = (ev[half:] if len(ev) % 2 == 0 else ev)
    d = (nxt[k + 1][0] - t) / 1e3 if k + 1 < len(nxt) else float("nan")
    print(f"{(t - base) / 1e3:8.1f} us  {nm:8s} -> next {d:6.1f} us")
