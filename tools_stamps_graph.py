"""Debug aid (-DGM_STAMPS build): phase stamps of CTA 0 of the last mp_edge launch of a CUDA-graph rollout step
(shows whether programmatic dependent launch overlaps the prologue with the previous kernel: long `pdl wait`)."""
import ctypes, sys
import numpy as np, torch
import gnn_manip_b200 as gm, scenes, bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
graph = (sys.argv[2] != "nograph") if len(sys.argv) > 2 else True
torch.manual_seed(0)
model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
obs, _ = scenes.make_state(n, seed=0, n_rigid=min(scenes.N_RIGID, n // 2))
rot, ty = bench.control_curve(20, 0)
plan = gm.Rollout(model, bench.make_graph_attr(gm), obs, rot=rot, ty=ty, ty_init=tuple(scenes.TY_INIT),
                  use_cuda_graph=graph, impl=gm._lib.GM_IMPL_TC)
plan.run(5)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); plan.run(10); b.record(); torch.cuda.synchronize()
st = (ctypes.c_ulonglong * 64)()
gm._lib.check(gm._lib.lib().gm_debug_stamps(st, 64), "gm_debug_stamps")
s = np.array(st[:], dtype=np.int64)
print(f"n={n} graph={graph} us/step={a.elapsed_time(b) * 100:.1f} edge pdl wait (cycles): {int(s[24] - s[23])}")
