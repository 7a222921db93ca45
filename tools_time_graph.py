"""Debug aid: live time of the graph build (row A1) alone on the C2 / C4 scenes."""
import sys, torch
import gnn_manip_b200 as gm, scenes
for n, ng in ((2000, 1), (2000, 8), (2000, 64), (100_000, 1)):
    obs, _ = scenes.make_state(n, seed=0, n_rigid=0)
    pos = obs[-1, :, 2:5].contiguous().cuda()
    if ng > 1:
        pos = pos.repeat(ng, 1).contiguous()
    for _ in range(3):
        g = gm.radius_graph(pos, scenes.CONN_R, scenes.MAX_NBRS, n_graphs=ng)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50):
        g = gm.radius_graph(pos, scenes.CONN_R, scenes.MAX_NBRS, n_graphs=ng)
    b.record(); torch.cuda.synchronize()
    print(f"n={n} graphs={ng}: {a.elapsed_time(b) / 50 * 1e3:.1f} us per build (eager launches, includes host launch cost)")
