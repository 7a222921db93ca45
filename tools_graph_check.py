"""Debug aid: the tiny-input graph path against the general path on awkward clouds (huge extents, batched graphs)."""
import os, sys, subprocess
import numpy as np, torch

def run(case):
    import gnn_manip_b200 as gm
    from oracle import scene
    g = torch.Generator().manual_seed(case)
    G, n = 5, 400
    pos = torch.cat([scene.lattice_positions(n, 0.005, seed=case * 10 + i)[0] for i in range(G)])
    if case >= 1:
        pos[::7, 1] += 1.2 * torch.randn((pos[::7].shape[0],), generator=g)       # far-away rigid particles
    if case >= 2:
        pos[::11] += 30.0 * torch.randn((pos[::11].shape[0], 3), generator=g)
    gr = gm.radius_graph(pos.cuda(), 0.015, 20, n_graphs=G)
    torch.cuda.synchronize()
    ei = gr.edge_index().cpu()
    print("case", case, "tiny", os.environ.get("GM_GRAPH_TINY", "1"), "edges", gr.n_edges, "checksum", int(ei.sum()), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(int(sys.argv[1]))
    else:
        for case in range(3):
            for tiny in ("1", "0"):
                env = dict(os.environ, GM_GRAPH_TINY=tiny, CUDA_LAUNCH_BLOCKING="1")
                r = subprocess.run([sys.executable, __file__, str(case)], env=env, capture_output=True, text=True)
                print(r.stdout.strip() or ("FAILED: " + r.stderr.strip().splitlines()[-1][:300]), flush=True)
