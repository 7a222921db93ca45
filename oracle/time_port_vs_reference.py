"""TEST INFRASTRUCTURE ONLY.  Build container only (needs /root/reference): wall time of the UNMODIFIED reference's
cma_objective loop against the oracle port (oracle/restate.rollout, what bench.py's CPU legs time) on the same cores and
the same C2 scene -- the ratio recorded in BASELINE.md section 3.   python -m oracle.time_port_vs_reference [steps]"""
import contextlib
import io
import os
import sys
import time

import numpy as np
import torch

from . import ref_shims, restate, scene


def main(steps=8):
    ref = ref_shims.load()
    torch.set_num_threads(os.cpu_count() or 1)
    n, n_rigid = 2000, scene.N_RIGID
    obs, nxt = scene.make_state(n, seed=0, n_rigid=n_rigid)
    st = np.load(os.path.join(ref_shims.REFERENCE_ROOT, "dataset", "sample_traj.npy"))
    rot, ty = np.deg2rad(st[:steps, 0]), st[:steps, 1] - st[0, 1]
    ga = ref.collate.GraphBoundedMultimaterialControl(scene.CONN_R, scene.make_stats(), scene.CART_IDX, [scene.MAT_IDX],
                                                      scene.CTRL_IDX, scene.make_bounds(), noise=None,
                                                      max_neighbours=scene.MAX_NBRS)
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        model = ref.EncProcDecGNN(25, 4, 3, 128, 2, 10)

    class Solver(ref.traj.CMAESolver):
        def interpolate_trajectory(self, x):
            return list(rot), list(ty)

        def compute_loss(self, *a, **k):
            return 0.0, None, None, None, None, None

    with contextlib.redirect_stdout(io.StringIO()):
        solver = Solver(model, ga, (obs.clone(), nxt.clone()), scene.RX_INIT_DEG, scene.TY_INIT, 0.05, 1.0, 1.0, 1.0, 1.0,
                        1.0, 1.0, "cpu", total_steps=steps, traj_points=1)
    torch.manual_seed(0)
    port = restate.EncProcDecGNN(25, 4, 3, 128, 2, 10)
    feats = restate.GraphFeatures(scene.CONN_R, scene.make_stats(), scene.make_bounds(), scene.CART_IDX, scene.MAT_IDX,
                                  scene.CTRL_IDX, scene.MAX_NBRS, connectivity=restate.get_connectivity_sklearn)
    traj = restate.rigid_body_trajectory(rot, ty, steps, scene.TY_INIT, obs[-1, :n_rigid][:, scene.CART_IDX])
    out = {}
    for rep in range(3):
        t0 = time.perf_counter()
        solver.cma_objective(np.zeros(2))
        t1 = time.perf_counter()
        restate.rollout(port, feats, (obs, None), traj, feats.stats, steps)
        t2 = time.perf_counter()
        out[rep] = ((t1 - t0) / steps, (t2 - t1) / steps)
        print(f"rep {rep}: reference {out[rep][0]:.3f} s/step, port {out[rep][1]:.3f} s/step, "
              f"reference / port = {out[rep][0] / out[rep][1]:.2f}  ({torch.get_num_threads()} threads, {steps} steps)")


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 8)
