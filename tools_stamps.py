"""Debug aid (needs a -DGM_STAMPS build: make -C gnn_manip_b200/csrc clean all EXTRA=-DGM_STAMPS):
prints the phase timestamps CTA 0 recorded during one forward on the C2 or C3 scene."""
import ctypes, sys
import numpy as np, torch
import gnn_manip_b200 as gm, scenes, bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
torch.manual_seed(0)
model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
obs, _ = scenes.make_state(n, seed=0, n_rigid=0)
per, E, N, _ = bench.profile_forward(gm, model, obs, reps=2, warm=1, flush=None)
st = (ctypes.c_ulonglong * 64)()
gm._lib.check(gm._lib.lib().gm_debug_stamps(st, 64), "gm_debug_stamps")
s = np.array(st[:], dtype=np.int64)
print("kernel ms", per, "E", E)
names16 = ["top", "ps landed", "GEMM1 done", "E1 published", "GEMM2 done", "E2 published", "GEMM3 done", "next operand", "LN pass 1",
           "LN exchange", "LN rows stored", "RED issued", "slot barrier", "sweeps done", "slot barrier", "next tile begun"]
print("mp_edge16 (default edge kernel) phase deltas:", {names16[i + 1]: int(s[i + 1] - s[i]) for i in range(15)}, "tile:", int(s[15] - s[0]))
print("edge compute stamps (cycles since stamp 0):", [int(x - s[0]) for x in s[:16]])
print("edge compute deltas:", [int(s[i + 1] - s[i]) for i in range(15)])
print("edge E1 fine stamps (since stamp 5: ld0 done, blk0, blk1, blk2, blk3, published, issued):", [int(x - s[5]) for x in s[16:23]])
print("edge LN fine stamps (since stamp 2: acc loaded, mean, rstd; stamp 10 = rows stored):", [int(x - s[2]) for x in (s[25], s[26], s[27], s[10])])
print("edge pdl wait (cycles):", int(s[24] - s[23]))
names_n = ["v rows -> operand", "publish + issue 1b", "GEMM 1b", "weights + pdl_wait", "agg rows -> operand", "publish + issue 1a", "GEMM 1a",
           "residual copy + epilogue 1", "publish + issue 2", "GEMM 2", "epilogue 2", "publish + issue 3", "GEMM 3", "LayerNorm + operand",
           "publish + issue Wr", "v' rows out", "GEMM Wr", "Pr staged", "publish + issue Ws", "Pr out + GEMM Ws", "Ps staged", "Ps out"]
sn = s[32:55]
print("mp_node16 phase deltas:", {names_n[i]: int(sn[i + 1] - sn[i]) for i in range(22)}, "tile:", int(sn[22] - sn[0]))
print("node_finish stamps (since 44: ln, a0, launch Wr, copy v, wait, stage Pr, launch Ws, wait, stage Ps, copy Ps):", [int(x - s[44]) for x in s[45:55]])
