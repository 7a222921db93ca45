"""Profiling aid: ONE training step (C5 batch) between cudaProfilerStart/Stop, for
ncu --profile-from-start off --metrics gpu__time_duration.sum ... python tools_train_launches.py"""
import torch
import gnn_manip_b200 as gm, scenes, bench_train as bt

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = gm.EncProcDecGNN(25, 4, 3, 128, 2, 10)
ga = gm.GraphBoundedMultimaterialControl(scenes.CONN_R, scenes.make_stats(), scenes.CART_IDX, [scenes.MAT_IDX], scenes.CTRL_IDX,
                                         scenes.make_bounds(), noise=bt.NOISE_STD, max_neighbours=scenes.MAX_NBRS)
n_nodes = bt.GRAPHS_PER_GPU * bt.PARTICLES
tr = gm.Trainer(model, n_nodes, n_nodes * scenes.MAX_NBRS, lr=1e-4, device=dev, bucketed_allreduce=False)
batch = gm.make_batch(ga, bt.make_samples(0), device=dev, seed=1)
for _ in range(2):
    tr.loss_batch(batch)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.loss_batch(batch)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("loss", float(tr.loss))
