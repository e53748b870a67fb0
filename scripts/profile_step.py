"""Runs ON THE GPU BOX under `ncu --profile-from-start off` (scripts/profile_round2.sh): every kernel of the hot path once,
inside ONE cudaProfilerStart/Stop region, at the benchmarked cfg2 size -- edge builders, index preparation, diffusion,
a 2-layer slice of the training step (forward, loss, backward, optimizer: kernels are the same for every layer), the
sampler step and COM projection, the structure-analysis kernels.  Numbers printed under ncu are never bench values."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from diffusion_model_b200 import evaluate as EV  # noqa: E402
from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN  # noqa: E402
from diffusion_model_b200.data import Data  # noqa: E402
from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess, remove_mean  # noqa: E402
from diffusion_model_b200.graph import GraphIndex, build_edges_cutoff, build_edges_fc  # noqa: E402
from diffusion_model_b200.optim import FlatAdam, RAdamScheduleFree  # noqa: E402
from diffusion_model_b200.sampler import sample_structures  # noqa: E402
from diffusion_model_b200.train_per_iretation import batch_loss  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
cfg = dict(bench.CFG)
if workload == "cfg4":
    cfg.update(bench.CFG4)
dev = torch.device("cuda", 0)
G, n, L = cfg["graphs"], cfg["atoms"], 2
pos, onehot, box = bench.synth_batch(G, n, cfg["seed"] + 1)
pos, onehot, box = pos.to(dev), onehot.to(dev), box.to(dev)
node_ptr = (torch.arange(G + 1) * n).to(torch.int32).to(dev)
torch.manual_seed(cfg["seed"])
net = EquivariantGNN(**bench.net_sizes(L, cfg["d"], cfg["H"])).to(dev)
net.precision = "bf16x3"
opt = FlatAdam(net, lr=1e-5)
dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
prm = dict(atom_type_size=2)


def step(ei):
    b = Data(x=onehot, pos=pos, edge_index=ei)
    b.node_ptr = node_ptr
    opt.zero_grad()
    loss, _, _ = batch_loss(net, b, dp, prm)
    loss.backward()
    opt.step()
    return loss


ei = build_edges_cutoff(pos, cfg["r_c"], node_ptr=node_ptr, box=box)
for _ in range(3):
    step(ei)
small = EquivariantGNN(**bench.net_sizes(2, 128, 3)).to(dev)
bench.tame_(small, 2)
small.precision = "bf16x3"
dps = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=4)
sample_structures(small, dps, 192, 64, use_cuda_graph=False)
torch.cuda.synchronize()

torch.cuda.profiler.start()
ei2 = build_edges_cutoff(pos, cfg["r_c"], node_ptr=node_ptr, box=box)          # cutoff_kernel<count/fill>, scan, fallback
GraphIndex(ei2, pos.shape[0], node_ptr=node_ptr)                               # convert_edges, boundaries, senders, sort_segments
step(ei2)                                                                      # diffuse, pack, edge/node/xupdate fwd, loss, bwd, reduce, adam
build_edges_fc([192] * 64, dev)                                                # fc_edges_kernel
sample_structures(small, dps, 192, 64, use_cuda_graph=False)                   # remove_mean, sampler_step, counter_add
x = torch.randn(pos.shape[0], 3, device=dev)
remove_mean(x, torch.repeat_interleave(torch.arange(G, device=dev), n))        # batch_to_i32, boundaries<batch>, remove_mean
ro = RAdamScheduleFree(list(net.parameters()), lr=1e-5)
ro.train()
for p in net.parameters():
    p.grad = torch.zeros_like(p)
for _ in range(6):
    ro.step()                                                                  # radam_sf_kernel (silent for the first steps)
ro.eval()                                                                      # lerp_kernel
P = pos.view(G, n, 3)
EV.RDF_batch(P)                                                                # rdf_local_kernel
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled region done")
