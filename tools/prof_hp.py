"""Profiling driver for the hp path (used under ncu): a few applies on a curved mixed-hp table file."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, d4est_b200
from d4est_b200 import tables
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
fx = tables.load(sys.argv[1] if len(sys.argv) > 1 else os.path.join(root, "oracle", "_ref", "fixtures", "cubed_sphere_hp3.fix"))
lvl = d4est_b200.MeshLevel(fx)
u = torch.rand(lvl.local_nodes, dtype=torch.float64, device="cuda"); Au = torch.empty_like(u)
for _ in range(4):
    lvl.apply_dev(u.data_ptr(), Au.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize(); print("ok", float(Au.abs().sum()))
