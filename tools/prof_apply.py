"""Small profiling driver (used under ncu): a few applies on a brick, nothing else."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, d4est_b200
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
deg = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lvl = d4est_b200.MeshLevel.brick(n, deg=deg)
u = torch.rand(lvl.local_nodes, dtype=torch.float64, device="cuda"); Au = torch.empty_like(u)
for _ in range(4):
    lvl.apply_dev(u.data_ptr(), Au.data_ptr(), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize(); print("ok", float(Au.abs().sum()))
