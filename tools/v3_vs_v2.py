"""Register-blocked kernel (v3) against the line-task kernel (v2) on uniform bricks: GDoF/s, roofline fraction, agreement."""
import os, sys, json
sys.path.insert(0, '/root/repo')
import numpy as np, torch, d4est_b200
out = {}
for deg, n in ((1, 128), (2, 96), (3, 80)):
    for tag, env in (("v3", {}), ("v2", {"D4B_NO_V3": "1"})):
        os.environ.update(env)
        try:
            lvl = d4est_b200.MeshLevel.brick(n, deg=deg)
            nd = lvl.local_nodes
            g = torch.Generator(device="cuda"); g.manual_seed(7)
            u = torch.rand(nd, dtype=torch.float64, device="cuda", generator=g); Au = torch.empty_like(u)
            st = torch.cuda.current_stream().cuda_stream
            for _ in range(3): lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20): lvl.apply_dev(u.data_ptr(), Au.data_ptr(), st)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            out[(deg, tag)] = (nd / ms / 1e6, Au.cpu().numpy().copy(), lvl.algorithmic_bytes / ms / 1e6 / 6541.5)
            lvl.close()
        finally:
            for k in env: del os.environ[k]
    a, b = out[(deg, "v3")][1], out[(deg, "v2")][1]
    print(json.dumps({"deg": deg, "n": n, "v3_gdofs": out[(deg, "v3")][0], "v3_frac": out[(deg, "v3")][2], "v2_gdofs": out[(deg, "v2")][0], "v2_frac": out[(deg, "v2")][2],
                      "rel_diff": float(np.linalg.norm(a - b) / np.linalg.norm(b))}), flush=True)
