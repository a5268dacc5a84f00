import sys, numpy as np, torch
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_b200 import abi, lib, scenes
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_trace import _brute_force_first_hit
lib.init(0)
nobj = int(sys.argv[1]) if len(sys.argv) > 1 else 160
b = scenes.scene_10m(width=64, height=36, spp=1, n_obj=nobj)
sc = lib.Scene(b)
print("tris", sc.info().n_tris, "nodes", sc.info().n_bvh_nodes)
pos_h, _, _, idx_h = b.arrays()
pos = torch.from_numpy(pos_h).cuda(); idx = torch.from_numpy(idx_h.astype(np.int64)).cuda()
n = 1 << 22
for kind in (0, 1):
    rays = torch.empty((n, 8), dtype=torch.float32, device="cuda"); hits = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    sc.gen_rays_device(kind, n, 1004, rays.data_ptr()); torch.cuda.synchronize()
    sc.trace_device(rays.data_ptr(), n, hits.data_ptr(), abi.TRACE_CLOSEST)
    prim = hits.view(torch.int32)[:, 0].clone(); t = hits[:, 3].clone()
    sel = torch.arange(0, n, n // 128, device="cuda")[:128]
    bt, bp = _brute_force_first_hit(pos, idx, rays[sel])
    gp, gt = prim[sel].long(), t[sel].double()
    bad = (bp != gp)
    print("kind", kind, "hit frac", (prim != -1).float().mean().item(), "disagree", bad.sum().item(), "of", len(sel))
    for k in torch.nonzero(bad)[:6, 0].tolist():
        print("   ray", sel[k].item(), "gpu prim", gp[k].item(), "t", gt[k].item(), " brute prim", bp[k].item(), "t", bt[k].item(), "ray", rays[sel[k]].tolist())
    hit = prim != -1
    r2 = rays.clone(); r2[:, 7] = torch.where(hit, t * (1 - 1e-4), r2[:, 7])
    h2 = torch.empty_like(hits); sc.trace_device(r2.data_ptr(), n, h2.data_ptr(), abi.TRACE_CLOSEST)
    p2 = h2.view(torch.int32)[:, 0]; bad2 = (p2 != -1) & hit
    print("retrace finds something:", bad2.sum().item())
    for k in torch.nonzero(bad2)[:5, 0].tolist():
        print("   ray", k, "first", prim[k].item(), t[k].item(), "retrace", p2[k].item(), h2[k, 3].item(), "maxt2", r2[k, 7].item(), "mint", rays[k,3].item())
