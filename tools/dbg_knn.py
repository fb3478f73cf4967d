import numpy as np, torch, sys
sys.path.insert(0, "/root/repo")
from memsurfer_b200 import point_normals_device
g = np.load("/root/repo/tests/golden/noisy_off_golden.npz")
v, box = g["verts"], g["box"]
keep = {}
point_normals_device(torch.from_numpy(v).cuda(), box, True, 0.2, 18, +1, keep=keep)
nbr = keep["nbr"].cpu().numpy(); want = g["orc_nbr"]; P = keep["points_all"].cpu().numpy().astype(np.float64)
bad = np.where((nbr != want).any(axis=1))[0]
print("rows", nbr.shape, "bad rows", len(bad))
for r in bad[:6]:
    d_got = ((P[nbr[r]] - P[r])**2).sum(1); d_want = ((P[want[r]] - P[r])**2).sum(1)
    print(r, "got ", nbr[r].tolist()); print("   want", want[r].tolist())
    print("   d_got ", np.round(d_got, 6).tolist()); print("   d_want", np.round(d_want, 6).tolist())
