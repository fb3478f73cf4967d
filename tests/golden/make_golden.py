"""Generates tests/golden/noisy_off_golden.npz (run in the BUILD container only).

Input: the reference's own example fixture /root/reference/examples/data/noisy.off (config 1,
examples/ex_simple.py:53-75: 10 201 vertices, box +-1.01, periodic).  Outputs stored:
  * float32 results of the reference's OWN C++ (oracle/_ref/libmsref.so, compiled in place from
    memsurfer/src/TriMesh.cpp + TriMesh_kde.cpp): need_normals, need_pointareas on the file's own
    faces, kde type 2 / sigma 0.2 / get_counts for 'all' and for the label '1' subset
    (ex_simple.py:116-120, labels seeded here because the script's are not);
  * the ghost-augmented point set size and neighbour lists / oriented normals of the fp64
    restatement (PARITY UNPINNED for these: CGAL is not available).
Nothing under /root/reference is read at test time; this file is the only consumer.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O          # noqa: E402
from memsurfer_b200 import synth        # noqa: E402

verts, faces = synth.read_off("/root/reference/examples/data/noisy.off")
verts = verts.astype(np.float32)
faces = faces.astype(np.uint32)
box = np.array([-1.01, -1.01, -1.01, 1.01, 1.01, 1.01], np.float32).reshape(2, 3)
labels = np.random.default_rng(1234).integers(0, 5, size=verts.shape[0]).astype(np.int32)
ids1 = np.where(labels == 1)[0].astype(np.int32)
bboxl = np.concatenate([verts.min(0), verts.max(0)]).astype(np.float32)

out = dict(verts=verts, faces=faces, box=box, labels=labels, bboxl=bboxl)
out["ref_normals"] = O.ref_need_normals(verts, faces)
out["ref_pareas"] = O.ref_need_pointareas(verts, faces)
out["ref_kde_all"] = O.ref_kde(verts, np.zeros(0, np.int32), 2, 0.2, True, bboxl)
out["ref_kde_l1"] = O.ref_kde(verts, ids1, 2, 0.2, True, bboxl)
out["ref_kde3_all_nonper"] = O.ref_kde(verts, np.zeros(0, np.int32), 3, 0.2, False, None)
pn, info = O.point_normals(verts, box, 0.2, 18, +1, return_all=True)
out["orc_m"] = np.int64(info["points"].shape[0])
out["orc_nbr"] = info["nbr"]
out["orc_pnormals"] = pn
out["orc_n_unoriented"] = np.int64(info["n_unoriented"])
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "noisy_off_golden.npz"), **out)
print({k: getattr(v, "shape", v) for k, v in out.items()})
