"""Synthetic inputs and the HOST stand-in for the reference's mesh producer.

The reference builds its membrane mesh on the host with pypoisson + CGAL
(memsurfer/membrane.py:160-234); by design that producer stays on the host and
is not part of the B200 hot path.  CGAL and a working pypoisson are absent in
this image, so this module supplies

  * the synthetic periodic undulating bilayer of SURVEY.md section 8(d), and
  * a scipy-based restatement of `TriMesh::periodicDelaunay`
    (memsurfer/src/TriMesh_cgal.cpp:160-281: keep every face of the periodic
    covering that has >= 1 original vertex) followed by a vectorised
    restatement of `TriMesh::trim_periodicDelaunay`
    (memsurfer/src/TriMesh.cpp:493-575),

so that the GPU path is fed arrays with exactly the reference layout
(`faces`, `_periodic_faces`, `_trimmed_faces`, `_duplicated_vidxs`,
`_duplicated_verts`; memsurfer/trimesh.py:421-436).  Connectivity is an INPUT
of the hot path; nothing here runs on the GPU.
"""
import numpy as np


def make_leaflet(npoints, seed=1234, z0=20.0, ntypes=3, fractions=None, bumps=0,
                 area_per_lipid=64.0, phase_shift=0.0):
    """Jittered sqrt(N) x sqrt(N) grid on an undulating periodic sheet (units: Angstrom).

    Returns (points (N,3) f32, labels (N,) int32, bbox (2,3) f32) with N = round(sqrt(npoints))**2.
    """
    rng = np.random.default_rng(seed)
    ns = int(round(np.sqrt(npoints)))
    n = ns * ns
    L = float(np.sqrt(area_per_lipid * n))
    h = L / ns
    i, j = np.meshgrid(np.arange(ns), np.arange(ns), indexing="xy")
    x = (i.reshape(-1) + 0.5 + rng.uniform(-0.3, 0.3, n)) * h
    y = (j.reshape(-1) + 0.5 + rng.uniform(-0.3, 0.3, n)) * h
    z = np.full(n, z0, dtype=np.float64)
    for m in range(4):
        p, q = int(rng.integers(0, 5)), int(rng.integers(0, 5))
        if p + q == 0:
            p = 1
        amp = rng.uniform(0.25, 1.0) * 0.02 * L / (p + q)
        amp = min(amp, 25.0)
        phi = rng.uniform(0, 2 * np.pi) + phase_shift * (m + 1)
        z += amp * np.sin(2 * np.pi * (p * x + q * y) / L + phi)
    for _ in range(int(bumps)):
        cx, cy = rng.uniform(0, L, 2)
        hgt, wid = rng.uniform(30, 60), rng.uniform(40, 80)
        dx = np.abs(x - cx); dx = np.minimum(dx, L - dx)
        dy = np.abs(y - cy); dy = np.minimum(dy, L - dy)
        z += hgt * np.exp(-(dx * dx + dy * dy) / (2 * wid * wid))
    pts = np.stack([x, y, z], axis=1).astype(np.float32)
    if fractions is None:
        fractions = np.full(ntypes, 1.0 / ntypes)
    labels = rng.choice(len(fractions), size=n, p=np.asarray(fractions) / np.sum(fractions)).astype(np.int32)
    bbox = np.array([[0.0, 0.0, -L / 2], [L, L, L / 2]], dtype=np.float32)
    return pts, labels, bbox


def make_leaflet_rect(nx, ny, seed=1234, z0=20.0, ntypes=3, fractions=None, area_per_lipid=64.0, jitter=0.3,
                      amplitude=1.0):
    """Rectangular-box variant of make_leaflet: a jittered nx x ny grid in a box Lx x Ly with Lx/Ly = nx/ny
    (`jitter` = 0 gives the exact lattice: with area_per_lipid = 64 every coordinate is a float32 integer + 4,
    so squared distances tie EXACTLY).  Same return values as make_leaflet."""
    rng = np.random.default_rng(seed)
    n = nx * ny
    h = float(np.sqrt(area_per_lipid))
    Lx, Ly = nx * h, ny * h
    i, j = np.meshgrid(np.arange(nx), np.arange(ny), indexing="xy")
    jx = rng.uniform(-jitter, jitter, n) if jitter > 0 else np.zeros(n)
    jy = rng.uniform(-jitter, jitter, n) if jitter > 0 else np.zeros(n)
    x = (i.reshape(-1) + 0.5 + jx) * h
    y = (j.reshape(-1) + 0.5 + jy) * h
    z = np.full(n, z0, dtype=np.float64)
    for m in range(4):
        p, q = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        if p + q == 0:
            p = 1
        amp = amplitude * rng.uniform(0.25, 1.0) * 0.02 * min(Lx, Ly) / (p + q)
        amp = min(amp, 25.0)
        phi = rng.uniform(0, 2 * np.pi)
        z += amp * np.sin(2 * np.pi * (p * x / Lx + q * y / Ly) + phi)
    pts = np.stack([x, y, z], axis=1).astype(np.float32)
    if fractions is None:
        fractions = np.full(ntypes, 1.0 / ntypes)
    labels = rng.choice(len(fractions), size=n, p=np.asarray(fractions) / np.sum(fractions)).astype(np.int32)
    Lz = max(Lx, Ly)
    bbox = np.array([[0.0, 0.0, -Lz / 2], [Lx, Ly, Lz / 2]], dtype=np.float32)
    return pts, labels, bbox


def make_leaflet_clustered(ns=80, seed=7, z0=20.0, ntypes=3, void_radius=100.0, cluster=300, area_per_lipid=64.0):
    """Hostile density: the left half of a square box at the nominal density, the right half at a quarter of it
    (every other row and column removed), a circular void of `void_radius` in the middle of the dense half and a
    micro-cluster of `cluster` points inside a 0.5 A disc in the sparse half.  Same return values as make_leaflet."""
    rng = np.random.default_rng(seed)
    h = float(np.sqrt(area_per_lipid))
    L = ns * h
    i, j = np.meshgrid(np.arange(ns), np.arange(ns), indexing="xy")
    i, j = i.reshape(-1), j.reshape(-1)
    keep = (i < ns // 2) | ((i % 2 == 0) & (j % 2 == 0))
    x = (i + 0.5 + rng.uniform(-0.3, 0.3, i.shape[0])) * h
    y = (j + 0.5 + rng.uniform(-0.3, 0.3, i.shape[0])) * h
    cx, cy = 0.25 * L, 0.5 * L
    keep &= (x - cx) ** 2 + (y - cy) ** 2 > void_radius ** 2
    x, y = x[keep], y[keep]
    if cluster > 0:
        r, t = 0.5 * np.sqrt(rng.uniform(0, 1, cluster)), rng.uniform(0, 2 * np.pi, cluster)
        x = np.concatenate([x, 0.75 * L + h / 3 + r * np.cos(t)])
        y = np.concatenate([y, 0.5 * L + h / 3 + r * np.sin(t)])
    z = z0 + 6.0 * np.sin(2 * np.pi * x / L + 0.3) + 4.0 * np.sin(2 * np.pi * (x + 2 * y) / L + 1.1)
    perm = rng.permutation(x.shape[0])              # no spatial order in the index either
    pts = np.stack([x, y, z], axis=1)[perm].astype(np.float32)
    labels = rng.integers(0, ntypes, pts.shape[0]).astype(np.int32)
    bbox = np.array([[0.0, 0.0, -L / 2], [L, L, L / 2]], dtype=np.float32)
    return pts, labels, bbox


def periodic_delaunay(xy, box0, box1, margin=None):
    """Host stand-in for TriMesh::periodicDelaunay (TriMesh_cgal.cpp:160-281).

    xy: (V,2) planar points inside [box0, box1).  Returns (dfaces (F,3) int64 original vertex
    ids, doffs (F,3,2) int8 periodic offsets); every face has >= 1 vertex with offset (0,0).
    Instead of the full 9-sheet covering only a `margin`-wide strip of images is triangulated,
    which yields the same faces around the original sheet when the strip is wider than the
    largest Delaunay circumcircle (default: 8 mean spacings, at least 5 % of the box).
    """
    from scipy.spatial import Delaunay
    xy = np.asarray(xy, dtype=np.float64)[:, :2]
    V = xy.shape[0]
    w = np.asarray(box1, np.float64)[:2] - np.asarray(box0, np.float64)[:2]
    if margin is None:
        margin = max(8.0 * np.sqrt(w[0] * w[1] / max(V, 1)), 0.05 * min(w))
    margin = min(margin, 0.999 * min(w))
    lo, hi = np.asarray(box0, np.float64)[:2], np.asarray(box1, np.float64)[:2]
    pts, orig, off = [xy], [np.arange(V)], [np.zeros((V, 2), np.int8)]
    for ox in (-1, 0, 1):
        for oy in (-1, 0, 1):
            if ox == 0 and oy == 0:
                continue
            q = xy + np.array([ox, oy]) * w
            sel = np.where((q[:, 0] >= lo[0] - margin) & (q[:, 0] < hi[0] + margin) &
                           (q[:, 1] >= lo[1] - margin) & (q[:, 1] < hi[1] + margin))[0]
            pts.append(q[sel]); orig.append(sel)
            off.append(np.tile(np.array([[ox, oy]], np.int8), (sel.shape[0], 1)))
    pts, orig, off = np.concatenate(pts), np.concatenate(orig), np.concatenate(off)
    tri = Delaunay(pts).simplices
    keep = (tri < V).any(axis=1)
    tri = tri[keep]
    # consistent counter-clockwise winding in the plane (CGAL faces are ccw)
    a, b, c = pts[tri[:, 0]], pts[tri[:, 1]], pts[tri[:, 2]]
    cw = ((b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0])) < 0
    tri[cw] = tri[cw][:, [0, 2, 1]]
    return orig[tri].astype(np.int64), off[tri].astype(np.int8)


def trim_periodic(verts, dfaces, doffs, bbox0, bbox1):
    """Vectorised restatement of TriMesh::trim_periodicDelaunay (TriMesh.cpp:493-575).

    verts (V,dim) f32; bbox0/bbox1: this mesh's own bbox (the reference re-trims each mesh with
    its own box, TriMesh.hpp:301-304).  Returns dict(faces, periodic_faces, trimmed_faces,
    duplicated_vidxs, duplicated_verts) with the dtypes of trimesh.py:421-436.
    """
    full = (doffs == 0).all(axis=2).all(axis=1)
    out = trim_partial(verts, dfaces[~full], doffs[~full], bbox0, bbox1)
    out["faces"] = dfaces[full].astype(np.uint32)
    return out


def trim_partial(verts, pf, po, bbox0, bbox1):
    """The part of trim_periodicDelaunay that deals with the faces crossing the box (pf (Fp,3) vertex ids,
    po (Fp,3,2) offsets, at least one corner outside the original sheet): duplicated vertices in
    first-appearance order of the face/corner scan and the faces re-indexed onto them."""
    verts = np.asarray(verts, dtype=np.float32)
    V, dim = verts.shape
    boxw = (np.asarray(bbox1, np.float32) - np.asarray(bbox0, np.float32)).astype(np.float32)
    pis = (po == 0).all(axis=2)                              # (Fp,3) corner in the original sheet
    # unique (orig, offx, offy) in first-appearance order of the face/corner scan
    key = (pf.astype(np.int64) * 9 + (po[..., 0].astype(np.int64) + 1) * 3 + (po[..., 1].astype(np.int64) + 1))
    flat_key = key.reshape(-1)
    flat_dup = ~pis.reshape(-1)
    dup_keys = flat_key[flat_dup]
    uniq, first = np.unique(dup_keys, return_index=True)
    order = np.argsort(first, kind="stable")
    uniq = uniq[order]
    rank = np.empty(uniq.shape[0], np.int64)
    sorter = np.argsort(uniq, kind="stable")
    rank[sorter] = np.arange(uniq.shape[0])
    pos = np.searchsorted(uniq[sorter], dup_keys)
    dup_id = sorter[pos]
    tf = pf.reshape(-1).astype(np.int64)
    tf[flat_dup] = V + dup_id
    dup_orig = (uniq // 9).astype(np.int64)
    dup_ox = ((uniq % 9) // 3 - 1).astype(np.float32)
    dup_oy = ((uniq % 3) - 1).astype(np.float32)
    dverts = verts[dup_orig].astype(np.float64)
    dverts[:, 0] += (dup_ox * boxw[0]).astype(np.float32)
    dverts[:, 1] += (dup_oy * boxw[1]).astype(np.float32)
    return dict(periodic_faces=pf.astype(np.uint32),
                trimmed_faces=tf.reshape(-1, 3).astype(np.uint32),
                duplicated_vidxs=dup_orig.astype(np.uint32),
                duplicated_verts=dverts.astype(np.float32))


def planar_delaunay(xy):
    """Non-periodic stand-in for TriMesh::delaunay (TriMesh_cgal.cpp:111-158)."""
    from scipy.spatial import Delaunay
    xy = np.asarray(xy, dtype=np.float64)[:, :2]
    tri = Delaunay(xy).simplices
    a, b, c = xy[tri[:, 0]], xy[tri[:, 1]], xy[tri[:, 2]]
    cw = ((b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0])) < 0
    tri[cw] = tri[cw][:, [0, 2, 1]]
    return tri.astype(np.uint32)


def read_off(path):
    """Minimal OFF reader (vertices, faces) for examples/data/noisy.off-style files."""
    with open(path) as fh:
        toks = fh.read().split()
    assert toks[0] == "OFF"
    nv, nf = int(toks[1]), int(toks[2])
    v = np.array(toks[4:4 + 3 * nv], dtype=np.float64).reshape(nv, 3)
    rest = toks[4 + 3 * nv:]
    f = np.array(rest[:4 * nf], dtype=np.int64).reshape(nf, 4)[:, 1:] if nf else np.zeros((0, 3), np.int64)
    return v, f
