"""Vertex neighbours and shells: the oracle against the compiled reference's
TriMesh::need_neighbors and against the Python loop of trimesh.py:392-404 on CPU; the CUDA path against the
oracle on the GPU (integer work: exact equality, list order included)."""
import numpy as np
import pytest


def _python_shells(counts_and_lists, nverts, ref_pts):
    """trimesh.py:378-404 restated verbatim in structure (list slicing + FIFO pop(0)); small meshes only."""
    ret_val = np.asarray(counts_and_lists).astype(int)
    nnbrs = ret_val[:nverts]
    nbrs, nidx = [], nverts
    for vidx in range(nverts):
        nbrs.append(ret_val[nidx:nidx + nnbrs[vidx]])
        nidx += nnbrs[vidx]
    shells = -1 * np.ones(nverts, dtype=int)
    for p in ref_pts:
        shells[p] = 0
    todo = list(ref_pts)
    while len(todo) > 0:
        cur = todo.pop(0)
        for nbr in nbrs[cur]:
            if shells[nbr] == -1:
                shells[nbr] = 1 + shells[cur]
                todo.append(nbr)
    return shells


def _mesh(n, seed):
    from memsurfer_b200 import synth
    pts, _, _ = synth.make_leaflet(n, seed=seed)
    return pts, synth.planar_delaunay(pts[:, :2]).astype(np.uint32)


def test_oracle_neighbors_anchor(oracle):
    # two triangles sharing the edge (1,2): the lists follow the face loop, first appearance wins
    faces = np.array([[0, 1, 2], [2, 1, 3]], np.uint32)
    off, lists = oracle.need_neighbors(faces, 5)
    assert off.tolist() == [0, 2, 5, 8, 10, 10]
    assert lists.tolist() == [1, 2, 2, 0, 3, 0, 1, 3, 2, 1]
    sh = oracle.shells(off, lists, [0])
    assert sh.tolist() == [0, 1, 1, 2, -1]                 # vertex 4 is isolated
    assert oracle.shells(off, lists, [3, 0, 3]).tolist() == [0, 1, 1, 0, -1]


def test_oracle_neighbors_vs_reference(oracle):
    if oracle.ref() is None or not hasattr(oracle.ref(), "msref_need_neighbors"):
        pytest.skip("compiled reference not built")
    pts, faces = _mesh(900, 5)
    rng = np.random.default_rng(0)
    faces = faces[rng.permutation(faces.shape[0])]          # the list order depends on the face order
    off, lists = oracle.need_neighbors(faces, pts.shape[0])
    lin = oracle.ref_need_neighbors(pts, faces)
    assert np.array_equal(lin[:pts.shape[0]], np.diff(off))
    assert np.array_equal(lin[pts.shape[0]:], lists)
    ref_pts = rng.choice(pts.shape[0], 7, replace=False)
    assert np.array_equal(oracle.shells(off, lists, ref_pts), _python_shells(lin, pts.shape[0], list(ref_pts)))


@pytest.mark.gpu
def test_neighbors_and_shells_gpu_vs_oracle(oracle):
    from memsurfer_b200 import TriMesh
    pts, faces = _mesh(10000, 9)
    rng = np.random.default_rng(1)
    faces = faces[rng.permutation(faces.shape[0])]
    V = pts.shape[0]
    mesh = TriMesh(pts, faces=faces)
    off, lists = oracle.need_neighbors(faces, V)
    lin = mesh.need_neighbors()
    assert lin.dtype == np.int32
    assert np.array_equal(lin[:V], np.diff(off)) and np.array_equal(lin[V:], lists)
    ref_pts = rng.choice(V, 25, replace=False)
    sh = mesh.compute_shells(list(ref_pts))
    assert sh.dtype == np.dtype(int) and sh.shape == (V,)
    assert np.array_equal(sh, oracle.shells(off, lists, ref_pts))
    assert mesh.compute_shells([0]) is sh                    # memoised under 'shell' like the reference
    # a vertex no face touches cannot be reached: the reference asserts
    pts2 = np.concatenate([pts, [[1e4, 1e4, 0.0]]]).astype(np.float32)
    with pytest.raises(AssertionError):
        TriMesh(pts2, faces=faces).compute_shells([0])


@pytest.mark.gpu
def test_shells_full_size_properties():
    """499 849-vertex leaflet mesh: shells are a graph distance (seeds 0, every other vertex has a neighbour one
    shell closer and none more than one shell away)."""
    import torch
    import bench
    lf = bench.make_frame_inputs(1000000, 0)[0]
    mesh = bench.new_mesh(lf, torch.device("cuda:0"))
    V = mesh.vertices.shape[0]
    ref_pts = np.flatnonzero(lf["labels"] == 2)[::997]
    sh = mesh.compute_shells(ref_pts)
    off, nbr = (t.cpu().numpy() for t in mesh.neighbors_device())
    assert sh.min() == 0 and np.array_equal(np.flatnonzero(sh == 0), np.sort(ref_pts))
    deg = np.diff(off)
    src = np.repeat(np.arange(V), deg)
    assert np.abs(sh[src] - sh[nbr]).max() <= 1
    closer = np.full(V, np.iinfo(np.int64).max)
    np.minimum.at(closer, src, sh[nbr])
    rest = sh > 0
    assert np.array_equal(closer[rest], sh[rest] - 1)
