#!/usr/bin/env python
"""bench.py -- membrane frames/sec of the per-frame hot path (normals + curvature + KDE).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One step = one frame: for each of the 2 leaflets of a synthetic periodic
undulating bilayer (BASELINE.json configs[2]: 1M lipids, ~500k points per leaflet)
compute_pnormals(k=18), compute_properties (vertex normals, point areas, mean + Gaussian
curvature) and compute_densities(types=[2], sigmas=[10,20,30,40], get_nlipdis=True) for 'all' and
each of the 3 lipid types.  The host mesh producer (Delaunay stand-in) runs once in set-up and is
not part of the timed region (it stays on the host by design).

 * `value`: frames/s with the inputs resident in HBM (C-ABI calls on device pointers).
 * `e2e`:   the same frame through the drop-in Membrane/TriMesh API from HOST arrays: host->device
            copies of points/mesh/ids and device->host copies of every result inside the timing.
 * `roofline`: the dominant kernel (KDE pair-sum) against the fp64-FMA peak measured in the same
            process by a DFMA probe; `roofline_geometry`: the fused mesh-properties kernel against
            the measured HBM copy bandwidth (MEASURED_PEAKS.json).
 * `cpu_baseline`: the reference's own C++ (oracle/_ref, single-threaded as shipped) for vertex
            normals, point areas and KDE, the fp64 restatement for the CGAL/VTK parts, timed on
            this box's host cores on a bounded sample (see `sample`).
N > 1: every rank runs the same pipeline on its own frame (weak scaling, frames sharded across
GPUs); the only collective is the all_gather of the per-frame summary records.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SIGMAS = [10.0, 20.0, 30.0, 40.0]
FRACTIONS = [0.4, 0.3, 0.3]
METRIC = "membrane_frames_per_sec"
# fp64 FLOP per evaluated 2-D periodic pair / algorithmic bytes per vertex
FLOP_PER_PAIR_2D = 32.0
FP64_INSTR_PER_PAIR = {"kde_pairs": 14, "kde_cells": 10, "kde_groups": 10}
GEOM_BYTES_PER_VERTEX = 84.0


def _load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except Exception:
        return {}


TRAFFIC = _load_traffic()


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--lipids", type=int, default=1000000, help="lipids per frame (both leaflets)")
    ap.add_argument("--kde-cutoff", type=float, default=7.5,
                    help="KDE cut-off in sigmas (the drop-in's default 7.5: dropped tail <= exp(-28.1) = 6.1e-13 relative; "
                         "0 = all pairs, the reference's literal sum)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------- inputs
def make_frame_inputs(lipids, frame):
    """Two leaflets of frame `frame` + their host-produced periodic meshes."""
    from memsurfer_b200 import synth, TriMesh
    leaflets = []
    for li, z0 in enumerate((20.0, -20.0)):
        pts, labels, bbox = synth.make_leaflet(lipids // 2, seed=1234 + li, z0=z0, ntypes=3, fractions=FRACTIONS,
                                               phase_shift=0.05 * frame)
        dfaces, doffs = synth.periodic_delaunay(pts[:, :2], bbox[0], bbox[1])
        tri = synth.trim_periodic(pts, dfaces, doffs, pts.min(axis=0), pts.max(axis=0))
        leaflets.append(dict(points=pts, labels=labels, bbox=bbox, tri=tri, lo=pts.min(axis=0), hi=pts.max(axis=0)))
    return leaflets


def new_mesh(lf, device):
    from memsurfer_b200 import TriMesh
    t = lf["tri"]
    mesh = TriMesh(lf["points"], periodic=True, label="memb_smooth", device=device)
    mesh.set_bbox(lf["lo"], lf["hi"])          # the producer's bbox (spoints min/max, membrane.py:214-217)
    mesh.set_periodic_triangulation(t["faces"], t["periodic_faces"], t["trimmed_faces"],
                                    t["duplicated_vidxs"], t["duplicated_verts"])
    return mesh


# ---------------------------------------------------------------------------------------------- clocks
class ClockSampler(object):
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------- CPU baseline
def cpu_reference_baseline(lf, kde_sample_verts=16384, all_threads=False):
    """Reference CPU path on a bounded sample of the frame workload -> (frames/s, details)."""
    from oracle import oracle as O
    import ctypes
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        ncores = os.cpu_count() or 1
        gomp.omp_set_num_threads(ncores if all_threads else 1)
    except OSError:
        ncores = 1
    pts, bbox, tri, labels = lf["points"], lf["bbox"], lf["tri"], lf["labels"]
    V = pts.shape[0]
    verts = np.ascontiguousarray(np.concatenate([pts, tri["duplicated_verts"]]), np.float32)
    faces = np.ascontiguousarray(np.concatenate([tri["faces"], tri["trimmed_faces"]]), np.uint32)
    R = O.ref()
    t = {}
    t0 = time.perf_counter(); O.point_normals(pts, bbox, 0.2, 18, +1); t["pnormals_port"] = time.perf_counter() - t0
    if R is not None:
        t0 = time.perf_counter(); O.ref_need_normals(verts, faces); O.ref_need_pointareas(verts, faces)
        t["normals_areas_reference"] = time.perf_counter() - t0
    else:
        t0 = time.perf_counter(); O.mesh_normals(verts, faces); O.pointareas(verts, faces)
        t["normals_areas_port"] = time.perf_counter() - t0
    bb = np.concatenate([pts.min(0), pts.max(0)]).astype(np.float32)
    xw, yw = float(np.float32(0.5) * (bb[3] - bb[0])), float(np.float32(0.5) * (bb[4] - bb[1]))
    t0 = time.perf_counter(); O.curvatures(verts, faces, (xw, yw)); t["curvatures_port"] = time.perf_counter() - t0
    # KDE: the reference evaluates all V x n_ids pairs; sample = a compact patch of the same leaflet
    ns = int(min(kde_sample_verts, V))
    side = int(round(np.sqrt(V)))
    s_side = int(np.sqrt(ns))
    idx = (np.arange(s_side)[:, None] * side + np.arange(s_side)[None, :]).reshape(-1)
    sub = np.ascontiguousarray(pts[idx]); ns = sub.shape[0]
    sbb = np.concatenate([sub.min(0), sub.max(0)]).astype(np.float32)
    t0 = time.perf_counter()
    if R is not None:
        O.ref_kde(sub, np.zeros(0, np.int32), 2, 20.0, True, sbb)
        kind = "reference"
    else:
        O.kde(sub, np.zeros(0, np.int32), 2, 20.0, True, sbb)
        kind = "port"
    t["kde_sample"] = time.perf_counter() - t0
    pairs_per_s = ns * ns / t["kde_sample"]
    nl = [int((labels == l).sum()) for l in np.unique(labels)]
    pairs_leaflet = len(SIGMAS) * (V * V + sum(V * n for n in nl))
    t_leaflet = t["pnormals_port"] + t.get("normals_areas_reference", t.get("normals_areas_port", 0.0)) \
        + t["curvatures_port"] + pairs_leaflet / pairs_per_s
    fps = 1.0 / (2.0 * t_leaflet)
    detail = {"seconds": {k: round(v, 4) for k, v in t.items()}, "kde_pairs_per_s": pairs_per_s,
              "kde_pairs_per_leaflet": pairs_leaflet, "extrapolated_frame_seconds": 2.0 * t_leaflet}
    sample = (f"one leaflet (V={V}): point normals [fp64 port of CGAL pca+mst], vertex normals + point areas "
              f"[{'reference C++' if R is not None else 'port'}], curvatures [fp64 port of vtkCurvatures] at full size; "
              f"KDE [{kind}] on a {ns}-vertex patch, type 2, sigma 20, all pairs ({ns * ns:.3g} pairs), extrapolated "
              f"linearly in pairs to the frame's {2 * pairs_leaflet:.3g} pairs; frame = 2 leaflets")
    return fps, kind, (ncores if all_threads else 1), sample, detail


# ---------------------------------------------------------------------------------------------- reference arm
def run_reference_arm(args, rank):
    if rank != 0:
        return
    leaflets = make_frame_inputs(args.lipids, 0)[:1]
    vals = []
    for i in range(args.warmup + args.steps):
        # each step = one bounded sample (smaller KDE patch during warm-up)
        fps, kind, cores, sample, detail = cpu_reference_baseline(leaflets[0], 4096 if i < args.warmup else 12288)
        if i >= args.warmup:
            vals.append(fps)
    v = float(np.mean(vals))
    V = leaflets[0]["points"].shape[0]
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (reference) / f64 (port)", "data": "synthetic",
            "config": workload_config(args, V),
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample,
                             "detail": detail},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "vertices_per_sec": v * 2 * V}
    print(json.dumps(line), flush=True)


def workload_config(args, V):
    return {"workload": f"configs[2]: synthetic periodic undulating bilayer, {2 * V} lipids (2 leaflets x {V} points), "
                        f"1 frame/step: pnormals(k=18) + vertex normals/areas/mean+Gauss curvature + KDE type 2, "
                        f"sigmas {SIGMAS}, labels all + {len(FRACTIONS)} types, get_nlipdis",
            "kde": "all pairs" if args.kde_cutoff <= 0 else f"cut-off {args.kde_cutoff} sigma (cell list)",
            "l2": "flushed between steps (512 MiB device write)",
            "parallelism": f"frames sharded over {args.gpus} GPU(s), 1 frame per rank per step",
            "mesh_producer": "host (scipy Delaunay stand-in for CGAL), untimed"}


# ---------------------------------------------------------------------------------------------- our arm
def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    from memsurfer_b200 import _lib, Membrane, TriMesh
    from memsurfer_b200.frame import LeafletFrame, gather_summaries
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    device = torch.device(f"cuda:{local_rank}")
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()
    Membrane.KDE_CUTOFF_SIGMAS = args.kde_cutoff

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # frame `rank` of the trajectory: every rank owns one frame per step (weak scaling)
    leaflets = make_frame_inputs(args.lipids, rank)
    V = leaflets[0]["points"].shape[0]
    frames = [LeafletFrame(lf["points"], lf["labels"], lf["bbox"], new_mesh(lf, device), device=device) for lf in leaflets]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)

    def step_device():
        recs = []
        for fr in frames:
            fr.run(SIGMAS, 18, args.kde_cutoff)
            recs.append(fr.summary())
        rec = torch.stack(recs)
        return gather_summaries(rec) if world > 1 else rec.unsqueeze(0)

    def timed(fn, nwarm, nsteps, profile=False):
        for _ in range(nwarm):
            fn()
        barrier()
        if profile:
            L.ms_profile_enable(1)
        L.ms_launch_count(1)
        total_ms = 0.0
        for _ in range(nsteps):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            total_ms += e0.elapsed_time(e1)
        launches = int(L.ms_launch_count(1))
        prof = _lib.profile_report() if profile else {}
        L.ms_profile_enable(0)
        barrier()
        t = torch.tensor([total_ms], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, prof

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    # nvidia-smi takes a few hundred ms to come up (NVML init stalls kernel submission while it does): keep
    # warming up until its first sample has arrived, so that the timed steps run next to a settled sampler
    # (step_device ends in a collective when world > 1: rank 0 decides, every rank follows)
    t_wait = time.perf_counter()
    while True:
        go = torch.tensor([int(rank == 0 and clocks.proc is not None and not clocks.lines
                               and time.perf_counter() - t_wait < 3.0)], dtype=torch.int32, device=device)
        if world > 1:
            dist.broadcast(go, 0)
        if int(go.item()) == 0:
            break
        step_device()
        torch.cuda.synchronize()
    total_ms, launches, _ = timed(step_device, args.warmup, args.steps, profile=False)
    clk = clocks.stop() if rank == 0 else None
    ms_per_step = total_ms / args.steps
    value = world * 1e3 / ms_per_step
    # second, shorter pass with an event around every kernel launch: the per-kernel times behind `roofline`
    psteps = max(1, min(args.steps, 5))
    ptotal_ms, _, prof = timed(step_device, 1, psteps, profile=True)

    # ---- end to end through the drop-in API, from host arrays
    e2e = None
    if not args.no_e2e:
        bytes_io = {"h2d": 0, "d2h": 0}

        def step_e2e():
            # one frame in the order of the reference's own driver (examples/ex_3lipid.py:165-175): build BOTH
            # membranes from host arrays (Membrane.compute: point normals + surface; the surface is the host
            # producer's output, injected), then properties and compute_densities([top, bottom], ...) for 'all'
            # and every lipid type; finally EVERY result is read back as numpy (downloads inside the timing)
            h2d = d2h = 0
            ms = []
            for lf in leaflets:
                m = Membrane(lf["points"], labels=lf["labels"], periodic=True, bbox=lf["bbox"], device=device)
                pn = m.compute_pnormals(18)
                h2d += m._pset_nbytes; d2h += pn.nbytes
                m.memb_smooth = new_mesh(lf, device)
                ms.append(m)
            for m, lf in zip(ms, leaflets):
                m.compute_properties('smooth')
                h2d += sum(a.nbytes for a in m.memb_smooth._dev['_host'])
                h2d += 4 * lf["labels"].shape[0]                # the per-vertex label index
            for l in ['all'] + list(range(len(FRACTIONS))):
                Membrane.compute_densities(ms, [2], SIGMAS, True, l)
            recs = []
            for m in ms:
                mesh = m.memb_smooth
                d2h += sum(mesh.attributes[k].nbytes for k in ('pnormal', 'parea', 'mean_curv', 'gaus_curv'))
                d2h += sum(v.nbytes for v in m.attributes.values())
                recs.append([float(mesh.attributes['parea'].sum()), float(mesh.attributes['mean_curv'].mean())])
            bytes_io["h2d"], bytes_io["d2h"] = h2d, d2h
            return recs

        esteps = max(1, min(args.steps, 10))
        e2e_ms, _, _ = timed(step_e2e, 2, esteps)
        e2e_ms /= esteps
        e2e = {"value": world * 1e3 / e2e_ms, "unit": "frames/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": bytes_io["h2d"], "d2h_bytes_per_step": bytes_io["d2h"],
               "api": "memsurfer_b200.Membrane / TriMesh (host numpy in, numpy out)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rooflines (rank 0): DFMA probe for the fp64 peak, measured HBM peak from MEASURED_PEAKS.json
    import ctypes
    probe = torch.empty(148 * 8 * 256, dtype=torch.float64, device=device)
    flop = ctypes.c_double(0)
    st = _lib.stream_ptr(device)
    best = 0.0
    for it in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); _lib.check(L.ms_dfma_probe(probe.data_ptr(), 200000, ctypes.byref(flop), st), "probe"); e1.record()
        torch.cuda.synchronize()
        best = max(best, flop.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    pairs_allpairs_equiv = sum(fr.pairs_per_run(SIGMAS) for fr in frames)
    pairs_step = sum(fr.evaluated_pairs(SIGMAS, cutoff_sigmas=args.kde_cutoff) for fr in frames)
    kname = max(("kde_groups", "kde_cells", "kde_pairs"), key=lambda k: prof.get(k, {"ms": 0.0})["ms"])
    kp = prof.get(kname, {"launches": 0, "ms": 0.0})
    roof = None
    if kp["launches"]:
        # fp64-pipe instructions per pair, counted in the SASS of the 2-D kernels (cuobjdump):
        # k_kde_pairs periodic 7 DADD + 4 DFMA + 2 DSETP + 1 DMUL = 14; k_kde_cells / k_kde_groups 5 DADD + 4 DFMA + 1 DMUL = 10
        instr = FP64_INSTR_PER_PAIR[kname]
        pairs_per_launch = pairs_step * psteps / kp["launches"]
        avg_ms = kp["ms"] / kp["launches"]
        pps = pairs_per_launch / (avg_ms * 1e-3)
        ach = pps * instr * 2.0 / 1e12
        roof = {"kernel": "k_" + kname, "bound": "fp64_fma", "achieved": ach, "peak": best, "unit": "TFLOP/s",
                "frac": ach / best if best > 0 else None, "traffic": TRAFFIC.get(kname),
                "definition": "evaluated pairs/s x fp64-pipe instructions per pair x 2 (every DADD/DMUL/DSETP/DFMA holds one "
                              "FMA slot of the fp64 pipe) / DFMA-probe peak",
                "peak_source": "ms_dfma_probe in this process (no fp64 figure in MEASURED_PEAKS.json)",
                "fp64_instr_per_pair": instr, "pairs_per_launch": pairs_per_launch, "avg_launch_ms": avg_ms,
                "launches": kp["launches"], "evaluated_pairs_per_s": pps,
                "allpairs_equivalent_pairs_per_s": pairs_allpairs_equiv * psteps / (kp["ms"] * 1e-3),
                "tflops_at_32_flop_per_pair_convention": pps * FLOP_PER_PAIR_2D / 1e12,
                "share_of_step": kp["ms"] / ptotal_ms}
    mesh_kernels = ("face_records", "edge_mean", "vertex_gather")
    mp_ms = sum(prof.get(k, {"ms": 0.0})["ms"] for k in mesh_kernels)
    mp_n = prof.get("vertex_gather", {"launches": 0})["launches"]
    roof_g = None
    if mp_n:
        avg_ms = mp_ms / mp_n
        ach = V * GEOM_BYTES_PER_VERTEX / (avg_ms * 1e-3) / 1e9
        roof_g = {"kernel": "ms_mesh_properties = k_face_records + k_edge_mean + k_vertex_gather", "bound": "hbm",
                  "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                  "traffic": TRAFFIC.get("mesh_properties"), "peak_source": hbm_src,
                  "bytes_per_vertex": GEOM_BYTES_PER_VERTEX, "avg_ms_per_mesh": avg_ms, "meshes": mp_n,
                  "share_of_step": mp_ms / ptotal_ms,
                  "note": "algorithmic bytes per SURVEY 8(d); the three passes are bound by fp64 sqrt/div/atan2 "
                          "latency and L2 gathers, not by HBM (DESIGN.md 4.2)"}
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, V),
            "vertices_per_sec": value * 2 * V, "gpu_launches": launches, "clocks": clk, "e2e": e2e,
            "roofline": roof, "roofline_geometry": roof_g,
            "kernel_ms_per_step": {k: round(v["ms"] / psteps, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}}
    if not args.no_e2e:
        # the mesh producer is NOT part of a frame; reported beside it: periodic Delaunay + trim of
        # one leaflet's planar points on the device, host arrays in, host arrays out
        from memsurfer_b200 import TriMesh
        lf = leaflets[0]
        try:
            t_prod = []
            for _ in range(4):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                pm = TriMesh(lf["points"][:, :2], periodic=True, device=device)
                pm.set_bbox(lf["bbox"][0, :2], lf["bbox"][1, :2])
                pm.compute_delaunay()
                t_prod.append((time.perf_counter() - t0) * 1e3)
            line["mesh_producer"] = {"ms_per_leaflet": min(t_prod[1:]), "faces": int(pm.faces.shape[0] + pm._trimmed_faces.shape[0]),
                                     "what": "TriMesh.compute_delaunay: ms_ghosts + ms_knn + ms_delaunay_periodic + ms_delaunay_split, "
                                             "duplicated-vertex bookkeeping on the host; not included in value / e2e"}
        except Exception as e:            # a side measurement must never cost the bench line
            line["mesh_producer"] = {"error": repr(e)}
    if world == 1 and not args.no_cpu_baseline:
        fps, kind, cores, sample, detail = cpu_reference_baseline(leaflets[0])
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample,
                                "detail": detail}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
