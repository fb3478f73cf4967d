"""memsurfer_b200.utils -- the O(N^2) guard of the reference flow, grid-based on the GPU.

`test_overlapping_points` mirrors memsurfer/utils.py:91-159 (same arguments, return value, warning text
and in-place `fix`): the reference builds scipy's condensed all-pairs distance vector (N(N-1)/2 doubles:
a 1M-lipid leaflet would need 1 TB), here the pairs closer than the threshold come from the kNN cell list
in HBM (include/memsurfer_b200.h: ms_overlapping_pairs).  No CPU fallback.
"""
import ctypes as C
import logging

import numpy as np
import torch

from . import _lib

LOGGER = logging.getLogger(__name__)
MAX_REPORTED_PAIRS = 1 << 20


def overlapping_pairs_device(points_dev, dthreshold):
    """(pairs (k,2) int32 sorted like the reference's condensed order, min distance above the threshold)."""
    _lib.require_cuda()
    L = _lib.lib()
    dev = points_dev.device
    n = int(points_dev.shape[0])
    pairs = torch.empty((MAX_REPORTED_PAIRS, 2), dtype=torch.int32, device=dev)
    ws = _lib.workspace(L.ms_overlap_workspace_bytes(n), dev)
    npairs, dmin = C.c_int64(0), C.c_double(0.0)
    _lib.check(L.ms_overlapping_pairs(points_dev.data_ptr(), n, float(dthreshold), pairs.data_ptr(), MAX_REPORTED_PAIRS,
                                      C.byref(npairs), C.byref(dmin), ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)),
               'ms_overlapping_pairs')
    k = int(npairs.value)
    if k > MAX_REPORTED_PAIRS:
        raise ValueError(f'{k} overlapping point pairs (more than {MAX_REPORTED_PAIRS}): the input is degenerate')
    if k == 0:
        return np.zeros((0, 2), np.int32), float(dmin.value)
    p = pairs[:k].cpu().numpy()
    p = p[np.lexsort((p[:, 1], p[:, 0]))]          # condensed pdist order: by row, then column
    return p, float(dmin.value)


class OverlapCheck(object):
    """The detection queued on the current stream without any read-back; `report()` fetches the verdict (and
    warns like the reference) once the caller synchronises anyway."""

    def __init__(self, points_dev, box4, dthreshold=0.00001, max_pairs=4096):
        _lib.require_cuda()
        L = _lib.lib()
        dev = points_dev.device
        n = int(points_dev.shape[0])
        self.pairs = torch.empty((max_pairs, 2), dtype=torch.int32, device=dev)
        self.result = torch.empty(2, dtype=torch.int64, device=dev)
        self.ws = _lib.workspace(L.ms_overlap_workspace_bytes(n), dev)
        self.max_pairs, self.done = max_pairs, False
        b = (C.c_float * 4)(*[float(x) for x in box4])
        _lib.check(L.ms_overlapping_pairs_async(points_dev.data_ptr(), n, float(dthreshold), b, self.pairs.data_ptr(), max_pairs,
                                                self.result.data_ptr(), self.ws.data_ptr(), self.ws.numel(), _lib.stream_ptr(dev)),
                   'ms_overlapping_pairs_async')
        self.event = torch.cuda.Event()           # report() may run on another stream than the one this was queued on
        self.event.record(torch.cuda.current_stream(dev))

    def report(self):
        """True if no two points overlap; else warn (memsurfer/utils.py:121) and return False."""
        if self.done:
            return self.ok
        self.event.synchronize()
        nzero = int(self.result[0].item())
        self.done, self.ok = True, nzero == 0
        if nzero:
            p = self.pairs[:min(nzero, self.max_pairs)].cpu().numpy()
            p = p[np.lexsort((p[:, 1], p[:, 0]))]
            LOGGER.warning(f'Found {nzero} overlapping points: {[(int(x), int(y)) for x, y in p]}')
        self.ws = self.pairs = None
        return self.ok


def test_overlapping_points(points, dthreshold=0.00001, fix=False, device='cuda:0', points_dev=None):
    """True if no two points are closer than dthreshold; else warn and return False, or (fix=True) move one point of
    every coincident pair in place exactly as memsurfer/utils.py:118-159 does.  `points_dev`: the same points
    already in HBM as an (n,3) float32 tensor (saves the upload)."""
    points = np.asarray(points) if not isinstance(points, np.ndarray) else points
    LOGGER.debug(f'Testing for overlapping points ({points.shape}, threshold={dthreshold}, fix={fix})')
    if points_dev is None:
        p3 = np.zeros((points.shape[0], 3), np.float32)
        p3[:, :points.shape[1]] = points
        points_dev = torch.from_numpy(p3).to(device)
    is_zero, dmin = overlapping_pairs_device(points_dev, dthreshold)
    nzero = is_zero.shape[0]
    if nzero == 0:
        return True
    is_zero = [(int(x), int(y)) for x, y in is_zero]
    LOGGER.warning(f'Found {nzero} overlapping points: {is_zero}')
    if not fix:
        return False

    # now, try to fix them: we want to move the fewest points possible (utils.py:124-159)
    dupls, dupls_cnts = np.unique([x for sublist in is_zero for x in sublist], return_counts=True)
    dupls = {dupls[i]: dupls_cnts[i] for i in range(len(dupls))}
    if not np.isfinite(dmin):
        raise ValueError('cannot separate overlapping points: no distinct neighbour found')
    max_move = 0.5 * dmin
    for x, y in is_zero:
        dist0 = np.linalg.norm(points[x] - points[y])
        if not np.isclose(dist0, 0.):      # we have moved either of these points before
            continue
        if dupls[x] == -1 or dupls[y] == -1:
            continue
        a, b = (x, y) if dupls[x] > dupls[y] else (y, x)      # move the one that impacts more points
        pa = f'{points[a]}'
        rad = np.random.rand() * max_move
        theta = np.random.rand() * np.pi
        points[a, 0] += rad * np.cos(theta)
        points[a, 1] += rad * np.sin(theta)
        dist1 = np.linalg.norm(points[x] - points[y])
        LOGGER.debug(f'Moved point {a}: {pa} --> {points[a]} : dist = {dist0} --> {dist1}')
        dupls[a] = -1          # mark this as moved


test_overlapping_points.__test__ = False      # not a pytest test


# ------------------------------------------------------------------------------ small host helpers of memsurfer/utils.py
def get_order_param(a2, a3, b2, b3, z=np.array([0, 0, 1])):
    """Lipid tail order parameter P2 = (3 cos^2(theta) - 1) / 2 per tail, theta between the bond a2-a3 (b2-b3)
    and the normal z (one normal or one per lipid); (n, 2) array, one column per tail (memsurfer/utils.py:163-204).
    O(n) host arithmetic on caller-supplied bead positions; not part of the timed path."""
    z = np.asarray(z)
    if z.ndim == 1:
        z = z[None, :]
    if not (a2.shape == a3.shape == b2.shape == b3.shape):
        raise AssertionError("Tail's shapes must match")
    if not (z.shape == (1, 3) or z.shape == a2.shape):
        raise AssertionError('Should provide a single normal or a normal per lipid')
    zn = np.linalg.norm(z, axis=1)

    def _p2(bond):
        c = (bond * z).sum(axis=1) / (np.linalg.norm(bond, axis=1) * zn)
        return 0.5 * (3.0 * c * c - 1)

    return np.stack((_p2(a2 - a3), _p2(b2 - b3)), axis=1)


class Timer(object):
    """Wall-clock timer with the reference's text form (memsurfer/utils.py:209-241)."""

    def __init__(self):
        self.stime = self.etime = None
        self.start()

    def start(self):
        import timeit
        self.stime = timeit.default_timer()

    def end(self, print_time=False):
        import timeit
        self.etime = timeit.default_timer()
        if print_time:
            print(self)

    def __str__(self):
        t = self.etime - self.stime
        if t < 1e-6:
            return '%.3f micro-sec.' % (t * 1e6)
        if t < 1e-3:
            return '%.3f milli-sec.' % (t * 1e3)
        if t < 60.0:
            return '%.3f sec.' % t
        m = int(t / 60.0)
        return '%d min. %.3f sec.' % (m, t - 60 * m)

    __repr__ = __str__


LFORMAT = '%(asctime)s [%(levelname)s] %(name)s:%(funcName)s - %(message)s'
_LEVELS = {1: logging.DEBUG, 2: logging.INFO, 3: logging.WARNING, 4: logging.ERROR, 5: logging.CRITICAL}


def create_logger(loglevel, logstdout, logfile, logpath, logfname):
    """Configure the root logger like memsurfer/utils.py:35-78: `loglevel` 1 (debug) ... 5 (critical), a stream
    handler when `logstdout == 1`, a file handler on `<logpath>/<logfname>[.log]` when `logfile == 1`."""
    import os
    if loglevel not in _LEVELS:
        raise AssertionError(f'loglevel must be 1..5, got {loglevel}')
    level = _LEVELS[loglevel]
    root = logging.getLogger()
    root.setLevel(level)
    handlers = []
    if logstdout == 1:
        handlers.append(logging.StreamHandler())
    if logfile == 1:
        logpath = os.path.expanduser(logpath)
        os.makedirs(logpath, exist_ok=True)
        if logfname[-4:] != '.log':
            logfname += '.log'
        handlers.append(logging.FileHandler(os.path.join(logpath, logfname)))
    for h in handlers:
        h.setLevel(level)
        h.setFormatter(logging.Formatter(LFORMAT))
        root.addHandler(h)
    return LOGGER
