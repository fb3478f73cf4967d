"""Export side of TriMesh: the data formats the reference writes after the hot path.

The reference goes through VTK (memsurfer/trimesh.py:441-567 `as_vtkpolydata`, :620-628 `write_vtp`) and pandas
(:570-617 `as_pddataframe`, :630-635 `write_pd`).  VTK is not in this image, so `PolyData` below is a plain
container with the same content a `vtkPolyData` would get (points, polys or verts, point/cell/field arrays chosen
by the reference's size rules) and `write_vtp` emits the VTK XML PolyData format itself (raw appended blocks or
ascii).  Host-side numpy only: nothing here is on the timed path.

The per-face Python loop of trimesh.py:484-499 (drop a face when one of its edges spans more than half the box
in x or y) is a vectorised comparison here; same float32 coordinates, same strict '>' test.
"""
import logging

import numpy as np

LOGGER = logging.getLogger(__name__)

_VTK_TYPES = {'float32': 'Float32', 'float64': 'Float64', 'int8': 'Int8', 'uint8': 'UInt8', 'int16': 'Int16',
              'uint16': 'UInt16', 'int32': 'Int32', 'uint32': 'UInt32', 'int64': 'Int64', 'uint64': 'UInt64'}


class PolyData(object):
    """What trimesh.py:450-567 puts into a vtkPolyData: `points` (n,3) float32; `polys` (m,3) int64 or None;
    `verts` (n,) int64 or None (one vtkVertex per point when the mesh has no faces); three name -> array dicts."""

    def __init__(self, points, polys=None, verts=None):
        self.points = points
        self.polys = polys
        self.verts = verts
        self.point_data = {}
        self.cell_data = {}
        self.field_data = {}

    def number_of_points(self):
        return self.points.shape[0]

    def number_of_cells(self):
        return self.polys.shape[0] if self.polys is not None else self.verts.shape[0]


def as_polydata(mesh, include_attributes, additional_attributes):
    """trimesh.py:441-567 without VTK."""
    include_periodic = mesh.periodic and (mesh._duplicated_vidxs is not None)
    if include_periodic:
        verts = np.concatenate((mesh.vertices, mesh._duplicated_verts), axis=0)
        faces = np.concatenate((mesh.faces, mesh._trimmed_faces), axis=0)
    else:
        verts, faces = mesh.vertices, mesh.faces

    points = np.zeros((verts.shape[0], 3), np.float32)       # 2-D vertices get z = 0 (trimesh.py:465-467)
    points[:, :verts.shape[1]] = verts

    if faces is not None:
        faces = np.asarray(faces).astype(np.int64).reshape(-1, 3)
        if include_periodic:                                  # trimesh.py:484-499
            xw = 0.5 * (mesh.bbox[1, 0] - mesh.bbox[0, 0])
            yw = 0.5 * (mesh.bbox[1, 1] - mesh.bbox[0, 1])
            p = points[faces].astype(np.float64)              # GetPoint returns doubles
            d = np.abs(p - np.roll(p, -1, axis=1))
            spans = ((d[:, :, 0] > xw) | (d[:, :, 1] > yw)).any(axis=1)
            faces = faces[~spans]
        poly = PolyData(points, polys=faces)
    else:
        poly = PolyData(points, verts=np.arange(mesh.vertices.shape[0], dtype=np.int64))

    def _add(_name, _data):
        if isinstance(_data, (int, float, str)):
            _data = np.array([_data])
        elif not isinstance(_data, np.ndarray):
            _data = np.array(_data)
        try:
            if include_periodic and _data.shape[0] == mesh.vertices.shape[0]:
                _data = np.concatenate((_data, _data[mesh._duplicated_vidxs]), axis=0)
            if _data.shape[0] == 1:
                poly.field_data[_name] = _data
            elif _data.shape[0] == poly.number_of_points():
                poly.point_data[_name] = _data
            elif _data.shape[0] == poly.number_of_cells():
                poly.cell_data[_name] = _data
            else:
                raise Exception('Failed to match size of data')
        except Exception as e:
            LOGGER.warning(f'Could not add "{_name}" with size {np.shape(_data)}. '
                           f'(nverts = {poly.number_of_points()}, faces = {poly.number_of_cells()}). Error = {e}')

    for att in include_attributes:
        if att in mesh.attributes.keys():
            _add(att, mesh.attributes[att])
        else:
            LOGGER.warning(f'Requested attribute "{att}" is not computed')
    for k, v in additional_attributes.items():
        _add(k, v)
    return poly


def as_pddataframe(mesh, include_attributes, additional_attributes):
    """trimesh.py:570-617: vertex positions + per-vertex attributes; duplicated vertices are not listed."""
    import pandas as pd
    df = pd.DataFrame()

    def _add(_name, _data):
        if not isinstance(_data, np.ndarray):
            _data = np.array([_data])
        if _data.ndim == 1:
            df[_name] = _data
        elif _data.ndim == 2 and _data.shape[1] == mesh.vertices.shape[1]:
            dims = ['x', 'y', 'z']
            for d in range(mesh.vertices.shape[1]):
                df[f'{_name}_{dims[d]}'] = _data[:, d]
        else:
            LOGGER.warning(f'Could not add "{_name}" with size {_data.shape}')

    _add('pos', np.asarray(mesh.vertices))
    for att in include_attributes:
        if att in mesh.attributes.keys():
            _add(att, mesh.attributes[att])
        else:
            LOGGER.warning(f'Requested attribute "{att}" is not computed')
    for k, v in additional_attributes.items():
        _add(k, v)
    return df


# ------------------------------------------------------------------------------------------ VTK XML PolyData
class _Blocks(object):
    """Raw appended section: every array is [uint64 byte count][bytes], addressed by its offset."""

    def __init__(self):
        self.chunks, self.size = [], 0

    def add(self, arr):
        off = self.size
        raw = np.ascontiguousarray(arr).tobytes()
        self.chunks += [np.uint64(len(raw)).tobytes(), raw]
        self.size += 8 + len(raw)
        return off


def _data_array(tag, name, arr, blocks, ascii):
    arr = np.asarray(arr)
    if arr.dtype.kind in 'US':          # vtkStringArray: character codes, each string 0-terminated
        codes = ' '.join(' '.join(str(b) for b in str(s).encode()) + ' 0' for s in arr.ravel())
        return f'<Array type="String" Name="{name}" NumberOfTuples="{arr.shape[0]}" format="ascii">{codes}</Array>\n'
    if arr.dtype.kind == 'b':
        arr = arr.astype(np.uint8)
    if arr.dtype.byteorder == '>':
        arr = arr.astype(arr.dtype.newbyteorder('<'))
    vt = _VTK_TYPES[arr.dtype.name]
    ncomp = 1 if arr.ndim == 1 else int(np.prod(arr.shape[1:]))
    head = f'<{tag} type="{vt}" Name="{name}" NumberOfComponents="{ncomp}"'
    if tag == 'Array':
        head += f' NumberOfTuples="{arr.shape[0]}"'
    if ascii:
        body = ' '.join(repr(x) for x in arr.ravel().tolist())
        return f'{head} format="ascii">{body}</{tag}>\n'
    return f'{head} format="appended" offset="{blocks.add(arr)}"/>\n'


def write_vtp(poly, filename, ascii=False):
    """VTK XML PolyData (version 1.0, little endian, UInt64 headers) of a `PolyData`.  `ascii=False` puts the arrays
    into one raw appended section (what ParaView/VTK read fastest), `ascii=True` writes them inline as text."""
    blocks = _Blocks()
    n = poly.number_of_points()
    npolys = poly.polys.shape[0] if poly.polys is not None else 0
    nverts = poly.verts.shape[0] if poly.verts is not None else 0
    x = ['<?xml version="1.0"?>\n',
         '<VTKFile type="PolyData" version="1.0" byte_order="LittleEndian" header_type="UInt64">\n<PolyData>\n']
    if poly.field_data:
        x.append('<FieldData>\n')
        x += [_data_array('Array', k, v, blocks, ascii) for k, v in poly.field_data.items()]
        x.append('</FieldData>\n')
    x.append(f'<Piece NumberOfPoints="{n}" NumberOfVerts="{nverts}" NumberOfLines="0" NumberOfStrips="0" '
             f'NumberOfPolys="{npolys}">\n')
    for sect, data in (('PointData', poly.point_data), ('CellData', poly.cell_data)):
        x.append(f'<{sect}>\n')
        for k, v in data.items():
            tag = 'Array' if np.asarray(v).dtype.kind in 'US' else 'DataArray'
            x.append(_data_array(tag, k, v, blocks, ascii))
        x.append(f'</{sect}>\n')
    x += ['<Points>\n', _data_array('DataArray', 'Points', poly.points, blocks, ascii), '</Points>\n']
    for sect, conn, offs in (('Verts', poly.verts, 1), ('Polys', poly.polys, 3)):
        x.append(f'<{sect}>\n')
        if conn is not None:
            conn = np.asarray(conn, np.int64)
            x.append(_data_array('DataArray', 'connectivity', conn.ravel(), blocks, ascii))
            x.append(_data_array('DataArray', 'offsets',
                                 np.arange(1, conn.shape[0] + 1, dtype=np.int64) * offs, blocks, ascii))
        x.append(f'</{sect}>\n')
    x.append('</Piece>\n</PolyData>\n')
    with open(filename, 'wb') as f:
        f.write(''.join(x).encode())
        if not ascii:
            f.write(b'<AppendedData encoding="raw">\n_')
            for c in blocks.chunks:
                f.write(c)
            f.write(b'\n</AppendedData>\n')
        f.write(b'</VTKFile>\n')


def read_vtp(filename):
    """Reader for the files `write_vtp` produces (both modes): {'Points', 'Polys', 'Verts', 'PointData',
    'CellData', 'FieldData'}.  For round-trip tests and for users without VTK; not a general VTK reader."""
    import re
    raw = open(filename, 'rb').read()
    cut = raw.find(b'<AppendedData')
    xml, app = (raw, b'') if cut < 0 else (raw[:cut], raw[raw.index(b'_', cut) + 1:])
    xml = xml.decode()
    inv = {v: k for k, v in _VTK_TYPES.items()}

    def arrays(section):
        m = re.search(rf'<{section}>(.*?)</{section}>', xml, re.S)
        out = {}
        if not m:
            return out
        for a in re.finditer(r'<(DataArray|Array)\s([^>]*?)(/>|>(.*?)</\1>)', m.group(1), re.S):
            at = dict(re.findall(r'(\w+)="([^"]*)"', a.group(2)))
            if at['type'] == 'String':
                codes = [int(c) for c in a.group(4).split()]
                strs, cur = [], []
                for c in codes:
                    if c == 0:
                        strs.append(bytes(cur).decode()); cur = []
                    else:
                        cur.append(c)
                out[at['Name']] = np.array(strs)
                continue
            dt = np.dtype(inv[at['type']])
            if at['format'] == 'appended':
                off = int(at['offset'])
                nb = int(np.frombuffer(app, np.uint64, 1, off)[0])
                v = np.frombuffer(app, dt, nb // dt.itemsize, off + 8).copy()
            else:
                v = np.array(a.group(4).split(), dtype=np.float64 if dt.kind == 'f' else np.int64).astype(dt) \
                    if dt.kind != 'u' else np.array([int(t) for t in a.group(4).split()], dtype=dt)
            nc = int(at.get('NumberOfComponents', 1))
            out[at['Name']] = v.reshape(-1, nc) if nc > 1 else v
        return out

    res = {'Points': arrays('Points')['Points'], 'PointData': arrays('PointData'), 'CellData': arrays('CellData'),
           'FieldData': arrays('FieldData'), 'Polys': None, 'Verts': None}
    p = arrays('Polys')
    if p:
        res['Polys'] = p['connectivity'].reshape(-1, 3)
    v = arrays('Verts')
    if v:
        res['Verts'] = v['connectivity']
    return res
