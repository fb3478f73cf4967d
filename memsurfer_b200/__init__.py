"""memsurfer_b200 -- B200-native (sm_100a) replacement of MemSurfer's per-frame membrane-property
hot path behind the reference's own Python API (`Membrane`, `TriMesh`)."""
from ._lib import pinned
from .membrane import Membrane, point_normals_device
from .trimesh import TriMesh

__all__ = ["Membrane", "TriMesh", "point_normals_device", "pinned"]
__version__ = "0.2.0"
