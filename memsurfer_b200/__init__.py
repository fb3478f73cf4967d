"""memsurfer_b200 -- B200-native (sm_100a) replacement of MemSurfer's per-frame membrane-property
hot path behind the reference's own Python API (`Membrane`, `TriMesh`)."""
from .membrane import Membrane, point_normals_device
from .trimesh import TriMesh

__all__ = ["Membrane", "TriMesh", "point_normals_device"]
__version__ = "0.1.0"
