"""blitzdg_b200: B200-native sw2d nodal-DG RHS + LSERK4 hot path behind the reference's setup API."""
from .bctypes import BCTag  # noqa: F401
from .dgcontext import DGContext2D  # noqa: F401
from .mesh_manager import MeshManager  # noqa: F401
from .triangle_nodes import TriangleNodesProvisioner  # noqa: F401
from .quad_nodes import QuadNodesProvisioner  # noqa: F401
from . import lserk4 as LSERK4  # noqa: F401
