"""rpt_b200 — B200-native light-transport core behind rpt's API.

Host-side mirror of neevparikh/rpt's public API (Scene / Object / Material / Medium / Light /
Camera / Renderer) that lowers a scene to the C ABI of include/rptx.h and runs it on the
CUDA library rpt_b200/librptx.so.  There is no CPU path in this package.
"""
from ._abi import LIB_PATH  # noqa: F401
from .io import load_mtl, load_obj, load_obj_with_mtl, load_stl  # noqa: F401
from .renderer import Buffer, Context, DeviceScene, Film, Filter, PhotonList, PhotonMap, PhotonRenderKind, Renderer, RptxError  # noqa: F401
from .scene import KdTree, Mesh, MonomialSurface, Triangle, cube, hex_color, monomial_surface, plane, polygon, sphere, vec3  # noqa: F401
from .world import Camera, Hdri, Light, Material, Medium, Object, Scene  # noqa: F401
