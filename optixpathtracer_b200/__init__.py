"""B200-native drop-in for the render hot path of bipul-mohanto/OptixPathTracer.

Host-side mirror of the reference interface: Model/TriangleMesh/loadOBJ/addBox, ProbeData, Camera,
SampleRenderer(resize/setCamera/setProbe/render/downloadPixels) over the C-ABI in include/b200pt.h.
"""
from .structs import Material, LaunchParams, Config, preset          # noqa: F401
from .model import Model, TriangleMesh, Texture, loadOBJ, addBox      # noqa: F401
from .probe import ProbeData, loadColor, loadProbe, read_hdr, write_hdr, analytic_sky                 # noqa: F401
from .camera import Camera, preset_camera                             # noqa: F401
from .renderer import SampleRenderer, host_params, set_camera         # noqa: F401
