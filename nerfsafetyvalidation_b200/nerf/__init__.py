from .renderer import NeRFRenderer  # noqa: F401
from .network_ff import NeRFNetwork  # noqa: F401
