"""``src/lib/lmvnerf/layers.py`` of the reference is identical to ``delta_ngf/layers.py``."""
from ..delta_ngf.layers import GraspReadout  # noqa: F401
