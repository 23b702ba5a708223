"""``src/lib/lmvnerf/grasp_optimizer.py`` of the reference is ``delta_ngf/grasp_optimizer.py`` for ``LanguageNeRF``."""
from ..delta_ngf.grasp_optimizer import DNGFOptimizer  # noqa: F401
