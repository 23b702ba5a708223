"""B200-native CLIP-NeRF volumetric render path (host side).

``render`` holds the differentiable render call, ``ops`` one thin wrapper per C-ABI entry
point, ``params`` the flat parameter layout <-> Keras variable names."""
from . import _abi  # noqa: F401
