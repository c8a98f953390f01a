"""aniso_b200: B200-native (sm_100a CUDA) implementation of the lowrank/aniso hot path, i.e. the black-box-FMM
matvec `mapping` of the anisotropic radiative-transfer integral operator and the GMRES solve that drives it.
The product is the C-ABI library libaniso_b200.so (include/aniso_b200.h); this package is the thin host-side
mirror of the reference's MEX/class interface (AnisoWrapper.cpp, class/@Aniso/Aniso.m)."""
from ._lib import AnisoError, LIB_PATH, load  # noqa: F401
from .api import Aniso, AnisoWrapper, BBFMMKernel  # noqa: F401
