"""covid-abm daily transmission cycle, B200-native (host side of the C ABI in include/cabm.h).

The directory name carries a hyphen (it mirrors the reference repository's name), so it is loaded
through the root-level shim `covid_abm_b200.py`; inside, modules use relative imports.
"""
from . import abi, params, replicates, sharded   # noqa: F401
from .community import Community, CabmError, Backend, cuda_backend, save_population, LIB_PATH   # noqa: F401
