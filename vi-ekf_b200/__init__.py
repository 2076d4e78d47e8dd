"""viekf_b200 -- host-side mirror of the reference's VIEKF interface over the B200 batch engine (C ABI in
include/viekf_batch.h, CUDA in csrc/).  Import through __graft_entry__.load_package() (the directory name
`vi-ekf_b200` is not a Python identifier)."""
from . import abi  # noqa: F401
from .batch import BatchVIEKF  # noqa: F401
from .tape import TapeBuilder  # noqa: F401
from . import batch, logfmt, refsim, shard, sim, tape  # noqa: F401,E402
