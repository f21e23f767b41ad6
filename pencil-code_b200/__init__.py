"""pencil-code_b200: B200-native solver core for the Pencil Code's RHS / RK3
hot path.  csrc/ holds the CUDA kernels and the C ABI (libpcb200.so),
fortran/ the drop-in `GPU` module for the reference host, and the Python
modules mirror that host interface for tests and the bench.

The directory name is not a Python identifier; load it through
`__graft_entry__.package()` (registers it as `pencil_code_b200`)."""
from . import abi, gpu, setups, snapshot  # noqa: F401
from .gpu import GPU, PCB200Error, load_library  # noqa: F401
