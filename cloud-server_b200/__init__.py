"""njucloud-b200: the B200-native hot path behind NJUCloud (view-sonic/Cloud-Server).

Import as ``cloud_server_b200``.  Everything numeric runs in hand-written sm_100a CUDA kernels
behind the C ABI declared in ``include/njucloud_b200.h``; there is no CPU fallback: using any
op without the built library or without a B200 raises.
"""
__version__ = "0.1.0"

from .build import build_library, library_path  # noqa: F401
