"""B200-native hybrid NeRF + mesh render path (drop-in for the render path of the instant-ngp fork
fnysalehi/Surface-Irradiance-Estimation-from-Neural-Radiance-Fields). The product is csrc/ (hand-written sm_100a CUDA behind
the C ABI of include/hybrid_nerf.h); this package is the thin host mirror of ``pyngp.Testbed`` plus synthetic inputs."""
from . import synthetic  # noqa: F401
from .testbed import EXPORTS, LIB_PATH, NOALLOC_MAX_POSITIONS, FrameStats, NetworkInfo, RenderMode, Testbed, camera_to_cols, decode_image_file  # noqa: F401

__all__ = ["Testbed", "NetworkInfo", "FrameStats", "RenderMode", "EXPORTS", "LIB_PATH", "NOALLOC_MAX_POSITIONS", "camera_to_cols", "decode_image_file", "synthetic"]
