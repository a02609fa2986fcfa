"""b2c — B200-native gradient synchronisation behind the ray_lightning accelerator API.

Drop-in surface (ray_lightning/__init__.py:1-4):
    from ray_lightning_accelerators_b200 import RayAccelerator, HorovodRayAccelerator
"""
__all__ = ["RayAccelerator", "HorovodRayAccelerator"]


def __getattr__(name):  # lazy: importing the package must not need CUDA
    if name == "RayAccelerator":
        from .ray_ddp import RayAccelerator
        return RayAccelerator
    if name == "HorovodRayAccelerator":
        from .ray_horovod import HorovodRayAccelerator
        return HorovodRayAccelerator
    raise AttributeError(name)
