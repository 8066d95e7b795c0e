"""cudapathtracer_b200 — ctypes binding of libnovum_b200.so (the C ABI in include/novum_b200.h).

The product is the shared library (hand-written sm_100a CUDA behind the reference's launch boundary); this
package only loads it for tests and bench.py. There is no CPU fallback: a missing library or a missing GPU
raises, loudly.
"""
from .binding import (  # noqa: F401
    NovumError, lib, lib_path, Scene, HashGrid, VcmSession, RenderConfig, Camera, RenderStats, Shard, Constants,
    make_scene, load_config, material_table, camera_make, device_count, DECLARED_SYMBOLS,
    FRAME_PROGRESSIVE, FRAME_FINAL, bmp_row_bytes, frame_finish, frame_finish_device, obj_parse, ObjDeviceMesh, accum_save, accum_info, accum_load, frame_save_bmp, frame_save_bmp_async, frame_save_wait, save_bmp, gamma_steps,
    hashgrid_build_static, STATIC_GRID_MAX_PHOTONS, LaunchCacheStats, TileRank,
)
