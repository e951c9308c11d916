"""owl_path_tracer_b200 — ctypes binding of the B200-native render loop.

Two shared libraries are bound, both built in-tree by ``make product``:

* ``lib/libb200pt.so``       the CUDA path behind the C ABI of ``include/b200pt.h``
* ``lib/libb200pt_scene.so`` the host stages that never touch the device (``include/b200pt_host.h``): settings /
  scene / OBJ / HDR loading, camera, material sweep, PNG output (no CUDA dependency)
* ``lib/libb200pt_host.so``  bind / render / test_loop on top of the two — the same code the ``bin/path_tracer``
  executable runs (reference: path_tracer/Main.cpp, path_tracer/src/application.cpp).

There is no CPU fallback here: ``Context()`` raises when the CUDA library is missing or no
sm_100 device is usable.  The scalar oracle lives under ``oracle/`` and is only ever
loaded by ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent
LIB_DIR = PKG_DIR / "lib"
ASSETS_DIR = REPO_DIR / "assets"

SHARD_SPP, SHARD_ROWS = 0, 1
FLAG_COUNTERS, FLAG_TIMING, FLAG_NEE = 1, 4, 8


class B200ptError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("device_ordinal", C.c_int32), ("rank", C.c_int32), ("world_size", C.c_int32),
                ("shard_mode", C.c_int32), ("flags", C.c_int32), ("streams_per_pixel", C.c_int32)]


MATERIAL_FIELDS = ["base_color_r", "base_color_g", "base_color_b", "subsurface", "metallic", "specular",
                   "specular_tint", "roughness", "anisotropic", "sheen", "sheen_tint", "clearcoat",
                   "clearcoat_gloss", "ior", "specular_transmission", "specular_transmission_roughness",
                   "emission"]
#: numpy view of b200pt_material (17 floats, device_global.hpp:19-36 order)
MATERIAL_DTYPE = np.dtype([(n, np.float32) for n in MATERIAL_FIELDS])
HIT_DTYPE = np.dtype([("mesh", np.int32), ("prim", np.int32), ("t", np.float32), ("u", np.float32),
                      ("v", np.float32)])


class Stats(C.Structure):
    _fields_ = [("render_ms", C.c_double), ("extend_ms", C.c_double), ("shade_ms", C.c_double),
                ("other_ms", C.c_double), ("build_ms", C.c_double), ("rays", C.c_uint64),
                ("paths", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("iterations", C.c_uint64), ("bvh_nodes", C.c_uint32),
                ("bvh_tris", C.c_uint32), ("bvh_bytes", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("shadow_ms", C.c_double), ("n_lights", C.c_uint32), ("reserved", C.c_uint32),
                ("refit_ms", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class PassInfo(C.Structure):
    """b200pt_pass_info (progressive passes)."""
    _fields_ = [("samples_done", C.c_int32), ("samples_in_pass", C.c_int32), ("complete", C.c_int32),
                ("stopped_early", C.c_int32), ("rel_variance", C.c_double), ("render_ms", C.c_double),
                ("rays", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


PASS_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(PassInfo), C.c_void_p, C.c_void_p)
TRANSFER_LINEAR, TRANSFER_GAMMA22 = 0, 1


class LoadOptions(C.Structure):
    _fields_ = [("scene", C.c_char_p), ("width", C.c_int32), ("height", C.c_int32),
                ("max_samples", C.c_int32), ("max_path_depth", C.c_int32), ("env_mode", C.c_int32),
                ("env_intensity", C.c_float), ("detail", C.c_int32), ("no_procedural", C.c_int32),
                ("env_float", C.c_int32)]


class SceneInfo(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_samples", C.c_int32),
                ("max_path_depth", C.c_int32), ("environment_use", C.c_int32),
                ("environment_auto", C.c_int32), ("environment_color", C.c_float * 3),
                ("environment_intensity", C.c_float), ("env_w", C.c_int32), ("env_h", C.c_int32),
                ("env_rgba", C.c_void_p), ("cam_origin", C.c_float * 3), ("cam_llc", C.c_float * 3),
                ("cam_horizontal", C.c_float * 3), ("cam_vertical", C.c_float * 3),
                ("n_entities", C.c_int32), ("n_materials", C.c_int32), ("n_notes", C.c_int32),
                ("n_triangles", C.c_int64), ("env_f32", C.c_void_p)]


class EntityView(C.Structure):
    _fields_ = [("name", C.c_char_p), ("verts", C.c_void_p), ("normals", C.c_void_p), ("uvs", C.c_void_p),
                ("nv", C.c_int32), ("idx", C.c_void_p), ("nt", C.c_int32), ("material_index", C.c_int32),
                ("tex_rgba", C.c_void_p), ("tex_w", C.c_int32), ("tex_h", C.c_int32)]


_libs = {}
_overrides = {}


def use_library(path, name="libb200pt.so"):
    """Bind a variant build of a library (tools/ab*.py) instead of lib/<name>.  Must be called before first use."""
    if name in _libs:
        raise B200ptError(f"{name} is already loaded")
    _overrides[name] = Path(path)


def _load(name: str) -> C.CDLL:
    if name not in _libs:
        path = LIB_DIR / name
        if name in _overrides:   # A/B builds: use_library() called explicitly by a tool, never the ambient environment
            path = _overrides[name]
        if not path.exists():
            raise B200ptError(f"{path} is missing: run `make product` (or __graft_entry__.build()); "
                              "there is no CPU fallback for the render path")
        _libs[name] = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
    return _libs[name]


def lib() -> C.CDLL:
    """libb200pt.so with argument types set."""
    L = _load("libb200pt.so")
    if getattr(L, "_typed", False):
        return L
    P = C.c_void_p
    L.b200pt_last_error.restype = C.c_char_p
    L.b200pt_create.argtypes = [C.POINTER(Config), C.POINTER(P)]
    L.b200pt_destroy.argtypes = [P]
    L.b200pt_add_mesh.argtypes = [P, P, P, P, C.c_int32, P, C.c_int32, C.c_int32, P, C.c_int32, C.c_int32,
                                  C.POINTER(C.c_int32)]
    L.b200pt_build_accel.argtypes = [P]
    L.b200pt_set_materials.argtypes = [P, P, C.c_int32]
    L.b200pt_set_environment.argtypes = [P, P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, P, C.c_float]
    L.b200pt_set_environment_f32.argtypes = [P, P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, P, C.c_float]
    L.b200pt_set_camera.argtypes = [P, P, P, P, P]
    L.b200pt_set_render.argtypes = [P, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    L.b200pt_render.argtypes = [P]
    L.b200pt_update_mesh.argtypes = [P, C.c_int32, P, P, C.c_int32]
    L.b200pt_refit_accel.argtypes = [P]
    L.b200pt_render_pass.argtypes = [P, C.c_int32, C.POINTER(PassInfo)]
    L.b200pt_render_restart.argtypes = [P]
    L.b200pt_pass_stats.argtypes = [P, C.POINTER(PassInfo)]
    L.b200pt_render_progressive.argtypes = [P, C.c_int32, C.c_double, PASS_FN, C.c_void_p, C.POINTER(PassInfo)]
    L.b200pt_set_output.argtypes = [P, C.c_int32]
    L.b200pt_accum_buffer.argtypes = [P, C.POINTER(P), C.POINTER(C.c_size_t)]
    L.b200pt_resolve.argtypes = [P]
    L.b200pt_comm_unique_id.argtypes = [P]
    L.b200pt_comm_init_rank.argtypes = [P, P]
    L.b200pt_comm_init_all.argtypes = [C.POINTER(P), C.c_int32]
    L.b200pt_reduce.argtypes = [P]
    L.b200pt_render_frame.argtypes = [P]
    L.b200pt_render_frame_all.argtypes = [C.POINTER(P), C.c_int32]
    L.b200pt_render_to_host.argtypes = [P, P, C.c_int32, P]
    L.b200pt_framebuffer_rgba8.argtypes = [P, C.POINTER(P)]
    L.b200pt_framebuffer_f32.argtypes = [P, C.POINTER(P)]
    L.b200pt_primary_hits.argtypes = [P, P]
    L.b200pt_trace.argtypes = [P, P, P, C.c_int32, C.c_float, C.c_float, P]
    L.b200pt_dedup_mesh.argtypes = [P, P, C.c_int32, P, C.c_int32, P, C.c_int32, P, P, P, C.c_int32, P, P, P, P, C.POINTER(C.c_int32)]
    L.b200pt_set_drain_callback.argtypes = [P, C.c_float, P, P]
    L.b200pt_pipeline_create.argtypes = [C.POINTER(P), C.c_int32, C.c_float, C.POINTER(P)]
    L.b200pt_pipeline_submit.argtypes = [P, P, C.c_int32, P]
    L.b200pt_pipeline_wait.argtypes = [P]
    L.b200pt_pipeline_destroy.argtypes = [P]
    L.b200pt_trace_static.argtypes = [P, P, P, C.c_int32, C.c_float, C.c_float, P]
    L.b200pt_get_stats.argtypes = [P, C.POINTER(Stats)]
    L.b200pt_sample_disney.argtypes = [P, C.c_int32] + [P] * 9
    L.b200pt_export_bvh.argtypes = [P, P, C.POINTER(C.c_uint32), P, C.POINTER(C.c_uint32)]
    L._typed = True
    return L


def host_lib() -> C.CDLL:
    """libb200pt_host.so (bind / render / test_loop; pulls in libb200pt.so and libb200pt_scene.so) with argument types
    set.  Every symbol of include/b200pt_host.h resolves through it."""
    lib()
    scene_lib()
    H = _load("libb200pt_host.so")
    if getattr(H, "_typed", False):
        return H
    _type_scene_api(H)
    P = C.c_void_p
    H.b200pt_scene_bind.argtypes = [P, P]
    H.b200pt_scene_test_loop.argtypes = [P, P, C.c_char_p]
    H.b200pt_obj_load_gpu.argtypes = [P, C.c_char_p, C.c_int32, C.POINTER(P)]
    H._typed = True
    return H


def scene_lib() -> C.CDLL:
    """libb200pt_scene.so: scene loading, the loaders stage by stage, image / OBJ writers.  Does NOT map the CUDA library,
    so CPU-only processes (bench.py --impl reference, the oracle tests) stay free of the product's device code."""
    S = _load("libb200pt_scene.so")
    if getattr(S, "_typed", False):
        return S
    _type_scene_api(S)
    S._typed = True
    return S


def _type_scene_api(H):
    P = C.c_void_p
    H.b200pt_host_last_error.restype = C.c_char_p
    H.b200pt_scene_load.argtypes = [C.c_char_p, C.POINTER(LoadOptions), C.POINTER(P)]
    H.b200pt_scene_free.argtypes = [P]
    H.b200pt_scene_info.argtypes = [P, C.POINTER(SceneInfo)]
    H.b200pt_scene_entity.argtypes = [P, C.c_int32, C.POINTER(EntityView)]
    H.b200pt_scene_materials.argtypes = [P, C.POINTER(C.c_int32)]
    H.b200pt_scene_materials.restype = P
    H.b200pt_scene_material_name.argtypes = [P, C.c_int32]
    H.b200pt_scene_material_name.restype = C.c_char_p
    H.b200pt_scene_note.argtypes = [P, C.c_int32]
    H.b200pt_scene_note.restype = C.c_char_p
    H.b200pt_scene_set_material_attr.argtypes = [P, C.c_char_p, C.c_char_p, C.c_float]
    H.b200pt_scene_set_material_color.argtypes = [P, C.c_char_p, P]
    H.b200pt_write_png.argtypes = [C.c_char_p, C.c_int32, C.c_int32, P]
    H.b200pt_write_pfm.argtypes = [C.c_char_p, C.c_int32, C.c_int32, P]
    H.b200pt_write_exr.argtypes = [C.c_char_p, C.c_int32, C.c_int32, P]
    H.b200pt_write_obj.argtypes = [P, C.c_char_p]
    H.b200pt_obj_load.argtypes = [C.c_char_p, C.c_int32, C.POINTER(P)]
    H.b200pt_obj_count.argtypes = [P]
    H.b200pt_obj_shape.argtypes = [P, C.c_int32, C.POINTER(EntityView)]
    H.b200pt_obj_free.argtypes = [P]
    H.b200pt_image_load.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(P)]
    H.b200pt_image_free.argtypes = [P]
    H.b200pt_parse_camera.argtypes = [C.c_char_p, C.c_int32, C.c_int32, P]
    H.b200pt_parse_settings.argtypes = [C.c_char_p, P]
    H.b200pt_parse_materials.argtypes = [C.c_char_p, C.POINTER(P)]
    H.b200pt_materials_count.argtypes = [P]
    H.b200pt_materials_name.argtypes = [P, C.c_int32]
    H.b200pt_materials_name.restype = C.c_char_p
    H.b200pt_materials_file.argtypes = [P, C.c_int32]
    H.b200pt_materials_file.restype = C.c_char_p
    H.b200pt_materials_data.argtypes = [P, C.c_int32]
    H.b200pt_materials_data.restype = P
    H.b200pt_materials_free.argtypes = [P]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _view(addr, shape, dtype):
    n = int(np.prod(shape))
    if not addr or n == 0:
        return None
    buf = (C.c_byte * (n * np.dtype(dtype).itemsize)).from_address(addr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


from ._scene import Scene  # noqa: E402
from ._context import Context, Pipeline  # noqa: E402

__all__ = ["Scene", "Context", "Pipeline", "Config", "Stats", "B200ptError", "MATERIAL_DTYPE", "HIT_DTYPE",
           "SHARD_SPP", "SHARD_ROWS", "FLAG_COUNTERS", "FLAG_TIMING", "FLAG_NEE", "lib", "host_lib", "scene_lib", "use_library", "ASSETS_DIR"]
