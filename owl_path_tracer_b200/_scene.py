"""Scene — the loaded program data (init_program_data, application.cpp:143-181)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import (ASSETS_DIR, B200ptError, EntityView, LoadOptions, MATERIAL_DTYPE, SceneInfo, _view, scene_lib)


class Scene:
    """Host-side scene: settings.json + <scene>.json + <scene>.obj.scene + environment.hdr.

    Keyword overrides default to the reference behaviour (everything from settings.json).
    ``env``: None | "color" | "auto" | "map".
    """

    def __init__(self, assets=None, scene=None, width=0, height=0, spp=0, depth=0, env=None,
                 env_intensity=-1.0, detail=0, allow_procedural=True, env_float=False):
        H = scene_lib()   # no CUDA library behind this one
        opt = LoadOptions()
        opt.scene = scene.encode() if scene else None
        opt.width, opt.height = int(width), int(height)
        opt.max_samples, opt.max_path_depth = int(spp), int(depth)
        opt.env_mode = {None: -1, "color": 0, "auto": 1, "map": 2}[env]
        opt.env_intensity = float(env_intensity)
        opt.detail = int(detail)
        opt.no_procedural = 0 if allow_procedural else 1
        opt.env_float = 1 if env_float else 0   # keep environment.hdr as float radiance (b200pt_set_environment_f32)
        self._h = C.c_void_p()
        self._H = H
        path = str(assets or ASSETS_DIR).encode()
        if H.b200pt_scene_load(path, C.byref(opt), C.byref(self._h)) != 0:
            raise B200ptError("scene load: " + H.b200pt_host_last_error().decode())
        self.info = SceneInfo()
        H.b200pt_scene_info(self._h, C.byref(self.info))

    def close(self):
        if getattr(self, "_h", None):
            self._H.b200pt_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- accessors (numpy views into the C++ vectors; valid while the Scene lives)
    @property
    def width(self):
        return self.info.width

    @property
    def height(self):
        return self.info.height

    @property
    def spp(self):
        return self.info.max_samples

    @property
    def depth(self):
        return self.info.max_path_depth

    @property
    def n_triangles(self):
        return int(self.info.n_triangles)

    @property
    def notes(self):
        return [self._H.b200pt_scene_note(self._h, i).decode() for i in range(self.info.n_notes)]

    @property
    def camera(self):
        i = self.info
        return (np.array(i.cam_origin[:], np.float32), np.array(i.cam_llc[:], np.float32),
                np.array(i.cam_horizontal[:], np.float32), np.array(i.cam_vertical[:], np.float32))

    @property
    def environment(self):
        i = self.info
        img = _view(i.env_rgba, (i.env_h, i.env_w, 4), np.uint8)
        f32 = _view(i.env_f32, (i.env_h, i.env_w, 3), np.float32) if i.env_f32 else None
        return dict(rgba=img, rgb_f32=f32, use=i.environment_use, auto=i.environment_auto,
                    color=np.array(i.environment_color[:], np.float32), intensity=float(i.environment_intensity))

    def entities(self):
        out = []
        for k in range(self.info.n_entities):
            v = EntityView()
            self._H.b200pt_scene_entity(self._h, k, C.byref(v))
            out.append(dict(
                name=v.name.decode(), material=v.material_index,
                verts=_view(v.verts, (v.nv, 3), np.float32), normals=_view(v.normals, (v.nv, 3), np.float32),
                uvs=_view(v.uvs, (v.nv, 2), np.float32), idx=_view(v.idx, (v.nt, 3), np.int32),
                tex=_view(v.tex_rgba, (v.tex_h, v.tex_w, 4), np.uint8)))
        return out

    def materials(self) -> np.ndarray:
        n = C.c_int32()
        p = self._H.b200pt_scene_materials(self._h, C.byref(n))
        return _view(p, (n.value,), MATERIAL_DTYPE).copy()

    def material_names(self):
        return [self._H.b200pt_scene_material_name(self._h, i).decode() for i in range(self.info.n_materials)]

    def set_material(self, material: str, attr: str, value):
        """modify_sbt (application.cpp:320-360), host side."""
        if attr == "base_color":
            rgb = np.asarray(value, np.float32)
            r = self._H.b200pt_scene_set_material_color(self._h, material.encode(), rgb.ctypes.data_as(C.c_void_p))
        else:
            r = self._H.b200pt_scene_set_material_attr(self._h, material.encode(), attr.encode(), float(value))
        if r != 0:
            raise B200ptError(f"unknown material {material!r}")

    def write_obj(self, path):
        if self._H.b200pt_write_obj(self._h, str(path).encode()) != 0:
            raise B200ptError(self._H.b200pt_host_last_error().decode())
