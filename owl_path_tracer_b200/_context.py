"""Context — one GPU's render context behind the C ABI (include/b200pt.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import B200ptError, Config, HIT_DTYPE, MATERIAL_DTYPE, PASS_FN, PassInfo, Stats, _ptr, _view, host_lib, lib


class Context:
    def __init__(self, device=0, rank=0, world_size=1, shard_mode=0, flags=0, streams=1):
        """streams: RNG streams per pixel on this GPU (b200pt_config.streams_per_pixel); 1 = the reference's order."""
        self._L = lib()
        cfg = Config(device, rank, world_size, shard_mode, flags, streams)
        self._h = C.c_void_p()
        self.rank, self.world_size = rank, world_size
        self._ck(self._L.b200pt_create(C.byref(cfg), C.byref(self._h)), "create")
        self.width = self.height = 0

    def _ck(self, status, what):
        if status != 0:
            raise B200ptError(f"b200pt_{what}: {self._L.b200pt_last_error().decode()} (status {status})")

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200pt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene upload, mirrors bind_sbt_data (application.cpp:184-294)
    def add_mesh(self, verts, normals, idx, material, uvs=None, tex=None):
        verts = np.ascontiguousarray(verts, np.float32)
        normals = np.ascontiguousarray(normals, np.float32)
        idx = np.ascontiguousarray(idx, np.int32)
        uvs = None if uvs is None else np.ascontiguousarray(uvs, np.float32)
        tex = None if tex is None else np.ascontiguousarray(tex, np.uint8)
        th, tw = (tex.shape[0], tex.shape[1]) if tex is not None else (0, 0)
        out = C.c_int32(-1)
        self._ck(self._L.b200pt_add_mesh(self._h, _ptr(verts), _ptr(normals), _ptr(uvs), verts.shape[0], _ptr(idx),
                                         idx.shape[0], int(material), _ptr(tex), tw, th, C.byref(out)), "add_mesh")
        return out.value

    def build_accel(self):
        self._ck(self._L.b200pt_build_accel(self._h), "build_accel")

    def update_mesh(self, mesh_index, verts, normals=None):
        """New vertex positions (and normals) of a mesh; follow with refit_accel() or build_accel()."""
        verts = np.ascontiguousarray(verts, np.float32)
        normals = None if normals is None else np.ascontiguousarray(normals, np.float32)
        self._ck(self._L.b200pt_update_mesh(self._h, int(mesh_index), _ptr(verts), _ptr(normals), verts.shape[0]), "update_mesh")

    def refit_accel(self):
        self._ck(self._L.b200pt_refit_accel(self._h), "refit_accel")

    def set_materials(self, mats):
        mats = np.ascontiguousarray(mats, MATERIAL_DTYPE)
        self._ck(self._L.b200pt_set_materials(self._h, _ptr(mats), mats.shape[0]), "set_materials")

    def set_environment(self, rgba=None, use=False, auto=False, color=(0, 0, 0), intensity=0.0):
        rgba = None if rgba is None else np.ascontiguousarray(rgba, np.uint8)
        h, w = (rgba.shape[0], rgba.shape[1]) if rgba is not None else (0, 0)
        col = np.asarray(color, np.float32)
        self._ck(self._L.b200pt_set_environment(self._h, _ptr(rgba), w, h, int(bool(use)), int(bool(auto)), _ptr(col),
                                                float(intensity)), "set_environment")

    def set_environment_f32(self, rgb, use=True, auto=False, color=(0, 0, 0), intensity=1.0):
        """float RGB environment map (h, w, 3), linear radiance (b200pt_set_environment_f32)"""
        rgb = np.ascontiguousarray(rgb, np.float32)
        col = np.asarray(color, np.float32)
        self._ck(self._L.b200pt_set_environment_f32(self._h, _ptr(rgb), rgb.shape[1], rgb.shape[0], int(bool(use)),
                                                    int(bool(auto)), _ptr(col), float(intensity)), "set_environment_f32")

    def set_camera(self, origin, llc, horizontal, vertical):
        a = [np.ascontiguousarray(x, np.float32) for x in (origin, llc, horizontal, vertical)]
        self._ck(self._L.b200pt_set_camera(self._h, *[_ptr(x) for x in a]), "set_camera")

    def set_render(self, width, height, spp, depth):
        self._ck(self._L.b200pt_set_render(self._h, int(width), int(height), int(spp), int(depth)), "set_render")
        self.width, self.height = int(width), int(height)

    def bind(self, scene):
        """bind_sbt_data through the C++ host layer (same code as bin/path_tracer)."""
        H = host_lib()
        if H.b200pt_scene_bind(scene._h, self._h) != 0:
            raise B200ptError("bind: " + H.b200pt_host_last_error().decode())
        self.width, self.height = scene.width, scene.height

    # ---- frames
    def render(self):
        self._ck(self._L.b200pt_render(self._h), "render")

    # ---- progressive surface (b200pt_render_pass / b200pt_render_progressive / b200pt_set_output)
    def render_pass(self, n_samples):
        """Next n_samples samples of every pixel; the frame is the mean over the samples so far.  Returns the pass info."""
        info = PassInfo()
        self._ck(self._L.b200pt_render_pass(self._h, int(n_samples), C.byref(info)), "render_pass")
        return info.as_dict()

    def pass_stats(self):
        """Figure of the frame behind accum_buffer (after the reduce on rank 0 of a multi-rank pass)."""
        info = PassInfo()
        self._ck(self._L.b200pt_pass_stats(self._h, C.byref(info)), "pass_stats")
        return info.as_dict()

    def render_restart(self):
        self._ck(self._L.b200pt_render_restart(self._h), "render_restart")

    def render_progressive(self, pass_samples, target_rel_variance=0.0, on_pass=None):
        """Passes until max_samples, until rel_variance <= target, or until on_pass(info, rgba, mean) returns True.
        rgba / mean are views that stay valid during the callback only."""
        last = PassInfo()
        H, W = self.height, self.width

        def tramp(_user, info, rgba, mean):
            if on_pass is None:
                return 0
            stop = on_pass(info.contents.as_dict(), _view(rgba, (H, W), np.uint32), _view(mean, (H, W, 4), np.float32))
            return 1 if stop else 0

        cb = PASS_FN(tramp)
        self._ck(self._L.b200pt_render_progressive(self._h, int(pass_samples), float(target_rel_variance), cb, None,
                                                   C.byref(last)), "render_progressive")
        return last.as_dict()

    def set_output(self, transfer):
        self._ck(self._L.b200pt_set_output(self._h, int(transfer)), "set_output")

    def resolve(self):
        self._ck(self._L.b200pt_resolve(self._h), "resolve")

    # ---- the frame's collective inside the library (NCCL; include/b200pt.h)
    @staticmethod
    def comm_unique_id() -> bytes:
        """128 bytes made on rank 0; ship them to the other ranks (any transport) for comm_init_rank."""
        buf = C.create_string_buffer(128)
        if lib().b200pt_comm_unique_id(buf) != 0:
            raise B200ptError("comm_unique_id: " + lib().b200pt_last_error().decode())
        return buf.raw

    def comm_init_rank(self, unique_id: bytes):
        self._ck(self._L.b200pt_comm_init_rank(self._h, C.create_string_buffer(unique_id, 128)), "comm_init_rank")

    def reduce(self):
        self._ck(self._L.b200pt_reduce(self._h), "reduce")

    def render_frame(self):
        """this rank's share + one ncclReduce of the float4 sums to rank 0 + (rank 0) resolve.  Collective."""
        self._ck(self._L.b200pt_render_frame(self._h), "render_frame")

    @staticmethod
    def comm_init_all(ctxs):
        arr = (C.c_void_p * len(ctxs))(*[c._h for c in ctxs])
        if lib().b200pt_comm_init_all(arr, len(ctxs)) != 0:
            raise B200ptError("comm_init_all: " + lib().b200pt_last_error().decode())

    @staticmethod
    def render_frame_all(ctxs):
        arr = (C.c_void_p * len(ctxs))(*[c._h for c in ctxs])
        if lib().b200pt_render_frame_all(arr, len(ctxs)) != 0:
            raise B200ptError("render_frame_all: " + lib().b200pt_last_error().decode())

    def accum_buffer(self):
        """(device pointer, number of floats) of the float4 sum buffer for the NCCL reduce."""
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._L.b200pt_accum_buffer(self._h, C.byref(p), C.byref(n)), "accum_buffer")
        return p.value, n.value

    def render_to_host(self, mats=None, out=None):
        if out is None:
            out = np.empty((self.height, self.width), np.uint32)
        mats = None if mats is None else np.ascontiguousarray(mats, MATERIAL_DTYPE)
        self._ck(self._L.b200pt_render_to_host(self._h, _ptr(mats), 0 if mats is None else mats.shape[0], _ptr(out)),
                 "render_to_host")
        return out

    def framebuffer_rgba8(self):
        p = C.c_void_p()
        self._ck(self._L.b200pt_framebuffer_rgba8(self._h, C.byref(p)), "framebuffer_rgba8")
        return _view(p.value, (self.height, self.width), np.uint32).copy()

    def framebuffer_f32(self):
        p = C.c_void_p()
        self._ck(self._L.b200pt_framebuffer_f32(self._h, C.byref(p)), "framebuffer_f32")
        return _view(p.value, (self.height, self.width, 4), np.float32).copy()

    def stats(self):
        s = Stats()
        self._ck(self._L.b200pt_get_stats(self._h, C.byref(s)), "get_stats")
        return s.as_dict()

    # ---- parity hooks
    def primary_hits(self):
        out = np.empty((self.height, self.width), HIT_DTYPE)
        self._ck(self._L.b200pt_primary_hits(self._h, _ptr(out)), "primary_hits")
        return out

    def trace(self, origins, dirs, tmin=1e-3, tmax=1e10):
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        out = np.empty(o.shape[0], HIT_DTYPE)
        self._ck(self._L.b200pt_trace(self._h, _ptr(o), _ptr(d), o.shape[0], tmin, tmax, _ptr(out)), "trace")
        return out

    def trace_static(self, origins, dirs, tmin=1e-3, tmax=1e10):
        """b200pt_trace_static: same result as trace(), no device allocation (at most 65536 rays per call)."""
        o = np.ascontiguousarray(origins, np.float32)
        d = np.ascontiguousarray(dirs, np.float32)
        out = np.empty(o.shape[0], HIT_DTYPE)
        self._ck(self._L.b200pt_trace_static(self._h, _ptr(o), _ptr(d), o.shape[0], tmin, tmax, _ptr(out)), "trace_static")
        return out

    def sample_disney(self, material, wo, lobe_in, rng_in):
        material = np.ascontiguousarray(material, np.int32)
        wo = np.ascontiguousarray(wo, np.float32)
        lobe_in = np.ascontiguousarray(lobe_in, np.int32)
        rng_in = np.ascontiguousarray(rng_in, np.uint32)
        n = material.shape[0]
        f, wi = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32)
        pdf, lobe, rng = np.empty(n, np.float32), np.empty(n, np.int32), np.empty(n, np.uint32)
        self._ck(self._L.b200pt_sample_disney(self._h, n, _ptr(material), _ptr(wo), _ptr(lobe_in), _ptr(rng_in), _ptr(f),
                                              _ptr(wi), _ptr(pdf), _ptr(lobe), _ptr(rng)), "sample_disney")
        return dict(f=f, wi=wi, pdf=pdf, lobe=lobe, rng=rng)

    def export_bvh(self):
        nn, nt = C.c_uint32(), C.c_uint32()
        self._ck(self._L.b200pt_export_bvh(self._h, None, C.byref(nn), None, C.byref(nt)), "export_bvh")
        nodes = np.empty((nn.value, 80), np.uint8)
        tris = np.empty((nt.value, 20), np.float32)   # 80-byte blocks of two triangles (pt_bvh8.h TriPair)
        self._ck(self._L.b200pt_export_bvh(self._h, _ptr(nodes), C.byref(nn), _ptr(tris), C.byref(nt)), "export_bvh")
        return nodes, tris


class Pipeline:
    """b200pt_pipeline_*: frames back to back on several contexts bound to the same scene; frame k+1 starts while frame k
    drains.  Every frame is bit-identical to the blocking render of its materials."""

    def __init__(self, contexts, start_fraction=0.0):
        self._L = lib()
        self._ctxs = list(contexts)
        arr = (C.c_void_p * len(self._ctxs))(*[c._h for c in self._ctxs])
        self._h = C.c_void_p()
        if self._L.b200pt_pipeline_create(arr, len(self._ctxs), float(start_fraction), C.byref(self._h)) != 0:
            raise B200ptError("pipeline_create: " + self._L.b200pt_last_error().decode())
        self._keep = []

    def submit(self, mats=None, out=None):
        """Queue one frame; `out` ((H, W) uint32, C-contiguous) is filled when the frame completes (after wait())."""
        m = None if mats is None else np.ascontiguousarray(mats, MATERIAL_DTYPE)
        self._keep.append((m, out))
        if self._L.b200pt_pipeline_submit(self._h, _ptr(m), 0 if m is None else m.shape[0], _ptr(out)) != 0:
            raise B200ptError("pipeline_submit: " + self._L.b200pt_last_error().decode())

    def wait(self):
        r = self._L.b200pt_pipeline_wait(self._h)
        self._keep.clear()
        if r != 0:
            raise B200ptError("pipeline: " + self._L.b200pt_last_error().decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200pt_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
