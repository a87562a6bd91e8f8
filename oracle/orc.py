"""ctypes front end of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  Never import this from renderengine_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

fp = C.POINTER(C.c_float)
u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


class OrcTexture(C.Structure):
    _fields_ = [("texels", u16p), ("width", C.c_uint32), ("height", C.c_uint32), ("srgb", C.c_uint32), ("pad", C.c_uint32)]


class OrcScene(C.Structure):
    _fields_ = [
        ("verts", fp), ("nVerts", C.c_uint64),
        ("indices", u32p), ("nIndices", C.c_uint64),
        ("transforms", fp), ("nNodes", C.c_uint32), ("nDraws", C.c_uint32),
        ("draws", u32p),
        ("materials", u8p), ("materialStride", C.c_uint32), ("nMaterials", C.c_uint32),
        ("lights", u8p), ("lightStride", C.c_uint32), ("nLights", C.c_uint32),
        ("view", u8p),
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("shadowSize", C.c_uint32), ("maxDescriptors", C.c_uint32),
        ("shadowView", fp), ("shadowProj", fp),
        ("textures", C.POINTER(OrcTexture)), ("nTextures", C.c_uint32), ("nThreads", C.c_uint32),
        ("rank", C.c_uint32), ("nRanks", C.c_uint32), ("bandH", C.c_uint32), ("shadingModel", C.c_uint32),
        ("skyFaces", u16p), ("skyW", C.c_uint32), ("skyH", C.c_uint32),
        ("translucentDraws", u32p), ("nTranslucentDraws", C.c_uint32),
    ]


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(HERE, "liboracle.so")
    src = [os.path.join(HERE, f) for f in ("oracle.cpp", "oracle.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src)
    if stale or force:
        subprocess.check_call(["make", "-s", "-C", HERE, os.path.join(HERE, "liboracle.so")])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_render_shadow.argtypes = [C.POINTER(OrcScene), fp, u64p]
        _LIB.orc_render_forward.argtypes = [C.POINTER(OrcScene), fp, fp, u32p, fp, u8p, u64p]
        _LIB.orc_phong_lighting.argtypes = [u8p, fp, fp, fp, fp, C.c_float, fp]
        _LIB.orc_compute_shadow.argtypes = [u8p, fp, fp, fp, fp, C.c_uint32]
        _LIB.orc_compute_shadow.restype = C.c_float
        _LIB.orc_render_forward_msaa4.argtypes = [C.POINTER(OrcScene), fp, fp, u32p, fp, u8p, u64p]
        _LIB.orc_sky_pixel.argtypes = [u8p, C.c_uint32, C.c_uint32, u16p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, fp]
        _LIB.orc_sky_lookup.argtypes = [u16p, C.c_uint32, C.c_uint32, fp, fp]
    return _LIB


def sky_lookup(faces, direction):
    """Appendix B.14 cube lookup (seamless bilinear) of one direction; faces = (6, h, w, 4) uint16."""
    f = np.ascontiguousarray(faces, np.uint16); d = np.ascontiguousarray(direction, np.float32); out = np.empty(4, np.float32)
    lib().orc_sky_lookup(_p(f, u16p), f.shape[2], f.shape[1], _p(d, fp), _p(out, fp))
    return out


def sky_pixel(view160, W, H, faces, x, y):
    f = np.ascontiguousarray(faces, np.uint16); v = np.ascontiguousarray(view160, np.uint8); out = np.empty(4, np.float32)
    lib().orc_sky_pixel(_p(v, u8p), W, H, _p(f, u16p), f.shape[2], f.shape[1], x, y, _p(out, fp))
    return out


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


class Oracle:
    """Renders a scene dict (see renderengine_b200.scene.make_scene) on the CPU."""

    def __init__(self, scene, n_threads=0, rank=0, n_ranks=1, band_h=64):
        self.scene = scene
        s = OrcScene()
        self._keep = []

        def hold(a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            self._keep.append(a)
            return a

        v = hold(scene["verts"], np.float32); s.verts = _p(v, fp); s.nVerts = v.shape[0]
        i = hold(scene["indices"], np.uint32); s.indices = _p(i, u32p); s.nIndices = i.size
        t = hold(scene["transforms"], np.float32); s.transforms = _p(t, fp); s.nNodes = t.shape[0]
        d = hold(scene["draws"], np.uint32); s.draws = _p(d, u32p); s.nDraws = d.shape[0]
        m = hold(scene["materials"], np.uint8); s.materials = _p(m, u8p); s.materialStride = m.shape[1]; s.nMaterials = m.shape[0]
        l = hold(scene["lights"], np.uint8); s.lights = _p(l, u8p); s.lightStride = l.shape[1]; s.nLights = l.shape[0]
        vw = hold(scene["view"], np.uint8); s.view = _p(vw, u8p)
        s.width, s.height = int(scene["width"]), int(scene["height"])
        s.shadowSize = int(scene.get("shadow_size", 0))
        s.maxDescriptors = int(scene.get("max_descriptors", 8192))
        if s.shadowSize:
            sv = hold(scene["shadow_view"], np.float32); sp = hold(scene["shadow_proj"], np.float32)
            s.shadowView = _p(sv, fp); s.shadowProj = _p(sp, fp)
        texs = scene.get("textures", [])
        if texs:
            arr = (OrcTexture * len(texs))()
            for k, tx in enumerate(texs):
                px = hold(tx["texels"], np.uint16)
                arr[k].texels = _p(px, u16p); arr[k].width = px.shape[1]; arr[k].height = px.shape[0]; arr[k].srgb = int(tx.get("srgb", 1))
            self._keep.append(arr)
            s.textures = arr; s.nTextures = len(texs)
        s.nThreads = n_threads
        s.rank, s.nRanks, s.bandH = rank, n_ranks, band_h
        s.shadingModel = int(scene.get("shading_model", 0))
        td = scene.get("translucent_draws")
        if td is not None and len(td):
            tdh = hold(td, np.uint32); s.translucentDraws = _p(tdh, u32p); s.nTranslucentDraws = tdh.shape[0]
        sky = scene.get("skybox")
        if sky is not None:
            f = hold(sky, np.uint16)                  # (6, h, w, 4) RGBA16, layers +X -X +Y -Y +Z -Z
            s.skyFaces = _p(f, u16p); s.skyH, s.skyW = f.shape[1], f.shape[2]
        self.s = s
        self.counters = np.zeros(4, np.uint64)

    def set_view(self, view160):
        vw = np.ascontiguousarray(view160, dtype=np.uint8)
        self._keep.append(vw)
        self.s.view = _p(vw, u8p)

    def render_shadow(self):
        S = self.s.shadowSize
        out = np.empty((S, S), np.float32)
        rc = lib().orc_render_shadow(C.byref(self.s), _p(out, fp), _p(self.counters, u64p))
        assert rc == 0
        return out

    def render_forward_msaa4(self, shadow_map=None):
        """Appendix B.15: 4 samples per pixel (Vulkan standard locations) + average resolve.  depth / tri_id: (4, H, W)."""
        W, H = self.s.width, self.s.height
        depth = np.empty((4, H, W), np.float32); tri = np.empty((4, H, W), np.uint32)
        cf = np.empty((H, W, 4), np.float32); c8 = np.empty((H, W, 4), np.uint8)
        sm = np.ascontiguousarray(shadow_map, np.float32) if shadow_map is not None else None
        rc = lib().orc_render_forward_msaa4(C.byref(self.s), _p(sm, fp), _p(depth, fp), _p(tri, u32p), _p(cf, fp), _p(c8, u8p), _p(self.counters, u64p))
        assert rc == 0
        return {"depth": depth, "tri_id": tri, "color_f32": cf, "color_srgb8": c8}

    def render_forward(self, shadow_map=None, color=True):
        W, H = self.s.width, self.s.height
        depth = np.empty((H, W), np.float32); tri = np.empty((H, W), np.uint32)
        cf = np.empty((H, W, 4), np.float32) if color else None
        c8 = np.empty((H, W, 4), np.uint8) if color else None
        sm = np.ascontiguousarray(shadow_map, np.float32) if shadow_map is not None else None
        rc = lib().orc_render_forward(C.byref(self.s), _p(sm, fp), _p(depth, fp), _p(tri, u32p), _p(cf, fp), _p(c8, u8p), _p(self.counters, u64p))
        assert rc == 0
        return {"depth": depth, "tri_id": tri, "color_f32": cf, "color_srgb8": c8}
