"""ctypes driver of the ORACLE (oracle/liborc_oracle.so) — test infrastructure only.

It consumes the same lowered scene (rpt_b200.world.lower_scene) as the CUDA library, so one
scene description feeds the checker and the product.  Nothing under rpt_b200/ imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from rpt_b200 import _abi as A
from rpt_b200.world import lower_scene, upload_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_DIR = os.path.join(ROOT, "oracle")
ORC_LIB = os.path.join(ORC_DIR, "liborc_oracle.so")

P, DP, U32P, U8P, U64P = C.c_void_p, A.DP, A.U32P, A.U8P, C.POINTER(C.c_uint64)
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", ORC_DIR])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORC_LIB):
            build()
        L = C.CDLL(ORC_LIB)
        sig = {
            "orc_scene_create": (P, []),
            "orc_scene_destroy": (None, [P]),
            "orc_scene_builtin": (P, [C.c_char_p, DP, DP, C.c_size_t]),
            "orc_last_error": (C.c_char_p, [P]),
            "orc_scene_add_mesh": (C.c_int, [P, DP, DP, C.c_size_t, U32P]),
            "orc_scene_add_object": (C.c_int, [P, C.POINTER(A.ShapeDesc), C.POINTER(A.MaterialDesc), U32P]),
            "orc_scene_add_light": (C.c_int, [P, C.POINTER(A.LightDesc)]),
            "orc_scene_set_medium": (C.c_int, [P, C.c_int32, C.c_double, C.c_double]),
            "orc_scene_set_environment_color": (C.c_int, [P, DP]),
            "orc_scene_set_environment_hdri": (C.c_int, [P, C.c_uint32, C.c_uint32, DP]),
            "orc_mesh_kd_stats": (C.c_int, [P, C.c_uint32, U64P]),
            "orc_render_begin": (P, [P, C.POINTER(A.CameraDesc), C.POINTER(A.Params), C.c_int]),
            "orc_film_destroy": (None, [P]),
            "orc_render_sample": (C.c_int, [P, C.c_uint32]),
            "orc_time_rows": (C.c_double, [P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, U64P]),
            "orc_film_last_seconds": (C.c_double, [P]),
            "orc_film_counters": (C.c_int, [P, U64P]),
            "orc_film_image_rgb8": (C.c_int, [P, U8P]),
            "orc_film_variance": (C.c_int, [P, DP]),
            "orc_film_mean": (C.c_int, [P, DP]),
            "orc_sample_radiance": (C.c_int, [P, U32P, U32P, U64P, C.c_size_t, DP]),
            "orc_dump_rays": (C.c_size_t, [P, C.c_uint32, C.c_size_t, DP, DP, U8P]),
            "orc_intersect_batch": (C.c_int, [P, DP, DP, C.c_size_t, C.c_double, DP, U32P, U32P, U32P, DP, C.c_int]),
            "orc_hex_color": (None, [C.c_uint32, DP]),
            "orc_color_bytes": (None, [DP, U8P]),
            "orc_rng_u64": (C.c_uint64, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32]),
            "orc_hardware_threads": (C.c_int, []),
            "orc_photons_emit": (P, [P, C.POINTER(A.CameraDesc), C.POINTER(A.Params), C.c_uint64, C.c_uint64, C.c_uint64, C.c_int32, C.c_int]),
            "orc_photons_destroy": (None, [P]),
            "orc_photons_count": (None, [P, U64P]),
            "orc_photons_rows": (None, [P, C.c_int32, DP]),
            "orc_photon_map_build": (P, [P, DP, C.c_size_t, DP, C.c_size_t, C.c_int32]),
            "orc_photon_map_destroy": (None, [P]),
            "orc_photon_map_knn": (C.c_int, [P, C.c_int32, DP, C.c_size_t, C.c_uint32, U64P, DP, U32P]),
            "orc_photon_map_spheres": (C.c_size_t, [P, DP, DP]),
            "orc_photon_map_gather": (C.c_int, [P, C.POINTER(A.CameraDesc), C.POINTER(A.Params), C.c_int, U8P, DP, DP]),
            "orc_photon_render": (C.c_int, [P, C.POINTER(A.CameraDesc), C.POINTER(A.Params), C.c_uint64, C.c_int32, C.c_int, U8P, DP, U64P, DP]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(DP)


class OracleScene:
    def __init__(self, scene):
        self.L = lib()
        self.h = self.L.orc_scene_create()
        self.low = lower_scene(scene)
        upload_scene(self.low, self)

    @classmethod
    def builtin(cls, name, teapot=None):
        """A BASELINE scene built inside the oracle from the example sources' literals (oracle/builtin_scenes.hpp): no
        Python lowering on the way.  teapot: (vertices [n,3,3], normals [n,3,3]) for the two teapot scenes."""
        self = cls.__new__(cls)
        self.L = lib()
        self.low = None
        if teapot is not None:
            v = np.ascontiguousarray(teapot[0], np.float64)
            n = np.ascontiguousarray(teapot[1], np.float64)
            self.h = self.L.orc_scene_builtin(name.encode(), _dp(v), _dp(n), len(v))
        else:
            self.h = self.L.orc_scene_builtin(name.encode(), None, None, 0)
        if not self.h:
            raise ValueError(f"unknown builtin scene {name}")
        return self

    def _ck(self, rc, what):
        if rc:
            raise RuntimeError(f"{what}: {self.L.orc_last_error(self.h).decode()}")

    def add_mesh(self, v, n):
        mid = C.c_uint32()
        v = np.ascontiguousarray(v, np.float64)
        nptr = None
        if n is not None:
            n = np.ascontiguousarray(n, np.float64)
            nptr = _dp(n)
        self._ck(self.L.orc_scene_add_mesh(self.h, _dp(v), nptr, len(v), C.byref(mid)), "add_mesh")
        return mid.value

    def add_object(self, sd, md):
        self._ck(self.L.orc_scene_add_object(self.h, C.byref(sd), C.byref(md), None), "add_object")

    def add_light(self, ld):
        self._ck(self.L.orc_scene_add_light(self.h, C.byref(ld)), "add_light")

    def set_medium(self, kind, a, s):
        self._ck(self.L.orc_scene_set_medium(self.h, kind, a, s), "set_medium")

    def set_environment_color(self, rgb):
        self._ck(self.L.orc_scene_set_environment_color(self.h, (C.c_double * 3)(*rgb)), "env")

    def set_environment_hdri(self, w, h, buf):
        buf = np.ascontiguousarray(buf, np.float64)
        self._ck(self.L.orc_scene_set_environment_hdri(self.h, w, h, _dp(buf)), "hdri")

    def commit(self):
        pass

    def intersect_batch(self, o, d, t_min=1e-12, threads=0):
        o = np.ascontiguousarray(o, np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, np.float64).reshape(-1, 3)
        n = len(o)
        t = np.empty(n, np.float64)
        obj, prim, inst = np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.uint32)
        nrm = np.empty((n, 3), np.float64)
        threads = threads or self.L.orc_hardware_threads()
        self._ck(self.L.orc_intersect_batch(self.h, _dp(o), _dp(d), n, t_min, _dp(t), obj.ctypes.data_as(U32P), prim.ctypes.data_as(U32P),
                                            inst.ctypes.data_as(U32P), _dp(nrm), threads), "intersect_batch")
        return t, obj, prim, inst, nrm

    def kd_stats(self, mesh_id):
        out = (C.c_uint64 * 3)()
        self._ck(self.L.orc_mesh_kd_stats(self.h, mesh_id, out), "kd_stats")
        return tuple(out)

    def photon_render(self, camera, params, photon_count, kind, threads=0):
        w, h = params.width, params.height
        img = np.empty((h, w, 3), np.uint8)
        mean = np.empty((h, w, 3), np.float64)
        counts = (C.c_uint64 * 2)()
        secs = (C.c_double * 3)()
        cam = camera.lower()
        self._ck(self.L.orc_photon_render(self.h, C.byref(cam), C.byref(params), photon_count, kind, threads, img.ctypes.data_as(U8P),
                                          _dp(mean), counts, secs), "photon_render")
        return img, mean, tuple(counts), tuple(secs)

    def emit_photons(self, camera, params, photon_count, kind, first=0, count=None, threads=0):
        """PASS 1 -> (surface rows [n,13], volume rows [n,13]) sorted by key."""
        count = photon_count - first if count is None else count
        cam = camera.lower()
        h = self.L.orc_photons_emit(self.h, C.byref(cam), C.byref(params), photon_count, first, count, kind, threads)
        if not h:
            raise RuntimeError(self.L.orc_last_error(self.h).decode())
        c = (C.c_uint64 * 2)()
        self.L.orc_photons_count(h, c)
        out = []
        for which in range(2):
            rows = np.empty((c[which], 13), np.float64)
            self.L.orc_photons_rows(h, which, _dp(rows))
            out.append(sort_rows(rows))
        self.L.orc_photons_destroy(h)
        return out

    def __del__(self):
        try:
            self.L.orc_scene_destroy(self.h)
        except Exception:
            pass


def sort_rows(rows):
    """Canonical order of photon rows: ascending u64 key (column 12)."""
    if len(rows) == 0:
        return rows
    return rows[np.argsort(np.ascontiguousarray(rows[:, 12]).view(np.uint64), kind="stable")]


class OraclePhotonMap:
    """PASS 2 + PASS 3 of the oracle over given photon rows."""

    def __init__(self, oscene, surface_rows, volume_rows, kind):
        self.L, self.oscene = oscene.L, oscene
        s = np.ascontiguousarray(surface_rows, np.float64).reshape(-1, 13)
        v = np.ascontiguousarray(volume_rows, np.float64).reshape(-1, 13)
        self.h = self.L.orc_photon_map_build(oscene.h, _dp(s), len(s), _dp(v), len(v), kind)

    def knn(self, which, q, k):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, 3)
        n = len(q)
        keys, d2, counts = np.empty((n, k), np.uint64), np.empty((n, k), np.float64), np.empty(n, np.uint32)
        self.L.orc_photon_map_knn(self.h, which, _dp(q), n, k, keys.ctypes.data_as(U64P), _dp(d2), counts.ctypes.data_as(U32P))
        return keys, d2, counts

    def spheres(self):
        n = self.L.orc_photon_map_spheres(self.h, None, None)
        c, r = np.empty((n, 3), np.float64), np.empty(n, np.float64)
        self.L.orc_photon_map_spheres(self.h, _dp(c), _dp(r))
        return c, r

    def gather(self, camera, params, threads=0):
        w, h = params.width, params.height
        mean = np.empty((h, w, 3), np.float64)
        img = np.empty((h, w, 3), np.uint8)
        sec = C.c_double()
        cam = camera.lower()
        self.L.orc_photon_map_gather(self.h, C.byref(cam), C.byref(params), threads, img.ctypes.data_as(U8P), _dp(mean), C.byref(sec))
        return img, mean, sec.value

    def __del__(self):
        try:
            self.L.orc_photon_map_destroy(self.h)
        except Exception:
            pass


class OracleFilm:
    def __init__(self, oscene, camera, params, threads=0):
        self.L, self.oscene = oscene.L, oscene
        self.w, self.h_px = params.width, params.height
        cam = camera.lower()
        self.h = self.L.orc_render_begin(oscene.h, C.byref(cam), C.byref(params), threads)

    def sample(self, iterations):
        self.L.orc_render_sample(self.h, iterations)

    def seconds(self):
        return self.L.orc_film_last_seconds(self.h)

    def time_rows(self, y0, y1, stride, iterations):
        ctr = (C.c_uint64 * 4)()
        sec = self.L.orc_time_rows(self.h, y0, y1, stride, iterations, ctr)
        return sec, tuple(ctr)

    def counters(self):
        out = (C.c_uint64 * 4)()
        self.L.orc_film_counters(self.h, out)
        return dict(extend_rays=out[0], shadow_rays=out[1], shadow_pass=out[2], shadow_near=out[3])

    def image(self):
        out = np.empty((self.h_px, self.w, 3), np.uint8)
        self.L.orc_film_image_rgb8(self.h, out.ctypes.data_as(U8P))
        return out

    def variance(self):
        v = C.c_double()
        self.L.orc_film_variance(self.h, C.byref(v))
        return v.value

    def mean(self):
        out = np.empty((self.h_px, self.w, 3), np.float64)
        self.L.orc_film_mean(self.h, _dp(out))
        return out

    def sample_radiance(self, px, py, sample):
        px, py = np.ascontiguousarray(px, np.uint32), np.ascontiguousarray(py, np.uint32)
        sample = np.ascontiguousarray(sample, np.uint64)
        out = np.empty((len(px), 3), np.float64)
        self.L.orc_sample_radiance(self.h, px.ctypes.data_as(U32P), py.ctypes.data_as(U32P), sample.ctypes.data_as(U64P), len(px), _dp(out))
        return out

    def dump_rays(self, iterations, cap):
        o, d = np.empty((cap, 3), np.float64), np.empty((cap, 3), np.float64)
        kind = np.empty(cap, np.uint8)
        n = self.L.orc_dump_rays(self.h, iterations, cap, _dp(o), _dp(d), kind.ctypes.data_as(U8P))
        return o[:n].copy(), d[:n].copy(), kind[:n].copy()

    def __del__(self):
        try:
            self.L.orc_film_destroy(self.h)
        except Exception:
            pass
