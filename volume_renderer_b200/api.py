"""Host-side mirror of the reference's MATLAB API (src/matlab/VolumeRender/*.m).

The reference's host language is MATLAB, which this image does not have; the classes below
keep the MATLAB names, argument meaning and error behaviour so a user script (or a parity
test) reads like `examples/example1.m`.  Everything below talks to the GPU through the C ABI
of include/vr_abi.h exactly as the mex gateway would:

    VolumeRender.m:180   volumeRender('new')            -> vr_create
    VolumeRender.m:185   volumeRender('delete', h)      -> vr_destroy
    VolumeRender.m:201   volumeRender('sync_volumes'..) -> vr_sync_volumes
    VolumeRender.m:227   volumeRender('mem_info', h)    -> vr_mem_info
    VolumeRender.m:565   volumeRender('render', ...)    -> vr_render
    HenyeyGreenstein(n)                                 -> vr_henyey_greenstein
    timestamp                                           -> vr_timestamp_ms

Arrays are numpy arrays in MATLAB (column-major) layout: a volume `Data` of shape (d1, d2, d3)
has d1 as the fastest axis; an image is (H, W, 3).
"""
import ctypes as C
import enum
import math
import threading
import time

import numpy as np

from . import _abi

__all__ = [
    "VolumeRender",
    "Volume",
    "DeviceVolume",
    "LightSource",
    "StereoRenderMode",
    "HenyeyGreenstein",
    "timestamp",
]

_ts_lock = threading.Lock()
_ts_last = 0


def timestamp():
    """timestamp mex (src/C/mex/timestamp.cpp): milliseconds since the epoch as uint64.

    Two deliberate differences, neither visible in rendered pixels: the value is not truncated
    to 32 bits (timestamp.cpp:25 writes through an int*), and it is strictly increasing inside
    a process so that two updates in the same millisecond are still ordered (the reference's
    dirty test is `last_update > timeLastMemSync`, volumeRender_kernel.cu:749-751).
    """
    global _ts_last
    with _ts_lock:
        now = time.time_ns() // 1000000 if _ts_clock is None else int(_ts_clock())
        if now <= _ts_last:
            now = _ts_last + 1
        _ts_last = now
        return now


_ts_clock = None  # test hook: callable returning ms


class StereoRenderMode(enum.Enum):
    """StereoRenderMode.m:1-40"""

    RedCyan = 0
    LeftRightHorizontal = 1


class _TrackedArray(np.ndarray):
    """ndarray whose in-place writes stamp the owning Volume, like MATLAB's indexed assignment
    into a property runs set.Data + the PostSet listener (Volume.m:82-90, 225-238)."""

    _owner = None

    def __setitem__(self, key, value):
        super().__setitem__(key, value)
        if self._owner is not None:
            self._owner._stamp()

    def __array_finalize__(self, obj):
        self._owner = None


class Volume:
    """Volume.m: handle class holding single-precision `Data` and `TimeLastUpdate`."""

    def __init__(self, data):
        self.TimeLastUpdate = 0
        self._data = None
        self.Data = data

    def _stamp(self):
        self.TimeLastUpdate = timestamp()

    @property
    def Data(self):
        return self._data

    @Data.setter
    def Data(self, data):
        a = np.asfortranarray(np.asarray(data), dtype=np.float32)  # set.Data: single(data), Volume.m:89
        if a.ndim < 2:
            a = np.asfortranarray(a.reshape((-1, 1) if a.ndim == 1 else (1, 1)))
        t = a.view(_TrackedArray)
        t._owner = self
        self._data = t
        self._stamp()  # PostSet listener, Volume.m:225-238

    def size(self):
        return tuple(self._data.shape)

    def resize(self, newsize):
        """Volume.m:92-106 uses imresize3/imresize (proprietary, cubic + antialiasing).  This
        mirror resamples with order-1 splines; results differ from MATLAB's, so the synthetic
        configs never depend on it for parity."""
        from scipy import ndimage

        a = np.asarray(self._data)
        if np.isscalar(newsize):
            zoom = [float(newsize)] * a.ndim
        else:
            zoom = [float(n) / s for n, s in zip(newsize, a.shape)]
        self.Data = ndimage.zoom(a, zoom, order=1, mode="nearest", grid_mode=True)

    def pad(self, padding, value):
        """Volume.m:113-125: pad all three dims by `padding` on both sides."""
        if self._data.ndim == 3:
            self.Data = np.pad(np.asarray(self._data), int(padding), mode="constant", constant_values=value)

    def mip(self):
        """Volume.m:127-131: max over dim 3 of permute(Data, [2 1 3])."""
        a = np.asarray(self._data)
        a = a if a.ndim == 3 else a[:, :, None]
        return np.transpose(a, (1, 0, 2)).max(axis=2)

    def mean(self):
        return float(np.asarray(self._data).mean())

    def max(self):
        return float(np.asarray(self._data).max())

    def min(self):
        return float(np.asarray(self._data).min())

    def grad(self, newsize=None):
        """Volume.m:181-205: [gx, gy, gz] = gradient(Data).  MATLAB's gradient returns the
        derivative along dim 2 first, then dim 1, then dim 3 (unit spacing, central differences
        inside, one-sided at the borders) - which is np.gradient per axis, reordered."""
        a = np.asarray(self._data, dtype=np.float32)
        if a.ndim == 3:
            g1, g2, g3 = np.gradient(a)
            out = [Volume(g2), Volume(g1), Volume(g3)]
        else:
            g1, g2 = np.gradient(a)
            out = [Volume(g2), Volume(g1), Volume(np.zeros_like(a))]
        if newsize is not None:
            for v in out:
                v.resize(newsize)
        return out

    def normalize(self, newMin, newMax):
        """Volume.m:207-222"""
        a = np.asarray(self._data)
        mx, mn = a.max(), a.min()
        self.Data = (a - mn) * np.float32((newMax - newMin) / (mx - mn)) + np.float32(newMin)


class DeviceVolume(Volume):
    """Extension: a volume that already lives in device memory (whole: _abi.MEM_DEVICE, or only this
    context's region plus halo: _abi.MEM_DEVICE_REGION).  `dims` are always the whole volume's."""

    def __init__(self, ptr, dims, dtype=_abi.DTYPE_F32, memspace=_abi.MEM_DEVICE, keepalive=None):
        self.TimeLastUpdate = 0
        self._keep = keepalive
        self._data = (int(ptr), tuple(int(d) for d in dims), int(dtype), int(memspace))
        self._stamp()

    @property
    def Data(self):
        return self._data

    def size(self):
        return self._data[1]


class LightSource:
    """LightSource.m: Position and Color as single 1x3."""

    def __init__(self, pos, col):
        self.Position = pos
        self.Color = col

    @staticmethod
    def _vec3(val, what):
        a = np.asarray(val, dtype=np.float32)
        if a.shape == (3,) or a.shape == (1, 3) or a.shape == (3, 1):
            return a.reshape(3).copy()
        raise ValueError("dimensions of %s must be [1,3]" % what)  # LightSource.m:56,62

    @property
    def Position(self):
        return self._pos

    @Position.setter
    def Position(self, val):
        self._pos = self._vec3(val, "position")

    @property
    def Color(self):
        return self._col

    @Color.setter
    def Color(self, val):
        self._col = self._vec3(val, "color")


def HenyeyGreenstein(n, g=0.8):
    """HenyeyGreenstein mex (src/C/mex/HenyeyGreenstein.cc:29-96): n x n x n single LUT of the
    Henyey-Greenstein phase function, index k = c*n^2 + a*n + b (dims (b, a, c), b fastest)."""
    n = int(n)
    if g > 1 or g < -1:
        raise ValueError("g must be in interval [-1,1]")  # HenyeyGreenstein.cc:44-45
    out = np.zeros(n * n * n, dtype=np.float32)
    _abi.check(_abi.lib().vr_henyey_greenstein(n, float(g), out.ctypes.data))
    return out.reshape((n, n, n), order="F")


class _CudaBackend:
    """The mex command protocol of render.cpp on top of the C ABI."""

    def new(self):
        h = C.c_void_p()
        _abi.check(_abi.lib().vr_create(C.byref(h)))
        return h

    def delete(self, h):
        return _abi.lib().vr_destroy(h)

    def mem_info(self, h):
        need = C.c_size_t(0)
        _abi.check(_abi.lib().vr_mem_info(h, None, 0, C.byref(need)), h)
        buf = C.create_string_buffer(need.value + 1)
        _abi.check(_abi.lib().vr_mem_info(h, buf, need.value + 1, C.byref(need)), h)
        return buf.value.decode()

    def sync_volumes(self, h, T, em, re, ab, gx=None, gy=None, gz=None):
        keep = []
        vols = [_abi.volume_struct(v.Data if isinstance(v, Volume) else v[0], v.TimeLastUpdate if isinstance(v, Volume) else v[1], keep) if v is not None else None for v in (em, re, ab, gx, gy, gz)]
        ptrs = [C.byref(v) if v is not None else None for v in vols]
        _abi.check(_abi.lib().vr_sync_volumes(h, int(T), *ptrs), h)

    def render(self, h, lights, lut, factors, es, res, rot, props, thr, color):
        H, W = int(res[0]), int(res[1])
        out = np.zeros((H, W, 3), dtype=np.float32, order="F")
        args = _render_args(lights, lut, factors, es, res, rot, props, thr, color)
        _abi.check(_abi.lib().vr_render(h, *args[0], out.ctypes.data), h)
        return out

    # extensions with no mex counterpart (include/vr_abi.h, "extensions")
    def render_batch(self, h, lights, lut, factors, es, res, rots, props_list, thr, color):
        """Frames that differ only in camera, in one launch (vr_render_batch)."""
        H, W = int(res[0]), int(res[1])
        n = len(rots)
        outs = [np.zeros((H, W, 3), dtype=np.float32, order="F") for _ in range(n)]
        args, keep = _render_args(lights, lut, factors, es, res, rots[0], props_list[0], thr, color)
        R = np.concatenate([np.asfortranarray(np.asarray(r, dtype=np.float32)).reshape(-1, order="F") for r in rots]).astype(np.float32)
        Pp = np.concatenate([np.asarray(p, dtype=np.float32).reshape(3) for p in props_list]).astype(np.float32)
        ptrs = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
        f32p = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
        _abi.check(_abi.lib().vr_render_batch(h, args[0], args[1], args[2], args[3], args[4], args[5], n, f32p(R), f32p(Pp),
                                              args[8], args[9], ptrs), h)
        return outs

    def render_stereo(self, h, lights, lut, factors, es, res_wide, rot, props_right, props_left, thr, color, delta, mode):
        """Both eyes in one launch and the crop + compose of VolumeRender.m:286-306 on the device: only the
        final image crosses PCIe."""
        lib = _abi.lib()
        H, Wp = int(res_wide[0]), int(res_wide[1])
        Wo = (Wp - delta) * (1 if mode == 0 else 2)
        nb = H * Wp * 3 * 4
        bufs = [C.c_void_p() for _ in range(3)]
        for b, size in zip(bufs, (nb, nb, H * Wo * 3 * 4)):
            _abi.check(lib.vr_device_alloc(h, size, C.byref(b)), h)
        try:
            args, keep = _render_args(lights, lut, factors, es, res_wide, rot, props_right, thr, color)
            R = np.tile(np.asfortranarray(np.asarray(rot, dtype=np.float32)).reshape(-1, order="F"), 2).astype(np.float32)
            Pp = np.concatenate([np.asarray(props_right, dtype=np.float32), np.asarray(props_left, dtype=np.float32)]).astype(np.float32)
            f32p = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
            outs = (C.c_void_p * 2)(bufs[0].value, bufs[1].value)  # right eye, left eye
            _abi.check(lib.vr_render_batch_device(h, args[0], args[1], args[2], args[3], args[4], args[5], 2, f32p(R), f32p(Pp), args[8], args[9], outs, None), h)
            rw = np.asarray(res_wide, dtype=np.uint64)
            _abi.check(lib.vr_compose_stereo_device(h, bufs[1], bufs[0], rw.ctypes.data_as(C.POINTER(C.c_uint64)), int(delta), int(mode), bufs[2], None), h)
            out = np.zeros((H, Wo, 3), dtype=np.float32, order="F")
            _abi.check(lib.vr_device_download(h, out.ctypes.data, bufs[2], out.nbytes), h)
            return out
        finally:
            for b in bufs:
                if b:
                    lib.vr_device_free(h, b)

    def stats(self, h):
        s = _abi.VrStats()
        _abi.check(_abi.lib().vr_get_stats(h, C.byref(s)), h)
        return s.as_dict()

    def set_option(self, h, option, value):
        _abi.check(_abi.lib().vr_set_option(h, int(option), int(value)), h)


def _render_args(lights, lut, factors, es, res, rot, props, thr, color):
    """Pack the 'render' arguments the way MATLAB lays them out in memory (render.cpp:142-240)."""
    keep = []
    if lights is not None and lut is not None:
        arr = (_abi.VrLight * len(lights))()
        for i, l in enumerate(lights):
            arr[i].position[:] = [float(x) for x in l.Position]
            arr[i].color[:] = [float(x) for x in l.Color]
        lutv = _abi.volume_struct(lut.Data, lut.TimeLastUpdate, keep)
        lp, ln, lv = arr, len(lights), C.byref(lutv)
        keep += [arr, lutv]
    else:
        lp, ln, lv = None, 0, None

    def f32(a, n):
        a = np.asfortranarray(np.asarray(a, dtype=np.float32))
        assert a.size == n, (a.shape, n)
        keep.append(a)
        return a.ctypes.data_as(C.POINTER(C.c_float))

    r = np.asarray(res, dtype=np.uint64)
    keep.append(r)
    packed = [
        lp, ln, lv,
        f32(factors, 3), f32(es, 3), r.ctypes.data_as(C.POINTER(C.c_uint64)),
        f32(rot, 9), f32(props, 3), C.c_float(float(thr)), f32(color, 3),
    ]
    return packed, keep


def _round_half_away(x):
    """MATLAB round()"""
    return int(math.floor(abs(x) + 0.5) * (1 if x >= 0 else -1))


def _cosd(a):
    return math.cos(math.radians(a))


def _sind(a):
    return math.sin(math.radians(a))


class VolumeRender:
    """VolumeRender.m: the renderer handle class (properties keep their MATLAB names)."""

    _backend_cls = _CudaBackend
    _volume_props = (
        "VolumeEmission", "VolumeAbsorption", "VolumeReflection",
        "VolumeGradientX", "VolumeGradientY", "VolumeGradientZ", "VolumeIllumination",
    )

    def __init__(self):
        object.__setattr__(self, "_vols", {})
        self.FocalLength = 0.0
        self.DistanceToObject = 0.0
        self.OpacityThreshold = 0.95
        self.LightSources = False
        self.Color = [1, 1, 1]
        self.FactorEmission = 1.0
        self.FactorReflection = 1.0
        self.FactorAbsorption = 1.0
        self.CameraXOffset = 0
        self.StereoOutput = StereoRenderMode.RedCyan
        self.ElementSizeUm = [1, 1, 1]
        self.RotationMatrix = np.eye(3)
        self.ImageResolution = [0, 0]
        self.TimeLastMemSync = 0
        self.BatchStereo = True  # extension: True = the stereo pair as one batch launch, "device" = also crop + compose on the GPU, False = two passes
        self._vols.update({k: False for k in self._volume_props})
        self._vols["VolumeReflection"] = Volume(1)  # VolumeRender.m:131
        self._backend = self._backend_cls()
        self.objectHandle = self._backend.new()  # VolumeRender.m:180

    # --- observable volume properties: PostSet stamps TimeLastUpdate (VolumeRender.m:723-740) ---
    def __getattr__(self, name):
        vols = object.__getattribute__(self, "_vols")
        if name in vols:
            return vols[name]
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in self._volume_props:
            if name == "VolumeAbsorption" and not isinstance(value, Volume):
                raise TypeError("VolumeAbsorption must be of type Volume")  # VolumeRender.m:491
            if not (isinstance(value, Volume) or value is False):
                raise TypeError("%s must be of type Volume" % name)
            self._vols[name] = value
            if isinstance(value, Volume):
                value.TimeLastUpdate = timestamp()
            return
        if name == "Color" or name == "ElementSizeUm":
            a = np.asarray(value, dtype=np.float64).reshape(-1)
            if a.size != 3:
                raise ValueError("dimensions of %s must be [1,3]" % name)
            value = a
        elif name == "ImageResolution":
            a = np.asarray(value).reshape(-1)
            if a.size != 2:
                raise ValueError("dimensions of ImageResolution must be [1,2]")
            value = np.array([int(a[0]), int(a[1])])
        elif name == "RotationMatrix":
            a = np.asarray(value, dtype=np.float64)
            if a.shape != (3, 3):
                raise ValueError("dimensions of RotationMatrix must be [3,3]")
            value = a.copy()
        elif name == "LightSources":
            if value is not False and isinstance(value, LightSource):
                value = [value]
        object.__setattr__(self, name, value)

    def delete(self):
        """VolumeRender.m:184-186"""
        if getattr(self, "objectHandle", None) is not None:
            self._backend.delete(self.objectHandle)
            object.__setattr__(self, "objectHandle", None)

    def __del__(self):
        try:
            self.delete()
        except Exception:
            pass

    def syncVolumes(self):
        """VolumeRender.m:188-219"""
        g = [self.VolumeGradientX, self.VolumeGradientY, self.VolumeGradientZ]
        if all(v is not False for v in g):
            self._backend.sync_volumes(self.objectHandle, self.TimeLastMemSync, self.VolumeEmission,
                                       self.VolumeReflection, self.VolumeAbsorption, *g)
        else:
            self._backend.sync_volumes(self.objectHandle, self.TimeLastMemSync, self.VolumeEmission,
                                       self.VolumeReflection, self.VolumeAbsorption)
        self.TimeLastMemSync = timestamp()

    ROLES = {"emission": 0, "absorption": 1, "reflection": 2, "gradientx": 3, "gradienty": 4, "gradientz": 5}

    def updateVolumeBox(self, role, lo, hi):
        """Extension (vr_update_box): the voxel box [lo, hi) of the volume that `role`
        ("emission", "absorption", "reflection", "gradientx|y|z") refers to has been rewritten in place on the host
        (the example3.m:135-161 pattern) - move only that box to the GPU instead of the whole volume.  The volume
        object is re-stamped and counts as synced afterwards."""
        attr = {0: "VolumeEmission", 1: "VolumeAbsorption", 2: "VolumeReflection", 3: "VolumeGradientX", 4: "VolumeGradientY",
                5: "VolumeGradientZ"}[self.ROLES[role.lower()]]
        vol = getattr(self, attr)
        if getattr(self, "TimeLastMemSync", 0) == 0:
            self.syncVolumes()  # nothing resident yet: the first upload is a whole one
            return
        vol._stamp()
        keep = []
        v = _abi.volume_struct(vol.Data, vol.TimeLastUpdate, keep)
        l3, h3 = (C.c_uint64 * 3)(*[int(x) for x in lo]), (C.c_uint64 * 3)(*[int(x) for x in hi])
        _abi.check(_abi.lib().vr_update_box(self.objectHandle, self.ROLES[role.lower()], C.byref(v), l3, h3), self.objectHandle)
        self.TimeLastMemSync = timestamp()

    def memInfo(self):
        """VolumeRender.m:221-228 (prints like mexPrintf, and returns the text)"""
        s = self._backend.mem_info(self.objectHandle)
        print(s, end="")
        return s

    def stats(self):
        """Counters of the last render (vr_get_stats); no MATLAB counterpart."""
        return self._backend.stats(self.objectHandle)

    def setOption(self, option, value):
        """vr_set_option; no MATLAB counterpart (defaults reproduce the reference as written)."""
        self._backend.set_option(self.objectHandle, option, value)

    def setDevices(self, device_ids):
        """vr_set_devices; no MATLAB counterpart in the reference (it renders on one device).  The gateway command is
        volumeRender('set_devices', handle, int32 ids): this object then renders on all the listed GPUs - volumes resident on
        each, tile columns dealt round-robin - and returns the same H x W x 3 image.  Call before the first render."""
        ids = (C.c_int * len(device_ids))(*[int(d) for d in device_ids])
        _abi.check(_abi.lib().vr_set_devices(self.objectHandle, len(device_ids), ids), self.objectHandle)

    def memClear(self):
        """VolumeRender.m:230-237: 'delete' on the handle; the later destructor hits a stale handle."""
        self._backend.delete(self.objectHandle)

    def rotate(self, alpha, beta, gamma):
        """VolumeRender.m:239-262"""
        Rx = np.array([[1, 0, 0], [0, _cosd(alpha), -_sind(alpha)], [0, _sind(alpha), _cosd(alpha)]])
        Ry = np.array([[_cosd(beta), 0, _sind(beta)], [0, 1, 0], [-_sind(beta), 0, _cosd(beta)]])
        Rz = np.array([[_cosd(gamma), -_sind(gamma), 0], [_sind(gamma), _cosd(gamma), 0], [0, 0, 1]])
        self.RotationMatrix = self.RotationMatrix @ Rx @ Ry @ Rz

    def stereo_delta(self):
        """VolumeRender.m:277-282"""
        base = self.CameraXOffset / 2
        fov = 2 * math.atan(1 / self.FocalLength)
        return base, _round_half_away((base * self.ImageResolution[1]) / (2 * self.FocalLength * math.tan(fov / 2)))

    def render(self):
        """VolumeRender.m:264-309"""
        flipped = [int(self.ImageResolution[1]), int(self.ImageResolution[0])]
        if self.CameraXOffset == 0:
            return self.p_render(np.float32(self.CameraXOffset), flipped)
        base, delta = self.stereo_delta()
        resolution = [flipped[0], flipped[1] + delta]
        if hasattr(self._backend, "render_stereo") and self.BatchStereo == "device":
            # both eyes in one launch, crop + compose on the device (vr_compose_stereo_device)
            self._check_volumes()
            self.syncVolumes()
            mode = 0 if self.StereoOutput == StereoRenderMode.RedCyan else 1
            img = self._backend.render_stereo(self.objectHandle, self.LightSources if self.LightSources is not False else None,
                                              self.VolumeIllumination if self.VolumeIllumination is not False else None,
                                              [self.FactorEmission, self.FactorReflection, self.FactorAbsorption], self.ElementSizeUm, resolution,
                                              np.flipud(np.asarray(self.RotationMatrix)), [base, self.FocalLength, self.DistanceToObject],
                                              [-base, self.FocalLength, self.DistanceToObject], self.OpacityThreshold, self.Color, delta, mode)
            return img.astype(np.float64) if mode == 0 else img
        if hasattr(self._backend, "render_batch") and self.BatchStereo:
            # both eyes in one launch (same pixels as two p_render passes; one work queue)
            rightImage, leftImage = self.p_render_batch([base, -base], resolution)
        else:
            rightImage = self.p_render(base, resolution)
            leftImage = self.p_render(-base, resolution)
        # imcrop(left, [delta+1 0 W' H]) keeps columns delta+1..W'; imcrop(right, [0 0 W'-delta H]) keeps 1..W'-delta
        leftImage = leftImage[:, delta:, :]
        rightImage = rightImage[:, : rightImage.shape[1] - delta, :]
        if self.StereoOutput == StereoRenderMode.RedCyan:
            image = np.zeros(leftImage.shape, dtype=np.float64, order="F")
            image[:, :, 0] = leftImage[:, :, 0]
            image[:, :, 1] = rightImage[:, :, 1]
            image[:, :, 2] = rightImage[:, :, 2]
            return image
        return np.concatenate([leftImage, rightImage], axis=1)

    def resetGradientVolumes(self):
        """VolumeRender.m:312-345"""
        self.VolumeGradientX = False
        self.VolumeGradientY = False
        self.VolumeGradientZ = False

    def p_render(self, CameraXOffset, resolution):
        """VolumeRender.m:497-583"""
        validate = [self.VolumeReflection is False, self.VolumeAbsorption is False, self.VolumeEmission is False]
        if self.VolumeIllumination is False:
            import warnings

            warnings.warn("VolumeIllumination is unset. Thus no lightning will be applied!")
        if all(validate):
            raise RuntimeError("Not all volumes are properly set!")
        if self.VolumeEmission is False or self.VolumeAbsorption is False or self.VolumeReflection is False:
            # the mex would dereference a logical as a Volume object here; fail cleanly instead
            raise RuntimeError("Not all volumes are properly set!")
        self.syncVolumes()
        factors = [self.FactorEmission, self.FactorReflection, self.FactorAbsorption]
        props = [CameraXOffset, self.FocalLength, self.DistanceToObject]
        matrix = np.flipud(np.asarray(self.RotationMatrix))  # flip(RotationMatrix): rows reversed
        lights = self.LightSources if self.LightSources is not False else None
        lut = self.VolumeIllumination if self.VolumeIllumination is not False else None
        return self._backend.render(self.objectHandle, lights, lut, factors, self.ElementSizeUm, resolution,
                                    matrix, props, self.OpacityThreshold, self.Color)

    def _check_volumes(self):
        if self.VolumeIllumination is False:
            import warnings

            warnings.warn("VolumeIllumination is unset. Thus no lightning will be applied!")
        if self.VolumeEmission is False or self.VolumeAbsorption is False or self.VolumeReflection is False:
            raise RuntimeError("Not all volumes are properly set!")

    def p_render_batch(self, offsets, resolution, rotations=None):
        """Extension: p_render for several cameras at once - `offsets` are CameraXOffsets, `rotations`
        (optional) RotationMatrix values per frame.  Same marshalling as p_render per frame."""
        self._check_volumes()
        self.syncVolumes()
        factors = [self.FactorEmission, self.FactorReflection, self.FactorAbsorption]
        rotations = rotations if rotations is not None else [self.RotationMatrix] * len(offsets)
        rots = [np.flipud(np.asarray(R)) for R in rotations]
        props = [[o, self.FocalLength, self.DistanceToObject] for o in offsets]
        lights = self.LightSources if self.LightSources is not False else None
        lut = self.VolumeIllumination if self.VolumeIllumination is not False else None
        return self._backend.render_batch(self.objectHandle, lights, lut, factors, self.ElementSizeUm, resolution, rots, props,
                                          self.OpacityThreshold, self.Color)

    # --- volume-space (sort-last) extension: this object owns one voxel box of every volume -------
    def setRegion(self, lo=None, hi=None, halo=0):
        """vr_set_region: volumes synced afterwards keep only [lo, hi) (+ halo) resident."""
        lib = _abi.lib()
        if lo is None:
            _abi.check(lib.vr_set_region(self.objectHandle, None, None, 0), self.objectHandle)
            return
        l3, h3 = (C.c_uint64 * 3)(*[int(v) for v in lo]), (C.c_uint64 * 3)(*[int(v) for v in hi])
        _abi.check(lib.vr_set_region(self.objectHandle, l3, h3, int(halo)), self.objectHandle)

    def _packed_mono_args(self):
        """p_render's marshalling (VolumeRender.m:497-583) for a mono frame, for the device entry points."""
        self._check_volumes()
        self.syncVolumes()
        factors = [self.FactorEmission, self.FactorReflection, self.FactorAbsorption]
        props = [np.float32(self.CameraXOffset), self.FocalLength, self.DistanceToObject]
        res = [int(self.ImageResolution[1]), int(self.ImageResolution[0])]
        lights = self.LightSources if self.LightSources is not False else None
        lut = self.VolumeIllumination if self.VolumeIllumination is not False else None
        return _render_args(lights, lut, factors, self.ElementSizeUm, res, np.flipud(np.asarray(self.RotationMatrix)), props,
                            self.OpacityThreshold, self.Color)

    def renderOpacity(self, d_opacity, stream=None, packed=None):
        """Pass 1 of a sort-last frame: this brick's own opacity per pixel into device memory (H*W floats).
        `packed` = the result of _packed_mono_args() when the caller packs once per frame for both passes."""
        args, keep = packed if packed is not None else self._packed_mono_args()
        _abi.check(_abi.lib().vr_render_opacity_device(self.objectHandle, *args, C.c_void_p(d_opacity), C.c_void_p(stream or 0)), self.objectHandle)

    def renderPartial(self, bricks, index, d_opacity_all, d_out, stream=None, packed=None):
        """Pass 2: this brick's additive share of the image (H*W*3 floats in device memory)."""
        from . import multigpu

        args, keep = packed if packed is not None else self._packed_mono_args()
        boxes = multigpu.boxes_array(bricks)
        _abi.check(_abi.lib().vr_render_partial_device(self.objectHandle, *args, len(bricks), int(index), boxes.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                       C.c_void_p(d_opacity_all), C.c_void_p(d_out), C.c_void_p(stream or 0)), self.objectHandle)

    def renderMoviePipelined(self, rotations, batch=8, out=None):
        """Extension: a long rotation movie (examples/example2.m:59-66) through vr_render_batch_begin / _end - batch k's
        frames travel to the host while batch k+1 is rendered.  `out` (optional) is a list of page-locked H x W x 3 arrays."""
        self._check_volumes()
        self.syncVolumes()
        H, W = int(self.ImageResolution[1]), int(self.ImageResolution[0])
        n = len(rotations)
        outs = out if out is not None else [np.zeros((H, W, 3), dtype=np.float32, order="F") for _ in range(n)]
        lib, h = _abi.lib(), self.objectHandle
        f32p = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
        keepalive = []
        try:
            for k in range(0, n, batch):
                nb = min(batch, n - k)
                args, keep = _render_args(self.LightSources if self.LightSources is not False else None,
                                          self.VolumeIllumination if self.VolumeIllumination is not False else None,
                                          [self.FactorEmission, self.FactorReflection, self.FactorAbsorption], self.ElementSizeUm, [H, W],
                                          np.flipud(np.asarray(rotations[k])), [np.float32(self.CameraXOffset), self.FocalLength, self.DistanceToObject],
                                          self.OpacityThreshold, self.Color)
                R = np.concatenate([np.asfortranarray(np.flipud(np.asarray(r)).astype(np.float32)).reshape(-1, order="F") for r in rotations[k:k + nb]]).astype(np.float32)
                Pp = np.tile(np.array([self.CameraXOffset, self.FocalLength, self.DistanceToObject], dtype=np.float32), nb)
                ptrs = (C.c_void_p * nb)(*[o.ctypes.data for o in outs[k:k + nb]])
                keepalive += [keep, R, Pp, ptrs]
                _abi.check(lib.vr_render_batch_begin(h, args[0], args[1], args[2], args[3], args[4], args[5], nb, f32p(R), f32p(Pp), args[8], args[9], ptrs), h)
        finally:
            _abi.check(lib.vr_render_batch_end(h), h)
        return outs

    def renderMovie(self, rotations):
        """Extension: the frames of a rotation movie (examples/example2.m:59-66) in one launch;
        `rotations` is a list of RotationMatrix values.  Mono cameras only."""
        flipped = [int(self.ImageResolution[1]), int(self.ImageResolution[0])]
        return self.p_render_batch([np.float32(self.CameraXOffset)] * len(rotations), flipped, rotations)

    @staticmethod
    def normalizeSequence(sequence):
        """VolumeRender.m:587-630"""
        sequence = np.asarray(sequence)
        if sequence.ndim < 4:
            raise ValueError("input must be a multiframe image (4D)")
        mx, mn = sequence.max(), sequence.min()
        out = np.zeros(sequence.shape)
        for i in range(sequence.shape[3]):
            out[:, :, :, i] = VolumeRender.normalizeImage(sequence[:, :, :, i], mn, mx)
        return out

    @staticmethod
    def normalizeImage(ImageRGB, *args):
        """VolumeRender.m:632-701 (including its `+ minValue` shift for negative minima)"""
        img = np.asarray(ImageRGB, dtype=np.float64)
        r, g, b = img[:, :, 0], img[:, :, 1], img[:, :, 2]
        mn = args[0] if len(args) >= 1 else min(r.min(), g.min(), b.min())
        mx = args[1] if len(args) >= 2 else max(r.max(), g.max(), b.max())
        if mn < 0:
            r, g, b = r + mn, g + mn, b + mn
            mx = mx + abs(mn)
        out = np.zeros(img.shape)
        out[:, :, 0], out[:, :, 1], out[:, :, 2] = r / mx, g / mx, b / mx
        return out
