"""ctypes binding of libzb200.so (include/zb200.h).  There is no CPU fallback: if the CUDA
library is missing or no GPU is visible, every compute entry point raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import params as P

# Concurrent contexts (one stream each) share the GPU's hardware work queues; with the default of 8 queues a
# stream that waits for its input copy holds up other pairs' kernels queued behind it (measured on B200, 24
# contexts: 227 -> 236 pairs/s device-resident, 198 -> 219 end to end with 32 queues).  Read by the CUDA
# runtime when the process creates its context, so it is set here, before the first CUDA call, unless the
# caller chose a value.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ZB_LIB") or os.path.join(_HERE, "libzb200.so")      # ZB_LIB: an experimental build (tools/)
ZB_NSTAGES = 12
MASK_DTYPES = {"uint8": 1, "bool": 1, "int16": 2, "int32": 4}

EXPORTS = ("zb_version", "zb_params_default", "zb_params_check", "zb_sizes_of", "zb_create", "zb_destroy",
           "zb_last_error", "zb_subtract", "zb_subtract_host", "zb_subtract_batch", "zb_last_timings", "zb_stage_name", "zb_launch_count", "zb_abi_sizeof",
           # row-tile sessions (bound by tiles.py)
           "zb_tile_plan", "zb_tile_begin", "zb_tile_begin_retry", "zb_tile_cells_export", "zb_tile_cells_import",
           "zb_tile_resolve", "zb_tile_research", "zb_tile_rows", "zb_tile_fill", "zb_tile_normal_export",
           "zb_tile_normal_import", "zb_tile_solve", "zb_tile_sig_export", "zb_tile_sig_import", "zb_tile_clip",
           "zb_tile_convolve", "zb_tile_finish",
           # wrapper pre/post steps (zuds.utils.quick_background_estimate, mask OR / boolean / bit 17, rms <-> weight)
           "zb_background_estimate", "zb_background_estimate_host", "zb_background_estimate_static",
           "zb_mask_combine", "zb_mask_flag_sentinel", "zb_rms_from_weight", "zb_weight_from_rms",
           # alignment of the reference onto the science grid (what zuds runs SWarp for)
           "zb_align_lanczos3", "zb_align_mask")


class ZbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("zb200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load():
    """Load libzb200.so (built in-tree by build.sh / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libzb200.so is not built (run ./build.sh); the subtraction path has no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.zb_version.restype = C.c_char_p
    lib.zb_params_default.argtypes = [C.POINTER(P.ZbParams)]
    lib.zb_params_check.argtypes = [C.POINTER(P.ZbParams), C.c_char_p, C.c_int]
    lib.zb_sizes_of.argtypes = [C.POINTER(P.ZbParams), C.POINTER(P.ZbSizes)]
    lib.zb_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.zb_destroy.argtypes = [vp]
    lib.zb_last_error.argtypes = [vp]
    lib.zb_last_error.restype = C.c_char_p
    sub_args = [vp, C.POINTER(P.ZbParams), vp, vp, vp, vp, vp, C.c_int, vp, C.c_int, vp, vp, vp,
                C.POINTER(P.ZbRegionInfo), vp, C.POINTER(P.ZbStampInfo)]
    lib.zb_subtract.argtypes = sub_args + [vp]
    lib.zb_subtract_host.argtypes = sub_args
    lib.zb_subtract_batch.argtypes = [vp, C.c_int, C.POINTER(P.ZbParams), C.POINTER(P.ZbHostPair), C.POINTER(P.ZbRegionInfo), vp,
                                      C.POINTER(P.ZbStampInfo), C.POINTER(C.c_int)]
    lib.zb_last_timings.argtypes = [vp, C.POINTER(C.c_float), C.c_int]
    lib.zb_stage_name.argtypes = [C.c_int]
    lib.zb_stage_name.restype = C.c_char_p
    lib.zb_launch_count.argtypes = [vp]
    lib.zb_launch_count.restype = C.c_longlong
    dp = C.POINTER(C.c_double)
    lib.zb_background_estimate.argtypes = [vp, vp, vp, C.c_int, C.c_longlong, dp, dp, vp]
    lib.zb_background_estimate_host.argtypes = [vp, vp, vp, C.c_int, C.c_longlong, dp, dp]
    lib.zb_background_estimate_static.argtypes = [vp, vp, C.c_int, C.c_longlong, vp, vp, C.c_char_p, C.c_int]
    lib.zb_mask_combine.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_longlong, C.c_int, vp, vp, vp]
    lib.zb_mask_flag_sentinel.argtypes = [vp, vp, vp, C.c_longlong, C.c_float, C.c_int, vp]
    lib.zb_rms_from_weight.argtypes = [vp, vp, vp, vp, C.c_float, C.c_float, C.c_longlong, vp, vp]
    lib.zb_weight_from_rms.argtypes = [vp, vp, vp, vp, C.c_float, C.c_longlong, vp, vp]
    wp = C.POINTER(P.ZbWcs)
    lib.zb_align_lanczos3.argtypes = [vp, vp, C.c_int, C.c_int, wp, C.c_int, C.c_int, wp, vp, vp, vp]
    lib.zb_align_mask.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, wp, C.c_int, C.c_int, wp, vp, C.c_int, vp]
    _lib = lib
    return lib


def sizes_of(p):
    s = P.ZbSizes()
    load().zb_sizes_of(C.byref(p), C.byref(s))
    return s


def check_params(p):
    buf = C.create_string_buffer(512)
    rc = load().zb_params_check(C.byref(p), buf, 512)
    return rc, buf.value.decode()


BKG_STATIC_MAX_PIXELS = 1 << 31        # ZB_BKG_STATIC_MAX_PIXELS (include/zb200.h)


def background_estimate_static(data, mask, out, n=None, stream=None):
    """zb_background_estimate_static: the background estimate without any allocation and without a context, on the
    current CUDA device.  ``data`` (float32), ``mask`` (uint8/int16/int32 or None) and ``out`` (float64, two elements)
    are CUDA tensors; only their data_ptr() crosses the C ABI and nothing is allocated here either.  ``n`` overrides
    the pixel count (to probe the size limit).  The result lands in ``out`` in stream order; no synchronisation."""
    import torch
    assert data.is_cuda and data.is_contiguous() and data.dtype == torch.float32
    assert out.is_cuda and out.is_contiguous() and out.dtype == torch.float64 and out.numel() >= 2
    dt = 0
    if mask is not None:
        assert mask.is_cuda and mask.is_contiguous() and mask.numel() == data.numel()
        dt = MASK_DTYPES[str(mask.dtype).replace("torch.", "")]
    s = torch.cuda.current_stream(data.device).cuda_stream if stream is None else stream.cuda_stream
    buf = C.create_string_buffer(256)
    rc = load().zb_background_estimate_static(C.c_void_p(data.data_ptr()), None if mask is None else C.c_void_p(mask.data_ptr()), dt,
                                              data.numel() if n is None else n, C.c_void_p(out.data_ptr()), C.c_void_p(s), buf, 256)
    if rc != 0:
        raise ZbError(rc, buf.value.decode())


class Result:
    """Outputs of one subtraction (arrays are torch tensors for device calls, numpy for host calls)."""


class OwnStream:
    """``stream=OwnStream()``: run on the context's private non-blocking stream (ZB_STREAM_OWN, include/zb200.h).
    The default, ``stream=None``, is torch's current stream; its handle 0 is CUDA's legacy default stream."""
    cuda_stream = -1


class Context:
    """One zb_ctx: owns the device workspace of one GPU; use from one thread at a time."""

    def __init__(self, device=0):
        self.lib = load()
        self.h = C.c_void_p()
        rc = self.lib.zb_create(int(device), C.byref(self.h))
        if rc != 0:
            raise ZbError(rc, "zb_create(device=%d) failed: no usable CUDA device" % device)
        self.device = int(device)

    def close(self):
        if self.h:
            self.lib.zb_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self, rc):
        raise ZbError(rc, (self.lib.zb_last_error(self.h) or b"").decode())

    def timings(self):
        ms = (C.c_float * ZB_NSTAGES)()
        self.lib.zb_last_timings(self.h, ms, ZB_NSTAGES)
        return {self.lib.zb_stage_name(i).decode(): ms[i] for i in range(ZB_NSTAGES) if self.lib.zb_stage_name(i)}

    def launch_count(self):
        return int(self.lib.zb_launch_count(self.h))

    def _outputs(self, p):
        sz = sizes_of(p)
        r = Result()
        r.sizes = sz
        r.regions = (P.ZbRegionInfo * sz.n_regions)()
        r.sol = np.zeros((sz.n_regions, sz.n_comp_total + 1), np.float64)
        r.stamps = (P.ZbStampInfo * (sz.n_regions * sz.n_stamps))()
        return r

    def background_estimate(self, data, mask=None):
        """zuds.utils.quick_background_estimate (zuds/utils.py:32-53) on the GPU: (median, 1.4826 x MAD) of
        the pixels whose mask value is zero, as Python floats.  ``data``/``mask``: numpy arrays (copied
        to the device inside the call) or torch CUDA tensors (only data_ptr() crosses the C ABI)."""
        bkg, std = C.c_double(), C.c_double()
        if isinstance(data, np.ndarray):
            d = np.ascontiguousarray(data, np.float32)
            m, dt = None, 0
            if mask is not None:
                m = np.ascontiguousarray(mask)
                if m.dtype == np.bool_:
                    m = m.view(np.uint8)
                if m.dtype.name not in MASK_DTYPES:
                    m = m.astype(np.int32)
                dt = MASK_DTYPES[m.dtype.name]
                assert m.size == d.size
            rc = self.lib.zb_background_estimate_host(self.h, d.ctypes.data_as(C.c_void_p),
                                                      None if m is None else m.ctypes.data_as(C.c_void_p), dt,
                                                      d.size, C.byref(bkg), C.byref(std))
        else:
            import torch
            assert data.is_cuda and data.is_contiguous() and data.dtype == torch.float32 and data.device.index == self.device
            dt = 0
            if mask is not None:
                assert mask.is_cuda and mask.is_contiguous() and mask.numel() == data.numel()
                dt = MASK_DTYPES[str(mask.dtype).replace("torch.", "")]
            s = torch.cuda.current_stream(data.device).cuda_stream
            rc = self.lib.zb_background_estimate(self.h, C.c_void_p(data.data_ptr()),
                                                 None if mask is None else C.c_void_p(mask.data_ptr()), dt,
                                                 data.numel(), C.byref(bkg), C.byref(std), C.c_void_p(s))
        if rc != 0:
            self._err(rc)
        return bkg.value, std.value

    # -- the other wrapper steps: torch CUDA tensors in and out, nothing returns to the host -------------
    def _dev(self, t, dtypes=None):
        import torch
        assert isinstance(t, torch.Tensor) and t.is_cuda and t.is_contiguous() and t.device.index == self.device, \
            "contiguous CUDA tensor on device %d expected" % self.device
        if dtypes is not None:
            assert t.dtype in dtypes, "dtype %s not in %s" % (t.dtype, dtypes)
        return C.c_void_p(t.data_ptr())

    @staticmethod
    def _mask_dt(t):
        return 0 if t is None else MASK_DTYPES[str(t.dtype).replace("torch.", "")]

    def _stream(self, ref_tensor, stream):
        import torch
        if stream is None:
            stream = torch.cuda.current_stream(ref_tensor.device)
        return C.c_void_p(stream.cuda_stream)

    def mask_combine(self, sci_mask, ref_mask, bad_sum, want_or=True, want_bool=True, stream=None):
        """``badpix = ref_mask | sci_mask`` (int32) and its boolean ``(badpix & bad_sum) > 0`` (uint8):
        zuds/subtraction.py:135 and zuds/mask.py:42-72.  Either mask may be None."""
        import torch
        some = sci_mask if sci_mask is not None else ref_mask
        o = torch.empty(some.shape, dtype=torch.int32, device=some.device) if want_or else None
        b = torch.empty(some.shape, dtype=torch.uint8, device=some.device) if want_bool else None
        rc = self.lib.zb_mask_combine(self.h, None if sci_mask is None else self._dev(sci_mask), self._mask_dt(sci_mask),
                                      None if ref_mask is None else self._dev(ref_mask), self._mask_dt(ref_mask),
                                      some.numel(), int(bad_sum), None if o is None else self._dev(o),
                                      None if b is None else self._dev(b), self._stream(some, stream))
        if rc != 0:
            self._err(rc)
        return o, b

    def mask_flag_sentinel(self, mask, diff, sentinel=1e-30, bit_value=2 ** 17, stream=None):
        """In place: ``mask |= bit_value`` where ``diff == float32(sentinel)`` (zuds/subtraction.py:167-177)."""
        import torch
        rc = self.lib.zb_mask_flag_sentinel(self.h, self._dev(mask, (torch.int32,)), self._dev(diff, (torch.float32,)),
                                            mask.numel(), float(np.float32(sentinel)), int(bit_value), self._stream(mask, stream))
        if rc != 0:
            self._err(rc)
        return mask

    def rms_from_weight(self, weight, bad=None, data=None, saturate=None, big_rms=None, stream=None):
        """zuds/image.py:179-190: ``1/sqrt(weight)``; ``big_rms`` where ``bad`` or ``data >= 0.9 * saturate``."""
        import torch
        from .synth import BIG_RMS
        out = torch.empty_like(weight)
        sat = float(np.float32(0.9 * saturate)) if (saturate is not None and data is not None) else float("inf")
        rc = self.lib.zb_rms_from_weight(self.h, self._dev(weight, (torch.float32,)),
                                         None if bad is None else self._dev(bad, (torch.uint8, torch.bool)),
                                         None if (data is None or saturate is None) else self._dev(data, (torch.float32,)), sat,
                                         float(np.float32(BIG_RMS if big_rms is None else big_rms)), weight.numel(),
                                         self._dev(out), self._stream(weight, stream))
        if rc != 0:
            self._err(rc)
        return out

    def weight_from_rms(self, rms, bad=None, data=None, saturate=None, stream=None):
        """zuds/image.py:146-149: ``1/rms**2``; 0 where ``bad`` or ``data >= 0.9 * saturate``."""
        import torch
        out = torch.empty_like(rms)
        sat = float(np.float32(0.9 * saturate)) if (saturate is not None and data is not None) else float("inf")
        rc = self.lib.zb_weight_from_rms(self.h, self._dev(rms, (torch.float32,)),
                                         None if bad is None else self._dev(bad, (torch.uint8, torch.bool)),
                                         None if (data is None or saturate is None) else self._dev(data, (torch.float32,)), sat,
                                         rms.numel(), self._dev(out), self._stream(rms, stream))
        if rc != 0:
            self._err(rc)
        return out

    def align_lanczos3(self, src, src_wcs, dst_shape, dst_wcs, stream=None):
        """Resample ``src`` (float32 CUDA tensor) onto the grid ``dst_shape`` = (ny, nx) with world coordinates
        ``dst_wcs``: (aligned image, weight map).  What zuds runs SWarp for (zuds/swarp.py:157-204)."""
        import torch
        sny, snx = src.shape
        dny, dnx = dst_shape
        out = torch.empty((dny, dnx), dtype=torch.float32, device=src.device)
        wgt = torch.empty((dny, dnx), dtype=torch.float32, device=src.device)
        rc = self.lib.zb_align_lanczos3(self.h, self._dev(src, (torch.float32,)), snx, sny, C.byref(src_wcs), dnx, dny, C.byref(dst_wcs),
                                        self._dev(out), self._dev(wgt), self._stream(src, stream))
        if rc != 0:
            self._err(rc)
        return out, wgt

    def align_mask(self, mask, src_wcs, dst_shape, dst_wcs, nodata_bit=2 ** 16, stream=None):
        """Nearest-neighbour transport of an integer mask onto the target grid, ``nodata_bit`` where the image's weight is 0."""
        import torch
        sny, snx = mask.shape
        dny, dnx = dst_shape
        out = torch.empty((dny, dnx), dtype=torch.int32, device=mask.device)
        rc = self.lib.zb_align_mask(self.h, self._dev(mask), self._mask_dt(mask), snx, sny, C.byref(src_wcs), dnx, dny, C.byref(dst_wcs),
                                    self._dev(out), int(nodata_bit), self._stream(mask, stream))
        if rc != 0:
            self._err(rc)
        return out

    def subtract_host(self, p, sci, ref, sci_rms=None, ref_rms=None, sci_mask=None, ref_mask=None,
                      want_mask=True, out=None):
        """numpy host arrays in, numpy host arrays out (H2D/D2H inside the call)."""
        ny, nx = sci.shape
        assert (p.nx, p.ny) == (nx, ny)
        r = self._outputs(p)
        if out is None:
            r.diff = np.empty((ny, nx), np.float32)
            r.noise = np.empty((ny, nx), np.float32)
            r.mask = np.empty((ny, nx), np.int32) if want_mask else None
        else:
            r.diff, r.noise, r.mask = out

        def ptr(a):
            return None if a is None else a.ctypes.data_as(C.c_void_p)

        def mk(a):
            if a is None:
                return None, 0
            a = np.ascontiguousarray(a)
            if a.dtype == np.bool_:
                a = a.view(np.uint8)
            return a, MASK_DTYPES[a.dtype.name]
        f = [None if a is None else np.ascontiguousarray(a, np.float32) for a in (sci, ref, sci_rms, ref_rms)]
        m1, d1 = mk(sci_mask)
        m2, d2 = mk(ref_mask)
        rc = self.lib.zb_subtract_host(self.h, C.byref(p), ptr(f[0]), ptr(f[1]), ptr(f[2]), ptr(f[3]),
                                       ptr(m1), d1, ptr(m2), d2, ptr(r.diff), ptr(r.noise), ptr(r.mask),
                                       r.regions, ptr(r.sol), r.stamps)
        if rc != 0:
            self._err(rc)
        return r

    def subtract_batch(self, params, inputs, outs=None, want_mask=False):
        """``zb_subtract_batch``: several pairs of one frame size from host (ideally pinned) numpy arrays, the copies of
        neighbouring pairs overlapping the computation inside this context.  ``params``: list of zb_params;
        ``inputs``: list of (sci, ref, sci_rms, ref_rms, sci_mask, ref_mask); ``outs``: list of (diff, noise, mask-or-None)
        arrays to fill (allocated when None).  Returns a list of Result."""
        n = len(params)
        assert n == len(inputs) and n > 0
        sz = sizes_of(params[0])
        ny, nx = inputs[0][0].shape
        pa = (P.ZbParams * n)(*params)
        hp = (P.ZbHostPair * n)()
        regions = (P.ZbRegionInfo * (n * sz.n_regions))()
        sol = np.zeros((n, sz.n_regions, sz.n_comp_total + 1), np.float64)
        stamps = (P.ZbStampInfo * (n * sz.n_regions * sz.n_stamps))()
        rcs = (C.c_int * n)()
        keep, res = [], []

        def ptr(a):
            return None if a is None else a.ctypes.data

        for i in range(n):
            f = [None if a is None else np.ascontiguousarray(a, np.float32) for a in inputs[i][:4]]
            ms = []
            for m in inputs[i][4:6]:
                if m is None:
                    ms.append((None, 0))
                else:
                    m = np.ascontiguousarray(m)
                    m = m.view(np.uint8) if m.dtype == np.bool_ else m
                    ms.append((m, MASK_DTYPES[m.dtype.name]))
            if outs is not None:
                d, nz, mk = outs[i]
            else:
                d, nz = np.empty((ny, nx), np.float32), np.empty((ny, nx), np.float32)
                mk = np.empty((ny, nx), np.int32) if want_mask else None
            keep.append((f, ms, d, nz, mk))
            hp[i].sci, hp[i].ref, hp[i].sci_rms, hp[i].ref_rms = ptr(f[0]), ptr(f[1]), ptr(f[2]), ptr(f[3])
            hp[i].sci_mask, hp[i].sci_mask_dtype = ptr(ms[0][0]), ms[0][1]
            hp[i].ref_mask, hp[i].ref_mask_dtype = ptr(ms[1][0]), ms[1][1]
            hp[i].diff, hp[i].noise, hp[i].out_mask = ptr(d), ptr(nz), ptr(mk)
        rc = self.lib.zb_subtract_batch(self.h, n, pa, hp, regions, sol.ctypes.data_as(C.c_void_p), stamps, rcs)
        if rc != 0:
            self._err(rc)
        for i in range(n):
            r = Result()
            r.sizes = sz
            r.regions = (P.ZbRegionInfo * sz.n_regions).from_buffer(regions, i * sz.n_regions * C.sizeof(P.ZbRegionInfo))
            r.sol = sol[i]
            r.stamps = (P.ZbStampInfo * (sz.n_regions * sz.n_stamps)).from_buffer(stamps, i * sz.n_regions * sz.n_stamps * C.sizeof(P.ZbStampInfo))
            r.diff, r.noise, r.mask = keep[i][2], keep[i][3], keep[i][4]
            res.append(r)
        return res

    def subtract(self, p, sci, ref, sci_rms=None, ref_rms=None, sci_mask=None, ref_mask=None,
                 want_mask=True, out=None, stream=None):
        """torch CUDA tensors in/out; only data_ptr() crosses the C ABI."""
        import torch
        ny, nx = sci.shape
        assert (p.nx, p.ny) == (nx, ny)
        for t in (sci, ref, sci_rms, ref_rms, sci_mask, ref_mask):
            if t is not None:
                assert t.is_cuda and t.is_contiguous() and t.device.index == self.device
        r = self._outputs(p)
        dev = sci.device
        if out is None:
            r.diff = torch.empty((ny, nx), dtype=torch.float32, device=dev)
            r.noise = torch.empty((ny, nx), dtype=torch.float32, device=dev)
            r.mask = torch.empty((ny, nx), dtype=torch.int32, device=dev) if want_mask else None
        else:
            r.diff, r.noise, r.mask = out

        def ptr(t):
            return None if t is None else C.c_void_p(t.data_ptr())

        def md(t):
            if t is None:
                return 0
            return MASK_DTYPES[str(t.dtype).replace("torch.", "")]
        if stream is None:
            stream = torch.cuda.current_stream(dev)
        s = C.c_void_p(stream.cuda_stream)
        rc = self.lib.zb_subtract(self.h, C.byref(p), ptr(sci), ptr(ref), ptr(sci_rms), ptr(ref_rms),
                                  ptr(sci_mask), md(sci_mask), ptr(ref_mask), md(ref_mask),
                                  ptr(r.diff), ptr(r.noise), ptr(r.mask), r.regions,
                                  r.sol.ctypes.data_as(C.c_void_p), r.stamps, s)
        if rc != 0:
            self._err(rc)
        return r
