"""libear_b200 -- Python harness over the C ABI of libear-b200 (include/earb200.h).

The product is the C-ABI shared library ``libear_b200/libearb200.so`` (hand-written
sm_100a CUDA kernels + C++ host code) and the C++ facade in
``libear_b200/include/ear``.  This module only binds the C ABI with ctypes so that the
parity tests and ``bench.py`` can call it; the class names mirror libear's
(``GainCalculatorObjects``, ``GainInterpolator``, ``BlockConvolver`` ...).

There is no CPU fallback: if the shared library is missing, importing this module
raises; if no CUDA device is present, creating any calculator raises ``CudaError``.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libearb200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make -C libear_b200` (or "
        "`python -c 'import __graft_entry__ as g; g.build()'`). libear-b200 has no CPU fallback.")

_lib = C.CDLL(LIB_PATH)
_lib.earb200_last_error.restype = C.c_char_p
_lib.earb200_version.restype = C.c_char_p
_lib.earb200_layout_name_at.restype = C.c_char_p
_lib.earb200_layout_channel_name.restype = C.c_char_p
for _f in ("earb200_objects_launch_count", "earb200_interp_launch_count",
           "earb200_convolver_launch_count", "earb200_direct_speakers_launch_count",
           "earb200_delay_launch_count", "earb200_delay_get_delay", "earb200_fft_launch_count",
           "earb200_fft_size", "earb200_objects_renderer_block_size", "earb200_conversion_launch_count",
           "earb200_objects_renderer_launch_count", "earb200_bus_result_stride", "earb200_bus_launch_count"):
    getattr(_lib, _f).restype = C.c_int64
_lib.earb200_bus_result.restype = C.c_void_p


# exception classes mirroring libear's (include/ear/exceptions.hpp:8-43)
class EarError(RuntimeError):
    pass


class NotImplementedFeature(EarError):      # ear::not_implemented
    pass


class InternalError(EarError):              # ear::internal_error
    pass


class AdmError(EarError, ValueError):       # ear::adm_error
    pass


class UnknownLayout(EarError, ValueError):  # ear::unknown_layout
    pass


class InvalidArgument(EarError, ValueError):  # ear::invalid_argument
    pass


class CudaError(EarError):
    pass


_ERRORS = {1: NotImplementedFeature, 2: InternalError, 3: AdmError, 4: UnknownLayout,
           5: InvalidArgument, 6: CudaError}


def _chk(rc):
    if rc != 0:
        raise _ERRORS.get(rc, EarError)(_lib.earb200_last_error().decode())


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))  # raw device pointer


def _f64(a, n=None):
    a = np.ascontiguousarray(np.broadcast_to(np.asarray(a, dtype=np.float64), (n,)) if n is not None
                             else np.asarray(a, dtype=np.float64))
    return a


def version():
    return _lib.earb200_version().decode()


def layout_names():
    return [_lib.earb200_layout_name_at(i).decode() for i in range(_lib.earb200_num_layouts())]


class Layout:
    """ear::Layout / ear::getLayout (include/ear/layout.hpp:62-96, src/bs2051.cpp:13-22)."""

    def __init__(self, name, real_positions=None, channels=None):
        """channels: dict(names, pos[, nominal, az_range, el_range, is_lfe]) for ear::Layout(name, channels)
        with arbitrary channels; a name outside BS.2051 makes the panner build its own convex hull."""
        self.h = C.c_void_p()
        self.name = name
        if channels is not None:
            names = [n.encode() for n in channels["names"]]
            arr = (C.c_char_p * len(names))(*names)
            pos = np.ascontiguousarray(channels["pos"], dtype=np.float64).reshape(-1, 3)
            opt = {k: (None if channels.get(k) is None else np.ascontiguousarray(channels[k], dtype=np.float64))
                   for k in ("nominal", "az_range", "el_range")}
            lfe = channels.get("is_lfe")
            lfe = None if lfe is None else np.ascontiguousarray(lfe, dtype=np.int32)
            _chk(_lib.earb200_layout_create_custom(name.encode(), len(names), arr, _ptr(pos), _ptr(opt["nominal"]),
                                                   _ptr(opt["az_range"]), _ptr(opt["el_range"]), _ptr(lfe),
                                                   C.byref(self.h)))
        elif real_positions is None:
            _chk(_lib.earb200_layout_create(name.encode(), C.byref(self.h)))
        else:
            rp = np.ascontiguousarray(real_positions, dtype=np.float64).reshape(-1, 3)
            _chk(_lib.earb200_layout_create_with_positions(name.encode(), len(rp), _ptr(rp),
                                                           C.byref(self.h)))
        n = _lib.earb200_layout_num_channels(self.h)
        self.channel_names = [_lib.earb200_layout_channel_name(self.h, i).decode() for i in range(n)]
        self.is_lfe = np.array([bool(_lib.earb200_layout_channel_is_lfe(self.h, i)) for i in range(n)])
        self.positions = np.zeros((n, 3))
        self.nominal_positions = np.zeros((n, 3))
        for i in range(n):
            _chk(_lib.earb200_layout_channel_position(self.h, i, _ptr(self.positions[i:i + 1]),
                                                      _ptr(self.nominal_positions[i:i + 1])))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_layout_destroy(self.h)
            self.h = None


def convex_hull(xyz):
    """ear::convex_hull (src/common/convex_hull.cpp:7-100): facets as sorted vertex index lists."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
    buf = np.zeros(64 * len(xyz) + 64, dtype=np.int32)
    n = C.c_int(0)
    _chk(_lib.earb200_convex_hull(len(xyz), _ptr(xyz), _ptr(buf), len(buf), C.byref(n)))
    out, k = [], 0
    while k < n.value:
        c = int(buf[k])
        out.append([int(v) for v in buf[k + 1:k + 1 + c]])
        k += 1 + c
    return out


def extent_cells(n_rows=37):
    """The extent core's clustering of the virtual-source grid into cells of 16 (host, setup-time):
    (grid_xyz [P][3], order [n_cells][16], centre [n_cells][3], radius [n_cells] in radians)."""
    cap = 1024
    order = np.full((cap, 16), -1, dtype=np.int32)
    centre = np.zeros((cap, 3), dtype=np.float32)
    radius = np.zeros(cap)
    grid = np.zeros((cap * 16, 3))
    nc, npts = C.c_int(0), C.c_int(0)
    _chk(_lib.earb200_extent_cells(int(n_rows), cap, _ptr(order), _ptr(centre), _ptr(radius), _ptr(grid),
                                   C.byref(nc), C.byref(npts)))
    return grid[:npts.value], order[:nc.value], centre[:nc.value], radius[:nc.value]


def selftest_fast_sqrt(x, device=0):
    """Number of elements of x where the gain epilogue's branch-free FP64 square root differs from
    sqrt() (self-test hook of the C ABI; must be 0)."""
    x = np.ascontiguousarray(x, dtype=np.float64).ravel()
    n = C.c_int64(0)
    _chk(_lib.earb200_selftest_fast_sqrt(int(device), C.c_int64(len(x)), _ptr(x), C.byref(n)))
    return n.value


def get_layout(name):
    return Layout(name)


def _as_layout(layout):
    return layout if isinstance(layout, Layout) else Layout(layout)


class PointSourcePanner:
    """ear::configurePolarPanner(layout.withoutLfe()) (src/common/point_source_panner.cpp:586-600)."""

    def __init__(self, layout, device=0):
        self.layout = _as_layout(layout)
        self.h = C.c_void_p()
        _chk(_lib.earb200_panner_create(self.layout.h, device, C.byref(self.h)))
        self.num_channels = _lib.earb200_panner_num_channels(self.h)
        self.num_regions = _lib.earb200_panner_num_regions(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_panner_destroy(self.h)
            self.h = None

    def handle(self, xyz, regions=False):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        n = len(xyz)
        gains = np.zeros((n, self.num_channels))
        reg = np.zeros(n, dtype=np.int32) if regions else None
        inner = np.zeros(n, dtype=np.int32) if regions else None
        _chk(_lib.earb200_panner_handle_host(self.h, C.c_int64(n), _ptr(xyz), _ptr(gains),
                                             _ptr(reg), _ptr(inner)))
        return (gains, reg, inner) if regions else gains


class _ObjectsSoA(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("azimuth", "elevation", "distance", "width", "height",
                                          "depth", "gain", "diffuse", "flags")]


class _ObjectsExt(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("divergence", "divergence_azimuth_range", "channel_lock_max_distance")]


# feature bits libear refuses with ear::not_implemented (earb200.h)
OBJ_CARTESIAN, OBJ_CHANNEL_LOCK, OBJ_SCREEN_REF = 1, 2, 4
OBJ_DIVERGENCE, OBJ_ZONE_EXCLUSION, OBJ_CARTESIAN_POSITION = 8, 16, 32


class GainCalculatorObjects:
    """ear::GainCalculatorObjects (include/ear/gain_calculators.hpp:37-64), batched."""

    def __init__(self, layout, device=0, grid="libear"):
        """grid: "libear" (1,652 points, the reference's) or "tdesign5200" (throughput shape only:
        no reference to compare with)."""
        self.layout = _as_layout(layout)
        self.h = C.c_void_p()
        if grid == "libear":
            _chk(_lib.earb200_objects_create(self.layout.h, device, C.byref(self.h)))
        else:
            kinds = {"tdesign5200": 1}
            if grid not in kinds:
                raise InvalidArgument("unknown extent grid %r" % (grid,))
            _chk(_lib.earb200_objects_create_with_grid(self.layout.h, device, kinds[grid], C.byref(self.h)))
        self.num_outputs = _lib.earb200_objects_num_outputs(self.h)
        self.grid_points = _lib.earb200_objects_grid_points(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_objects_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_objects_launch_count(self.h)

    def set_profiling(self, enable=True):
        _chk(_lib.earb200_objects_set_profiling(self.h, int(enable)))

    def kernel_time(self):
        """(ms, launches, pass_slots) of the extent core kernel since the last query."""
        ms, n, p = C.c_double(0), C.c_int64(0), C.c_int64(0)
        _chk(_lib.earb200_objects_kernel_time(self.h, C.byref(ms), C.byref(n), C.byref(p)))
        return ms.value, n.value, p.value

    def calculate(self, azimuth, elevation, distance=1.0, width=0.0, height=0.0, depth=0.0,
                  gain=1.0, diffuse=0.0, flags=None, dtype=np.float32, regions=False):
        """Host buffers in, host buffers out: (direct, diffuse[, region_idx])."""
        az = np.atleast_1d(np.asarray(azimuth, dtype=np.float64))
        n = len(az)
        arrs = [_f64(a, n) for a in (az, elevation, distance, width, height, depth, gain, diffuse)]
        fl = None if flags is None else np.ascontiguousarray(np.broadcast_to(flags, (n,)), dtype=np.int32)
        soa = _ObjectsSoA(*[a.ctypes.data for a in arrs], fl.ctypes.data if fl is not None else None)
        direct = np.zeros((n, self.num_outputs), dtype=dtype)
        diffuse_out = np.zeros((n, self.num_outputs), dtype=dtype)
        reg = np.zeros(n, dtype=np.int32) if regions else None
        _chk(_lib.earb200_objects_calculate_host(self.h, C.c_int64(n), C.byref(soa),
                                                 int(np.dtype(dtype) == np.float64), _ptr(direct),
                                                 _ptr(diffuse_out), _ptr(reg)))
        return (direct, diffuse_out, reg) if regions else (direct, diffuse_out)

    def calculate_device(self, n, d_soa, d_direct, d_diffuse, d_regions=None, out_is_double=False,
                         stream=0):
        """Device pointers (ints) in d_soa = 8 device addresses in field order; asynchronous."""
        soa = _ObjectsSoA(*[int(p) for p in d_soa], None)
        _chk(_lib.earb200_objects_calculate_device(self.h, C.c_int64(n), C.byref(soa), int(out_is_double),
                                                   _ptr(d_direct), _ptr(d_diffuse), _ptr(d_regions),
                                                   C.c_void_p(int(stream))))

    STATIC_MAX = 4096  # EARB200_OBJECTS_STATIC_MAX

    def calculate_device_static(self, n, d_soa, d_direct, d_diffuse, d_regions=None, out_is_double=False,
                                stream=0):
        """calculate_device without any allocation (scratch in a file-scope device array that every
        call on the device shares): n <= STATIC_MAX; calls that overlap on the GPU race."""
        soa = _ObjectsSoA(*[int(p) for p in d_soa], None)
        _chk(_lib.earb200_objects_calculate_device_static(self.h, C.c_int64(n), C.byref(soa), int(out_is_double),
                                                          _ptr(d_direct), _ptr(d_diffuse), _ptr(d_regions),
                                                          C.c_void_p(int(stream))))

    def enable_extensions(self, divergence=False, channel_lock=False):
        """OPT-IN features beyond libear (parity unpinned; libear throws not_implemented for both)."""
        _chk(_lib.earb200_objects_enable_extensions(self.h, C.c_uint32(int(divergence) | 2 * int(channel_lock))))

    def calculate_ext_device(self, n, d_soa, d_direct, d_diffuse, d_divergence=None, d_azimuth_range=None,
                             d_lock_max_distance=None, out_is_double=False, stream=0):
        """calculate_device with objectDivergence / channelLock arrays (device addresses or None)."""
        soa = _ObjectsSoA(*[int(p) for p in d_soa], None)
        ext = _ObjectsExt(*[None if p is None else int(p) for p in (d_divergence, d_azimuth_range, d_lock_max_distance)])
        _chk(_lib.earb200_objects_calculate_ext_device(self.h, C.c_int64(n), C.byref(soa), C.byref(ext),
                                                       int(out_is_double), _ptr(d_direct), _ptr(d_diffuse),
                                                       C.c_void_p(int(stream))))

    def extent(self, xyz, width, height, depth):
        """ear::PolarExtent::handle (src/object_based/polar_extent.cpp:281-302): [n][S] double."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        n = len(xyz)
        w, h, d = (_f64(a, n) for a in (width, height, depth))
        S = int((~self.layout.is_lfe).sum())
        out = np.zeros((n, S))
        _chk(_lib.earb200_objects_extent_host(self.h, C.c_int64(n), _ptr(xyz), _ptr(w), _ptr(h), _ptr(d),
                                              _ptr(out)))
        return out


class GainInterpolator:
    """ear::dsp::GainInterpolator<LinearInterpMatrix> (include/ear/dsp/gain_interpolator.hpp).

    ``interp_points`` is a list of (sample_index, gains[n_in][n_out]); call
    ``set_points`` after changing it.  N = 1 gives LinearInterpVector, N = M = 1
    LinearInterpSingle.
    """

    def __init__(self, n_in, n_out, device=0):
        self.n_in, self.n_out = n_in, n_out
        self.h = C.c_void_p()
        _chk(_lib.earb200_interp_create(n_in, n_out, device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_interp_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_interp_launch_count(self.h)

    def set_profiling(self, enable=True):
        _chk(_lib.earb200_interp_set_profiling(self.h, int(enable)))

    def kernel_time(self):
        ms, n = C.c_double(0), C.c_int64(0)
        _chk(_lib.earb200_interp_kernel_time(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def set_accumulate(self, enable=True):
        """Extension: process() adds into the output bus instead of overwriting it."""
        _chk(_lib.earb200_interp_set_accumulate(self.h, int(enable)))

    def set_points(self, sample_idx, gains):
        s = np.ascontiguousarray(sample_idx, dtype=np.int64)
        g = np.ascontiguousarray(gains, dtype=np.float32).reshape(len(s), self.n_in, self.n_out)
        _chk(_lib.earb200_interp_set_points(self.h, C.c_int64(len(s)), _ptr(s), _ptr(g)))

    def set_points_device(self, sample_idx, d_gains, stream=0):
        """d_gains: device address of float32 [n_points][n_in][n_out] (no host round trip)."""
        s = np.ascontiguousarray(sample_idx, dtype=np.int64)
        _chk(_lib.earb200_interp_set_points_device(self.h, C.c_int64(len(s)), _ptr(s), _ptr(d_gains),
                                                   C.c_void_p(int(stream))))

    def process(self, block_start, inp, out=None):
        """inp: float32 [n_in][nsamples] host array -> out float32 [n_out][nsamples]; with
        set_accumulate(True), `out` is the bus the result is added to."""
        inp = np.ascontiguousarray(inp, dtype=np.float32).reshape(self.n_in, -1)
        n = inp.shape[1]
        if out is None:
            out = np.zeros((self.n_out, n), dtype=np.float32)
        else:
            out = np.array(out, dtype=np.float32).reshape(self.n_out, n)
        _chk(_lib.earb200_interp_process_host(self.h, C.c_int64(block_start), C.c_int64(n), _ptr(inp),
                                              C.c_int64(n), _ptr(out), C.c_int64(n)))
        return out

    def process_device_bus(self, block_start, nsamples, d_in, in_stride, bus, bus_t0, d_partial=None,
                           partial_stride=0, stream=0):
        """process() of this rank's objects into the node-wide bus (Bus)."""
        _chk(_lib.earb200_interp_process_device_bus(self.h, C.c_int64(block_start), C.c_int64(nsamples), _ptr(d_in),
                                                    C.c_int64(in_stride), _ptr(d_partial), C.c_int64(partial_stride),
                                                    bus.h, C.c_int64(bus_t0), C.c_void_p(int(stream))))

    def process_device(self, block_start, nsamples, d_in, in_stride, d_out, out_stride, stream=0):
        _chk(_lib.earb200_interp_process_device(self.h, C.c_int64(block_start), C.c_int64(nsamples),
                                                _ptr(d_in), C.c_int64(in_stride), _ptr(d_out),
                                                C.c_int64(out_stride), C.c_void_p(int(stream))))


class Bus:
    """The loudspeaker bus shared by the GPUs of one node (earb200_bus_*, include/earb200.h): the
    exchange step of object-sharded mixing over peer memory.  One instance per rank."""

    def __init__(self, n_channels, n_samples, rank=0, n_ranks=1, device=0):
        self.n_channels, self.n_samples, self.rank, self.n_ranks = n_channels, n_samples, rank, n_ranks
        self.h = C.c_void_p()
        _chk(_lib.earb200_bus_create(device, rank, n_ranks, n_channels, C.c_int64(n_samples), C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_bus_destroy(self.h)
            self.h = None

    def ipc_handle(self):
        buf = C.create_string_buffer(_lib.earb200_bus_ipc_handle_bytes())
        _chk(_lib.earb200_bus_get_ipc_handle(self.h, buf))
        return buf.raw

    def open_ipc(self, handles):
        """handles: the ipc_handle() of every rank, in rank order."""
        _chk(_lib.earb200_bus_open_ipc(self.h, b"".join(handles)))

    @staticmethod
    def connect_local(buses):
        """All ranks in this process: wire them up with direct peer access."""
        arr = (C.c_void_p * len(buses))(*[b.h for b in buses])
        _chk(_lib.earb200_bus_connect_local(arr, len(buses)))

    def slice(self, rank=None):
        t0, t1 = C.c_int64(0), C.c_int64(0)
        _chk(_lib.earb200_bus_slice(self.h, self.rank if rank is None else rank, C.byref(t0), C.byref(t1)))
        return t0.value, t1.value

    def push(self, d_partial, partial_stride, t0, n, stream=0):
        _chk(_lib.earb200_bus_push(self.h, _ptr(d_partial), C.c_int64(partial_stride), C.c_int64(t0), C.c_int64(n),
                                   C.c_void_p(int(stream))))

    def publish(self, stream=0):
        _chk(_lib.earb200_bus_publish(self.h, C.c_void_p(int(stream))))

    def wait_result(self, stream=0):
        _chk(_lib.earb200_bus_wait_result(self.h, C.c_void_p(int(stream))))

    def reduce(self, d_out=None, out_stride=0, gather=False, stream=0):
        _chk(_lib.earb200_bus_reduce(self.h, _ptr(d_out), C.c_int64(out_stride), int(gather), C.c_void_p(int(stream))))

    def copy_result(self, d_dst, dst_stride, t0, n, stream=0):
        _chk(_lib.earb200_bus_copy_result(self.h, _ptr(d_dst), C.c_int64(dst_stride), C.c_int64(t0), C.c_int64(n),
                                          C.c_void_p(int(stream))))

    @property
    def result_ptr(self):
        return _lib.earb200_bus_result(self.h)

    @property
    def result_stride(self):
        return _lib.earb200_bus_result_stride(self.h)

    @property
    def launch_count(self):
        return _lib.earb200_bus_launch_count(self.h)


class BlockConvolver:
    """A bank of ear::dsp::block_convolver::BlockConvolver (include/ear/dsp/block_convolver.hpp)."""

    def __init__(self, block_size, n_channels, num_blocks, device=0):
        self.block_size, self.n_channels, self.num_blocks = block_size, n_channels, num_blocks
        self.h = C.c_void_p()
        _chk(_lib.earb200_convolver_create(C.c_int64(block_size), n_channels, C.c_int64(num_blocks),
                                           device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_convolver_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_convolver_launch_count(self.h)

    def set_profiling(self, enable=True):
        _chk(_lib.earb200_convolver_set_profiling(self.h, int(enable)))

    def kernel_time(self):
        ms, n = C.c_double(0), C.c_int64(0)
        _chk(_lib.earb200_convolver_kernel_time(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def add_filter(self, taps):
        taps = np.ascontiguousarray(taps, dtype=np.float32)
        fid, nb = C.c_int(0), C.c_int(0)
        _chk(_lib.earb200_convolver_add_filter(self.h, C.c_int64(len(taps)), _ptr(taps), C.byref(fid),
                                               C.byref(nb)))
        return fid.value, nb.value

    def set_filter(self, filter_id, channel=-1):
        _chk(_lib.earb200_convolver_set_filter(self.h, channel, -1 if filter_id is None else filter_id))

    def crossfade_filter(self, filter_id, channel=-1):
        _chk(_lib.earb200_convolver_crossfade_filter(self.h, channel,
                                                     -1 if filter_id is None else filter_id))

    def process(self, x, n_blocks=None):
        """x: float32 [n_channels][n_blocks*B] host array, or None for silence."""
        if x is None:
            out = np.zeros((self.n_channels, n_blocks * self.block_size), dtype=np.float32)
            _chk(_lib.earb200_convolver_process_host(self.h, C.c_int64(n_blocks), None, C.c_int64(0),
                                                     _ptr(out), C.c_int64(out.shape[1])))
            return out
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(self.n_channels, -1)
        nb = x.shape[1] // self.block_size
        out = np.zeros((self.n_channels, nb * self.block_size), dtype=np.float32)
        _chk(_lib.earb200_convolver_process_host(self.h, C.c_int64(nb), _ptr(x), C.c_int64(x.shape[1]),
                                                 _ptr(out), C.c_int64(out.shape[1])))
        return out

    def process_device(self, n_blocks, d_in, in_stride, d_out, out_stride, stream=0):
        _chk(_lib.earb200_convolver_process_device(self.h, C.c_int64(n_blocks), _ptr(d_in),
                                                   C.c_int64(in_stride), _ptr(d_out),
                                                   C.c_int64(out_stride), C.c_void_p(int(stream))))


class ObjectsRenderer:
    """The Objects signal flow of libear's docs (docs/dsp.rst:40-71) as one device pipeline:
    interpolated direct / diffuse buses, decorrelators on the diffuse bus, compensation delay on
    the direct bus, sum."""

    def __init__(self, layout, n_objects, device=0):
        self.layout = _as_layout(layout)
        self.n_objects = n_objects
        self.h = C.c_void_p()
        _chk(_lib.earb200_objects_renderer_create(self.layout.h, n_objects, device, C.byref(self.h)))
        self.num_outputs = _lib.earb200_objects_renderer_num_outputs(self.h)
        self.block_size = _lib.earb200_objects_renderer_block_size(self.h)
        self.delay = _lib.earb200_objects_renderer_delay(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_objects_renderer_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_objects_renderer_launch_count(self.h)

    def set_gains(self, sample_idx, direct, diffuse):
        s = np.ascontiguousarray(sample_idx, dtype=np.int64)
        shape = (len(s), self.n_objects, self.num_outputs)
        d = np.ascontiguousarray(direct, dtype=np.float32).reshape(shape)
        f = np.ascontiguousarray(diffuse, dtype=np.float32).reshape(shape)
        _chk(_lib.earb200_objects_renderer_set_gains(self.h, C.c_int64(len(s)), _ptr(s), _ptr(d), _ptr(f)))

    def set_gains_device(self, sample_idx, d_direct, d_diffuse, stream=0):
        s = np.ascontiguousarray(sample_idx, dtype=np.int64)
        _chk(_lib.earb200_objects_renderer_set_gains_device(self.h, C.c_int64(len(s)), _ptr(s), _ptr(d_direct),
                                                            _ptr(d_diffuse), C.c_void_p(int(stream))))

    def process(self, block_start, x):
        """x: float32 [n_objects][n_blocks * 512] host array -> [C_out][n_blocks * 512]."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(self.n_objects, -1)
        T = x.shape[1]
        if T % self.block_size:
            raise InvalidArgument("the renderer processes whole blocks of %d samples" % self.block_size)
        out = np.zeros((self.num_outputs, T), dtype=np.float32)
        _chk(_lib.earb200_objects_renderer_process_host(self.h, C.c_int64(block_start), C.c_int64(T // self.block_size),
                                                        _ptr(x), C.c_int64(T), _ptr(out), C.c_int64(T)))
        return out

    def process_device(self, block_start, n_blocks, d_in, in_stride, d_out, out_stride, stream=0):
        _chk(_lib.earb200_objects_renderer_process_device(self.h, C.c_int64(block_start), C.c_int64(n_blocks),
                                                          _ptr(d_in), C.c_int64(in_stride), _ptr(d_out),
                                                          C.c_int64(out_stride), C.c_void_p(int(stream))))


class FFT:
    """Batched r2c / c2r transforms behind ear::FFTImpl / FFTPlan (include/ear/fft.hpp:27-62)."""

    def __init__(self, n_fft, device=0):
        self.n_fft = n_fft
        self.h = C.c_void_p()
        _chk(_lib.earb200_fft_create(C.c_int64(n_fft), device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_fft_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_fft_launch_count(self.h)

    def forward(self, x):
        """x float32 [batch][n_fft] -> complex64 [batch][n_fft/2+1]."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, self.n_fft)
        out = np.zeros((len(x), self.n_fft // 2 + 1), dtype=np.complex64)
        _chk(_lib.earb200_fft_forward_host(self.h, C.c_int64(len(x)), _ptr(x), _ptr(out)))
        return out

    def reverse(self, X):
        """X complex64 [batch][n_fft/2+1] -> float32 [batch][n_fft], unnormalised."""
        X = np.ascontiguousarray(X, dtype=np.complex64).reshape(-1, self.n_fft // 2 + 1)
        out = np.zeros((len(X), self.n_fft), dtype=np.float32)
        _chk(_lib.earb200_fft_reverse_host(self.h, C.c_int64(len(X)), _ptr(X), _ptr(out)))
        return out


class DelayBuffer:
    """ear::dsp::DelayBuffer (include/ear/dsp/delay_buffer.hpp:12-29); delay memory on the device."""

    def __init__(self, n_channels, delay, device=0):
        self.n_channels, self.delay = n_channels, delay
        self.h = C.c_void_p()
        _chk(_lib.earb200_delay_create(n_channels, C.c_int64(delay), device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_delay_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_delay_launch_count(self.h)

    def get_delay(self):
        return _lib.earb200_delay_get_delay(self.h)

    def process(self, x):
        """x: float32 [n_channels][nsamples] host array -> delayed copy."""
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(self.n_channels, -1)
        out = np.zeros_like(x)
        _chk(_lib.earb200_delay_process_host(self.h, C.c_int64(x.shape[1]), _ptr(x), C.c_int64(x.shape[1]),
                                             _ptr(out), C.c_int64(x.shape[1])))
        return out

    def process_device(self, nsamples, d_in, in_stride, d_out, out_stride, stream=0):
        _chk(_lib.earb200_delay_process_device(self.h, C.c_int64(nsamples), _ptr(d_in), C.c_int64(in_stride),
                                               _ptr(d_out), C.c_int64(out_stride), C.c_void_p(int(stream))))


class Conversion:
    """ear::conversion (include/ear/conversion.hpp:14-71) for a batch of objects: positions
    [n][3] (X, Y, Z <-> azimuth, elevation, distance) and optional extents [n][3]
    (width, height, depth)."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        _chk(_lib.earb200_conversion_create(device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_conversion_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_conversion_launch_count(self.h)

    def _run(self, fn, pos, extent):
        pos = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1, 3)
        pos_out = np.zeros_like(pos)
        ext_out = None
        if extent is not None:
            extent = np.ascontiguousarray(extent, dtype=np.float64).reshape(-1, 3)
            if extent.shape != pos.shape:
                raise InvalidArgument("extent must have one row per position")
            ext_out = np.zeros_like(extent)
        _chk(fn(self.h, C.c_int64(len(pos)), _ptr(pos), _ptr(extent), _ptr(pos_out), _ptr(ext_out)))
        return (pos_out, ext_out) if extent is not None else pos_out

    def cart_to_polar(self, pos, extent=None):
        """pointCartToPolar (extent None) / extentCartToPolar per row."""
        return self._run(_lib.earb200_conversion_cart_to_polar_host, pos, extent)

    def polar_to_cart(self, pos, extent=None):
        """pointPolarToCart (extent None) / extentPolarToCart per row."""
        return self._run(_lib.earb200_conversion_polar_to_cart_host, pos, extent)

    def cart_to_polar_device(self, n, d_pos, d_extent, d_pos_out, d_extent_out, stream=0):
        _chk(_lib.earb200_conversion_cart_to_polar_device(self.h, C.c_int64(n), _ptr(d_pos), _ptr(d_extent),
                                                          _ptr(d_pos_out), _ptr(d_extent_out),
                                                          C.c_void_p(int(stream))))

    def polar_to_cart_device(self, n, d_pos, d_extent, d_pos_out, d_extent_out, stream=0):
        _chk(_lib.earb200_conversion_polar_to_cart_device(self.h, C.c_int64(n), _ptr(d_pos), _ptr(d_extent),
                                                          _ptr(d_pos_out), _ptr(d_extent_out),
                                                          C.c_void_p(int(stream))))

    def check(self, stream=0):
        _chk(_lib.earb200_conversion_check(self.h, C.c_void_p(int(stream))))


WARN_HOA_SCREENREF, WARN_HOA_NFCREFDIST = 4, 5


class GainCalculatorHOA:
    """ear::GainCalculatorHOA (include/ear/gain_calculators.hpp:67-82)."""

    def __init__(self, layout, device=0):
        self.layout = _as_layout(layout)
        self.h = C.c_void_p()
        _chk(_lib.earb200_hoa_create(self.layout.h, device, C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_hoa_destroy(self.h)
            self.h = None

    def calculate(self, orders, degrees, normalization="SN3D", nfc_ref_dist=0.0, screen_ref=False):
        """Returns (decode[n_channels][C_out], [warning codes])."""
        o = np.ascontiguousarray(orders, dtype=np.int32)
        d = np.ascontiguousarray(degrees, dtype=np.int32)
        if len(o) != len(d):
            raise InvalidArgument("orders and degrees must be the same size")
        n_out = len(self.layout.channel_names)
        D = np.zeros((len(o), n_out))
        w = C.c_uint32(0)
        _chk(_lib.earb200_hoa_calculate(self.h, len(o), _ptr(o), _ptr(d), normalization.encode(),
                                        C.c_double(nfc_ref_dist), int(screen_ref), _ptr(D), C.byref(w)))
        return D, [c for c in range(32) if w.value >> c & 1]


WARN_FREQ_SPEAKERLABEL_LFE_MISMATCH, WARN_FREQ_NOT_LFE = 1, 2


class _DirectSpeakersMetadata(C.Structure):
    _fields_ = [("speaker_labels", C.POINTER(C.c_char_p)), ("n_speaker_labels", C.c_int32),
                ("cartesian_position", C.c_int32), ("position", C.c_double * 3),
                ("bounds", C.c_double * 6), ("has_bounds", C.c_uint32), ("screen_edge_lock", C.c_uint32),
                ("low_pass", C.c_double), ("high_pass", C.c_double), ("has_frequency", C.c_uint32),
                ("reserved", C.c_uint32), ("audio_pack_format_id", C.c_char_p)]


_BOUND_KEYS = ["azMin", "azMax", "elMin", "elMax", "distMin", "distMax"]


class GainCalculatorDirectSpeakers:
    """ear::GainCalculatorDirectSpeakers (include/ear/gain_calculators.hpp:18-34).  Routing is
    host-side; only channels that fall through to the point-source panner touch the device."""

    def __init__(self, layout, additional_substitutions=None, device=0):
        self.layout = _as_layout(layout)
        subs = dict(additional_substitutions or {})
        frm = (C.c_char_p * max(1, len(subs)))(*[k.encode() for k in subs])
        to = (C.c_char_p * max(1, len(subs)))(*[v.encode() for v in subs.values()])
        self.h = C.c_void_p()
        _chk(_lib.earb200_direct_speakers_create(self.layout.h, device, len(subs), frm, to, C.byref(self.h)))
        self.num_outputs = _lib.earb200_direct_speakers_num_outputs(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            if _lib is not None:  # None while the interpreter shuts down
                _lib.earb200_direct_speakers_destroy(self.h)
            self.h = None

    @property
    def launch_count(self):
        return _lib.earb200_direct_speakers_launch_count(self.h)

    @staticmethod
    def _flatten(md, keep):
        m = _DirectSpeakersMetadata()
        labels = [x.encode() for x in md.get("labels", [])]
        arr = (C.c_char_p * max(1, len(labels)))(*labels)
        keep.append(arr)
        m.speaker_labels = C.cast(arr, C.POINTER(C.c_char_p))
        m.n_speaker_labels = len(labels)
        m.cartesian_position = int(bool(md.get("cartesian", False)))
        m.position[:] = [md.get("az", 0.0), md.get("el", 0.0), md.get("dist", 1.0)]
        for i, k in enumerate(_BOUND_KEYS):
            if md.get("bounds") and k in md["bounds"]:
                m.bounds[i] = md["bounds"][k]
                m.has_bounds |= 1 << i
        m.screen_edge_lock = int(md.get("screen_edge_lock", 0))
        if md.get("low_pass") is not None:
            m.low_pass, m.has_frequency = md["low_pass"], m.has_frequency | 1
        if md.get("high_pass") is not None:
            m.high_pass, m.has_frequency = md["high_pass"], m.has_frequency | 2
        m.audio_pack_format_id = md["pack"].encode() if md.get("pack") else None
        return m

    def calculate_batch(self, channels):
        """channels: list of dicts (labels, az, el, dist, bounds, low_pass, high_pass, pack,
        cartesian, screen_edge_lock).  Returns (gains[n][C_out], [[warning codes] per channel])."""
        keep = []
        mds = (_DirectSpeakersMetadata * max(1, len(channels)))(*[self._flatten(c, keep) for c in channels])
        gains = np.zeros((len(channels), self.num_outputs))
        w = np.zeros(max(1, len(channels)), dtype=np.uint32)
        _chk(_lib.earb200_direct_speakers_calculate(self.h, C.c_int64(len(channels)), mds, _ptr(gains), _ptr(w)))
        return gains, [[c for c in range(32) if int(x) >> c & 1] for x in w[:len(channels)]]

    def calculate(self, labels, az=0.0, el=0.0, dist=1.0, bounds=None, low_pass=None, high_pass=None,
                  pack=None, **kw):
        g, w = self.calculate_batch([dict(labels=labels, az=az, el=el, dist=dist, bounds=bounds,
                                          low_pass=low_pass, high_pass=high_pass, pack=pack, **kw)])
        return g[0], w[0]


def design_decorrelators(layout, dtype=np.float32):
    """ear::designDecorrelators<T> (src/decorrelate.cpp:83-90); host-side setup."""
    layout = _as_layout(layout)
    out = np.zeros((len(layout.channel_names), _lib.earb200_decorrelator_size()), dtype=dtype)
    if np.dtype(dtype) == np.float32:
        _chk(_lib.earb200_design_decorrelators(layout.h, _ptr(out), None))
    else:
        _chk(_lib.earb200_design_decorrelators(layout.h, None, _ptr(out)))
    return out


def decorrelator_compensation_delay():
    return _lib.earb200_decorrelator_compensation_delay()
