"""TEST INFRASTRUCTURE ONLY -- ctypes binding to oracle/_ref/libear_ref.so.

libear_ref.so is the UNMODIFIED libear reference compiled from /root/reference by
oracle/Makefile plus the extern "C" wrapper oracle/ref_capi.cpp.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product (libear_b200) never does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libear_ref.so")

_lib = None


class RefError(RuntimeError):
    """Raised when the reference throws; .code maps onto the ear:: exception class."""

    CODES = {1: "not_implemented", 2: "internal_error", 3: "adm_error",
             4: "unknown_layout", 5: "invalid_argument", 6: "std::exception", 7: "none"}

    def __init__(self, code, msg):
        super().__init__(f"{self.CODES.get(code, code)}: {msg}")
        self.code = code
        self.kind = self.CODES.get(code, str(code))


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.ref_last_error.restype = C.c_char_p
        _lib.ref_extent_isa.restype = C.c_char_p
    return _lib


def _chk(rc):
    if rc != 0:
        raise RefError(rc, lib().ref_last_error().decode())


def _p(a, t=None):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def layout_names():
    L = lib()
    out = []
    for i in range(L.ref_num_layouts()):
        buf = C.create_string_buffer(32)
        L.ref_layout_name(i, buf, 32)
        out.append(buf.value.decode())
    return out


def layout_info(name):
    L = lib()
    n = C.c_int(0)
    _chk(L.ref_layout_info(name.encode(), C.byref(n), None, None, None, None, None))
    n = n.value
    is_lfe = np.zeros(n, dtype=np.int32)
    pos = np.zeros((n, 3))
    nom = np.zeros((n, 3))
    rng = np.zeros((n, 4))
    names = C.create_string_buffer(16 * n)
    nn = C.c_int(0)
    _chk(L.ref_layout_info(name.encode(), C.byref(nn), _p(is_lfe), _p(pos), _p(nom), _p(rng), names))
    nm = [names.raw[16 * i:16 * i + 16].split(b"\0")[0].decode() for i in range(n)]
    return dict(names=nm, is_lfe=is_lfe.astype(bool), pos=pos, nominal=nom, ranges=rng)


def facets(name):
    buf = np.zeros(4096, dtype=np.int32)
    n = C.c_int(0)
    _chk(lib().ref_facets(name.encode(), _p(buf), len(buf), C.byref(n)))
    out, k = [], 0
    while k < n.value:
        c = int(buf[k])
        out.append([int(v) for v in buf[k + 1:k + 1 + c]])
        k += 1 + c
    return out


class PointSourcePanner:
    """ear::configurePolarPanner(layout.withoutLfe()) + region-index probe."""

    def __init__(self, layout):
        self.h = C.c_void_p()
        if isinstance(layout, dict):  # custom layout, see custom_layout()
            _chk(lib().ref_psp_new_custom(*_custom_args(layout), C.byref(self.h)))
        else:
            _chk(lib().ref_psp_new(layout.encode(), C.byref(self.h)))
        self.S = lib().ref_psp_num_channels(self.h)
        self.n_regions = lib().ref_psp_num_regions(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_psp_free(self.h)

    def handle(self, xyz, regions=True):
        xyz = _f64(xyz).reshape(-1, 3)
        n = len(xyz)
        g = np.zeros((n, self.S))
        reg = np.zeros(n, dtype=np.int32) if regions else None
        inner = np.zeros(n, dtype=np.int32) if regions else None
        _chk(lib().ref_psp_handle(self.h, C.c_long(n), _p(xyz), _p(g), _p(reg), _p(inner)))
        return (g, reg, inner) if regions else g


class Extent:
    """ear::PolarExtent::handle. core: 0 xsimd-dispatched, 1 scalar, 2 double test reference."""

    def __init__(self, layout, core=0):
        self.h = C.c_void_p()
        _chk(lib().ref_extent_new(layout.encode(), core, C.byref(self.h)))
        self.S = len([x for x in layout_info(layout)["is_lfe"] if not x])

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_extent_free(self.h)

    def handle(self, xyz, width, height, depth):
        xyz = _f64(xyz).reshape(-1, 3)
        n = len(xyz)
        w, h, d = (_f64(np.broadcast_to(x, (n,))) for x in (width, height, depth))
        out = np.zeros((n, self.S))
        _chk(lib().ref_extent_handle(self.h, C.c_long(n), _p(xyz), _p(w), _p(h), _p(d), _p(out)))
        return out


def extent_isa():
    return lib().ref_extent_isa().decode()


class Objects:
    """ear::GainCalculatorObjects."""

    def __init__(self, layout):
        self.h = C.c_void_p()
        if isinstance(layout, dict):
            _chk(lib().ref_objects_new_custom(*_custom_args(layout), C.byref(self.h)))
            self.n_out = len(layout["names"])
        else:
            _chk(lib().ref_objects_new(layout.encode(), C.byref(self.h)))
            self.n_out = len(layout_info(layout)["names"])

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_objects_free(self.h)

    def calculate(self, az, el, dist, width, height, depth, gain, diffuse,
                  flags=None, dtype=np.float64):
        az = _f64(np.atleast_1d(az))
        n = len(az)
        arrs = [_f64(np.broadcast_to(x, (n,))) for x in (el, dist, width, height, depth, gain, diffuse)]
        fl = None if flags is None else np.ascontiguousarray(np.broadcast_to(flags, (n,)), dtype=np.int32)
        direct = np.zeros((n, self.n_out), dtype=dtype)
        diff = np.zeros((n, self.n_out), dtype=dtype)
        _chk(lib().ref_objects_calc(self.h, C.c_long(n), _p(az), *[_p(a) for a in arrs], _p(fl),
                                    self.n_out, int(dtype == np.float32), _p(direct), _p(diff)))
        return direct, diff


def hoa_calc(layout, orders, degrees, norm="SN3D", nfc_ref_dist=0.0, screen_ref=False):
    orders = np.ascontiguousarray(orders, dtype=np.int32)
    degrees = np.ascontiguousarray(degrees, dtype=np.int32)
    Cn = len(orders)
    n_out = len(layout_info(layout)["names"])
    D = np.zeros((Cn, n_out))
    nw = C.c_int(0)
    _chk(lib().ref_hoa_calc(layout.encode(), Cn, _p(orders), _p(degrees), norm.encode(),
                            C.c_double(nfc_ref_dist), int(screen_ref), n_out, _p(D), C.byref(nw)))
    return D, nw.value


def hoa_points():
    n = lib().ref_hoa_num_points()
    p = np.zeros((n, 3))
    lib().ref_hoa_points(_p(p))
    return p


def direct_speakers_calc(layout, labels, az=0.0, el=0.0, dist=1.0, bounds=None,
                         low_pass=None, high_pass=None, pack_id=None, substitutions=None,
                         cartesian=False, screen_edge_lock=0, codes=False):
    """bounds: dict with any of azMin, azMax, elMin, elMax, distMin, distMax.
    Returns (gains, number of warnings) or, with codes=True, (gains, sorted warning codes)."""
    n_out = len(layout_info(layout)["names"])
    keys = ["azMin", "azMax", "elMin", "elMax", "distMin", "distMax"]
    b = np.zeros(6)
    mask = 0
    for i, k in enumerate(keys):
        if bounds and k in bounds:
            b[i] = bounds[k]
            mask |= 1 << i
    fmask = (1 if low_pass is not None else 0) | (2 if high_pass is not None else 0)
    g = np.zeros(n_out)
    nw = C.c_int(0)
    _chk(lib().ref_direct_speakers_calc(
        layout.encode(), "\n".join(labels).encode(), C.c_double(az), C.c_double(el), C.c_double(dist),
        _p(b), mask, C.c_double(low_pass or 0.0), C.c_double(high_pass or 0.0), fmask,
        (pack_id or "").encode(), n_out, _p(g), C.byref(nw),
        "\n".join(f"{k}={v}" for k, v in (substitutions or {}).items()).encode(),
        int(cartesian), int(screen_edge_lock)))
    if codes:
        return g, [c for c in range(1, 8) if nw.value >> (8 + c) & 1]
    return g, nw.value & 255


def decorrelators(layout, dtype=np.float32):
    n = len(layout_info(layout)["names"])
    out = np.zeros((n, 512), dtype=dtype)
    if dtype == np.float32:
        _chk(lib().ref_decorrelators(layout.encode(), n, _p(out), None))
    else:
        _chk(lib().ref_decorrelators(layout.encode(), n, None, _p(out)))
    return out


def decorrelator_delay():
    return lib().ref_decorrelator_delay()


def interp_matrix(inp, point_sample, point_gains, block_start0=0, call_size=None):
    """GainInterpolator<LinearInterpMatrix>::process; inp [N][T], gains [P][N][M]."""
    inp = np.ascontiguousarray(inp, dtype=np.float32)
    N, T = inp.shape
    pg = np.ascontiguousarray(point_gains, dtype=np.float32)
    P, N2, M = pg.shape
    assert N2 == N
    ps = np.ascontiguousarray(point_sample, dtype=np.int64)
    out = np.zeros((M, T), dtype=np.float32)
    _chk(lib().ref_interp_matrix(N, M, C.c_long(T), _p(inp), _p(out), P, _p(ps), _p(pg),
                                 C.c_long(block_start0), C.c_long(call_size or T)))
    return out


def interp_vector(inp, point_sample, point_gains, block_start0=0, call_size=None):
    inp = np.ascontiguousarray(inp, dtype=np.float32).reshape(-1)
    T = len(inp)
    pg = np.ascontiguousarray(point_gains, dtype=np.float32)
    P, M = pg.shape
    ps = np.ascontiguousarray(point_sample, dtype=np.int64)
    out = np.zeros((M, T), dtype=np.float32)
    _chk(lib().ref_interp_vector(M, C.c_long(T), _p(inp), _p(out), P, _p(ps), _p(pg),
                                 C.c_long(block_start0), C.c_long(call_size or T)))
    return out


def interp_single(inp, point_sample, point_gains, block_start0=0, call_size=None):
    inp = np.ascontiguousarray(inp, dtype=np.float32).reshape(-1)
    T = len(inp)
    pg = np.ascontiguousarray(point_gains, dtype=np.float32).reshape(-1)
    ps = np.ascontiguousarray(point_sample, dtype=np.int64)
    out = np.zeros(T, dtype=np.float32)
    _chk(lib().ref_interp_single(C.c_long(T), _p(inp), _p(out), len(pg), _p(ps), _p(pg),
                                 C.c_long(block_start0), C.c_long(call_size or T)))
    return out


class Convolver:
    """ear::dsp::block_convolver::BlockConvolver with the KISS float FFT."""

    def __init__(self, block_size, num_blocks):
        self.h = C.c_void_p()
        self.block_size = block_size
        _chk(lib().ref_conv_new(C.c_long(block_size), C.c_long(num_blocks), C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_conv_free(self.h)

    def add_filter(self, taps):
        taps = np.ascontiguousarray(taps, dtype=np.float32)
        fid = C.c_int(0)
        nb = C.c_int(0)
        _chk(lib().ref_conv_add_filter(self.h, C.c_long(len(taps)), _p(taps), C.byref(fid), C.byref(nb)))
        return fid.value, nb.value

    def set_filter(self, fid):
        _chk(lib().ref_conv_set_filter(self.h, -1 if fid is None else fid))

    def crossfade_filter(self, fid):
        _chk(lib().ref_conv_crossfade_filter(self.h, -1 if fid is None else fid))

    def process(self, x, n_blocks=None):
        """x: float32 [n_blocks*B] or None (silence; n_blocks required)."""
        if x is None:
            out = np.zeros(n_blocks * self.block_size, dtype=np.float32)
            _chk(lib().ref_conv_process(self.h, C.c_long(n_blocks), None, _p(out)))
            return out
        x = np.ascontiguousarray(x, dtype=np.float32)
        nb = len(x) // self.block_size
        out = np.zeros(nb * self.block_size, dtype=np.float32)
        _chk(lib().ref_conv_process(self.h, C.c_long(nb), _p(x), _p(out)))
        return out


def delay_process(x, delay, call_size=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    nch, T = x.shape
    out = np.zeros_like(x)
    _chk(lib().ref_delay_process(nch, C.c_long(delay), C.c_long(T), C.c_long(call_size or T), _p(x), _p(out)))
    return out


def fft_kiss(x, forward=True, n_fft=None):
    """libear's own KISS transforms (src/fft_kiss.cpp:52-99): forward float32 [batch][n_fft] ->
    complex64 [batch][n_fft/2+1]; reverse complex64 -> float32 [batch][n_fft], unnormalised."""
    if forward:
        x = np.ascontiguousarray(x, dtype=np.float32)
        n_fft = x.shape[-1]
        x = x.reshape(-1, n_fft)
        out = np.zeros((len(x), n_fft // 2 + 1), dtype=np.complex64)
        _chk(lib().ref_fft_kiss(C.c_long(n_fft), C.c_long(len(x)), 1, _p(x), _p(out)))
        return out
    x = np.ascontiguousarray(x, dtype=np.complex64)
    n_fft = n_fft or 2 * (x.shape[-1] - 1)
    x = x.reshape(-1, n_fft // 2 + 1)
    out = np.zeros((len(x), n_fft), dtype=np.float32)
    _chk(lib().ref_fft_kiss(C.c_long(n_fft), C.c_long(len(x)), 0, _p(x), _p(out)))
    return out


def vbs_process(block_size, x, n_out, sizes):
    """libear's VariableBlockSizeAdapter over calls of the given sizes around the probe process
    function of ref_capi.cpp::ref_vbs_process.  x: float32 [n_in][sum(sizes)] -> ([n_out][T], delay)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n_in, T = x.shape
    sizes = np.ascontiguousarray(sizes, dtype=np.int64)
    assert int(sizes.sum()) == T
    out = np.zeros((n_out, T), dtype=np.float32)
    delay = C.c_int(0)
    _chk(lib().ref_vbs_process(C.c_long(block_size), n_in, n_out, len(sizes),
                               sizes.ctypes.data_as(C.POINTER(C.c_long)), _p(x), _p(out), C.byref(delay)))
    return out, delay.value


def conversion(pos, ext=None, to_polar=True):
    """ear::conversion::extentCartToPolar / extentPolarToCart (point* when ext is None) per row."""
    pos = _f64(pos).reshape(-1, 3)
    pos_out = np.zeros_like(pos)
    if ext is not None:
        ext = _f64(ext).reshape(-1, 3)
        ext_out = np.zeros_like(ext)
    else:
        ext_out = None
    _chk(lib().ref_conversion(C.c_long(len(pos)), int(bool(to_polar)), _p(pos), _p(ext), _p(pos_out), _p(ext_out)))
    return pos_out, ext_out


def custom_layout(name, names, pos, nominal=None, az_range=None, el_range=None, is_lfe=None):
    """Flat description of an ear::Layout(name, channels); defaults as ear::Channel."""
    pos = _f64(pos).reshape(-1, 3)
    nominal = pos.copy() if nominal is None else _f64(nominal).reshape(-1, 3)
    az_range = np.stack([pos[:, 0], pos[:, 0]], 1) if az_range is None else _f64(az_range).reshape(-1, 2)
    el_range = np.stack([pos[:, 1], pos[:, 1]], 1) if el_range is None else _f64(el_range).reshape(-1, 2)
    is_lfe = np.zeros(len(pos), np.int32) if is_lfe is None else np.ascontiguousarray(is_lfe, dtype=np.int32)
    return dict(name=name, names=list(names), pos=pos, nominal=nominal, az_range=np.ascontiguousarray(az_range),
                el_range=np.ascontiguousarray(el_range), is_lfe=is_lfe)


def _custom_args(lay):
    return (lay["name"].encode(), len(lay["names"]), "\n".join(lay["names"]).encode(), _p(lay["pos"]),
            _p(lay["nominal"]), _p(lay["az_range"]), _p(lay["el_range"]), _p(lay["is_lfe"]))


def convex_hull(xyz):
    """ear::convex_hull: list of facets (sorted vertex index lists) in the reference's order."""
    xyz = _f64(xyz).reshape(-1, 3)
    buf = np.zeros(8 * len(xyz) * 8 + 64, dtype=np.int32)
    n = C.c_int(0)
    _chk(lib().ref_convex_hull(len(xyz), _p(xyz), _p(buf), len(buf), C.byref(n)))
    out, k = [], 0
    while k < n.value:
        c = int(buf[k])
        out.append([int(v) for v in buf[k + 1:k + 1 + c]])
        k += 1 + c
    return out
