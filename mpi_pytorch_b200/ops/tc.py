"""Tensor-core (tcgen05 / TMEM / TMA) GEMM and implicit-GEMM convolution front end.

Python face of ``csrc/tc_gemm.cu``.  ``*_ok`` predicates say whether a call can run on the
hand-written sm_100a kernel; anything else is counted as a library fallback by ``ops.cuda_impl``.
Conventions: activations NHWC (channels-last) bf16, conv weights logical (K,C,R,S) with physical
KRSC (channels-last), fp32 accumulation everywhere; weight gradients come back in fp32.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

import torch

from . import scratch as _scratch
from ..build import lib_path

ENABLED = os.environ.get("MPB_TC", "1").lower() not in ("0", "false", "off")
_LIB = ctypes.CDLL(lib_path())
LAUNCHES = {"n": 0}
c_void_p, c_int, c_ll = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
SMS = 148


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return c_void_p(0 if t is None else t.data_ptr())


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError("{} failed with code {}".format(what, rc))


# ----------------------------------------------------------------------------------------------- GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, m: int, n: int, k: int, a_mn: bool, b_mn: bool,
         bias: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16, splitk: int = 1,
         stats: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """D[m,n] = A[m,k] @ B[n,k]^T.  ``a`` is stored [m][k] (a_mn False) or [k][m] (True); same for ``b``.
    ``out`` (fp32 [m,n], contiguous): the product is ADDED to it (fp32 RED) -- a slot of the gradient arena."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous()
    if out is not None:
        assert out.dtype == torch.float32 and out.is_contiguous() and tuple(out.shape) == (m, n)
        d, mode = out, 2
    elif splitk > 1:
        d = torch.zeros((m, n), dtype=torch.float32, device=a.device)
        mode = 2
    else:
        d = torch.empty((m, n), dtype=out_dtype, device=a.device)
        mode = 0 if out_dtype == torch.bfloat16 else 1
    rc = _LIB.mpb_tc_gemm(_p(a), _p(b), _p(d), _p(bias), _p(stats), c_int(m), c_int(n), c_int(k), c_ll(a.stride(0)),
                          c_ll(b.stride(0)), c_ll(n), c_int(int(a_mn)), c_int(int(b_mn)), c_int(mode), c_int(splitk),
                          c_void_p(_stream()))
    LAUNCHES["n"] += 1
    _check(rc, "mpb_tc_gemm")
    return d if splitk == 1 or out_dtype == torch.float32 else d.to(out_dtype)


def gemm_supported(x: torch.Tensor, w: torch.Tensor) -> bool:
    if not ENABLED or x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 2:
        return False
    out_f, in_f = w.shape
    return in_f % 8 == 0 and out_f % 8 == 0 and x.shape[0] % 8 == 0 and w.is_contiguous()


def linear_fwd(x, w, bias):
    return gemm(x, w, x.shape[0], w.shape[0], w.shape[1], False, False, None if bias is None else bias.float())


def linear_dx(gy, w):
    """dX[M, in] = gy[M, out] @ W[out, in].  A wide head (64,500 classes) makes this a long reduction over very few
    output tiles (128 x 512 -> 4 CTAs, 1,008 k-blocks each): split K across CTAs so the GPU is filled."""
    m, n, k = gy.shape[0], w.shape[1], w.shape[0]
    tiles = ((m + 127) // 128) * ((n + 127) // 128)
    splitk = 1
    if tiles < 74 and k >= 4096:
        splitk = max(1, min(k // 512, 296 // tiles))
    return gemm(gy, w, m, n, k, False, True, splitk=splitk)


def linear_dw(gy, x, out=None, rows=None):
    """dW[out, in] = gy^T x (fp32).  ``out``: accumulate into that fp32 buffer (the weight's slot of the gradient
    arena); ``rows``: number of valid output rows when gy carries zero-padded columns."""
    m = gy.shape[1] if rows is None else rows
    return gemm(gy, x, m, x.shape[1], gy.shape[0], True, True, out_dtype=torch.float32, out=out)


# ----------------------------------------------------------------------------------------------- conv
def _geom(x_shape, w_shape, stride, padding):
    n, c, h, w = x_shape
    k, _, r, s = w_shape
    p = (h + 2 * padding[0] - r) // stride[0] + 1
    q = (w + 2 * padding[1] - s) // stride[1] + 1
    return n, c, h, w, k, r, s, p, q


def _base_ok(x, w, stride, padding, dilation, groups) -> bool:
    if not ENABLED or groups != 1 or tuple(dilation) != (1, 1) or stride[0] != stride[1]:
        return False
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 4:
        return False
    if not x.is_contiguous(memory_format=torch.channels_last) or not w.is_contiguous(memory_format=torch.channels_last):
        return False
    return x.data_ptr() % 16 == 0 and w.data_ptr() % 16 == 0


def fprop_ok(x, w, stride, padding, dilation, groups) -> bool:
    # channel tails are handled by TMA out-of-bounds zero fill; 16-byte global strides need C, K multiples of 8
    return _base_ok(x, w, stride, padding, dilation, groups) and x.shape[1] % 8 == 0 and w.shape[0] % 8 == 0


def dgrad_ok(x, w, stride, padding, dilation, groups) -> bool:
    return (_base_ok(x, w, stride, padding, dilation, groups) and x.shape[1] % 8 == 0
            and w.shape[0] % 8 == 0 and padding[0] <= w.shape[2] - 1 and padding[1] <= w.shape[3] - 1)


def wgrad_ok(x, w, stride, padding, dilation, groups) -> bool:
    return _base_ok(x, w, stride, padding, dilation, groups) and x.shape[1] % 8 == 0 and w.shape[0] % 8 == 0


def conv_supported(x, w, stride, padding, dilation, groups) -> bool:
    return fprop_ok(x, w, stride, padding, dilation, groups)


_LIB.mpb_tc_halo_efficiency.restype = ctypes.c_float
HALO = os.environ.get("MPB_TC_HALO", "1").lower() not in ("0", "false", "off")
HALO_MIN_EFF = float(os.environ.get("MPB_TC_HALO_MIN_EFF", "0.6"))


def halo_ok(out_h: int, out_w: int, r: int, s: int, stride, ci: int = 0, co: int = 0) -> bool:
    """Stride-1 filters larger than 1x1 whose padded output rows tile the 128 GEMM rows well enough run on the
    halo-tile kernel (activation tile fetched once per channel block instead of once per tap)."""
    if not HALO or stride[0] != 1 or stride[1] != 1 or r * s == 1:
        return False
    if float(_LIB.mpb_tc_halo_efficiency(c_int(out_h), c_int(out_w), c_int(r), c_int(s))) < HALO_MIN_EFF:
        return False
    return bool(_LIB.mpb_tc_halo_persistent_ok(c_int(ci), c_int(co), c_int(r), c_int(s), c_int(out_h), c_int(out_w)))


WGRAD_HALO = os.environ.get("MPB_TC_WGRAD_HALO", "1") != "0"


def wgrad_halo_ok(c: int, k: int, r: int, s: int, out_h: int, out_w: int, stride) -> bool:
    """Stride-1 filters larger than 1x1 with C == 64 or C % 128 == 0 (and K % 64 == 0) take the persistent halo-tile
    weight-gradient kernel (tc_wgrad_halo.cu): x tile fetched once per filter row instead of once per tap."""
    if not WGRAD_HALO or stride[0] != 1 or stride[1] != 1 or r * s == 1:
        return False
    return bool(_LIB.mpb_tc_wgrad_halo_ok(c_int(c), c_int(k), c_int(r), c_int(s), c_int(out_h), c_int(out_w)))


def conv_fprop(x, w, bias, stride, padding, want_stats=False, relu=False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    n, c, h, wd, k, r, s, p, q = _geom(x.shape, w.shape, stride, padding)
    y = torch.empty((n, k, p, q), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    stats = _scratch.take(2 * k, x.device)[0] if want_stats else None
    if halo_ok(p, q, r, s, stride, c, k):
        rc = _LIB.mpb_tc_conv_fprop_halo_act(_p(x), _p(w), _p(y), _p(None if bias is None else bias.float()), _p(stats),
                                             c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k), c_int(r), c_int(s),
                                             c_int(p), c_int(q), c_int(padding[0]), c_int(padding[1]), c_int(int(relu)),
                                             c_void_p(_stream()))
        LAUNCHES["n"] += 1 + int(want_stats)
        _check(rc, "mpb_tc_conv_fprop_halo")
        return y, stats
    rc = _LIB.mpb_tc_conv_fprop_act(_p(x), _p(w), _p(y), _p(None if bias is None else bias.float()), _p(stats), c_int(n),
                                    c_int(h), c_int(wd), c_int(c), c_int(k), c_int(r), c_int(s), c_int(p), c_int(q),
                                    c_int(stride[0]), c_int(padding[0]), c_int(padding[1]), c_int(1), c_int(int(relu)),
                                    c_void_p(_stream()))
    LAUNCHES["n"] += 1 + int(want_stats)
    _check(rc, "mpb_tc_conv_fprop")
    return y, stats


def conv_dgrad(gy, w, x_shape, stride, padding) -> torch.Tensor:
    n, c, h, wd, k, r, s, p, q = _geom(x_shape, w.shape, stride, padding)
    dx = torch.empty(tuple(x_shape), dtype=torch.bfloat16, device=gy.device, memory_format=torch.channels_last)
    if stride[0] > 1:
        rc = _LIB.mpb_tc_conv_dgrad_strided(_p(gy), _p(w), _p(dx), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k),
                                            c_int(r), c_int(s), c_int(p), c_int(q), c_int(stride[0]), c_int(padding[0]),
                                            c_int(padding[1]), c_void_p(_stream()))
        LAUNCHES["n"] += 1 + stride[0] * stride[0]
        _check(rc, "mpb_tc_conv_dgrad_strided")
        return dx
    if halo_ok(h, wd, r, s, stride, k, c):
        rc = _LIB.mpb_tc_conv_dgrad_halo(_p(gy), _p(w), _p(dx), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k), c_int(r),
                                         c_int(s), c_int(p), c_int(q), c_int(padding[0]), c_int(padding[1]),
                                         c_void_p(_stream()))
        LAUNCHES["n"] += 1
        _check(rc, "mpb_tc_conv_dgrad_halo")
        return dx
    rc = _LIB.mpb_tc_conv_dgrad(_p(gy), _p(w), _p(dx), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k), c_int(r),
                                c_int(s), c_int(p), c_int(q), c_int(padding[0]), c_int(padding[1]), c_void_p(_stream()))
    LAUNCHES["n"] += 1
    _check(rc, "mpb_tc_conv_dgrad")
    return dx


def conv_wgrad(gy, x, w_shape, stride, padding, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 weight gradient.  With ``out`` (a zero-initialised or partially accumulated fp32 buffer in
    physical KRSC order, e.g. the parameter's slot in the gradient arena) the split-K partial sums are
    atomically accumulated straight into it -- no temporary, no separate add."""
    n, c, h, wd, k, r, s, p, q = _geom(x.shape, w_shape, stride, padding)
    dw = torch.zeros((k, r, s, c), dtype=torch.float32, device=gy.device) if out is None else out
    if wgrad_halo_ok(c, k, r, s, p, q, stride):
        rc = _LIB.mpb_tc_conv_wgrad_halo(_p(gy), _p(x), _p(dw), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k),
                                         c_int(r), c_int(s), c_int(p), c_int(q), c_int(padding[0]), c_int(padding[1]),
                                         c_void_p(_stream()))
        LAUNCHES["n"] += 2 if out is None else 1
        _check(rc, "mpb_tc_conv_wgrad_halo")
        return out if out is not None else dw.permute(0, 3, 1, 2)
    bn = 128 if c > 64 else 64
    tiles = ((k + 127) // 128) * r * s * ((c + bn - 1) // bn)
    num_kb = (n * p * q + 63) // 64
    # split-K: enough CTAs to fill the machine (>= 2 per SM) but >= ~96 K-blocks each, so the fp32 atomic
    # epilogue (128 x BN RED ops per CTA) stays a small fraction of the CTA's life (measured sweep)
    lo, hi = (2 * SMS + tiles - 1) // tiles, (8 * SMS + tiles - 1) // tiles
    splitk = max(1, min(hi, max(lo, num_kb // 96)))
    if "MPB_WGRAD_CTAS" in os.environ:
        splitk = max(1, (int(os.environ["MPB_WGRAD_CTAS"]) + tiles - 1) // tiles)
    splitk = min(splitk, max(1, num_kb))
    rc = _LIB.mpb_tc_conv_wgrad(_p(gy), _p(x), _p(dw), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(k), c_int(r),
                                c_int(s), c_int(p), c_int(q), c_int(stride[0]), c_int(padding[0]), c_int(padding[1]),
                                c_int(1), c_int(splitk), c_void_p(_stream()))
    LAUNCHES["n"] += 2 if out is None else 1
    _check(rc, "mpb_tc_conv_wgrad")
    if out is not None:
        return out
    return dw.permute(0, 3, 1, 2)          # logical (K,C,R,S), physical KRSC


# ------------------------------------------------------------------------ explicit-im2col path (C % 8 != 0)
def _kpad(c: int, r: int, s: int) -> int:
    return (c * r * s + 7) // 8 * 8


def im2col_ok(x, w, stride, padding, dilation, groups) -> bool:
    """3-channel stems and other inputs the TMA im2col mode cannot address (pixel rows < 16 bytes)."""
    if not ENABLED or groups != 1 or tuple(dilation) != (1, 1) or stride[0] != stride[1]:
        return False
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or x.dim() != 4 or w.shape[0] % 8 != 0:
        return False
    return x.is_contiguous(memory_format=torch.channels_last) and w.shape[2] < 256 and w.shape[3] < 256


def im2col(x, w_shape, stride, padding) -> torch.Tensor:
    n, c, h, wd, k, r, s, p, q = _geom(x.shape, w_shape, stride, padding)
    kp = _kpad(c, r, s)
    out = torch.empty((n * p * q, kp), dtype=torch.bfloat16, device=x.device)
    rc = _LIB.mpb_im2col_nhwc(_p(x), _p(out), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(r), c_int(s), c_int(p),
                              c_int(q), c_int(stride[0]), c_int(padding[0]), c_int(padding[1]), c_int(kp),
                              c_void_p(_stream()))
    LAUNCHES["n"] += 1
    _check(rc, "mpb_im2col_nhwc")
    return out


def conv_fprop_im2col(x, w, bias, stride, padding, want_stats=False):
    """fprop as im2col + tensor-core GEMM.  Returns (y NHWC, stats, xcol) -- xcol is kept for the wgrad GEMM."""
    n, c, h, wd, k, r, s, p, q = _geom(x.shape, w.shape, stride, padding)
    kp = _kpad(c, r, s)
    xcol = im2col(x, w.shape, stride, padding)
    wmat = torch.zeros((k, kp), dtype=torch.bfloat16, device=x.device)
    wmat[:, : c * r * s] = w.permute(0, 2, 3, 1).reshape(k, c * r * s)          # KRSC flattened
    stats = _scratch.take(2 * k, x.device)[0] if want_stats else None
    y2d = gemm(xcol, wmat, n * p * q, k, kp, False, False, None if bias is None else bias.float(), stats=stats)
    y = y2d.view(n, p, q, k).permute(0, 3, 1, 2)                                  # logical NCHW, physical NHWC
    return y, stats, xcol


def conv_wgrad_im2col(gy, xcol, w_shape) -> torch.Tensor:
    k, c, r, s = w_shape
    m, kp = xcol.shape
    gy2d = gy.permute(0, 2, 3, 1).reshape(m, k)
    splitk = max(1, min(8 * SMS // max(1, (kp + 127) // 128), (m + 63) // 64 // 96))
    dw = gemm(gy2d, xcol, k, kp, m, True, True, out_dtype=torch.float32, splitk=max(2, splitk))
    return dw[:, : c * r * s].reshape(k, r, s, c).permute(0, 3, 1, 2)            # logical (K,C,R,S)


# ------------------------------------------------------------------------------------------- space-to-depth stems
STEM_S2D = os.environ.get("MPB_TC_STEM_S2D", "1") != "0"
_S2D_IDX = {}


def stem_s2d_ok(x, w, stride, padding, dilation, groups) -> bool:
    """7x7 stride-2 convolutions over <= 4 input channels (ResNet / DenseNet / SqueezeNet stems) run as a 4x1 stride-1
    convolution with 64 'channels' over a space-to-depth copy of the input (no im2col matrix); models.py:33,68,81."""
    if not STEM_S2D or x.dtype != torch.bfloat16 or groups != 1 or tuple(dilation) != (1, 1) or tuple(stride) != (2, 2):
        return False
    k, c, r, s = w.shape
    return (r == 7 and s == 7 and 4 * c <= 16 and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0 and k % 64 == 0
            and padding[0] == padding[1] and 0 <= padding[0] <= 3)


def _s2d_index(c: int, pad: int, device):
    """Gather maps between the KRSC filter (7*7*c per output channel, + one zero slot) and its s2d form [4][64]."""
    key = (c, pad, str(device))
    if key not in _S2D_IDX:
        off, zero = pad % 2, 49 * c
        fwd = torch.full((4 * 64,), -1, dtype=torch.int64)      # -1: a zero column (mpb_gather_cols)
        inv = torch.zeros(49 * c, dtype=torch.int64)
        for a in range(4):
            for a2 in range(4):
                for b in range(2):
                    for b2 in range(2):
                        r, s = 2 * a + b - off, 2 * a2 + b2 - off
                        if 0 <= r < 7 and 0 <= s < 7:
                            for ch in range(c):
                                dst = a * 64 + a2 * 16 + (b * 2 + b2) * c + ch
                                fwd[dst] = (r * 7 + s) * c + ch
                                inv[(r * 7 + s) * c + ch] = dst
        _S2D_IDX[key] = (fwd.to(device), inv.to(device))
    return _S2D_IDX[key]


def _s2d_geom(x_shape, w_shape, padding):
    n, c, h, wd = x_shape
    k = w_shape[0]
    pad = padding[0]
    p, q = (h + 2 * pad - 7) // 2 + 1, (wd + 2 * pad - 7) // 2 + 1
    lpad = (pad + 1) // 2                              # rows / pixels of zero padding in s2d space (top / left)
    wp = max(q + 3, lpad + wd // 2)
    return n, c, h, wd, k, p, q, lpad, wp


def conv_fprop_stem(x, w, bias, padding, want_stats=False):
    """-> (y, stats, xs2d).  xs2d ([N, H/2, Wp, 16] bf16) is kept for the wgrad."""
    n, c, h, wd, k, p, q, lpad, wp = _s2d_geom(x.shape, w.shape, padding)
    xs = torch.empty((n, h // 2, wp, 16), dtype=torch.bfloat16, device=x.device)
    rc = _LIB.mpb_s2d_stem(_p(x), _p(xs), c_int(n), c_int(h), c_int(wd), c_int(c), c_int(wp), c_int(lpad),
                           c_void_p(_stream()))
    _check(rc, "mpb_s2d_stem")
    fwd, _ = _s2d_index(c, padding[0], x.device)
    wk = w.permute(0, 2, 3, 1)                       # KRSC view: free for a channels-last filter (the bf16 shadow)
    if not wk.is_contiguous():
        wk = wk.contiguous()
    w2 = torch.empty((k, 4 * 64), dtype=torch.bfloat16, device=x.device)                # [K][4 taps][64]
    rc = _LIB.mpb_gather_cols(_p(wk), _p(fwd), _p(w2), c_int(k), c_int(49 * c), c_int(4 * 64),
                              c_int(0 if wk.dtype == torch.float32 else 1), c_int(1), c_int(0), c_void_p(_stream()))
    _check(rc, "mpb_gather_cols")
    y = torch.empty((n, k, p, q), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    stats = _scratch.take(2 * k, x.device)[0] if want_stats else None
    ll = ctypes.c_longlong
    rc = _LIB.mpb_tc_conv_fprop_halo_strided(_p(xs), _p(w2), _p(y), _p(None if bias is None else bias.float()), _p(stats),
                                             c_int(n), c_int(h // 2), c_int(wp - 3), c_int(64), c_int(k), c_int(4),
                                             c_int(1), c_int(p), c_int(q), c_int(lpad), c_int(0), ll(16), ll(wp * 16),
                                             ll((h // 2) * wp * 16), c_void_p(_stream()))
    LAUNCHES["n"] += 2 + int(want_stats)
    _check(rc, "mpb_tc_conv_fprop_halo_strided")
    return y, stats, xs


def conv_wgrad_stem(gy, xs, x_shape, w_shape, padding, out=None) -> torch.Tensor:
    """Filter gradient of the space-to-depth stem.  With ``out`` (the fp32 KRSC slot of the gradient arena, logical
    shape (K,C,7,7)) the result is ADDED there by the gather kernel and ``out`` is returned."""
    n, c, h, wd, k, p, q, lpad, wp = _s2d_geom(x_shape, w_shape, padding)
    dw2 = _scratch.take(k * 4 * 64, gy.device)[0].view(k, 4 * 64)
    ll = ctypes.c_longlong
    rc = _LIB.mpb_tc_conv_wgrad_halo_strided(_p(gy), _p(xs), _p(dw2), c_int(n), c_int(h // 2), c_int(wp - 3), c_int(64),
                                             c_int(k), c_int(4), c_int(1), c_int(p), c_int(q), c_int(lpad), c_int(0),
                                             ll(16), ll(wp * 16), ll((h // 2) * wp * 16), c_void_p(_stream()))
    LAUNCHES["n"] += 1
    _check(rc, "mpb_tc_conv_wgrad_halo_strided")
    _, inv = _s2d_index(c, padding[0], gy.device)
    acc = out is not None
    dst = out.permute(0, 2, 3, 1) if acc else torch.empty((k, 7, 7, c), dtype=torch.float32, device=gy.device)
    rc = _LIB.mpb_gather_cols(_p(dw2), _p(inv), _p(dst), c_int(k), c_int(4 * 64), c_int(49 * c), c_int(0), c_int(0),
                              c_int(int(acc)), c_void_p(_stream()))
    LAUNCHES["n"] += 1
    _check(rc, "mpb_gather_cols")
    return out if acc else dst.permute(0, 3, 1, 2)                                  # logical (K,C,R,S)
