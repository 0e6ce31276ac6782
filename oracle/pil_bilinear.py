"""Oracle (test infrastructure): numpy restatement of the image preprocess path.

Follows
  * ``/root/reference/src/python/examples/image_client.py:154-193`` (``preprocess``; the same
    body is at ``grpc_image_client.py:171-210``), and
  * Pillow 12.2.0's 8-bit BILINEAR resampler, which the reference reaches through
    ``Image.resize((w, h), Image.BILINEAR)`` at ``image_client.py:166``. Pillow is a
    third-party dependency that is absent from ``/root/reference`` and unpinned by it; the
    arithmetic below restates its published algorithm (``src/libImaging/Resample.c``:
    ``precompute_coeffs``, ``normalize_coeffs_8bpc``, ``ImagingResampleHorizontal_8bpc``,
    ``ImagingResampleVertical_8bpc``; ``src/libImaging/Convert.c``: ``rgb2l``; and the
    Python-level pass-order rule in ``PIL/Image.py::Image.resize``).

Two implementations of ``preprocess`` live here:
  * ``preprocess_pil``  - the reference's function body, calling the installed Pillow (this is
    exactly what the reference executes; used as the timed CPU baseline and as a checker
    when Pillow is importable).
  * ``preprocess_restated`` - no Pillow: the resampler arithmetic restated in numpy. It
    documents the integer arithmetic the CUDA kernel must reproduce bit for bit.
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2  # Pillow Resample.c: 22-bit fixed point coefficients

# model_config.proto ModelInput.Format (un-vendored in the reference; values from
# image_client.py:185 usage and SURVEY.md 8a): FORMAT_NONE=0, FORMAT_NHWC=1, FORMAT_NCHW=2
FORMAT_NONE, FORMAT_NHWC, FORMAT_NCHW = 0, 1, 2


def triton_to_np_dtype(dtype):
    """Restates tritonclient/utils/__init__.py:163-190."""
    return {
        "BOOL": bool, "INT8": np.int8, "INT16": np.int16, "INT32": np.int32,
        "INT64": np.int64, "UINT8": np.uint8, "UINT16": np.uint16, "UINT32": np.uint32,
        "UINT64": np.uint64, "FP16": np.float16, "FP32": np.float32, "BF16": np.float32,
        "FP64": np.float64, "BYTES": np.object_,
    }.get(dtype)


def _tri(x):
    # Resample.c bilinear_filter
    if x < 0.0:
        x = -x
    if x < 1.0:
        return 1.0 - x
    return 0.0


def precompute_coeffs(in_size, out_size):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the BILINEAR filter (support 1.0)
    over the full axis (box = [0, in_size)).

    Returns (bounds[out,2] int32 = (xmin, xmax_count), kk[out,ksize] int32, ksize).
    All floating point is IEEE double evaluated in the same order as the C code
    (sequential accumulation of ww, multiply by the reciprocal ss, then divide by ww).
    """
    scale = float(in_size) / float(out_size)
    filterscale = scale
    if filterscale < 1.0:
        filterscale = 1.0
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        ww = 0.0
        xmin = int(center - support + 0.5)  # C (int) cast: truncation toward zero
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [0.0] * xmax
        for x in range(xmax):
            v = _tri((x + xmin - center + 0.5) * ss)
            w[x] = v
            ww += v
        for x in range(xmax):
            if ww != 0.0:
                w[x] /= ww
            # normalize_coeffs_8bpc: (int)(+-0.5 + w * 2^22)
            if w[x] < 0:
                kk[xx, x] = int(-0.5 + w[x] * (1 << PRECISION_BITS))
            else:
                kk[xx, x] = int(0.5 + w[x] * (1 << PRECISION_BITS))
        bounds[xx, 0] = xmin
        bounds[xx, 1] = xmax
    return bounds, kk, ksize


def _clip8_shift(acc):
    # Resample.c clip8(): lookup table indexed by (acc >> PRECISION_BITS), i.e. an arithmetic
    # shift followed by a clamp to [0, 255].
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def resample_axis(src, out_size, axis):
    """One separable 8-bit pass (ImagingResampleHorizontal_8bpc for axis=1,
    ImagingResampleVertical_8bpc for axis=0) over src[H, W, C] uint8."""
    in_size = src.shape[axis]
    bounds, kk, _ = precompute_coeffs(in_size, out_size)
    out_shape = list(src.shape)
    out_shape[axis] = out_size
    out = np.empty(out_shape, dtype=np.uint8)
    s = src.astype(np.int64)
    for xx in range(out_size):
        xmin, n = int(bounds[xx, 0]), int(bounds[xx, 1])
        k = kk[xx, :n].astype(np.int64)
        if axis == 1:
            acc = (1 << (PRECISION_BITS - 1)) + np.tensordot(s[:, xmin:xmin + n, :], k, axes=([1], [0]))
            out[:, xx, :] = _clip8_shift(acc)
        else:
            acc = (1 << (PRECISION_BITS - 1)) + np.tensordot(s[xmin:xmin + n, :, :], k, axes=([0], [0]))
            out[xx, :, :] = _clip8_shift(acc)
    return out


def vertical_first(in_w, in_h, out_w, out_h):
    """PIL/Image.py::Image.resize (Pillow 12.2.0): tall-narrow images are reduced vertically
    first. Because the intermediate is rounded to uint8 the order changes results by +-1."""
    return in_h > in_w * 100 and out_h < in_h


def resize_bilinear_u8(src, out_w, out_h):
    """Image.resize((out_w, out_h), Image.BILINEAR) for an 8-bit multi-band image held as
    src[H, W, C] uint8 (C in {1, 3}). Resample.c ImagingResampleInner: horizontal pass first
    (skipped when the width is unchanged), uint8 intermediate, then the vertical pass
    (skipped when the height is unchanged); same size -> copy."""
    assert src.dtype == np.uint8 and src.ndim == 3
    in_h, in_w = src.shape[0], src.shape[1]
    if (in_w, in_h) == (out_w, out_h):
        return src.copy()
    if vertical_first(in_w, in_h, out_w, out_h):
        tmp = resample_axis(src, out_h, 0)            # (in_w, out_h); height changes by rule
        return resample_axis(tmp, out_w, 1) if out_w != in_w else tmp
    tmp = resample_axis(src, out_w, 1) if out_w != in_w else src
    return resample_axis(tmp, out_h, 0) if out_h != in_h else tmp.copy()


def rgb_to_l(rgb):
    """Convert.c rgb2l: L = (R*19595 + G*38470 + B*7471 + 0x8000) >> 16."""
    r = rgb[..., 0].astype(np.uint32)
    g = rgb[..., 1].astype(np.uint32)
    b = rgb[..., 2].astype(np.uint32)
    return ((r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16).astype(np.uint8)


def convert_u8(src, c):
    """img.convert("L" if c == 1 else "RGB") (image_client.py:161-164) for src[H, W, Cin] uint8
    with Cin in {1 (L), 3 (RGB), 4 (RGBA)}."""
    cin = src.shape[2]
    if c == 1:
        if cin == 1:
            return src
        return rgb_to_l(src[..., :3])[..., None]
    if cin == 1:
        return np.repeat(src, 3, axis=2)       # Convert.c l2rgb: replicate
    return np.ascontiguousarray(src[..., :3])  # rgba2rgb drops alpha; rgb->rgb copies


def scaling_lut(dtype, c, scaling):
    """The 256-entry per-channel table of image_client.py:171-182 evaluated WITH NUMPY ITSELF
    on every possible uint8 value: astype(npdtype) then INCEPTION / VGG / NONE. Returns
    lut[c, 256] whose dtype is whatever numpy promotion yields (FP32->f32, FP16->f16 with
    numpy's double rounding, UINT8+INCEPTION->f64, UINT8+VGG wraps mod 256, ...)."""
    npdtype = triton_to_np_dtype(dtype)
    resized = np.broadcast_to(np.arange(256, dtype=np.uint8)[:, None, None], (256, 1, c)).copy()
    typed = resized.astype(npdtype)
    if scaling == "INCEPTION":
        scaled = (typed / 127.5) - 1
    elif scaling == "VGG":
        if c == 1:
            scaled = typed - np.asarray((128,), dtype=npdtype)
        else:
            scaled = typed - np.asarray((123, 117, 104), dtype=npdtype)
    else:
        scaled = typed
    return np.ascontiguousarray(scaled[:, 0, :].T)  # [c, 256]


def preprocess_restated(src, format, dtype, c, h, w, scaling, protocol=None):
    """image_client.py:154-193 without Pillow: src is the image as uint8 [H, W, Cin]."""
    if src.ndim == 2:
        src = src[:, :, None]
    sample = convert_u8(src, c)
    resized = resize_bilinear_u8(sample, w, h)
    npdtype = triton_to_np_dtype(dtype)
    typed = resized.astype(npdtype)
    if scaling == "INCEPTION":
        scaled = (typed / 127.5) - 1
    elif scaling == "VGG":
        if c == 1:
            scaled = typed - np.asarray((128,), dtype=npdtype)
        else:
            scaled = typed - np.asarray((123, 117, 104), dtype=npdtype)
    else:
        scaled = typed
    if format == FORMAT_NCHW:
        ordered = np.transpose(scaled, (2, 0, 1))
    else:
        ordered = scaled
    return ordered


def preprocess_pil(img, format, dtype, c, h, w, scaling, protocol=None):
    """The reference's preprocess body (image_client.py:154-193) with the installed Pillow:
    this is what the reference itself executes on the CPU."""
    from PIL import Image

    if c == 1:
        sample_img = img.convert("L")
    else:
        sample_img = img.convert("RGB")
    resized_img = sample_img.resize((w, h), Image.BILINEAR)
    resized = np.array(resized_img)
    if resized.ndim == 2:
        resized = resized[:, :, np.newaxis]
    npdtype = triton_to_np_dtype(dtype)
    typed = resized.astype(npdtype)
    if scaling == "INCEPTION":
        scaled = (typed / 127.5) - 1
    elif scaling == "VGG":
        if c == 1:
            scaled = typed - np.asarray((128,), dtype=npdtype)
        else:
            scaled = typed - np.asarray((123, 117, 104), dtype=npdtype)
    else:
        scaled = typed
    if format == FORMAT_NCHW:
        ordered = np.transpose(scaled, (2, 0, 1))
    else:
        ordered = scaled
    return ordered
