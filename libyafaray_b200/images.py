"""Image loading for image textures: decodes a file into the float RGBA texels `Image::getColor(x, y)` returns in
the reference, i.e. after the loader's colour-space conversion and the storage rounding of the image's
optimisation mode.  Host-side data preparation for `y4gpu_pt_set_textures` (include/y4gpu.h); the lookups
themselves run on the device (csrc/y4_pt_tex.cuh).

Reference code restated (numpy fp32/fp64 with the reference's operation order):
  Image::factory                        src/image/image.cc:38-101   (HDR files force LinearRGB + no optimisation)
  TgaFormat::loadFromFile / read*Image  src/format/format_tga.cc:92-135,277-395
  HdrFormat::loadFromFile / readArle    src/format/format_hdr.cc:37-310, include/format/format_hdr_util.h:98-107
  Rgb::linearRgbFromColorSpace          include/color/color.h:338-380 (math::pow is the FAST_MATH exp2(log2(a) * b))
  Rgba1010108 / Rgb101010 storage       include/image/image_buffers.h:236-296

Formats: the two loaders libYafaRay carries without external libraries (TGA true-colour / grey, Radiance HDR).  A build
of the reference without libpng/libjpeg/libtiff/OpenEXR (the default configure on this image, oracle/Makefile) cannot
load the other formats: createImage fails there, the texture and every material node that uses it are dropped and the
material falls back to its plain colour (NodeMaterial::loadNodes, src/material/material_node.cc:102-169).  `load_image`
raises ImageLoadError for them and the Interface mirrors that fall-back.
"""
from __future__ import annotations

import os

import numpy as np

f32 = np.float32


class ImageLoadError(RuntimeError):
    pass


# ---- FAST_MATH pow (include/math/math.h:102-179) -------------------------------------------------------
def _polyexp(x):
    return x * (x * (x * (x * (x * f32(1.8775767e-3) + f32(8.9893397e-3)) + f32(5.5826318e-2)) + f32(2.4015361e-1)) + f32(6.9315308e-1)) + f32(9.9999994e-1)


def _polylog(x):
    return x * (x * (x * (x * (x * f32(-3.4436006e-2) + f32(3.1821337e-1)) + f32(-1.2315303)) + f32(2.5988452)) + f32(-3.3241990)) + f32(3.1157899)


def fast_log2(x: np.ndarray) -> np.ndarray:
    i = np.ascontiguousarray(x, f32).view(np.int32)
    e = (((i & 0x7F800000) >> 23) - 127).astype(f32)
    m = ((i & 0x7FFFFF) | np.float32(1.0).view(np.int32)).view(f32)
    return _polylog(m) * (m - f32(1.0)) + e


def fast_exp2(x: np.ndarray) -> np.ndarray:
    x = np.maximum(np.minimum(np.asarray(x, f32), f32(129.0)), f32(-126.99999))
    ipart = (x - f32(0.5)).astype(np.int32)              # static_cast<int>: truncation
    fpart = x - ipart.astype(f32)
    expipart = ((ipart + 127) << 23).astype(np.int32).view(f32)
    return expipart * _polyexp(fpart)


def fast_pow(a: np.ndarray, b: float) -> np.ndarray:
    return fast_exp2(fast_log2(a) * f32(b))


def linear_from_srgb(v: np.ndarray) -> np.ndarray:
    """Rgb::linearRgbFromSRgb (color.h:338-343)."""
    v = np.asarray(v, f32)
    with np.errstate(all="ignore"):
        hi = fast_pow((v + f32(0.055)) / f32(1.055), 2.4)
    return np.where(v <= f32(0.04045), v / f32(12.92), hi).astype(f32)


def to_linear(rgba: np.ndarray, color_space: str, gamma: float) -> np.ndarray:
    """Rgba::linearRgbFromColorSpace: alpha is not converted."""
    out = np.array(rgba, f32, copy=True)
    if color_space == "sRGB":
        out[..., :3] = linear_from_srgb(out[..., :3])
    elif color_space == "XYZ":
        m = np.array([[3.2406255, -1.537208, -0.4986286], [-0.9689307, 1.8757561, 0.0415175], [0.0557101, -0.2040211, 1.0569959]], f32)
        r, g, b = out[..., 0].copy(), out[..., 1].copy(), out[..., 2].copy()
        for k in range(3):
            out[..., k] = (m[k, 0] * r + m[k, 1] * g) + m[k, 2] * b
    elif color_space == "LinearRGB":
        pass
    elif f32(gamma) != f32(1.0):         # Raw_Manual_Gamma (also the default for unknown names)
        with np.errstate(all="ignore"):
            out[..., :3] = fast_pow(out[..., :3], gamma)
    return out


def _roundf(x):
    """C roundf: halves away from zero (the inputs here are non-negative)."""
    return np.floor(np.asarray(x, f32) + f32(0.5))


def store(rgba: np.ndarray, has_alpha: bool, optimization: str) -> np.ndarray:
    """What getColor returns after setColor into the image buffer of this type / optimisation (image.cc:103-140)."""
    rgba = np.asarray(rgba, f32)
    out = np.empty_like(rgba)
    if optimization == "none":
        out[...] = rgba
        if not has_alpha:
            out[..., 3] = 1.0
        return out
    if optimization == "optimized":      # Rgba1010108 / Rgb101010
        q = _roundf(rgba[..., :3] * f32(1023.0)).astype(np.uint16).astype(f32)       # (uint16_t) roundf(c * 1023.f)
        out[..., :3] = q / f32(1023.0)
        if has_alpha:
            out[..., 3] = _roundf(rgba[..., 3] * f32(255.0)).astype(np.uint8).astype(f32) / f32(255.0)
        else:
            out[..., 3] = 1.0
        return out
    raise ImageLoadError(f"image_optimization '{optimization}' is not supported by the GPU back end")


# ---- TGA (format_tga.cc) ------------------------------------------------------------------------------
def _load_tga(data: bytes):
    if len(data) < 18:
        raise ImageLoadError("TGA: short file")
    id_len, cm_type, img_type = data[0], data[1], data[2]
    cm_entry_bits = data[7]
    width, height = int.from_bytes(data[12:14], "little"), int.from_bytes(data[14:16], "little")
    bits, desc = data[16], data[17]
    alpha_bits = desc & 0x0F
    if img_type in (1, 9) or cm_type:
        raise ImageLoadError("TGA: colour-mapped files are not supported by the GPU back end")
    if img_type not in (2, 3, 10, 11):
        raise ImageLoadError("TGA: no image data")
    is_gray, is_rle = img_type in (3, 11), img_type in (10, 11)
    if (is_gray and bits != 8) or (not is_gray and bits not in (24, 32)):
        raise ImageLoadError(f"TGA: bit depth {bits} is not supported by the GPU back end")
    if bits == 32 and alpha_bits != 8:
        raise ImageLoadError("TGA: invalid alpha bit depth for a 32 bit colour image")
    bpp = bits // 8
    n = width * height
    body = np.frombuffer(data, np.uint8, offset=18 + id_len)
    if is_rle:
        px = np.empty((n, bpp), np.uint8)
        i = pos = 0
        while i < n:
            d = int(body[pos]); pos += 1
            rep = (d & 0x7F) + 1
            if d & 0x80:
                px[i:i + rep] = body[pos:pos + bpp]; pos += bpp
            else:
                px[i:i + rep] = body[pos:pos + rep * bpp].reshape(rep, bpp); pos += rep * bpp
            i += rep
    else:
        px = body[:n * bpp].reshape(n, bpp)
    inv255 = 1.0 / 255.0                                            # static constexpr double inv_max_8_bit_
    rgba = np.ones((n, 4), f32)
    if is_gray:
        rgba[:, :] = (px[:, 0].astype(np.float64) * inv255).astype(f32)[:, None]      # Rgba(float): all four channels
    else:
        for k, src in enumerate((2, 1, 0)):                          # file order b, g, r, [a]
            rgba[:, k] = (px[:, src].astype(np.float64) * inv255).astype(f32)
        if bpp == 4:
            rgba[:, 3] = (px[:, 3].astype(np.float64) * inv255).astype(f32)
    img = rgba.reshape(height, width, 4)
    if not (desc & 0x20):                # from_top == false: rows arrive bottom-up
        img = img[::-1]
    if desc & 0x10:                      # "from_left" bit set: columns arrive right to left
        img = img[:, ::-1]
    has_alpha = alpha_bits != 0 or cm_entry_bits == 32 or bits == 32      # Image::Type::ColorAlpha rule of loadFromFile
    return np.ascontiguousarray(img), has_alpha


# ---- Radiance HDR (format_hdr.cc) ----------------------------------------------------------------------
def _load_hdr(data: bytes):
    pos = data.find(b"\n")
    if pos < 0 or b"#?" not in data[:pos]:
        raise ImageLoadError("HDR: not a Radiance RGBE image")
    pos += 1
    while True:
        end = data.find(b"\n", pos)
        if end < 0:
            raise ImageLoadError("HDR: truncated header")
        line = data[pos:end]
        pos = end + 1
        if line == b"":
            break
        if b"FORMAT=" in line and b"32-bit_rle_rgbe" not in line:
            raise ImageLoadError("HDR: only RGBE images are supported")
    end = data.find(b"\n", pos)
    tok = data[pos:end].decode("ascii", "replace").split()
    pos = end + 1
    if len(tok) != 4 or "Y" not in tok[0]:
        raise ImageLoadError("HDR: only row-major (Y first) files are supported by the GPU back end")
    height, width = int(tok[1]), int(tok[3])
    from_left, from_top = "+" in tok[2], "-" in tok[0]
    body = np.frombuffer(data, np.uint8, offset=pos)
    rgbe = np.empty((height, width, 4), np.uint8)
    p = 0
    for y in range(height):
        if 8 <= width <= 0x7FFF and body[p] == 2 and body[p + 1] == 2 and ((int(body[p + 2]) << 8) | int(body[p + 3])) < 0x8000:
            if ((int(body[p + 2]) << 8) | int(body[p + 3])) != width:
                raise ImageLoadError("HDR: invalid ARLE scanline width")
            p += 4
            for ch in range(4):
                j = 0
                while j < width:
                    c = int(body[p]); p += 1
                    if c > 128:
                        c &= 0x7F
                        rgbe[y, j:j + c, ch] = body[p]; p += 1
                    else:
                        rgbe[y, j:j + c, ch] = body[p:p + c]; p += c
                    j += c
        else:                               # flat pixels (old-style RLE runs are not produced by current writers)
            row = body[p:p + 4 * width].reshape(width, 4)
            if ((row[:, 0] == 1) & (row[:, 1] == 1) & (row[:, 2] == 1)).any():
                raise ImageLoadError("HDR: old-style RLE scanlines are not supported by the GPU back end")
            rgbe[y] = row; p += 4 * width
    e = rgbe[..., 3].astype(np.int32)
    f = np.ldexp(f32(1.0), e - (128 + 8)).astype(f32)                 # math::ldexp(1.f, e - (128 + 8))
    rgba = np.ones((height, width, 4), f32)
    for k in range(3):
        rgba[..., k] = np.where(e != 0, f * rgbe[..., k].astype(f32), f32(0.0))
    if not from_top:
        rgba = rgba[::-1]
    if not from_left:
        rgba = rgba[:, ::-1]
    return np.ascontiguousarray(rgba), True


LOADERS = {"tga": _load_tga, "tpic": _load_tga, "hdr": _load_hdr, "pic": _load_hdr}


def empty_image(params: dict) -> np.ndarray:
    """Image::factory for an image without a file (image.cc:38-60,103-140): width x height texels of the cleared buffer."""
    t = str(params.get("type", "ColorAlpha"))
    if t not in ("ColorAlpha", "Color"):
        raise ImageLoadError(f"in-memory image type '{t}' is not supported by the GPU back end (ColorAlpha / Color)")
    w, h = int(params.get("width", 100)), int(params.get("height", 100))
    img = np.zeros((h, w, 4), f32)
    if t == "Color":
        img[..., 3] = 1.0
    elif _optimization(params) == "none":
        img[..., 3] = 0.0        # RgbAlpha buffers start as Rgba{0.f}; the optimised ones as all-zero bits (alpha 0 as well)
    return img


def _optimization(params: dict) -> str:
    o = str(params.get("image_optimization", "optimized"))
    return o if o in ("none", "optimized", "compressed") else "optimized"


def storage_of(params: dict, texels: np.ndarray) -> tuple:
    """(has alpha channel, optimisation) of the buffer createImage made: what setImageColor rounds through.
    `params` is the dict load_image saw (it records the file's alpha rule under "_has_alpha")."""
    fname = str(params.get("filename", ""))
    ext = os.path.splitext(fname)[1].lower().lstrip(".")
    if ext in ("hdr", "pic"):
        return True, "none"
    if fname:
        return bool(params.get("_has_alpha", True)), _optimization(params)
    return str(params.get("type", "ColorAlpha")) == "ColorAlpha", _optimization(params)


def load_image(params: dict, search_dir: str = ".") -> np.ndarray:
    """Image::factory(params) for an image with a `filename` -> [height, width, 4] float32 texels (row y, column x)."""
    filename = str(params.get("filename", ""))
    if not filename:
        raise ImageLoadError("images without a file are not supported by the GPU back end")
    path = filename if os.path.isabs(filename) else os.path.join(search_dir, filename)
    ext = os.path.splitext(filename)[1].lower().lstrip(".")
    if ext not in LOADERS:
        raise ImageLoadError(f"no loader for '{ext}' files (libYafaRay built without the optional image libraries)")
    try:
        with open(path, "rb") as fh:
            data = fh.read()
    except OSError as exc:
        raise ImageLoadError(f"cannot open {path}: {exc}") from exc
    color_space, gamma = str(params.get("color_space", "Raw_Manual_Gamma")), float(params.get("gamma", 1.0))
    optimization = str(params.get("image_optimization", "optimized"))
    if optimization not in ("none", "optimized", "compressed"):
        optimization = "optimized"
    raw, has_alpha = LOADERS[ext](data)
    if ext in ("hdr", "pic"):
        color_space, optimization = "LinearRGB", "none"
    params["_has_alpha"] = has_alpha          # for setImageColor on a loaded image (storage_of)
    return np.ascontiguousarray(store(to_linear(raw, color_space, gamma), has_alpha, optimization), f32)
