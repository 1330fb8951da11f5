"""The reference's internal film file ("YAF_FILMv4_0_0"): save, load, and the merge of every film file of a folder that a
`film_load_save_mode = load-save` render starts from.  This is how libYafaRay checkpoints / resumes a render and how
several computers ("nodes") or GPUs combine their samples, so files written here are read by the reference and vice versa.

Reference code restated:
  ImageFilm::imageFilmSave              src/render/imagefilm.cc:1020-1102  (layout below)
  ImageFilm::imageFilmLoad              src/render/imagefilm.cc:827-936    (header checks, in this order)
  ImageFilm::imageFilmLoadAllInFolder   src/render/imagefilm.cc:938-1018   (sorted file list, sums, max of the offsets)
  ImageFilm::getFilmPath                src/render/imagefilm.cc:817-825    ("<path> - node NNNN.film")
  File::append(std::string) / read      src/common/file.cc:169-194         (bytes + one 0x00)

Layout (native little-endian): "YAF_FILMv4_0_0\\0", u32 computer_node, u32 base_sampling_offset, u32 sampling_offset,
i32 width, height, cx0, cx1, cy0, cy1, i32 number of layers, float weights[h][w], then per layer float rgba[h][w][4].
Layers are stored in LayerDef::Type order.  The reference always defines two: Combined and the internal
DebugSamplingFactor layer (scene.cc:732-735), a grey image of the hit material's `samplingfactor` that scales the sample
count of adaptive passes (integrator_tiled.cc:295-310).  The GPU back end renders Combined only and supports
samplingfactor = 1 with background_resampling = true (the defaults), for which that scale is exactly 1 for every pixel;
it writes the second layer as "factor 1 for every sample" (weights in rgb, Combined's alpha) so that the reference
accepts the file (it checks the layer count) and computes the same scale from it.  Both one- and two-layer files load.
"""
from __future__ import annotations

import os

import numpy as np

MAGIC = b"YAF_FILMv4_0_0"


class FilmFileError(RuntimeError):
    pass


def film_path(path: str, computer_node: int = 0) -> str:
    return "%s - node %04d.film" % (path, computer_node)


def save_film(filename: str, weights: np.ndarray, rgba: np.ndarray, computer_node: int = 0, base_sampling_offset: int = 0,
              sampling_offset: int = 0, cx0: int = 0, cy0: int = 0, sampling_factor_layer: bool = True) -> None:
    h, w = weights.shape
    if rgba.shape != (h, w, 4):
        raise FilmFileError("film colour buffer does not match the weights")
    with open(filename, "wb") as f:
        f.write(MAGIC + b"\0")
        f.write(np.array([computer_node, base_sampling_offset, sampling_offset], np.uint32).tobytes())
        f.write(np.array([w, h, cx0, cx0 + w, cy0, cy0 + h, 2 if sampling_factor_layer else 1], np.int32).tobytes())
        f.write(np.ascontiguousarray(weights, np.float32).tobytes())
        f.write(np.ascontiguousarray(rgba, np.float32).tobytes())
        if sampling_factor_layer:
            sf = np.empty((h, w, 4), np.float32)
            sf[..., :3] = np.asarray(weights, np.float32)[..., None]
            sf[..., 3] = rgba[..., 3]
            f.write(sf.tobytes())


def load_film(filename: str, width: int, height: int, cx0: int = 0, cy0: int = 0, num_layers=(1, 2)) -> dict:
    """-> {"weights", "layers" [num_layers][h][w][4], "computer_node", "base_sampling_offset", "sampling_offset"};
    raises FilmFileError where imageFilmLoad returns false (same checks, same order)."""
    try:
        data = open(filename, "rb").read()
    except OSError:
        raise FilmFileError(f"imageFilm file '{filename}' not found")
    if not data.startswith(MAGIC + b"\0"):
        raise FilmFileError(f"imageFilm file '{filename}' does not contain a valid YafaRay image file")
    pos = len(MAGIC) + 1
    node, base_offs, offs = (int(v) for v in np.frombuffer(data, np.uint32, 3, pos))
    pos += 12
    fw, fh, fcx0, fcx1, fcy0, fcy1, n_layers = (int(v) for v in np.frombuffer(data, np.int32, 7, pos))
    pos += 28
    for what, got, want in (("Image width", fw, width), ("Image height", fh, height), ("Border cx0", fcx0, cx0), ("Border cx1", fcx1, cx0 + width),
                            ("Border cy0", fcy0, cy0), ("Border cy1", fcy1, cy0 + height), ("Number of image layers", n_layers, num_layers)):
        if (got not in want) if isinstance(want, tuple) else (got != want):
            raise FilmFileError(f"imageFilm: loading/reusing film check failed. {what}, expected={want}, in reused/loaded film={got}")
    npx = width * height
    if len(data) < pos + 4 * npx * (1 + 4 * n_layers):
        raise FilmFileError(f"imageFilm file '{filename}' is truncated")
    weights = np.frombuffer(data, np.float32, npx, pos).reshape(height, width).copy()
    pos += 4 * npx
    layers = np.frombuffer(data, np.float32, npx * 4 * n_layers, pos).reshape(n_layers, height, width, 4).copy()
    return {"weights": weights, "layers": layers, "computer_node": node, "base_sampling_offset": base_offs, "sampling_offset": offs}


def film_files_in_folder(path: str) -> list:
    """The files imageFilmLoadAllInFolder picks: extension "film", base name starting with the output's base name, sorted."""
    d = os.path.dirname(path) or "."
    base = os.path.splitext(os.path.basename(path))[0]
    out = []
    for name in os.listdir(d):
        stem, ext = os.path.splitext(name)
        if ext == ".film" and stem.startswith(base) and os.path.isfile(os.path.join(d, name)):
            out.append(os.path.join(d, name))
    return sorted(out)


def load_all_in_folder(path: str, width: int, height: int, cx0: int = 0, cy0: int = 0, base_sampling_offset: int = 0):
    """-> (weights, rgba, base_sampling_offset, sampling_offset, n_loaded): the sums over every loadable film file in the
    folder of `path` (fp32, file by file in sorted order) and the largest offsets; files that fail a check are skipped."""
    weights = np.zeros((height, width), np.float32)
    rgba = np.zeros((height, width, 4), np.float32)
    offs, n = 0, 0
    for fn in film_files_in_folder(path):
        try:
            f = load_film(fn, width, height, cx0, cy0)
        except FilmFileError:
            continue
        weights += f["weights"]
        rgba += f["layers"][0]
        offs = max(offs, f["sampling_offset"])
        base_sampling_offset = max(base_sampling_offset, f["base_sampling_offset"])
        n += 1
    return weights, rgba, base_sampling_offset, offs, n
