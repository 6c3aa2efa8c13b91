"""AMReX plotfiles: the reader the render path uses (ctypes over
amrvr_plotfile_*) and a writer for the synthetic inputs of
the tests and benchmarks (there is no AMReX on this machine to produce them).

On-disk layout, as amrex::WriteMultiLevelPlotfile / VisMF::Write lay it down
and amrex::PlotFileData (used at VolumeRenderer/VolumeRenderer.cpp:599-679)
reads it back:

    <dir>/Header                 text: names, geometry, per-level grid tables
    <dir>/Level_n/Cell_H         text: VisMF header (BoxArray + FabOnDisk table)
    <dir>/Level_n/Cell_D_00000   FAB records: one text line, then the reals of
                                 all components, component-major, i fastest
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from pathlib import Path

import numpy as np

from . import ctypes_api as api

_P = C.POINTER

_REAL_DESCRIPTORS = {
    # (precision, big_endian) -> (RealDescriptor text, numpy dtype)
    ("double", False): ("((8, (64 11 52 0 1 12 0 1023)),(8, (8 7 6 5 4 3 2 1)))", "<f8"),
    ("double", True): ("((8, (64 11 52 0 1 12 0 1023)),(8, (1 2 3 4 5 6 7 8)))", ">f8"),
    ("single", False): ("((8, (32 8 23 0 1 9 0 127)),(4, (4 3 2 1)))", "<f4"),
    ("single", True): ("((8, (32 8 23 0 1 9 0 127)),(4, (1 2 3 4)))", ">f4"),
}


@dataclass
class LevelSpec:
    """One AMR level to write: boxes = [(lo[3], data)], data shaped
    (ncomp, nz, ny, nx) or (nz, ny, nx) -- i fastest, as in a FAB."""

    domain_lo: tuple
    domain_hi: tuple
    cell_size: tuple
    boxes: list


def _box_text(lo, hi) -> str:
    return f"(({lo[0]},{lo[1]},{lo[2]}) ({hi[0]},{hi[1]},{hi[2]}) (0,0,0))"


def write_plotfile(path, levels, var_names=("density",), prob_lo=(0.0, 0.0, 0.0),
                   prob_hi=(1.0, 1.0, 1.0), ref_ratio=2, time=0.0, precision="double",
                   big_endian=False, nghost=0, fabs_per_file=8) -> Path:
    """Write `levels` (list of LevelSpec, coarsest first) as an AMReX plotfile.
    nghost > 0 stores every FAB grown by that many cells (filled with 1e30), as a
    MultiFab written with ghost cells would be."""
    path = Path(path)
    path.mkdir(parents=True, exist_ok=True)
    ncomp = len(var_names)
    nlev = len(levels)
    ratios = [ref_ratio] * max(nlev - 1, 0) if np.isscalar(ref_ratio) else list(ref_ratio)
    desc, dtype = _REAL_DESCRIPTORS[(precision, bool(big_endian))]

    h = ["HyperCLaw-V1.1", str(ncomp), *var_names, "3", repr(float(time)), str(nlev - 1),
         " ".join(repr(float(v)) for v in prob_lo) + " ",
         " ".join(repr(float(v)) for v in prob_hi) + " ",
         " ".join(str(int(r)) for r in ratios) + " ",
         " ".join(_box_text(L.domain_lo, L.domain_hi) for L in levels) + " ",
         " ".join("0" for _ in levels) + " "]
    for L in levels:
        h.append(" ".join(repr(float(v)) for v in L.cell_size) + " ")
    h += ["0", "0"]

    for li, L in enumerate(levels):
        ldir = path / f"Level_{li}"
        ldir.mkdir(exist_ok=True)
        h.append(f"{li} {len(L.boxes)} {float(time)!r}")
        h.append("0")
        table, mins, maxs, spans = [], [], [], []
        handle, current = None, None
        for bi, (lo, data) in enumerate(L.boxes):
            data = np.asarray(data, dtype=np.float64)
            if data.ndim == 3:
                data = data[None]
            assert data.shape[0] == ncomp, "component count differs from var_names"
            nz, ny, nx = data.shape[1:]
            hi = (lo[0] + nx - 1, lo[1] + ny - 1, lo[2] + nz - 1)
            spans.append((lo, hi))
            for d, n in enumerate((nx, ny, nz)):
                h.append(f"{prob_lo[d] + lo[d] * L.cell_size[d]!r} "
                         f"{prob_lo[d] + (lo[d] + n) * L.cell_size[d]!r}")
            if nghost:
                g = nghost
                grown = np.full((ncomp, nz + 2 * g, ny + 2 * g, nx + 2 * g), 1e30)
                grown[:, g:g + nz, g:g + ny, g:g + nx] = data
                disk, dlo, dhi = grown, tuple(v - g for v in lo), tuple(v + g for v in hi)
            else:
                disk, dlo, dhi = data, lo, hi
            fname = f"Cell_D_{bi // max(fabs_per_file, 1):05d}"
            if fname != current:
                if handle is not None:
                    handle.close()
                handle = open(ldir / fname, "wb")
                current = fname
            table.append((fname, handle.tell()))
            handle.write(f"FAB {desc}{_box_text(dlo, dhi)} {ncomp}\n".encode())
            handle.write(np.ascontiguousarray(disk).astype(dtype).tobytes())
            mins.append(data.reshape(ncomp, -1).min(axis=1))
            maxs.append(data.reshape(ncomp, -1).max(axis=1))
        if handle is not None:
            handle.close()
        h.append(f"Level_{li}/Cell")

        c = ["1", "1", str(ncomp), str(nghost), f"({len(L.boxes)} 0"]
        c += [_box_text(lo, hi) for lo, hi in spans]
        c += [")", str(len(L.boxes))]
        c += [f"FabOnDisk: {name} {off}" for name, off in table]
        c.append("")
        for block in (mins, maxs):
            c.append(f"{len(L.boxes)},{ncomp}")
            c += ["".join(f"{v:.17e}," for v in row) for row in block]
            c.append("")
        (ldir / "Cell_H").write_text("\n".join(c) + "\n")
    (path / "Header").write_text("\n".join(h) + "\n")
    return path


class PlotFile:
    """amrvr_plotfile handle; mirrors the amrex::PlotFileData calls the
    reference makes (spaceDim, varNames, finestLevel, probLo, cellSize,
    refRatio, boxArray, get)."""

    def __init__(self, path: os.PathLike | str):
        self.lib = api.load_library()
        self._h = C.c_void_p()
        status = self.lib.amrvr_plotfile_open(os.fsencode(str(path)), C.byref(self._h))
        if status != 0:
            msg = self.lib.amrvr_last_error(None).decode()
            if status == 1:
                raise ValueError(msg)  # std::invalid_argument
            raise api.AmrvrError(status, msg)
        info = api.PlotfileInfo()
        self._check(self.lib.amrvr_plotfile_get_info(self._h, C.byref(info)))
        self.info = info
        self.space_dim = int(info.space_dim)
        self.finest_level = int(info.finest_level)
        self.prob_lo = tuple(info.prob_lo)
        self.prob_hi = tuple(info.prob_hi)
        self.var_names = [self.lib.amrvr_plotfile_var_name(self._h, c).decode()
                          for c in range(info.ncomp)]

    def _check(self, status):
        if status != 0:
            raise api.AmrvrError(status, self.lib.amrvr_last_error(None).decode())

    def close(self):
        if self._h:
            self.lib.amrvr_plotfile_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def var_index(self, name: str) -> int:
        return int(self.lib.amrvr_plotfile_var_index(self._h, name.encode()))

    def level(self, level: int) -> api.PlotfileLevel:
        out = api.PlotfileLevel()
        self._check(self.lib.amrvr_plotfile_get_level(self._h, level, C.byref(out)))
        return out

    def boxes(self, level: int) -> np.ndarray:
        """int32[n, 6]: lo[3], hi[3] of every box of the level's BoxArray."""
        n = self.level(level).nboxes
        out = np.zeros((max(n, 1), 6), np.int32)
        self._check(self.lib.amrvr_plotfile_level_boxes(
            self._h, level, out.ctypes.data_as(_P(C.c_int32)), n))
        return out[:n]

    def read_fab(self, level: int, box: int, component: int = 0) -> np.ndarray:
        b = self.boxes(level)[box]
        nx, ny, nz = (int(b[3 + d] - b[d] + 1) for d in range(3))
        out = np.empty((nz, ny, nx), np.float64)
        self._check(self.lib.amrvr_plotfile_read_fab(self._h, level, box, component,
                                                     out.ctypes.data_as(C.c_void_p), out.size))
        return out

    def level_share(self, level: int, rank: int = 0, nranks: int = 1, rule: int = 0):
        """(box indices, element offsets incl. the total as last entry) of the
        FABs read_level() hands to `rank`."""
        n = self.level(level).nboxes
        idx = np.zeros(n + 1, np.int32)
        off = np.zeros(n + 1, np.int64)
        got = C.c_int32(0)
        self._check(self.lib.amrvr_plotfile_read_level(
            self._h, level, 0, rank, nranks, rule, 1, None, 0,
            idx.ctypes.data_as(_P(C.c_int32)), off.ctypes.data_as(_P(C.c_int64)), n + 1,
            C.byref(got)))
        k = int(got.value)
        return idx[:k].copy(), off[:k + 1].copy()

    def read_level(self, level: int, component: int, dst: np.ndarray, rank: int = 0,
                   nranks: int = 1, rule: int = 0, threads: int = 8):
        """This rank's FABs of one level, densely one after another into dst
        (float64, typically pinned).  Returns (box indices, offsets)."""
        idx, off = self.level_share(level, rank, nranks, rule)
        assert dst.dtype == np.float64 and dst.flags.c_contiguous and dst.size >= int(off[-1])
        got = C.c_int32(0)
        self._check(self.lib.amrvr_plotfile_read_level(
            self._h, level, component, rank, nranks, rule, threads,
            dst.ctypes.data_as(C.c_void_p), dst.size, None, None, 0, C.byref(got)))
        return idx, off


# ------------------------------------------------------- synthetic hierarchies
def synthetic_hierarchy(base_cells=32, box_cells=16, nlevels=3, seed=1003, ncomp=1,
                        component=0, extent=(1.0, 1.0, 1.0)):
    """A properly nested AMR hierarchy as a simulation code would dump it, i.e.
    NOT convex: every level covers its whole refined region, the coarse data
    under finer levels included.  Level 0 = base_cells^3 over `extent`; level
    l+1 refines the central half of level l by 2.  Field = SURVEY.md 8(d)'s two
    Gaussians + 5 % seeded noise, in `component`; other components hold junk.
    Returns (list of LevelSpec, prob_lo, prob_hi)."""
    rng = np.random.default_rng(seed)
    levels = []
    lo_cells = np.zeros(3, np.int64)
    n_cells = np.array([base_cells] * 3, np.int64)
    for l in range(nlevels):
        dx = [extent[d] / (base_cells * 2 ** l) for d in range(3)]
        boxes = []
        per_axis = [int(n_cells[d] // box_cells) for d in range(3)]
        assert all(p * box_cells == n_cells[d] for d, p in enumerate(per_axis))
        for kz in range(per_axis[2]):
            for jy in range(per_axis[1]):
                for ix in range(per_axis[0]):
                    lo = (int(lo_cells[0] + ix * box_cells), int(lo_cells[1] + jy * box_cells),
                          int(lo_cells[2] + kz * box_cells))
                    ax = [(lo[d] + np.arange(box_cells) + 0.5) * dx[d] / extent[d] for d in range(3)]
                    pz, py, px = np.meshgrid(ax[2], ax[1], ax[0], indexing="ij")
                    f = (0.6 * np.exp(-((px - .5) ** 2 + (py - .5) ** 2 + (pz - .5) ** 2) / 0.02)
                         + 0.4 * np.exp(-((px - .62) ** 2 + (py - .4) ** 2 + (pz - .55) ** 2) / 0.005)
                         + 0.05 * rng.random(px.shape))
                    data = np.full((ncomp,) + f.shape, -7.0)
                    data[component] = f
                    boxes.append((lo, data))
        levels.append(LevelSpec(tuple(int(v) for v in (0, 0, 0)),
                                tuple(int(base_cells * 2 ** l - 1) for _ in range(3)),
                                tuple(dx), boxes))
        # next level: central half of this level's region, refined by 2
        lo_cells = (lo_cells + n_cells // 4) * 2
        n_cells = (n_cells // 2) * 2
    return levels, (0.0, 0.0, 0.0), tuple(float(e) for e in extent)
