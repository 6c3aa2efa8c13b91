"""oracle/amr_io_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

numpy restatement of the two un-vendored AMReX pieces the reference calls on
the way into the render path (AMReX 26.04, pinned at CMakeLists.txt:44-52 of
the reference; its source is NOT under /root/reference):

  * amrex::PlotFileData  (call site VolumeRenderer/VolumeRenderer.cpp:599-679)
  * amrex::convexify     (call sites VolumeRenderer.cpp:669-670,
                          VolumeRendererApi.cpp:97-98)

PARITY UNPINNED: the reference holds no test, fixture or golden vector for
either, and AMReX cannot be built here, so these restate the published
on-disk format and the set semantics (coarse cells under a finer level are
dropped) only.  The exact box cuts and box->rank map AMReX would produce are
not reproduced; the checks built on this file are therefore set-level
(cell masks) and byte-level (FAB payloads), not box-for-box.
Only tests/ may import this module.
"""
from __future__ import annotations

import re
from pathlib import Path

import numpy as np

_BOX = re.compile(r"\(\((-?\d+),(-?\d+),(-?\d+)\) \((-?\d+),(-?\d+),(-?\d+)\)(?: \((\d+),(\d+),(\d+)\))?\)")


def _boxes(text: str):
    return [tuple(int(v) for v in m.groups()[:6]) for m in _BOX.finditer(text)]


def read_plotfile(path):
    """Plain-Python reading of Header / Cell_H / Cell_D_*: returns a dict with
    names, prob_lo, prob_hi, ref_ratio, and per level: cell_size, domain, boxes
    [(lo, hi)], fabs [ndarray (ncomp, nz, ny, nx) float64 of the VALID box]."""
    path = Path(path)
    lines = (path / "Header").read_text().split("\n")
    it = iter(lines)
    next(it)
    ncomp = int(next(it))
    names = [next(it).strip() for _ in range(ncomp)]
    dim = int(next(it))
    time = float(next(it))
    finest = int(next(it))
    prob_lo = [float(v) for v in next(it).split()]
    prob_hi = [float(v) for v in next(it).split()]
    ratios = [int(v) for v in next(it).split()]
    domains = _boxes(next(it))
    next(it)  # level steps
    cell = [[float(v) for v in next(it).split()] for _ in range(finest + 1)]
    next(it)
    next(it)
    levels = []
    for l in range(finest + 1):
        _lev, ngrids, _t = next(it).split()
        next(it)
        for _ in range(int(ngrids) * dim):
            next(it)
        rel = next(it).strip()
        head = (path / (rel + "_H")).read_text()
        hl = head.split("\n")
        assert hl[0].strip() == "1"
        file_ncomp = int(hl[2])
        assert hl[4].startswith("(")
        n = int(hl[4][1:].split()[0])
        boxes = _boxes(head)[:n]
        fod = re.findall(r"FabOnDisk: (\S+) (\d+)", head)
        assert len(fod) == n == int(ngrids)
        fabs = []
        for (lo_hi, (fname, off)) in zip(boxes, fod):
            raw = (path / Path(rel).parent / fname).read_bytes()
            off = int(off)
            end = raw.index(b"\n", off)
            line = raw[off:end].decode()
            m = re.match(r"FAB \(\((\d+), \(([\d ]+)\)\),\((\d+), \(([\d ]+)\)\)\)", line)
            bits = int(m.group(2).split()[0])
            order = [int(v) for v in m.group(4).split()]
            fb = _boxes(line)[0]
            nc = int(line.split()[-1])
            assert nc == file_ncomp
            little = order[0] == len(order)
            dt = np.dtype(("<" if little else ">") + ("f8" if bits == 64 else "f4"))
            fx, fy, fz = (fb[3 + d] - fb[d] + 1 for d in range(3))
            arr = np.frombuffer(raw, dt, nc * fx * fy * fz, end + 1).reshape(nc, fz, fy, fx)
            sl = tuple(slice(lo_hi[d] - fb[d], lo_hi[3 + d] - fb[d] + 1) for d in (2, 1, 0))
            fabs.append(arr[(slice(None),) + sl].astype(np.float64))
        levels.append({"cell_size": cell[l], "domain": domains[l], "boxes": boxes, "fabs": fabs})
    return {"names": names, "dim": dim, "time": time, "finest_level": finest, "prob_lo": prob_lo,
            "prob_hi": prob_hi, "ref_ratio": ratios, "levels": levels}


def uncovered_mask(domain, coarse_boxes, fine_boxes, ratio):
    """Boolean mask over `domain` (lo[3], hi[3]) of the coarse cells that belong
    to a coarse box and lie under NO fine box (fine boxes coarsened by floor
    division, as amrex::coarsen)."""
    lo, hi = np.array(domain[:3]), np.array(domain[3:])
    shape = tuple((hi - lo + 1)[::-1])  # (z, y, x)
    have = np.zeros(shape, bool)
    for b in coarse_boxes:
        have[b[2] - lo[2]:b[5] - lo[2] + 1, b[1] - lo[1]:b[4] - lo[1] + 1,
             b[0] - lo[0]:b[3] - lo[0] + 1] = True
    for b in fine_boxes:
        c = [int(np.floor_divide(b[d], ratio[d % 3])) for d in range(6)]
        c = [max(c[d], lo[d]) for d in range(3)] + [min(c[3 + d], hi[d]) for d in range(3)]
        if all(c[3 + d] >= c[d] for d in range(3)):
            have[c[2] - lo[2]:c[5] - lo[2] + 1, c[1] - lo[1]:c[4] - lo[1] + 1,
                 c[0] - lo[0]:c[3] - lo[0] + 1] = False
    return have


def pieces_mask(domain, pieces):
    """(mask, overlap count max) painted from a list of piece boxes."""
    lo, hi = np.array(domain[:3]), np.array(domain[3:])
    count = np.zeros(tuple((hi - lo + 1)[::-1]), np.int32)
    for b in pieces:
        count[b[2] - lo[2]:b[5] - lo[2] + 1, b[1] - lo[1]:b[4] - lo[1] + 1,
              b[0] - lo[0]:b[3] - lo[0] + 1] += 1
    return count > 0, int(count.max()) if count.size else 0
