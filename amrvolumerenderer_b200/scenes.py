"""Synthetic AMR scenes at the SceneGeometry / AmrBox level.

The reference gets its boxes from ``amrex::convexify`` + ``BuildSceneGeometry``
(VolumeRenderer/SceneBuilder.cpp:113-425).  convexify and the plotfile reader
live in un-vendored AMReX, so benchmarks and parity tests feed ALREADY-CONVEX
box lists with an explicit box->rank map; this module generates them (numpy
only, seeded) for configs C1..C4 and C2' and computes the same derived
quantities BuildSceneGeometry would: global rescale so the shortest domain
edge is 1 (:229-254), 5 % padded bounds (:303-313) and the normalise-to-data-
range ScalarTransform (:354-421).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import ctypes_api as api

DEFAULT_EYE = (1.9, 1.3, 2.4)
DEFAULT_LOOK_AT = (0.5, 0.5, 0.5)
DEFAULT_UP = (0.0, 1.0, 0.0)


@dataclass
class BoxSpec:
    level: int
    lo: tuple            # index-space small end at this level
    dims: tuple          # (nx, ny, nz)
    min_corner: np.ndarray  # float64[3], after global rescale
    max_corner: np.ndarray
    data: np.ndarray     # float64, shape (nz, ny, nx): i fastest == AMReX Fortran order
    owner: int = 0
    gen_index: int = 0   # position in the generator's (level, k, j, i) order


@dataclass
class Scene:
    name: str
    boxes: list          # owner-major, owner-local order (= MPI_Allgatherv order)
    transform: api.ScalarTransform
    scalar_range: tuple
    bounds_min: np.ndarray
    bounds_max: np.ndarray
    nranks: int = 1
    levels: list = field(default_factory=list)  # [(prob_lo[3], cell_size[3])]

    # -------------------------------------------------------------- ctypes views
    def owners(self) -> np.ndarray:
        return np.array([b.owner for b in self.boxes], dtype=np.int32)

    def c_boxes(self, only_rank: int | None = None):
        """(ctypes array of amrvr_box, keepalive) for all boxes or one rank's."""
        sel = [b for b in self.boxes if only_rank is None or b.owner == only_rank]
        # the array is rebuilt only when a box's storage changed (a host application
        # keeps its amrvr_box table between frames the same way)
        key = (only_rank, tuple(id(b.data) for b in sel))
        cache = self.__dict__.setdefault("_cbox_cache", {})
        if key in cache:
            return cache[key], len(sel), sel
        arr = (api.Box * max(len(sel), 1))()
        for i, b in enumerate(sel):
            nx, ny, nz = b.dims
            cb = arr[i]
            for d in range(3):
                cb.min_corner[d] = float(b.min_corner[d])
                cb.max_corner[d] = float(b.max_corner[d])
                cb.dims[d] = int(b.dims[d])
                cb.lo[d] = int(b.lo[d])
                cb.begin[d] = int(b.lo[d])
            assert b.data.dtype == np.float64 and b.data.flags.c_contiguous
            cb.data = b.data.ctypes.data
            cb.jstride = nx
            cb.kstride = nx * ny
            cb.nstride = nx * ny * nz
            cb.component = 0
        cache.clear()
        cache[key] = arr
        return arr, len(sel), sel

    def records(self):
        """Global box table (amrvr_box_record[]) in owner-major order."""
        arr = (api.BoxRecord * max(len(self.boxes), 1))()
        local = {}
        for i, b in enumerate(self.boxes):
            r = arr[i]
            for d in range(3):
                r.min_corner[d] = float(b.min_corner[d])
                r.max_corner[d] = float(b.max_corner[d])
                r.dims[d] = int(b.dims[d])
            r.owner = int(b.owner)
            r.local_index = local.get(b.owner, 0)
            local[b.owner] = r.local_index + 1
        return arr, len(self.boxes)

    def total_cells(self) -> int:
        return int(sum(int(np.prod(b.dims)) for b in self.boxes))

    def with_owners(self, nranks: int, rule: str = "chunks") -> "Scene":
        """Re-assign box->rank: 'chunks' = contiguous ceil(n/P) chunks of the
        (level, k, j, i) order; 'round_robin' = box i -> i mod P; 'skewed' = box i -> (i + i div P) mod P.  The result
        is re-sorted owner-major (stable), the order an allgather produces."""
        base = sorted(self.boxes, key=lambda b: b.gen_index)
        n = len(base)
        per = -(-n // nranks)
        out = []
        for i, b in enumerate(base):
            if rule == "chunks":
                owner = i // per
            elif rule == "skewed":
                # round-robin whose start advances by one every P boxes: boxes of one
                # grid column no longer land on one rank, so every rank gets near and
                # far boxes alike (sample counts within 3-9 % at P = 8 from any side;
                # plain round-robin is 20-38 % off)
                owner = (i + i // nranks) % nranks
            else:
                owner = i % nranks
            nb = BoxSpec(b.level, b.lo, b.dims, b.min_corner, b.max_corner, b.data, owner,
                         b.gen_index)
            out.append(nb)
        out.sort(key=lambda b: b.owner)  # stable
        return Scene(self.name, out, self.transform, self.scalar_range, self.bounds_min,
                     self.bounds_max, nranks, self.levels)


def c_levels(scene: "Scene", only_rank: int | None = None, ghost: int = 0, ncomp: int = 1,
             component: int = 0):
    """The scene as api::AmrData would hand it over: one amrvr_level per AMR level
    with the FABs of `only_rank` (or of all ranks) in generation order.  ghost > 0
    pads every FAB with that many ghost cells (filled with 1e30, never part of
    the valid box) so the Array4 strides differ from the valid extents; ncomp > 1
    stores the field in `component` and junk in the others.
    Returns (Level array, nlevels, keepalive)."""
    nlevels = len(scene.levels)
    per = [[] for _ in range(nlevels)]
    for b in sorted(scene.boxes, key=lambda b: b.gen_index):
        if only_rank is None or b.owner == only_rank:
            per[b.level].append(b)
    levels = (api.Level * max(nlevels, 1))()
    keep = []
    for li in range(nlevels):
        prob_lo, cs = scene.levels[li]
        fabs = (api.Fab * max(len(per[li]), 1))()
        for fi, b in enumerate(per[li]):
            nx, ny, nz = b.dims
            gx, gy, gz = nx + 2 * ghost, ny + 2 * ghost, nz + 2 * ghost
            if ghost == 0 and ncomp == 1:
                arr = b.data
            else:
                arr = np.full((ncomp, gz, gy, gx), 1e30, dtype=np.float64)
                arr[component, ghost:ghost + nz, ghost:ghost + ny, ghost:ghost + nx] = b.data
            keep.append(arr)
            f = fabs[fi]
            for d in range(3):
                f.lo[d] = int(b.lo[d])
                f.hi[d] = int(b.lo[d] + b.dims[d] - 1)
                f.begin[d] = int(b.lo[d]) - ghost
            f.data = arr.ctypes.data
            f.jstride = gx
            f.kstride = gx * gy
            f.nstride = gx * gy * gz
        keep.append(fabs)
        for d in range(3):
            levels[li].prob_lo[d] = float(prob_lo[d])
            levels[li].cell_size[d] = float(cs[d])
        levels[li].nfabs = len(per[li])
        levels[li].fabs = fabs
    return levels, nlevels, keep


def make_camera(eye=DEFAULT_EYE, look_at=DEFAULT_LOOK_AT, up=DEFAULT_UP, fov_y=45.0,
                near=0.1, far=10.0) -> api.Camera:
    cam = api.Camera()
    for d in range(3):
        cam.eye[d] = float(eye[d])
        cam.look_at[d] = float(look_at[d])
        cam.up[d] = float(up[d])
    cam.fov_y_degrees = float(fov_y)
    cam.near_plane = float(near)
    cam.far_plane = float(far)
    return cam


def identity_transform() -> api.ScalarTransform:
    t = api.ScalarTransform()
    t.log_scale_input = 0
    t.normalize_to_unit_range = 0
    t.positive_floor = 0.0
    t.processed_min = 0.0
    t.processed_max = 1.0
    t.inverse_processed_span = 1.0
    t.normalization_min = 0.0
    t.normalization_max = 1.0
    t.inverse_normalization_span = 1.0
    return t


def data_range_transform(vmin: float, vmax: float, log_scale: bool = False,
                         positive_floor: float = 0.0):
    """ScalarTransform + scalarRange as BuildSceneGeometry with
    normalizeToDataRange=true leaves them (SceneBuilder.cpp:354-421, 427-443)."""
    pmin, pmax = float(vmin), float(vmax)
    if log_scale:
        pmin = float(np.log(np.float64(positive_floor)))
        pmax = float(np.log(np.float64(max(vmax, positive_floor))))
    if pmin == pmax:
        pmax = pmin + 1.0
    t = api.ScalarTransform()
    t.log_scale_input = 1 if log_scale else 0
    t.positive_floor = float(positive_floor) if log_scale else 0.0
    t.processed_min = pmin
    t.processed_max = pmax
    t.inverse_processed_span = 1.0 / (pmax - pmin)
    t.normalize_to_unit_range = 1
    t.normalization_min = pmin
    t.normalization_max = pmax
    t.inverse_normalization_span = 1.0 / (pmax - pmin)
    return t, (0.0, 1.0)


def _gauss_field(px, py, pz, rng, shape):
    c1 = (0.5, 0.5, 0.5)
    c2 = (0.62, 0.4, 0.55)
    r1 = (px - c1[0]) ** 2 + (py - c1[1]) ** 2 + (pz - c1[2]) ** 2
    r2 = (px - c2[0]) ** 2 + (py - c2[1]) ** 2 + (pz - c2[2]) ** 2
    u = rng.random(shape)
    return 0.6 * np.exp(-r1 / 0.02) + 0.4 * np.exp(-r2 / 0.005) + 0.05 * u


def _finish(name, raw_boxes, levels, extent_phys, field_kind, seed, nranks, rule):
    """raw_boxes: [(level, lo(3), dims(3))] in generation order."""
    rng = np.random.default_rng(seed)
    # BuildSceneGeometry: corners = probLo + index*cellSize, then *globalScale
    gmin = np.full(3, np.inf)
    gmax = np.full(3, -np.inf)
    corners = []
    for (lvl, lo, dims) in raw_boxes:
        prob_lo, cs = levels[lvl]
        mn = np.array([prob_lo[d] + float(lo[d]) * cs[d] for d in range(3)])
        mx = np.array([prob_lo[d] + float(lo[d] + dims[d]) * cs[d] for d in range(3)])
        corners.append((mn, mx))
        gmin = np.minimum(gmin, mn)
        gmax = np.maximum(gmax, mx)
    ext = np.abs(gmax - gmin)
    min_extent = float(np.min(ext[ext > 0]))
    scale = 1.0 / min_extent
    boxes = []
    vmin, vmax = np.inf, -np.inf
    for gi, ((lvl, lo, dims), (mn, mx)) in enumerate(zip(raw_boxes, corners)):
        nx, ny, nz = dims
        prob_lo, cs = levels[lvl]
        # cell centres in [0,1]^3 "field" coordinates
        ci = (prob_lo[0] + (lo[0] + np.arange(nx) + 0.5) * cs[0]) / extent_phys[0]
        cj = (prob_lo[1] + (lo[1] + np.arange(ny) + 0.5) * cs[1]) / extent_phys[1]
        ck = (prob_lo[2] + (lo[2] + np.arange(nz) + 0.5) * cs[2]) / extent_phys[2]
        if field_kind == "gauss":
            data = _gauss_field(ci[None, None, :], cj[None, :, None], ck[:, None, None],
                                rng, (nz, ny, nx))
        elif field_kind == "hash":
            # C5: one constant per box, h = hash(box id) in [0.5, 1]
            h = 0.5 + 0.5 * (((gi + 1) * 2654435761) % (1 << 32)) / float(1 << 32)
            data = np.full((nz, ny, nx), h)
        else:
            raise ValueError(field_kind)
        data = np.ascontiguousarray(data, dtype=np.float64)
        vmin = min(vmin, float(data.min()))
        vmax = max(vmax, float(data.max()))
        if scale != 1.0:
            mn = mn * scale
            mx = mx * scale
        b = BoxSpec(lvl, tuple(int(v) for v in lo), tuple(int(v) for v in dims), mn, mx, data,
                    0, gi)
        boxes.append(b)
    smin = np.min([b.min_corner for b in boxes], axis=0)
    smax = np.max([b.max_corner for b in boxes], axis=0)
    max_extent = float(np.max(smax - smin))
    pad = max_extent * 0.05 if max_extent > 0 else 1.0
    t, rng_ = data_range_transform(vmin, vmax)
    scene = Scene(name, boxes, t, rng_, smin - pad, smax + pad, 1, levels)
    return scene.with_owners(nranks, rule)


def _grid(level, origin_cells, nboxes_axis, box_cells, skip=None):
    out = []
    for k in range(nboxes_axis):
        for j in range(nboxes_axis):
            for i in range(nboxes_axis):
                if skip is not None and skip(i, j, k):
                    continue
                lo = (origin_cells[0] + i * box_cells, origin_cells[1] + j * box_cells,
                      origin_cells[2] + k * box_cells)
                out.append((level, lo, (box_cells,) * 3))
    return out


def _central(i, j, k):
    return 1 <= i <= 2 and 1 <= j <= 2 and 1 <= k <= 2


def scene_c2(box_cells=64, nranks=1, seed=1001, rule="chunks"):
    """C2: single level, 4x4x4 boxes of box_cells^3 (64 -> 256^3)."""
    n = 4 * box_cells
    levels = [((0.0, 0.0, 0.0), (1.0 / n,) * 3)]
    return _finish("C2", _grid(0, (0, 0, 0), 4, box_cells), levels, (1.0, 1.0, 1.0),
                   "gauss", seed, nranks, rule)


def scene_c3(box_cells=64, nranks=1, seed=1002, rule="chunks"):
    """C3: L0 4^3 boxes minus the central 2^3 (56) + L1 4^3 boxes over [.25,.75]^3."""
    n0 = 4 * box_cells
    levels = [((0.0,) * 3, (1.0 / n0,) * 3), ((0.0,) * 3, (1.0 / (2 * n0),) * 3)]
    raw = _grid(0, (0, 0, 0), 4, box_cells, _central)
    raw += _grid(1, (2 * box_cells,) * 3, 4, box_cells)
    return _finish("C3", raw, levels, (1.0, 1.0, 1.0), "gauss", seed, nranks, rule)


def scene_c4(box_cells=64, nranks=1, seed=1003, rule="chunks"):
    """C4: three levels, 56 + 56 + 64 = 176 boxes (1024^3 effective at 64)."""
    n0 = 4 * box_cells
    levels = [((0.0,) * 3, (1.0 / n0,) * 3), ((0.0,) * 3, (1.0 / (2 * n0),) * 3),
              ((0.0,) * 3, (1.0 / (4 * n0),) * 3)]
    raw = _grid(0, (0, 0, 0), 4, box_cells, _central)
    raw += _grid(1, (2 * box_cells,) * 3, 4, box_cells, _central)
    # L2: 4^3 boxes over [.375,.625]^3 -> origin at 0.375 * 4*n0 = 6*box_cells
    raw += _grid(2, (6 * box_cells,) * 3, 4, box_cells)
    return _finish("C4", raw, levels, (1.0, 1.0, 1.0), "gauss", seed, nranks, rule)


def scene_c5(n=4, box_cells=2, nranks=1, seed=1005, rule="round_robin"):
    """C5 (compositor sweep): n^3 bricks tiling the unit cube, n in {4, 8, 16} ->
    64 / 512 / 4096 layers.  Each brick holds one constant h = hash(box id) in
    [0.5, 1], so a layer is the brick's projected footprint filled with one
    colour and its entry depth; the march is a few samples per ray and the frame
    time is the composite's.  Boxes are dealt round-robin."""
    cells = n * box_cells
    levels = [((0.0,) * 3, (1.0 / cells,) * 3)]
    return _finish(f"C5-{n}", _grid(0, (0, 0, 0), n, box_cells), levels, (1.0, 1.0, 1.0),
                   "hash", seed, nranks, rule)


def scene_c2prime(box_cells=50, nranks=1, seed=1011, rule="chunks"):
    """C2': 3x3x3 boxes over a domain of physical size 1.0 x 1.37 x 0.9
    (non-power-of-two spacing, globalScale != 1, inexact step recurrence)."""
    n = 3 * box_cells
    phys = (1.0, 1.37, 0.9)
    levels = [((0.0,) * 3, tuple(p / n for p in phys))]
    raw = []
    for k in range(3):
        for j in range(3):
            for i in range(3):
                raw.append((0, (i * box_cells, j * box_cells, k * box_cells), (box_cells,) * 3))
    return _finish("C2prime", raw, levels, phys, "gauss", seed, nranks, rule)


def scene_c1(nranks=1):
    """C1: Examples/RenderFromMultiFab.cpp:20-46 -- 32^3 cells in 8 boxes of
    16^3 (maxSize(16)), field x^2+y^2+z^2 with x = i/(n-1)."""
    n, bc = 32, 16
    levels = [((0.0,) * 3, (1.0 / n,) * 3)]
    raw = _grid(0, (0, 0, 0), 2, bc)
    boxes = []
    vmin, vmax = np.inf, -np.inf
    for gi, (lvl, lo, dims) in enumerate(raw):
        i = (lo[0] + np.arange(bc, dtype=np.float64)) / (n - 1)
        j = (lo[1] + np.arange(bc, dtype=np.float64)) / (n - 1)
        k = (lo[2] + np.arange(bc, dtype=np.float64)) / (n - 1)
        x, y, z = i[None, None, :], j[None, :, None], k[:, None, None]
        data = np.ascontiguousarray(x * x + y * y + z * z, dtype=np.float64)
        vmin, vmax = min(vmin, float(data.min())), max(vmax, float(data.max()))
        mn = np.array([0.0 + float(lo[d]) * (1.0 / n) for d in range(3)])
        mx = np.array([0.0 + float(lo[d] + bc) * (1.0 / n) for d in range(3)])
        b = BoxSpec(0, lo, dims, mn, mx, data, 0, gi)
        boxes.append(b)
    t, rng_ = data_range_transform(vmin, vmax)
    pad = 1.0 * 0.05
    scene = Scene("C1", boxes, t, rng_, np.zeros(3) - pad, np.ones(3) + pad, 1, levels)
    return scene.with_owners(nranks, "chunks")


def auto_camera_draws(seed: int = 91021):
    """The two std::uniform_real_distribution<float> draws of
    VolumeRenderer.cpp:992-997 over std::mt19937(seed), as libstdc++ computes
    them (generate_canonical<float,24> takes ONE 32-bit draw:
    (x / 2^32) in float, clamped below 1, then scaled a + (b-a)*u)."""
    rs = np.random.RandomState()
    rs.seed(seed)  # MT19937 init_genrand(seed), identical to std::mt19937(seed)

    def canonical():
        x = np.float32(rs.randint(0, 2 ** 32, dtype=np.uint64))
        u = np.float32(x / np.float32(4294967296.0))
        if u >= np.float32(1.0):
            u = np.nextafter(np.float32(1.0), np.float32(0.0))
        return u

    two_pi = np.float32(2.0) * np.float32(3.14159265358979323846)
    max_alt = np.float32(3.14159265358979323846) * np.float32(0.25)
    az = np.float32(canonical() * (two_pi - np.float32(0.0)) + np.float32(0.0))
    alt = np.float32(canonical() * (max_alt - (-max_alt)) + (-max_alt))
    return float(az), float(alt)
