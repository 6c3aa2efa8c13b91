"""Writes <out>/base (a valid two-level plotfile) and <out>/mNNN (400 seeded
corruptions of it, the mutations of tests/fuzz_plotfile_worker.py)."""
import pathlib
import random
import shutil
import sys

ROOT = pathlib.Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from fuzz_plotfile_worker import corrupt  # noqa: E402
from amrvolumerenderer_b200.plotfile import synthetic_hierarchy, write_plotfile  # noqa: E402

out = pathlib.Path(sys.argv[1])
shutil.rmtree(out, ignore_errors=True)
out.mkdir(parents=True)
levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=2, ncomp=2)
base = write_plotfile(out / "base", levels, ("rho", "p"), plo, phi, nghost=1)
rng = random.Random(99)
for k in range(400):
    d = out / f"m{k:03d}"
    shutil.copytree(base, d)
    files = sorted(p for p in d.rglob("*") if p.is_file())
    headers = [p for p in files if p.name in ("Header", "Cell_H")]
    victim = rng.choice(headers if rng.random() < 0.75 else files)
    victim.write_bytes(corrupt(victim.read_bytes(), rng))
