"""Child process of test_plotfile_reader_survives_corrupted_files: corrupts one
file of a small two-level plotfile per iteration and reads everything the C ABI
offers from it.  Prints "<opened> <refused>"; any other exception, or a crash,
fails the parent test.  usage: fuzz_plotfile_worker.py <scratch dir> <seed> <n>"""
import os
import pathlib
import random
import shutil
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

from amrvolumerenderer_b200.plotfile import PlotFile, synthetic_hierarchy, write_plotfile  # noqa: E402

JUNK = [b"", b"-1", b"99999999999", b"1e400", b"abc", b"((0,0,0) (7,7,7) (0,0,0))"]


def corrupt(raw: bytes, rng: random.Random) -> bytes:
    b = bytearray(raw)
    mode = rng.randrange(4)
    if mode == 0:  # truncate
        return bytes(b[:rng.randrange(len(b) + 1)])
    if mode == 1:  # overwrite 1-3 bytes with header-ish characters
        for _ in range(rng.randrange(1, 4)):
            b[rng.randrange(len(b))] = rng.choice(b"0123456789 -\n()x,.e")
        return bytes(b)
    lines = bytes(b).split(b"\n")
    i = rng.randrange(len(lines))
    if mode == 2:  # replace or double a line
        lines[i] = rng.choice(JUNK + [lines[i] + lines[i]])
    else:  # drop a line
        del lines[i]
    return b"\n".join(lines)


def read_everything(path) -> None:
    with PlotFile(path) as pf:
        for level in range(pf.finest_level + 1):
            boxes = pf.boxes(level)
            for k in range(min(len(boxes), 3)):
                pf.read_fab(level, k, 0)
            cells = int(sum(np.prod((b[3:] - b[:3] + 1).astype(np.int64)) for b in boxes))
            pf.read_level(level, 0, np.empty(cells, np.float64), 0, 2, 1, threads=3)


def main(scratch: str, seed: int, count: int) -> None:
    scratch = pathlib.Path(scratch)
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=2, ncomp=2)
    base = write_plotfile(scratch / "base", levels, ("rho", "p"), plo, phi, nghost=1)
    read_everything(base)
    rng = random.Random(seed)
    opened = refused = 0
    for _ in range(count):
        work = scratch / "mutated"
        shutil.rmtree(work, ignore_errors=True)
        shutil.copytree(base, work)
        files = sorted(p for p in work.rglob("*") if p.is_file())
        headers = [p for p in files if p.name in ("Header", "Cell_H")]
        victim = rng.choice(headers if rng.random() < 0.75 else files)
        victim.write_bytes(corrupt(victim.read_bytes(), rng))
        try:
            read_everything(work)
            opened += 1
        except (RuntimeError, ValueError):  # AmrvrError is a RuntimeError
            refused += 1
    print(opened, refused)


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
