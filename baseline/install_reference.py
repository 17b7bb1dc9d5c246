"""Places an UNMODIFIED copy of the reference's scripts under baseline/_ref (git-ignored).

`pip install /root/reference` is not possible: the reference is four script directories with
no setup.py / pyproject.toml (pip: "Directory '/root/reference' is not installable").  The
copy made here is byte-identical to the source tree; `python baseline/install_reference.py
--verify` re-checks that against the recorded digests (baseline/reference.sha256).
"""
import argparse
import hashlib
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
DST = HERE / "_ref"
DIGESTS = HERE / "reference.sha256"
FILES = ["part1/main.py", "part1/model.py", "part2/part2a/main.py", "part2/part2a/model.py",
         "part2/part2b/main.py", "part2/part2b/model.py", "part3/main.py", "part3/model.py", "README.md"]


def sha(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default="/root/reference")
    ap.add_argument("--verify", action="store_true")
    a = ap.parse_args()
    if a.verify:
        want = dict(line.split() [::-1] for line in DIGESTS.read_text().splitlines())
        bad = [f for f in FILES if not (DST / f).exists() or sha(DST / f) != want[f]]
        print("reference copy OK" if not bad else f"MISMATCH: {bad}")
        return 1 if bad else 0
    src = Path(a.src)
    lines = []
    for f in FILES:
        (DST / f).parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(src / f, DST / f)
        lines.append(f"{sha(DST / f)}  {f}")
    DIGESTS.write_text("\n".join(lines) + "\n")
    print(f"copied {len(FILES)} files to {DST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
