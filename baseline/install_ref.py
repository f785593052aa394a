"""Place the UNMODIFIED reference under baseline/_ref (git-ignored, travels to the GPU box).

`pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref
/root/reference` fails because the reference has no setup.py / pyproject.toml ("Directory is not
installable") — it is four scripts, not a package.  The scripts are therefore copied verbatim;
bench.py --impl reference imports them from there and drives their own train() functions.
"""
import hashlib
import shutil
import sys
from pathlib import Path

SRC = Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
DST = Path(__file__).resolve().parent / "_ref"


def main() -> int:
    if not SRC.exists():
        print(f"reference not found at {SRC}")
        return 1
    DST.mkdir(parents=True, exist_ok=True)
    lines = []
    for f in sorted(SRC.glob("*.py")) + [SRC / "README.md", SRC / "LICENSE"]:
        if f.exists():
            shutil.copyfile(f, DST / f.name)
            lines.append(f"{hashlib.sha256(f.read_bytes()).hexdigest()}  {f.name}")
    (DST / "SHA256SUMS").write_text("\n".join(lines) + "\n")
    print(f"copied {len(lines)} files to {DST}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
