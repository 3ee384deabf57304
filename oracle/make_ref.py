"""Recipe that makes the reference's own CPU implementation available to the CPU arm of bench.py.

TEST / MEASUREMENT INFRASTRUCTURE ONLY (never imported by pocket_tts_b200/).

kyutai-labs/pocket-tts is a pure-Python package (pocket_tts/: .py files and the YAML model configs).  The offline
install the contract names (`pip install --no-index --no-build-isolation --target ... /root/reference`) fails in this
image because its build backend (hatchling) is not in the wheelhouse; what that wheel build would do is copy the
package directory, so this recipe does exactly that and nothing else: /root/reference/pocket_tts -> oracle/_ref/pocket_tts,
files unmodified.  oracle/_ref/ is git-ignored (no reference source enters the history) but travels to the GPU box with
the snapshot, like the built .so files.

    python oracle/make_ref.py            # run in the build container (where /root/reference exists)
"""
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = Path("/root/reference/pocket_tts")
DST = HERE / "_ref" / "pocket_tts"


def make_ref(force: bool = False) -> Path | None:
    if not SRC.exists():
        return DST if DST.exists() else None
    if DST.exists() and not force:
        newest_src = max(p.stat().st_mtime for p in SRC.rglob("*") if p.is_file())
        if (DST / ".stamp").exists() and (DST / ".stamp").stat().st_mtime >= newest_src:
            return DST
    if DST.exists():
        for q in [DST, *DST.rglob("*")]:     # the copy inherits the read-only modes of /root/reference
            q.chmod(0o755 if q.is_dir() else 0o644)
        shutil.rmtree(DST)
    DST.parent.mkdir(parents=True, exist_ok=True)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "static"))
    for q in [DST, *DST.rglob("*")]:
        q.chmod(0o755 if q.is_dir() else 0o644)
    (DST / ".stamp").write_text("copied unmodified from /root/reference/pocket_tts by oracle/make_ref.py\n")
    return DST


if __name__ == "__main__":
    out = make_ref(force="--force" in sys.argv)
    print(out if out else "reference not present and no vendored copy")
