"""Installs the UNMODIFIED reference into the git-ignored ``baseline/_ref`` (idempotent).

Called from ``__graft_entry__.build()`` (so the install exists in the working tree before it is
shipped to a GPU box) and, on demand, from ``bench.py --impl reference``.

The documented command is used as is::

    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
        --no-deps --target baseline/_ref <copy of /root/reference>

``/root/reference`` is read-only and setuptools writes ``build/`` + ``*.egg-info`` into the source
tree, hence the temporary copy.  ``--no-deps``: the reference's requirements (cyy_*, hydra, gevent,
torch_geometric, ...) exist neither in the image nor in the wheelhouse; ``baseline/shims`` stands in
for them at run time.
"""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "baseline", "_ref")
PACKAGE = os.path.join(REF_DIR, "distributed_learning_simulator")
SOURCES = [os.environ.get("DLS_REFERENCE_SRC", ""), "/root/reference"]


class ReferenceUnavailable(RuntimeError):
    pass


def is_installed() -> bool:
    return os.path.isfile(os.path.join(PACKAGE, "training.py"))


def ensure_installed(verbose: bool = False) -> str:
    """Returns the install directory; raises ``ReferenceUnavailable`` with the reason otherwise."""
    if is_installed():
        return REF_DIR
    src = next((s for s in SOURCES if s and os.path.isfile(os.path.join(s, "setup.py"))), None)
    if src is None:
        raise ReferenceUnavailable(
            "baseline/_ref is not installed and no reference source tree (/root/reference) exists on this machine")
    tmp = tempfile.mkdtemp(prefix="dls_refcopy_")
    try:
        copy = os.path.join(tmp, "reference")
        shutil.copytree(src, copy, symlinks=True, ignore=shutil.ignore_patterns(".git"))
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
               "/opt/wheelhouse", "--no-deps", "--upgrade", "--target", REF_DIR, copy]
        proc = subprocess.run(cmd, capture_output=True, text=True, cwd=tmp)
        if verbose:
            print(proc.stdout[-2000:])
        if proc.returncode != 0 or not is_installed():
            raise ReferenceUnavailable("pip install of the reference failed: " + (proc.stderr or proc.stdout)[-400:])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return REF_DIR


if __name__ == "__main__":
    try:
        print(ensure_installed(verbose=True))
    except ReferenceUnavailable as e:
        print(f"reference unavailable: {e}", file=sys.stderr)
        sys.exit(1)
