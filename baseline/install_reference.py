#!/usr/bin/env python
"""Install the UNMODIFIED reference into ``baseline/_ref`` for the ``bench.py --impl reference`` arm.

The reference tree has no ``setup.py`` / ``pyproject.toml`` (it is a loose ``src`` package plus scripts), so
``pip install /root/reference`` fails with "not installable".  This script therefore
  1. copies the tree to a temp dir (the mount is read-only),
  2. adds ONLY packaging metadata there (a 10-line ``pyproject.toml`` naming the ``src`` package) -
     every ``.py`` file is byte-identical to the reference, which ``verify()`` checks by hash,
  3. runs the prescribed offline install:
     ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --no-deps --target baseline/_ref``,
  4. writes two import stand-ins next to it for packages that do not exist on this box and cannot be
     installed: ``structlog`` (a logging adapter; skipped if the real one is importable) and an empty
     ``smdistributed.dataparallel.torch.torch_smddp`` (the reference imports it unconditionally in parallel
     mode, src/trainer.py:43-44, then the harness passes backend='nccl').
``baseline/_ref`` is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
TARGET = os.path.join(HERE, "_ref")
REFERENCE = os.environ.get("MLT_REFERENCE_DIR", "/root/reference")

PYPROJECT = """[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"

[project]
name = "ml-trainer-reference"
version = "0.0.0"

[tool.setuptools]
packages = ["src", "src.utils"]
"""

STRUCTLOG_SHIM = '''"""Stand-in for the missing ``structlog`` package (bench harness only; not part of the reference)."""
import logging

class _L:
    def __init__(self, name):
        self._l = logging.getLogger(name)
    def _fmt(self, ev, kw):
        return str(ev) + ("" if not kw else " " + " ".join(f"{k}={v!r}" for k, v in kw.items()))
    def info(self, ev, **kw): self._l.info(self._fmt(ev, kw))
    def debug(self, ev, **kw): self._l.debug(self._fmt(ev, kw))
    def warning(self, ev, **kw): self._l.warning(self._fmt(ev, kw))
    warn = warning
    def error(self, ev, **kw): self._l.error(self._fmt(ev, kw))

def get_logger(name=None, *a, **k):
    return _L(name or "structlog")
'''


def _sha(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def verify() -> bool:
    """True when every installed reference .py is byte-identical to the mounted reference."""
    ok = os.path.exists(os.path.join(TARGET, "src", "trainer.py"))
    if ok and os.path.isdir(REFERENCE):
        for rel in ("src/trainer.py", "src/model.py", "src/dataloader.py", "src/utils/functions.py", "src/utils/utils.py"):
            ok = ok and _sha(os.path.join(REFERENCE, rel)) == _sha(os.path.join(TARGET, rel))
    return ok


def install(force: bool = False) -> str:
    if verify() and not force:
        return "already installed"
    if not os.path.isdir(REFERENCE):
        raise RuntimeError(f"reference tree {REFERENCE} not mounted; cannot install")
    tmp = tempfile.mkdtemp(prefix="mlt_ref_")
    try:
        work = os.path.join(tmp, "reference")
        shutil.copytree(REFERENCE, work)
        with open(os.path.join(work, "pyproject.toml"), "w") as f:
            f.write(PYPROJECT)
        shutil.rmtree(TARGET, ignore_errors=True)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
               "/opt/wheelhouse", "--no-deps", "--target", TARGET, work]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"pip install of the reference failed:\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
        for name in ("main.py", "test.py"):          # scripts are not part of the package; keep them beside it
            shutil.copy(os.path.join(REFERENCE, name), os.path.join(TARGET, name))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    shims = os.path.join(TARGET, "_shims")
    os.makedirs(os.path.join(shims, "smdistributed", "dataparallel", "torch"), exist_ok=True)
    for sub in ("smdistributed", "smdistributed/dataparallel", "smdistributed/dataparallel/torch"):
        open(os.path.join(shims, sub, "__init__.py"), "w").close()
    with open(os.path.join(shims, "smdistributed", "dataparallel", "torch", "torch_smddp.py"), "w") as f:
        f.write('"""Empty stand-in: the SageMaker-only smddp backend is not used; the harness passes backend=nccl."""\n')
    os.makedirs(os.path.join(shims, "structlog"), exist_ok=True)
    with open(os.path.join(shims, "structlog", "__init__.py"), "w") as f:
        f.write(STRUCTLOG_SHIM)
    if not verify():
        raise RuntimeError("installed reference differs from the mounted tree")
    return "installed"


if __name__ == "__main__":
    print(install(force="--force" in sys.argv))
