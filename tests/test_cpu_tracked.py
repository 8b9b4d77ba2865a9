"""Release hygiene: everything the package imports must be in git (a clean ``git clone`` has to import and run).
Round 1 shipped a tree whose ``dba_mod_b200/data`` package was swallowed by an unanchored ``data/`` ignore pattern;
this test fails on that class of mistake."""
import importlib
import os
import pkgutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _git(*args):
    return subprocess.run(["git", "-C", ROOT, *args], capture_output=True, text=True)


def _sources():
    out = []
    for top in ("dba_mod_b200", "tests", "tools", "baseline", "csrc", "include", "utils"):
        for dp, dn, fn in os.walk(os.path.join(ROOT, top)):
            dn[:] = [d for d in dn if d not in ("__pycache__", "_ref")]
            out += [os.path.relpath(os.path.join(dp, f), ROOT) for f in fn
                    if f.endswith((".py", ".cu", ".cuh", ".h", ".yaml", ".sh"))]
    return out + ["main.py", "bench.py", "__graft_entry__.py"]


def test_every_module_imports():
    import dba_mod_b200
    names = [m.name for m in pkgutil.walk_packages(dba_mod_b200.__path__, "dba_mod_b200.")]
    assert "dba_mod_b200.data.image" in names and "dba_mod_b200.fl.simulator" in names
    for n in names:
        if n.endswith("__main__") or n.rsplit(".", 1)[-1].startswith("lib"):
            continue                     # entry scripts; the ctypes library (not a Python extension module)
        importlib.import_module(n)


def test_no_source_file_is_untracked_or_ignored():
    if _git("rev-parse", "--is-inside-work-tree").returncode != 0:
        pytest.skip("not a git checkout (e.g. the snapshot shipped to the GPU box)")
    files = _sources()
    ignored = _git("check-ignore", *files).stdout.split()
    assert ignored == [], f"source files hidden by .gitignore: {ignored}"
    tracked = set(_git("ls-files").stdout.split("\n"))
    missing = [f for f in files if f not in tracked]
    assert missing == [], f"source files not in git (git add them): {missing}"
