#!/usr/bin/env python3
"""Build the UNMODIFIED reference (tenpy/cyten) from the sources where they lie.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package ``cyten_b200`` may import
``oracle/``; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs use what this script produces.

What it does
------------
* compiles every translation unit listed in ``/root/reference/src/CMakeLists.txt`` and
  ``/root/reference/pybind/CMakeLists.txt`` with ``/usr/bin/g++ -std=c++23`` directly (a ninja
  file written by this script; the reference's own cmake build system is NOT run),
* links ``oracle/_ref/cyten/_core.cpython-*.so`` against the venv's libtorch (same image on the
  GPU box, so the result travels with the gpurun snapshot),
* stages the thin ``cyten/*.py`` shim package next to it (with the three Python-3.14-only syntax
  spots adapted for the 3.12 interpreter of this image),
* outputs ONLY into ``oracle/_ref/`` (git-ignored; never committed).

Usage:  python oracle/build_ref.py [-j N]
"""
from __future__ import annotations

import argparse
import os
import re
import shutil
import subprocess
import sys
import sysconfig
from pathlib import Path

REF = Path(os.environ.get("CYTEN_REFERENCE", "/root/reference"))
HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
OBJ = OUT / "_obj"


def _sources(cmakelists: Path, base: Path) -> list[Path]:
    txt = cmakelists.read_text()
    return [base / m for m in re.findall(r"^\s+([\w/]+\.cpp)\s*$", txt, flags=re.M)]


def _stage_python_package() -> None:
    dst = OUT / "cyten"
    so_keep = list(dst.glob("_core*.so")) if dst.exists() else []
    tmp = OUT / "_so_keep"
    tmp.mkdir(parents=True, exist_ok=True)
    for so in so_keep:
        shutil.move(str(so), tmp / so.name)
    if dst.exists():
        shutil.rmtree(dst)
    shutil.copytree(REF / "cyten", dst)
    for so in tmp.glob("_core*.so"):
        shutil.move(str(so), dst / so.name)
    shutil.rmtree(tmp)
    # python-3.14-only syntax -> 3.12 (copies only)
    pat = re.compile(r"^(\s*except) ([A-Za-z_][\w\.]*(?:, [A-Za-z_][\w\.]*)+):", flags=re.M)
    for py in dst.rglob("*.py"):
        s = py.read_text()
        s2 = pat.sub(lambda m: f"{m.group(1)} ({m.group(2)}):", s)
        if py.name == "sparse.py" and "from __future__ import annotations" not in s2:
            # must come after the module docstring
            m = re.match(r'\s*(""".*?"""|\'\'\'.*?\'\'\')\s*', s2, flags=re.S)
            k = m.end() if m else 0
            s2 = s2[:k] + "\nfrom __future__ import annotations\n" + s2[k:]
        if s2 != s:
            py.write_text(s2)
    (dst / "_version.py").write_text(
        "__version__ = '0.0.1+ref'\n__version_tuple__ = (0, 0, 1)\n__commit_id__ = 'reference'\n"
        "version = __version__\nversion_tuple = __version_tuple__\ncommit_id = __commit_id__\n"
    )
    # reference helpers needed to drive its own tests / toy codes (the GPU box has no /root/reference: the reference's
    # pytest suite travels with the built artefact so that it can be run THROUGH the plugin there, see
    # integration/pytest_cyten_b200.py and tests/test_plugin.py)
    for name in ("toycodes", "tests/python_tests"):
        if (REF / name).exists():
            if (OUT / name).exists():
                shutil.rmtree(OUT / name)
            (OUT / name).parent.mkdir(parents=True, exist_ok=True)
            shutil.copytree(REF / name, OUT / name)
    for name in ("conftest.py", "pyproject.toml"):
        if (REF / name).exists():
            shutil.copy(REF / name, OUT / name)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("-j", type=int, default=os.cpu_count() or 8)
    ap.add_argument("--stage-only", action="store_true")
    args = ap.parse_args()
    if not REF.exists():
        print(f"reference not present at {REF}; keeping prebuilt oracle/_ref as is")
        return 0
    OUT.mkdir(exist_ok=True)
    OBJ.mkdir(exist_ok=True)
    _stage_python_package()
    if args.stage_only:
        return 0

    import pybind11
    import torch
    from torch.utils import cpp_extension as ce

    srcs = _sources(REF / "src/CMakeLists.txt", REF / "src") + _sources(
        REF / "pybind/CMakeLists.txt", REF / "pybind"
    )
    missing = [s for s in srcs if not s.exists()]
    if missing:
        print("missing sources:", missing)
        return 1
    incs = [REF / "include", REF / "src", REF / "pybind", pybind11.get_include(),
            sysconfig.get_paths()["include"], *ce.include_paths()]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cflags = (
        "-std=c++23 -O2 -g0 -fPIC -fvisibility=hidden -fvisibility-inlines-hidden -DNDEBUG "
        f"-D_GLIBCXX_USE_CXX11_ABI={abi} -Wno-deprecated-declarations "
        + " ".join(f"-I{p}" for p in incs)
    )
    torch_lib = ce.library_paths()[0]
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    target = OUT / "cyten" / f"_core{ext}"
    ldflags = (
        f"-shared -L{torch_lib} -Wl,-rpath,{torch_lib} -ltorch -ltorch_cpu -lc10 "
        "-Wl,--no-as-needed"
    )
    lines = [
        "cxx = /usr/bin/g++",
        f"cflags = {cflags}",
        f"ldflags = {ldflags}",
        "rule cc",
        "  command = $cxx $cflags -MMD -MF $out.d -c $in -o $out",
        "  depfile = $out.d",
        "  deps = gcc",
        "  description = CXX $in",
        "rule link",
        "  command = $cxx $in $ldflags -o $out",
        "  description = LINK $out",
    ]
    objs = []
    for s in srcs:
        rel = s.relative_to(REF)
        o = OBJ / (str(rel).replace("/", "__") + ".o")
        objs.append(o)
        lines.append(f"build {o}: cc {s}")
    lines.append(f"build {target}: link {' '.join(map(str, objs))}")
    lines.append(f"default {target}")
    (OBJ / "build.ninja").write_text("\n".join(lines) + "\n")
    r = subprocess.run(["ninja", "-C", str(OBJ), f"-j{args.j}"])
    return r.returncode


if __name__ == "__main__":
    sys.exit(main())
