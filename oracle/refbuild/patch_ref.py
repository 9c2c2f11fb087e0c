#!/usr/bin/env python3
"""Patch a THROW-AWAY COPY of the reference sources so g++ 13 accepts them.

TEST INFRASTRUCTURE ONLY (see oracle/README.md).  The reference
(Thomi6York/pbrt-v2-skin) is maintained for MSVC 2012 only (README.md:42-43);
g++ rejects about twenty MSVC-isms.  None of the edits below touches arithmetic
on the subsurface hot path -- they are spelling fixes (missing `template<>`,
`__declspec`, `_strtoui64`, header case, PCH reliance).  The list follows
SURVEY.md section 8(c).

Usage: patch_ref.py <copy-of-reference-src-dir>
The copy lives under a mktemp dir and is deleted by build_ref.sh after the
build; only objects/binaries land in oracle/_ref/.
"""
import re
import sys
from pathlib import Path


def sub(path: Path, old: str, new: str, count: int = 0, must: bool = True):
    s = path.read_text(errors="surrogateescape")
    if old not in s:
        if must:
            raise SystemExit(f"patch_ref: pattern not found in {path}: {old!r}")
        return
    s = s.replace(old, new, count) if count else s.replace(old, new)
    path.write_text(s, errors="surrogateescape")


def main(src: Path):
    core = src / "core"
    # (2) the PCH header is what makes most fork headers self-sufficient
    sub(core / "stdafx.h", "#if defined(_MSC_VER)", "#if 1")
    (core / "StdAfx.h").write_text('#include "stdafx.h"\n')
    # (3) dependent-name lookups in RayDifferentialBase
    g = core / "geometry.h"
    s = g.read_text()
    m = re.search(r"class RayDifferentialBase\s*:\s*public RayBase<scalar>\s*\{\s*public:", s)
    if not m:
        raise SystemExit("patch_ref: RayDifferentialBase not found")
    s = s[:m.end()] + ("\n    typedef ScalarTraits<scalar> Traits;\n"
                       "    using RayBase<scalar>::o; using RayBase<scalar>::d;\n") + s[m.end():]
    g.write_text(s)
    # (4) explicit specialisations need template<>
    gc = core / "geometry.cpp"
    sub(gc, "void BBox::BoundingSphere", "template<> void BBox::BoundingSphere")
    sub(gc, "bool BBox::IntersectP", "template<> bool BBox::IntersectP")
    # (5) memory.cpp
    mc = core / "memory.cpp"
    s = mc.read_text()
    s = s.replace("__declspec(thread)", "__thread")
    s = s.replace("auto& iter = mapSelfishAllocators.find", "auto iter = mapSelfishAllocators.find")
    mc.write_text(s)
    # (6) parallel.cpp: sys/sysctl.h is gone from glibc
    sub(core / "parallel.cpp", "#include <sys/sysctl.h>", "", must=False)
    # (7) KahanSum copy-initialisation from double
    sub(core / "reflection.cpp", "KahanSum<Spectrum> r = 0.;", "KahanSum<Spectrum> r(Spectrum(0.f));")
    # (8) _strtoui64
    for f in ("materials/layeredskin.cpp", "renderers/mcprofile.cpp", "renderers/batchmcprofile.cpp"):
        sub(src / f, "_strtoui64", "strtoull", must=False)
    # (9) ARRAYSIZE
    sk = src / "materials/skincoeffs.cpp"
    s = sk.read_text()
    if "define ARRAYSIZE" not in s:
        idx = s.find("#include")
        s = s[:idx] + "#ifndef ARRAYSIZE\n#define ARRAYSIZE(a) (sizeof(a)/sizeof((a)[0]))\n#endif\n" + s[idx:]
    sk.write_text(s)
    # (10) typeinfo / bad_cast
    for f in ("renderers/surfacepoints.cpp", "core/primitive.cpp"):
        p = src / f
        s = p.read_text()
        idx = s.find("#include")
        s = s[:idx] + "#include <typeinfo>\n" + s[idx:]
        s = s.replace("catch (std::bad_cast())", "catch (std::bad_cast&)")
        s = s.replace("catch (std::bad_cast)", "catch (std::bad_cast&)")
        p.write_text(s)
    # (11) wide string literal
    li = src / "integrators/layeredintegrator.cpp"
    s = li.read_text()
    s = re.sub(r'\bL"', '"', s)
    li.write_text(s)
    # (12) multipole DLL: portable stdafx + temporaries bound to non-const refs
    mp = src / "multipole/MultipoleProfileCalculator"
    (mp / "stdafx.h").write_text(
        "#pragma once\n#include <cstring>\n#include <cstdlib>\n#include <cmath>\n#include <math.h>\n"
        "#include <cassert>\n#include <algorithm>\n#include <utility>\n#include <functional>\n"
        "using std::min; using std::max; using std::make_pair;\n")
    m = mp / "MultipoleProfileCalculator.cpp"
    s = m.read_text()
    for op in ("+=", "-=", "*=", "/="):
        s = s.replace(f"return kiss_fft_cpx(cpx) {op} other;",
                      f"kiss_fft_cpx tmp_ = cpx; tmp_ {op} other; return tmp_;")
    m.write_text(s)
    mt = mp / "mpc-types.h"
    s = mt.read_text()
    s = s.replace("__declspec(dllexport)", "").replace("__declspec(dllimport)", "")
    mt.write_text(s)
    print("patch_ref: ok")


if __name__ == "__main__":
    main(Path(sys.argv[1]))
