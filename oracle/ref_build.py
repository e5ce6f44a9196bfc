"""TEST INFRASTRUCTURE — not product code.

Builds oracle/_ref/libgsref.so: the reference's OWN shader bodies
(/root/reference/shaders/gau_vert.glsl, gau_frag.glsl), compiled as C++ against oracle/glsl_shim.hpp
and driven by oracle/ref_driver.inc.  The shader text is read from /root/reference at build time,
passed to g++ on stdin and never written into this repository; only the .so lands in the
git-ignored oracle/_ref/.

The only textual edits made to the shader sources (none touches an arithmetic expression):
  * the `#version` line is dropped;
  * `layout(...)` qualifiers are dropped; the two std430 `buffer` blocks become plain pointers
    (`const float* g_data`, `const int* gi`);
  * `uniform` -> nothing (a namespace-scope variable), `in`/`out` -> `thread_local`;
  * `void main()` -> `void shader_main()`;
  * `discard;` -> `{ gl_discarded = true; return; }`.

Usage:  python oracle/ref_build.py [--reference /root/reference]
Returns 0 and prints the .so path; returns 3 when the reference tree is not present (GPU box).
"""
from __future__ import annotations

import argparse
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "libgsref.so")


def _strip_common(src: str) -> str:
    src = re.sub(r"^\s*#version[^\n]*\n", "\n", src, flags=re.M)
    # std430 buffer blocks -> removed (pointers are declared by the wrapper)
    src = re.sub(r"layout\s*\(\s*std430[^)]*\)\s*buffer\s+\w+\s*\{[^}]*\}\s*;", "", src, flags=re.S)
    src = re.sub(r"layout\s*\([^)]*\)\s*", "", src)
    src = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", src)
    src = re.sub(r"\bdiscard\s*;", "{ gl_discarded = true; return; }", src)
    return src


def make_translation_unit(ref_root: str) -> str:
    with open(os.path.join(ref_root, "shaders", "gau_vert.glsl")) as f:
        vert = _strip_common(f.read())
    with open(os.path.join(ref_root, "shaders", "gau_frag.glsl")) as f:
        frag = _strip_common(f.read())
    parts = [
        '#include "glsl_shim.hpp"\n',
        "#define uniform\n#define in thread_local\n#define out thread_local\n",
        "namespace glsl { namespace gau_vert {\n",
        "const float* g_data; const int* gi;\n",
        "thread_local int gl_InstanceID; thread_local vec4 gl_Position;\n",
        '#line 1 "gau_vert.glsl"\n',
        vert,
        "\n} }\n",
        "namespace glsl { namespace gau_frag {\n",
        "thread_local bool gl_discarded;\n",
        '#line 1 "gau_frag.glsl"\n',
        frag,
        "\n} }\n",
        "#undef uniform\n#undef in\n#undef out\n",
        '#include "ref_driver.inc"\n',
    ]
    return "".join(parts)


def build(ref_root: str = "/root/reference", verbose: bool = False) -> str | None:
    if not os.path.isfile(os.path.join(ref_root, "shaders", "gau_vert.glsl")):
        return None
    os.makedirs(OUT_DIR, exist_ok=True)
    tu = make_translation_unit(ref_root)
    cmd = [
        "g++", "-x", "c++", "-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp",
        "-fPIC", "-shared", "-I", HERE, "-o", OUT_SO, "-",
    ]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    res = subprocess.run(cmd, input=tu.encode(), capture_output=True)
    if res.returncode != 0:
        sys.stderr.write(res.stderr.decode())
        raise RuntimeError("oracle/_ref build failed")
    return OUT_SO


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    so = build(a.reference, a.verbose)
    if so is None:
        print("reference tree not present; oracle/_ref not rebuilt", file=sys.stderr)
        sys.exit(3)
    print(so)
