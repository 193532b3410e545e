#!/usr/bin/env python3
"""glsl2cpp.py — makes the reference's OWN shader sources compile as C++ (test infrastructure).

Usage: glsl2cpp.py <reference shader dir> <output dir>

Reads pt_common.glsl, raygen.rgen, rayhit.rchit, miss.rmiss and shadowmiss.rmiss WHERE THEY LIE under
/root/reference/src/GPUSpectral/assets/shaders and writes <name>.cpp files into the (temporary) output directory;
oracle/ref/ref_glue.cpp includes each of them inside its own namespace against the reference's external/glm.
Nothing of the reference is copied into the repository: the output directory is a build scratch directory and
only the linked oracle/_ref/libgsref.so (git-ignored) is kept.

The rewrite is purely syntactic — every arithmetic expression, constant and statement of the shaders is compiled
as written.  Each rule and why it does not change the meaning:

 1. `#version`, `#pragma shader_stage`, `#extension` lines are dropped (no C++ meaning).
 2. `#include "pt_common.glsl"` -> `#include "pt_common.glsl.cpp"` (the rewritten copy); `#include "BSDF.inc"` is
    left to the C preprocessor, which reads BSDF.inc straight from the reference directory (-I).
 3. floating literals without a suffix get an `f`: GLSL literals are 32-bit floats, C++ ones would be doubles.
 4. `out T x` / `inout T x` parameters -> `T& x`.
 5. swizzles `.xyz` / `.xy` -> glm's swizzle functions `.xyz()` / `.xy()` (GLM_FORCE_SWIZZLE).
 6. `layout(buffer_reference, ...) readonly buffer NAME { T field[]; };` -> `REF_BUFFER(NAME, T, field)`: a struct with a
    `const T* field` constructed from the uvec2 device address, which is what a buffer reference is.
 7. interface variables (`rayPayloadEXT`, `rayPayloadInEXT`, `hitAttributeEXT`, `uniform ...`) are dropped; the glue
    provides `prd`, `shadowed`, `attribs`, `topLevelAS`, `image`, `renderState` and the `gl_*` built-ins.
 8. `void main()` -> `void shader_main()`.
 9. `uint rngState;` -> `thread_local uint rngState;` (one shader invocation per host thread).
10. `vec2(randUniform(), randUniform())` -> `vec2{randUniform(), randUniform()}`: GLSL evaluates constructor arguments
    left to right (GLSL 4.60 spec, section 6.1.1 / 5.9 "order of evaluation"), C++ only guarantees that order for
    braced initialisers.  The script refuses to run if any OTHER expression holds two RNG calls.
11. raygen.rgen's `#define MAX_DEPTH 50` -> `#define REF_MAX_DEPTH_LITERAL 50` + `#define MAX_DEPTH (gsref::g_maxDepth)`: the
    bound becomes a run-time integer that STARTS OUT as the shader's literal (ref_glue.cpp: g_maxDepth), so that
    BASELINE.json's "max depth 8" configurations can be run through the shader text (ref_set_max_depth); every pinning
    test runs with the literal.
"""
import os
import re
import sys

FILES = ["pt_common.glsl", "raygen.rgen", "rayhit.rchit", "miss.rmiss", "shadowmiss.rmiss"]

FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")
BUFFER_REF = re.compile(
    r"layout\s*\(\s*buffer_reference[^)]*\)\s*readonly\s+buffer\s+([\w#]+)\s*\\?\s*\{\s*\\?\s*([\w#]+)\s+(\w+)\[\];\s*\\?\s*\};")
UNIFORM_BLOCK = re.compile(r"layout\s*\([^)]*\)\s*uniform\s+_RenderState\s*\{[^}]*\};")
INTERFACE_VAR = re.compile(
    r"^\s*(?:layout\s*\([^)]*\)\s*)?(?:rayPayloadInEXT|rayPayloadEXT|hitAttributeEXT|uniform)\s+\w+\s+\w+\s*;\s*$", re.M)


def strip_comments_for_check(text):
    text = re.sub(r"//[^\n]*", "", text)
    return re.sub(r"/\*.*?\*/", "", text, flags=re.S)


def convert(name, text):
    out = []
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("#version") or s.startswith("#pragma") or s.startswith("#extension"):
            continue
        out.append(line)
    text = "\n".join(out) + "\n"
    text = text.replace('#include "pt_common.glsl"', '#include "pt_common.glsl.cpp"')
    # rule 10 first (and its guard): no other statement may hold two RNG calls
    text = text.replace("vec2(randUniform(), randUniform())", "vec2{randUniform(), randUniform()}")
    for stmt in strip_comments_for_check(text).split(";"):
        if len(re.findall(r"(?<!float )(?<!uint )\brand(?:Uniform|Pcg)\s*\(\s*\)", stmt)) > 1 and "vec2{randUniform(), randUniform()}" not in stmt:
            raise SystemExit(f"{name}: an expression with two RNG calls needs an evaluation-order rule: {stmt.strip()[:120]}")
    text, n_depth = re.subn(r"^#define\s+MAX_DEPTH\s+(\d+)\s*$", r"#define REF_MAX_DEPTH_LITERAL \1\n#define MAX_DEPTH (gsref::g_maxDepth)", text, flags=re.M)
    if name == "raygen.rgen" and n_depth != 1:
        raise SystemExit("raygen.rgen: expected exactly one '#define MAX_DEPTH <int>'")
    text = BUFFER_REF.sub(lambda m: f"REF_BUFFER({m.group(1)}, {m.group(2)}, {m.group(3)})", text)
    text = UNIFORM_BLOCK.sub("", text)
    text = INTERFACE_VAR.sub("", text)
    text = FLOAT_LIT.sub(lambda m: m.group(1) + "f", text)
    text = re.sub(r"\b(?:in)?out\s+(\w+)\s+(\w+)", r"\1& \2", text)
    text = re.sub(r"\.(xyz|xy)\b(?!\s*\()", r".\1()", text)
    text = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", text)
    text = re.sub(r"^uint rngState;", "thread_local uint rngState;", text, flags=re.M)
    for leftover in ("layout(", "rayPayload", "hitAttributeEXT", "inout "):
        if leftover in strip_comments_for_check(text):
            raise SystemExit(f"{name}: unconverted GLSL construct '{leftover}'")
    return text


def main():
    src, dst = sys.argv[1], sys.argv[2]
    os.makedirs(dst, exist_ok=True)
    for f in FILES:
        with open(os.path.join(src, f)) as fh:
            text = fh.read()
        with open(os.path.join(dst, f + ".cpp"), "w") as fh:
            fh.write(convert(f, text))


if __name__ == "__main__":
    main()
