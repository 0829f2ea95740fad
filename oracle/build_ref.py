#!/usr/bin/env python3
"""Build oracle/_ref/libref_shader.so: the reference's gpu_shader.comp compiled as C++.

TEST INFRASTRUCTURE.  Reads the shader where it lies under the reference tree
(default /root/reference, read-only), applies a mechanical text transform and
writes ONLY into oracle/_ref/ (git-ignored; the built .so travels to the GPU
box with the snapshot).  Nothing is copied into tracked files.

Transform:
  * drop `#version` and the `layout(...)` interface block (-> REF_BUFFERS, which
    the glue defines as plain statics of the same names);
  * suffix every floating literal with `f` (GLSL literals are fp32);
  * `out float` parameters -> `float&`;
  * `value.xyz = accumulatedColor;` -> vec4 constructor (no swizzle stores in GLM);
  * `ivec2(gl_GlobalInvocationID.xy)` -> a thread-local ivec2.
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def transform(src: str) -> str:
    lines = src.split("\n")
    out = []
    in_layout = False
    emitted = False
    for ln in lines:
        s = ln.strip()
        if s.startswith("#version"):
            continue
        if s.startswith("layout("):
            in_layout = not s.endswith(";")
            if not emitted:
                out.append("REF_BUFFERS")
                emitted = True
            continue
        if in_layout:
            if s.startswith("};") or s == "}":
                in_layout = False
            continue
        out.append(ln)
    body = "\n".join(out)
    lit = re.compile(r"(?<![A-Za-z0-9_.])((?:[0-9]+\.[0-9]*|\.[0-9]+)(?:[eE][-+]?[0-9]+)?|[0-9]+[eE][-+]?[0-9]+)(?![0-9A-Za-z_.])")
    body = lit.sub(lambda m: m.group(1) + "f", body)
    body = body.replace("out float", "float&")
    body = body.replace("value.xyz = accumulatedColor;", "value = vec4(accumulatedColor, value.w);")
    body = body.replace("ivec2(gl_GlobalInvocationID.xy)", "gl_GlobalInvocationID_xy")
    return body


def cut(src: str, head_re: str) -> str:
    """The definition that starts at the line matching head_re, through its matching closing brace (plus a trailing
    declarator such as `} wheels[4];`).  Text is taken verbatim."""
    for m in re.finditer(head_re, src, re.M):
        i = src.index("{", m.start())
        if ";" not in src[m.start():i]:          # a prototype (ends in ';' before any brace) is not the definition
            break
    else:
        raise SystemExit("build_ref: %r not found in main.cpp" % head_re)
    depth, j = 0, i
    while True:
        c = src[j]
        if c == "{":
            depth += 1
        elif c == "}":
            depth -= 1
            if depth == 0:
                break
        j += 1
    k = src.index("\n", j)
    return src[m.start():k] + "\n"


def extract_main(ref: str, outdir: str) -> None:
    """src/main.cpp -> oracle/_ref/ref_main_{node,wheel,funcs}.inc (verbatim definitions, see ref_bvh_glue.cpp)."""
    with open(os.path.join(ref, "src", "main.cpp"), "r", encoding="utf-8", errors="replace") as f:
        src = f.read()
    with open(os.path.join(outdir, "ref_main_node.inc"), "w") as f:
        f.write(cut(src, r"^class Node\b"))
    with open(os.path.join(outdir, "ref_main_wheel.inc"), "w") as f:
        f.write(cut(src, r"^struct Wheel\b"))
    heads = [r"^void serializeBVH\(std::vector<FlatNode>& nodes, std::vector<int>& indices\) \{", r"^void updateScene\(FlatScene& flatScene, GLuint ssbo\)\s*$",
             r"^FlatShape serializeShape\(const std::unique_ptr<Shape>& shape\)\s*$", r"^void updateBVH\(\) \{",
             r"^void bounceSphere\(", r"^void updateWheelAnimations\(float elapsedTime\) \{",
             r"^void split\(std::unique_ptr<Node>& parentNode, int depth\) \{", r"^int buildBVH\(int maxDepth\) \{"]
    with open(os.path.join(outdir, "ref_main_funcs.inc"), "w") as f:
        for h in heads:
            f.write(cut(src, h) + "\n")


def main() -> int:
    ref = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("RT_REFERENCE_DIR", "/root/reference")
    shader = os.path.join(ref, "src", "shaders", "gpu_shader.comp")
    glm = os.path.join(ref, "glm")
    if not os.path.isfile(shader) or not os.path.isdir(glm):
        print(f"build_ref: reference not present at {ref}; keeping any prebuilt oracle/_ref", file=sys.stderr)
        return 0
    outdir = os.path.join(HERE, "_ref")
    os.makedirs(outdir, exist_ok=True)
    with open(shader, "r", encoding="utf-8", errors="replace") as f:
        body = transform(f.read())
    with open(os.path.join(outdir, "gpu_shader_body.inc"), "w") as f:
        f.write(body)
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-w", "-fPIC", "-shared", "-pthread",
           "-I", glm, "-I", HERE, os.path.join(HERE, "ref_shader_glue.cpp"),
           "-o", os.path.join(outdir, "libref_shader.so")]
    print(" ".join(cmd))
    rc = subprocess.call(cmd)
    if rc != 0:
        return rc
    # the same shader text with min / max / clamp / mix as the GLSL specification words them (NaN operands: see the glue)
    cmd = cmd[:-2] + ["-DREF_GLSL_BUILTINS=1", "-o", os.path.join(outdir, "libref_shader_glsl.so")]
    print(" ".join(cmd))
    rc = subprocess.call(cmd)
    if rc != 0:
        return rc
    # the reference's own C++ shape classes, for the primitive-level cross-check
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-w", "-fPIC", "-shared",
           "-I", os.path.join(ref, "src"), "-I", ref, "-I", glm, "-I", os.path.join(HERE, "stubs"),
           os.path.join(HERE, "ref_shapes_glue.cpp"), "-o", os.path.join(outdir, "libref_shapes.so")]
    print(" ".join(cmd))
    rc = subprocess.call(cmd)
    if rc != 0:
        return rc
    # the reference's wheel-animation expression on its vendored GLM (pins host/rt_anim_math.h)
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-w", "-fPIC", "-shared", "-I", glm,
           os.path.join(HERE, "ref_anim_glue.cpp"), "-o", os.path.join(outdir, "libref_anim.so")]
    print(" ".join(cmd))
    rc = subprocess.call(cmd)
    if rc != 0:
        return rc
    # the reference's own Camera class + serializeCamera's field copies (pins the headless host's rh_camera_*); the one OpenGL
    # type camera.hpp names (GLboolean) comes from oracle/stubs/glad/glad.h
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-w", "-fPIC", "-shared",
           "-I", os.path.join(ref, "src"), "-I", glm, "-I", os.path.join(HERE, "stubs"),
           os.path.join(HERE, "ref_camera_glue.cpp"), "-o", os.path.join(outdir, "libref_camera.so")]
    print(" ".join(cmd))
    rc = subprocess.call(cmd)
    if rc != 0:
        return rc
    # the reference's own builder / serialiser / animation step cut out of main.cpp (pins host builder, K6, K7, K8)
    extract_main(ref, outdir)
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-w", "-fPIC", "-shared",
           "-I", os.path.join(ref, "src"), "-I", ref, "-I", glm, "-I", os.path.join(HERE, "stubs"), "-I", HERE,
           os.path.join(HERE, "ref_bvh_glue.cpp"), "-o", os.path.join(outdir, "libref_bvh.so")]
    print(" ".join(cmd))
    return subprocess.call(cmd)


if __name__ == "__main__":
    sys.exit(main())
