#!/usr/bin/env python
"""transpile_wgsl.py — a second, MECHANICAL derivation of the oracle (TEST INFRASTRUCTURE).

The reference's hot path exists only as WGSL (compute.wgsl, funcs.wgsl, sort.wgsl, shaders/jump_flood.wgsl) and cannot be
compiled or run here (Rust + wgpu + a graphics API: none on this box). oracle/fluid_oracle.cpp restates it by hand; this
script instead rewrites the shader TEXT, token by token, into C++ that g++ compiles, so that the hand-written restatement
can be checked bit for bit against code nobody re-typed (tests/test_wgsl_ref.py).

    python oracle/transpile_wgsl.py [--reference /root/reference]
        reads   <reference>/compute.wgsl (+ `#include funcs.wgsl`, spliced as src/shader.rs:16-33 does), sort.wgsl,
                shaders/jump_flood.wgsl                       — where they lie; nothing is copied into the repository
        writes  oracle/_ref/wgsl_compute.inc, wgsl_sort.inc, wgsl_jfa.inc   (generated C++; oracle/_ref/ is git-ignored)
        builds  oracle/_ref/libwgsl_ref.so from oracle/wgsl_ref_driver.cpp (the dispatch loops = what src/simulation.rs:459-539
                records) + oracle/wgsl_prelude.h (what the WGSL tokens mean in C++)

What the rewriter handles (everything the four files use, nothing more): struct declarations; `@group/@binding var<...>`
globals (uniform / storage array / rg32uint storage texture); `const NAME: T = v;`; `fn f(a: T, ...) -> R {`;
`@compute @workgroup_size(..) fn f(@builtin(global_invocation_id) id: vec3<u32>)` entry points; `let` / `var` with and
without a type; `if c {`, `while true {`, C-style `for`; the scalar types f32/u32/i32 and vec2/3/4 of them;
`ptr<function, T>` parameters with `&x` / `*p`; value constructors `T(x)` (-> wgsl::to_T, WGSL's conversion rules) and
`vecN<T>(...)` / `vecN(...)` / `array<T, N>(...)` (-> brace initialisers: C++ evaluates them left to right like WGSL
does, which matters for `vec2<f32>(rand_f32(&s), rand_f32(&s))`, compute.wgsl:212); `.xy` swizzles; abstract-float
literals (-> f32 literals); the built-ins dot length normalize sqrt pow abs sign floor max select all textureLoad
textureStore; `/* */` comments. One semantic pin is applied textually: inside an entry point, a store to
`in_particles[...]` goes to `out_particles[...]` — the driver points both at the same array for every pass except
move_particle, where the reference has a read/write race on neighbour velocity (compute.wgsl:155 vs :288) and SURVEY.md
§3.3 pins Jacobi semantics (every read sees the state before the pass).
"""
import argparse
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
BUILTINS = ["dot", "length", "normalize", "sqrt", "pow", "abs", "sign", "floor", "max", "select", "all", "textureLoad", "textureStore"]
SCALARS = {"f32": "float", "u32": "uint32_t", "i32": "int32_t", "bool": "bool"}


def strip_block_comments(src):
    return re.sub(r"/\*.*?\*/", "", src, flags=re.S)


def splice_includes(src, base_dir):
    """`#include <path>` lines, path relative to the crate root (src/shader.rs:22-26)."""
    def repl(m):
        with open(os.path.join(base_dir, m.group(1).strip())) as f:
            return splice_includes(strip_block_comments(f.read()), base_dir)
    return re.sub(r"^#include\s+(\S+)\s*$", repl, src, flags=re.M)


def map_type(t):
    t = t.strip()
    m = re.fullmatch(r"ptr<\s*function\s*,\s*(\w+)\s*>", t)
    if m:
        return map_type(m.group(1)) + "*"
    m = re.fullmatch(r"array<\s*(.+?)\s*,\s*(\d+)\s*>", t)
    if m:
        return f"std::array<{map_type(m.group(1))}, {m.group(2)}>"
    m = re.fullmatch(r"array<\s*(.+)\s*>", t)
    if m:
        return f"wgsl::Storage<{map_type(m.group(1))}>"
    m = re.fullmatch(r"(vec[234])<\s*(\w+)\s*>", t)
    if m:
        return f"{m.group(1)}<{SCALARS[m.group(2)]}>"
    if t.startswith("texture_storage_2d"):
        return "wgsl::Texture2D"
    return SCALARS.get(t, t)


def convert_constructors(src):
    """vecN<T>( ... ) / vecN( ... ) / array<T, N>( ... )  ->  brace initialisers (left-to-right evaluation)."""
    pat = re.compile(r"\b(array<\s*(?:vec[234]<\w+>|\w+)\s*,\s*\d+\s*>|vec[234](?:<\w+>)?)\(")
    pos = 0
    while True:
        m = pat.search(src, pos)
        if not m:
            return src
        open_i = m.end() - 1
        depth, j = 0, open_i
        while True:
            c = src[j]
            if c == "(":
                depth += 1
            elif c == ")":
                depth -= 1
                if depth == 0:
                    break
            j += 1
        head = m.group(1)
        if head.startswith("array<"):
            head = map_type(head)
        src = src[:m.start()] + head + "{" + src[open_i + 1:j] + "}" + src[j + 1:]
        pos = m.start() + len(head) + 1


def translate(src, namespace):
    src = strip_block_comments(src)
    protos, entry_points = [], []

    # ---- globals -------------------------------------------------------------------------------------------------
    def global_decl(m):
        name, typ = m.group(2), map_type(m.group(3))
        return f"static {typ} {name};"
    src = re.sub(r"@group\(\d+\)\s*@binding\(\d+\)\s*var(?:<([^>]*)>)?\s+(\w+)\s*:\s*([^;]+);", global_decl, src)
    src = re.sub(r"^const\s+(\w+)\s*:\s*(\w+)\s*=\s*([^;]+);", lambda m: f"static const {map_type(m.group(2))} {m.group(1)} = {m.group(3)};", src, flags=re.M)

    # ---- structs ---------------------------------------------------------------------------------------------------
    def struct_decl(m):
        body = re.sub(r"(\w+)\s*:\s*([^,\n]+),", lambda f: f"{map_type(f.group(2))} {f.group(1)};", m.group(2))
        return f"struct {m.group(1)} {{{body}}};"
    src = re.sub(r"\bstruct\s+(\w+)\s*\{(.*?)\}", struct_decl, src, flags=re.S)

    # ---- functions ---------------------------------------------------------------------------------------------------
    def params(p):
        out = []
        for part in [x for x in re.split(r",(?![^<]*>)", p) if x.strip()]:
            name, typ = part.split(":", 1)
            out.append(f"{map_type(typ)} {name.strip()}")
        return ", ".join(out)

    def entry(m):
        entry_points.append((m.group(2), m.group(1)))
        sig = f"static void {m.group(2)}(vec3<uint32_t> {m.group(3)})"
        protos.append(sig + ";")
        return sig + " { /*ENTRY*/"
    src = re.sub(r"@compute\s+@workgroup_size\(([^)]*)\)\s*fn\s+(\w+)\s*\(\s*@builtin\(global_invocation_id\)\s*(\w+)\s*:\s*vec3<u32>\s*,?\s*\)\s*\{",
                 entry, src, flags=re.S)

    def func(m):
        ret = map_type(m.group(3)) if m.group(3) else "void"
        sig = f"static {ret} {m.group(1)}({params(m.group(2))})"
        protos.append(sig + ";")
        return sig + " {"
    src = re.sub(r"\bfn\s+(\w+)\s*\(([^)]*)\)\s*(?:->\s*([^{]+?))?\s*\{", func, src)

    # the Jacobi pin: stores to in_particles[...] at statement start inside entry points go to out_particles[...]
    lines, in_entry, depth = [], False, 0
    for line in src.split("\n"):
        if "/*ENTRY*/" in line:
            in_entry, depth = True, 0
        if in_entry:
            if re.match(r"\s*in_particles\[[^\]]*\](\.\w+)?\s*=[^=]", line):
                line = line.replace("in_particles[", "out_particles[", 1)
            depth += line.count("{") - line.count("}")
            if depth <= 0 and "/*ENTRY*/" not in line:
                in_entry = False
        lines.append(line.replace(" /*ENTRY*/", ""))
    src = "\n".join(lines)

    # ---- statements -----------------------------------------------------------------------------------------------
    src = re.sub(r"\bvar\s+(\w+)\s*:\s*([\w<>, ]+?)\s*=", lambda m: f"{map_type(m.group(2))} {m.group(1)} =", src)
    src = re.sub(r"\blet\s+", "const auto ", src)
    src = re.sub(r"\bvar\s+", "auto ", src)
    src = re.sub(r"\bwhile\s+true\s*\{", "while (true) {", src)
    src = re.sub(r"\bif\s+([^{;]+?)\s*\{", r"if (\1) {", src)
    src = convert_constructors(src)
    src = re.sub(r"\.xy\b", ".xy()", src)
    for s_, fn in (("f32", "to_f32"), ("u32", "to_u32"), ("i32", "to_i32")):
        src = re.sub(rf"(?<![\w<]){s_}\(", f"wgsl::{fn}(", src)
    for s_, c in SCALARS.items():
        src = re.sub(rf"\b{s_}\b", c, src)
    for b in BUILTINS:
        src = re.sub(rf"(?<![\w.:]){b}\(", f"wgsl::{b}(", src)
    # abstract-float literals -> f32
    src = re.sub(r"(?<![\w.])(\d+\.\d*(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])", r"\1f", src)

    # forward declarations go after the last struct / global (all of them precede the first function in these files)
    first_fn = min((src.find(p[:-1]) for p in protos if src.find(p[:-1]) >= 0), default=0)
    extra = "static wgsl::Storage<ParticleInstance> out_particles;\n" if "in_particles" in src else ""
    src = src[:first_fn] + extra + "\n".join(protos) + "\n" + src[first_fn:]
    head = (f"// GENERATED by oracle/transpile_wgsl.py from the reference's WGSL — never committed (oracle/_ref/ is git-ignored)\n"
            f"namespace {namespace} {{\nusing wgsl::vec2; using wgsl::vec3; using wgsl::vec4;\n")
    return head + src + f"\n}}  // namespace {namespace}\n", entry_points


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--no-build", action="store_true")
    a = ap.parse_args()
    ref = a.reference
    os.makedirs(OUT, exist_ok=True)
    jobs = [("compute.wgsl", "wgsl_compute"), ("sort.wgsl", "wgsl_sort"), (os.path.join("shaders", "jump_flood.wgsl"), "wgsl_jfa")]
    for rel, ns in jobs:
        with open(os.path.join(ref, rel)) as f:
            src = splice_includes(strip_block_comments(f.read()), ref)
        out, eps = translate(src, ns)
        with open(os.path.join(OUT, ns + ".inc"), "w") as f:
            f.write(out)
        print(f"{rel}: {len(src.splitlines())} lines of WGSL -> {ns}.inc, entry points {[e for e, _ in eps]}")
    if a.no_build:
        return
    cmd = ["/usr/bin/g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-std=c++17", "-shared",
           "-I", HERE, "-I", OUT, "-o", os.path.join(OUT, "libwgsl_ref.so"), os.path.join(HERE, "wgsl_ref_driver.cpp")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise SystemExit("g++ failed on the generated code")
    print("built", os.path.join(OUT, "libwgsl_ref.so"))


if __name__ == "__main__":
    main()
