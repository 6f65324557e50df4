#!/usr/bin/env python
"""make_ref.py — TEST INFRASTRUCTURE. Builds oracle/_ref/ from the reference's own shader text.

    python oracle/make_ref.py [--wgsl /root/reference/shaders/ray_tracer.wgsl] [--out oracle/_ref]

The reference's path exists only as WGSL (shaders/ray_tracer.wgsl) and nothing in this image can
run WGSL.  This script is the mechanical stand-in for a WGSL compiler: it READS the shader from
/root/reference at build time (the text is never copied into the repository; oracle/_ref/ is
git-ignored), rewrites it token by token into C++ against oracle/wgsl_builtins.h, appends the
generic dispatch driver oracle/wgsl_ref_driver.inc and compiles

    oracle/_ref/librt_wgsl_ref.so        -DWGSL_CONTRACT  (the arithmetic contract of DESIGN.md)
    oracle/_ref/librt_wgsl_ref_plain.so  WGSL-spec formulas, nothing fused, glibc log/cos/pow

Both export wgslref_dispatch() with the argument list of the hand-written oracle's rto_dispatch().
The hand oracle (oracle/rt_oracle.cpp) and the CUDA path are tested against these libraries, so
loop bounds, RNG draw order, branch structure, the closest-hit tie rule and the record layouts
are taken from the reference's text, not from anybody's reading of it.

The rewrite knows WGSL syntax, not this shader:
  * module scope: `struct`, `const`, `@group/@binding var<...>`, `fn` (attributes skipped);
    declarations are re-ordered (structs by dependency, then constants, resources, prototypes,
    bodies) because WGSL allows use before declaration and C++ does not;
  * types: f32 u32 i32 bool vecN<T> array<T> ptr<function,T> texture_storage_2d<rgba32float,..>;
  * every struct also gets wgsl::layout<S> with offsets from the WGSL alignment rules, so
    buffers are decoded from the caller's raw bytes as the shader would read them;
  * statements: `let` -> const auto, `var x: T;` -> zero-initialised, `var x: T = e` / `var x = e`,
    `if cond {` -> `if (cond) {`; abstract-float literals get an `f` suffix (they are f32 wherever
    this shader uses them); value conversions u32()/i32()/f32() -> wgsl::to_*() (saturating);
  * a local that shadows a parameter (legal WGSL, an error in C++) renames the parameter;
  * WGSL evaluates arguments left to right, C++ in unspecified order: a vecN constructor whose arguments hold
    several side-effecting calls (functions with a `ptr` parameter) is written with braces, which C++
    sequences left to right; any other statement with two such calls is refused;
  * identifiers that are reserved in C++ (`main`) get a trailing underscore;
  * every function is wrapped with entry / return hooks (WGSL_HOOK_ENTER / WGSL_HOOK_RETURN) that
    the driver uses for counters and the primary-hit AOV; they do not touch the computation.
"""
import argparse
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_WGSL = "/root/reference/shaders/ray_tracer.wgsl"

TOKEN_RE = re.compile(r"""
    (?P<ws>\s+)
  | (?P<comment>//[^\n]*|/\*.*?\*/)
  | (?P<num>0[xX][0-9a-fA-F]*\.?[0-9a-fA-F]*(?:[pP][+-]?[0-9]+)?[fhiu]?
          | (?:[0-9]+\.[0-9]*|\.[0-9]+|[0-9]+)(?:[eE][+-]?[0-9]+)?[fhiu]?)
  | (?P<ident>[A-Za-z_][A-Za-z0-9_]*)
  | (?P<op>->|<=|>=|==|!=|&&|\|\||\+=|-=|\*=|/=|%=|<<|>>|[-+*/%<>=!&|^~.,;:(){}\[\]@])
""", re.X | re.S)

CPP_RESERVED = {"main", "class", "new", "delete", "template", "typename", "namespace", "this", "operator", "union",
                "public", "private", "protected", "virtual", "friend", "inline", "register", "signed", "unsigned",
                "short", "long", "int", "float", "double", "char", "void", "auto", "goto", "try", "catch", "throw",
                "and", "or", "not", "xor", "near", "far"}
SCALARS = {"f32", "u32", "i32", "bool"}
SCALAR_LAYOUT = {"f32": (4, 4), "u32": (4, 4), "i32": (4, 4)}


class Tok:
    __slots__ = ("kind", "text", "pre")

    def __init__(self, kind, text, pre=""):
        self.kind, self.text, self.pre = kind, text, pre    # pre: whitespace/comments before the token

    def __repr__(self):
        return "%s:%r" % (self.kind, self.text)


def tokenize(src):
    toks, pre, pos = [], "", 0
    while pos < len(src):
        m = TOKEN_RE.match(src, pos)
        if not m:
            raise SyntaxError("make_ref: cannot tokenize at offset %d: %r" % (pos, src[pos:pos + 20]))
        pos = m.end()
        kind = m.lastgroup
        if kind in ("ws", "comment"):
            pre += m.group()
            continue
        toks.append(Tok(kind, m.group(), pre))
        pre = ""
    toks.append(Tok("eof", "", pre))
    return toks


def cname(name):
    return name + "_" if name in CPP_RESERVED else name


# ---- types ---------------------------------------------------------------------------------
class Type:
    """kind: scalar | vec | array | ptr | texture | struct"""

    def __init__(self, kind, name, elem=None, n=0):
        self.kind, self.name, self.elem, self.n = kind, name, elem, n

    def cpp(self):
        if self.kind == "scalar":
            return self.name
        if self.kind == "vec":
            return "vec%d<%s>" % (self.n, self.elem.cpp())
        if self.kind == "array":
            return "storage_array<%s>" % self.elem.cpp()
        if self.kind == "ptr":
            return self.elem.cpp() + "*"
        if self.kind == "texture":
            return "texture_storage_2d_" + self.name
        return cname(self.name)

    def deps(self):
        if self.kind == "struct":
            return {self.name}
        return self.elem.deps() if self.elem is not None else set()


class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self, k=0):
        return self.t[self.i + k]

    def next(self):
        tok = self.t[self.i]
        self.i += 1
        return tok

    def expect(self, text):
        tok = self.next()
        if tok.text != text:
            raise SyntaxError("make_ref: expected %r, found %r (token %d)" % (text, tok.text, self.i))
        return tok

    def accept(self, text):
        if self.peek().text == text:
            return self.next()
        return None

    def skip_attributes(self):
        attrs = []
        while self.peek().text == "@":
            self.next()
            name = self.next().text
            args = []
            if self.accept("("):
                depth = 1
                while depth:
                    tok = self.next()
                    if tok.text == "(":
                        depth += 1
                    elif tok.text == ")":
                        depth -= 1
                    if depth:
                        args.append(tok.text)
            attrs.append((name, "".join(args)))
        return attrs

    def parse_type(self):
        name = self.next().text
        if name in SCALARS:
            return Type("scalar", name)
        m = re.fullmatch(r"vec([234])", name)
        if m:
            self.expect("<")
            elem = self.parse_type()
            self.expect(">")
            return Type("vec", name, elem, int(m.group(1)))
        if name == "array":
            self.expect("<")
            elem = self.parse_type()
            self.expect(">")
            return Type("array", name, elem)
        if name == "ptr":
            self.expect("<")
            self.next()                      # address space
            self.expect(",")
            elem = self.parse_type()
            if self.accept(","):
                self.next()                  # access mode
            self.expect(">")
            return Type("ptr", name, elem)
        if name == "texture_storage_2d":
            self.expect("<")
            fmt = self.next().text
            self.expect(",")
            self.next()
            self.expect(">")
            return Type("texture", fmt)
        return Type("struct", name)


# ---- module-scope parse --------------------------------------------------------------------
class Module:
    def __init__(self):
        self.structs = {}        # name -> [(field, Type)]
        self.struct_order = []
        self.consts = []         # (name, Type|None, [tokens])
        self.resources = []      # (group, binding, name, Type, space)
        self.functions = []      # dict(name, params[(name, Type)], ret Type|None, body [tokens], attrs)


def parse_module(toks):
    p, mod = Parser(toks), Module()
    while p.peek().kind != "eof":
        attrs = p.skip_attributes()
        tok = p.next()
        if tok.text == ";":
            continue
        if tok.text == "struct":
            name = p.next().text
            p.expect("{")
            fields = []
            while not p.accept("}"):
                p.skip_attributes()
                fname = p.next().text
                p.expect(":")
                ftype = p.parse_type()
                fields.append((fname, ftype))
                if not (p.accept(",") or p.accept(";")):
                    if p.peek().text != "}":
                        raise SyntaxError("make_ref: struct %s: expected , or }" % name)
            p.accept(";")
            mod.structs[name] = fields
            mod.struct_order.append(name)
        elif tok.text in ("const", "let") and p.peek(1).text in (":", "="):
            name = p.next().text
            ctype = None
            if p.accept(":"):
                ctype = p.parse_type()
            p.expect("=")
            expr = []
            while p.peek().text != ";":
                expr.append(p.next())
            p.expect(";")
            mod.consts.append((name, ctype, expr))
        elif tok.text == "var":
            space = ""
            if p.accept("<"):
                parts = []
                while not p.accept(">"):
                    parts.append(p.next().text)
                space = "".join(parts)
            name = p.next().text
            p.expect(":")
            vtype = p.parse_type()
            p.expect(";")
            a = dict(attrs)
            mod.resources.append((int(a.get("group", "0")), int(a.get("binding", "-1")), name, vtype, space))
        elif tok.text == "fn":
            name = p.next().text
            p.expect("(")
            params = []
            while not p.accept(")"):
                p.skip_attributes()
                pname = p.next().text
                p.expect(":")
                params.append((pname, p.parse_type()))
                p.accept(",")
            ret = None
            if p.accept("->"):
                p.skip_attributes()
                ret = p.parse_type()
            p.expect("{")
            depth, body = 1, []
            while depth:
                tok = p.next()
                if tok.kind == "eof":
                    raise SyntaxError("make_ref: fn %s: unterminated body" % name)
                if tok.text == "{":
                    depth += 1
                elif tok.text == "}":
                    depth -= 1
                if depth:
                    body.append(tok)
            mod.functions.append(dict(name=name, params=params, ret=ret, body=body, attrs=attrs, tail=tok.pre))
        else:
            raise SyntaxError("make_ref: unexpected %r at module scope" % tok.text)
    return mod


# ---- WGSL host-shareable layout (spec 13.4.1: alignment and size) ----------------------------
def layout_of(mod, ty, cache):
    if ty.kind == "scalar":
        return SCALAR_LAYOUT[ty.name]
    if ty.kind == "vec":
        es = layout_of(mod, ty.elem, cache)[0]
        return {2: (2 * es, 2 * es), 3: (3 * es, 4 * es), 4: (4 * es, 4 * es)}[ty.n]
    if ty.kind == "struct":
        return struct_layout(mod, ty.name, cache)[:2]
    raise ValueError("make_ref: no host-shareable layout for %s" % ty.name)


def struct_layout(mod, name, cache):
    if name in cache:
        return cache[name]
    off, align, offsets = 0, 1, []
    for fname, ftype in mod.structs[name]:
        fs, fa = layout_of(mod, ftype, cache)
        off = (off + fa - 1) // fa * fa
        offsets.append(off)
        off += fs
        align = max(align, fa)
    size = (off + align - 1) // align * align
    cache[name] = (size, align, offsets)
    return cache[name]


def host_shareable(mod, name):
    for _, ftype in mod.structs[name]:
        if ftype.kind == "scalar" and ftype.name == "bool":
            return False
        if ftype.kind == "struct" and not host_shareable(mod, ftype.name):
            return False
        if ftype.kind not in ("scalar", "vec", "struct"):
            return False
    return True


# ---- expression / statement token rewriting ------------------------------------------------
def is_float_literal(text):
    if text[-1] in "fhiu":
        return False
    if text[:2].lower() == "0x":
        return "." in text or "p" in text.lower()
    return "." in text or "e" in text.lower()


def emit_tokens(toks):
    return "".join(t.pre + t.text for t in toks)


def rewrite_expr_tokens(toks):
    """literal suffixes, conversions, reserved identifiers; type syntax vecN<T>(...) is valid C++ as is."""
    out = []
    for k, t in enumerate(toks):
        text = t.text
        if t.kind == "num" and is_float_literal(text):
            text += "f"
        elif t.kind == "ident":
            nxt = toks[k + 1].text if k + 1 < len(toks) else ""
            prev = toks[k - 1].text if k else ""
            if text in ("u32", "i32", "f32") and nxt == "(":
                text = "to_" + text
            elif prev != ".":
                text = cname(text)
        out.append(Tok(t.kind, text, t.pre))
    return out


def rewrite_body(fn):
    """Statement-level rewriting of one function body (a flat token list)."""
    toks = fn["body"]
    params = {n for n, _ in fn["params"]}
    renamed = {}                                   # parameter -> index of the token where the shadowing local is declared
    # pass 1: find locals that shadow parameters
    for k, t in enumerate(toks):
        if t.kind == "ident" and t.text in ("var", "let") and toks[k + 1].kind == "ident" and toks[k + 1].text in params:
            name = toks[k + 1].text
            if name not in renamed:
                end = k
                while toks[end].text != ";":
                    end += 1
                renamed[name] = (k + 1, end)       # uses up to the end of this statement (except the declared name) mean the parameter
    out, k, n = [], 0, len(toks)
    p = Parser(toks + [Tok("eof", "")])
    while k < n:
        t = toks[k]
        if t.kind == "ident" and t.text in ("var", "let") and toks[k + 1].kind == "ident" and (k == 0 or toks[k - 1].text != "."):
            is_let = t.text == "let"
            name_tok = toks[k + 1]
            p.i = k + 2
            vtype = None
            if p.peek().text == ":":
                p.next()
                vtype = p.parse_type()
            has_init = p.peek().text == "="
            decl = ("const " if is_let else "") + (vtype.cpp() if vtype is not None else "auto") + " " + cname(name_tok.text)
            if not has_init:
                decl += "{}"
            out.append(Tok("raw", decl, t.pre))
            k = p.i
            continue
        if t.kind == "ident" and t.text == "if" and toks[k + 1].text != "(":
            # `if cond {` -> `if (cond) {`
            out.append(Tok("ident", "if", t.pre))
            out.append(Tok("op", "(", " "))
            k += 1
            depth = 0
            while not (toks[k].text == "{" and depth == 0):
                if toks[k].text in "([":
                    depth += 1
                elif toks[k].text in ")]":
                    depth -= 1
                out.append(toks[k])
                k += 1
            out.append(Tok("op", ")", ""))
            continue
        if t.kind == "ident" and t.text in renamed and (k == 0 or toks[k - 1].text != "."):
            decl_at, stmt_end = renamed[t.text]
            if k < stmt_end and k != decl_at:
                out.append(Tok("ident", t.text + "_param", t.pre))
                k += 1
                continue
        out.append(t)
        k += 1
    return rewrite_expr_tokens(sequence_arguments(out, fn["impure"])), set(renamed)


def sequence_arguments(toks, impure):
    """WGSL evaluates call arguments left to right; C++ leaves the order unspecified.  Where one statement holds
    two or more calls of functions with side effects (a `ptr` parameter), they must be sibling arguments of a
    vecN constructor, which is then written with braces (C++ sequences a braced list left to right); any other
    shape is refused rather than mistranslated."""
    n, k, start = len(toks), 0, 0
    out = list(toks)
    while k <= n:
        if k == n or toks[k].text in (";", "{", "}"):
            calls = [m for m in range(start, k) if toks[m].kind == "ident" and toks[m].text in impure
                     and m + 1 < k and toks[m + 1].text == "(" and (m == 0 or toks[m - 1].text != ".")]
            if len(calls) >= 2:
                chains = []
                for m in calls:
                    depth, q, chain = 0, m - 1, []
                    while q >= start:                       # every unmatched "(" to the left, innermost first
                        if toks[q].text == ")":
                            depth += 1
                        elif toks[q].text == "(":
                            if depth == 0:
                                chain.append(q)
                            else:
                                depth -= 1
                        q -= 1
                    chains.append(chain)
                common = [q for q in chains[0] if all(q in c for c in chains[1:])]
                if not common:
                    raise SyntaxError("make_ref: several side-effecting calls in one statement outside one constructor: " + emit_tokens(toks[start:k]))
                op = common[0]                              # innermost group holding all of them
                args = []
                for m in calls:                             # each must sit in its own top-level argument of that group
                    depth, arg = 0, 0
                    for q in range(op + 1, m):
                        if toks[q].text in "([":
                            depth += 1
                        elif toks[q].text in ")]":
                            depth -= 1
                        elif toks[q].text == "," and depth == 0:
                            arg += 1
                    args.append(arg)
                if len(set(args)) != len(args):
                    raise SyntaxError("make_ref: several side-effecting calls inside one argument: " + emit_tokens(toks[start:k]))
                q = op - 1
                if toks[q].text == ">":                    # vecN<T>( ... )
                    while toks[q].text != "<":
                        q -= 1
                    q -= 1
                if not re.fullmatch(r"vec[234]", toks[q].text):
                    raise SyntaxError("make_ref: several side-effecting calls as arguments of one call: " + emit_tokens(toks[start:k]))
                depth, cl = 0, op
                while True:
                    if toks[cl].text == "(":
                        depth += 1
                    elif toks[cl].text == ")":
                        depth -= 1
                        if depth == 0:
                            break
                    cl += 1
                out[op] = Tok("op", "{", toks[op].pre)
                out[cl] = Tok("op", "}", toks[cl].pre)
            start = k + 1
        k += 1
    return out


# ---- code generation -----------------------------------------------------------------------
def generate(src, wgsl_path):
    mod = parse_module(tokenize(src))
    cache = {}
    o = []
    w = o.append
    w("// GENERATED by oracle/make_ref.py from %s — do not edit, do not commit.\n" % wgsl_path)
    w("// A token-level C++ rewrite of the reference's compute shader; see make_ref.py for the rules.\n")
    for fn in mod.functions:
        w("#define WGSL_USER_%s 1\n" % fn["name"])
    w('#include "wgsl_builtins.h"\n\nnamespace wgsl { namespace ref {\n\n')

    # function ids for the hooks
    w("enum fn_id {\n")
    for k, fn in enumerate(mod.functions):
        w("    FN_%s = %d,\n" % (fn["name"], k))
    w("    FN_COUNT = %d\n};\n" % len(mod.functions))
    w("static const char* const fn_names[] = {%s};\n\n" % ", ".join('"%s"' % fn["name"] for fn in mod.functions))
    w('} }\n#include "wgsl_ref_hooks.h"\nnamespace wgsl { namespace ref {\n\n')

    # structs, dependency order
    done, order = set(), []

    def visit(name):
        if name in done:
            return
        done.add(name)
        for _, ftype in mod.structs[name]:
            for d in ftype.deps():
                visit(d)
        order.append(name)
    for name in mod.struct_order:
        visit(name)
    for name in order:
        w("struct %s {\n" % cname(name))
        for fname, ftype in mod.structs[name]:
            w("    %s %s{};\n" % (ftype.cpp(), cname(fname)))
        w("};\n")
    w("\n} }\n\nnamespace wgsl {\n")
    for name in order:
        if not host_shareable(mod, name):
            continue
        size, align, offsets = struct_layout(mod, name, cache)
        w("template <> struct layout<ref::%s> { enum { size = %d, align = %d };\n" % (cname(name), size, align))
        w("    static ref::%s load(const uint8_t* p) { ref::%s s;" % (cname(name), cname(name)))
        for (fname, ftype), off in zip(mod.structs[name], offsets):
            w(" s.%s = layout<%s>::load(p + %d);" % (cname(fname), ("ref::" if ftype.kind == "struct" else "") + ftype.cpp(), off))
        w(" return s; } };\n")
    w("}\n\nnamespace wgsl { namespace ref {\n\n")

    # layout table for the driver's checks
    w("struct layout_entry { const char* name; unsigned size, align; };\nstatic const layout_entry layouts[] = {\n")
    for name in order:
        if host_shareable(mod, name):
            size, align, _ = struct_layout(mod, name, cache)
            w('    {"%s", %d, %d},\n' % (name, size, align))
    w("    {nullptr, 0, 0}};\n\n")

    for name, ctype, expr in mod.consts:
        w("static const %s %s =%s;\n" % (ctype.cpp() if ctype else "auto", cname(name), emit_tokens(rewrite_expr_tokens(expr))))
    w("\n")
    for group, binding, name, vtype, space in mod.resources:
        w("static %s %s;   // @group(%d) @binding(%d) var<%s>\n" % (vtype.cpp(), cname(name), group, binding, space))
    w("\n")

    def signature(fn, name, pnames):
        ret = fn["ret"].cpp() if fn["ret"] is not None else "void"
        return "static %s %s(%s)" % (ret, name, ", ".join("%s %s" % (t.cpp(), n) for (_, t), n in zip(fn["params"], pnames)))

    for fn in mod.functions:
        w(signature(fn, cname(fn["name"]), [cname(n) for n, _ in fn["params"]]) + ";\n")
    w("\n")
    impure = {fn["name"] for fn in mod.functions if any(t.kind == "ptr" for _, t in fn["params"])}
    for fn in mod.functions:
        fn["impure"] = impure
        body, shadowed = rewrite_body(fn)
        inner = [cname(n) + ("_param" if n in shadowed else "") for n, _ in fn["params"]]
        outer = [cname(n) for n, _ in fn["params"]]
        w(signature(fn, cname(fn["name"]) + "_body", inner) + " {")
        w(emit_tokens(body))
        w(fn["tail"] + "}\n")
        args = ", ".join(outer)
        w(signature(fn, cname(fn["name"]), outer) + " {\n")
        w("    WGSL_HOOK_ENTER(FN_%s%s);\n" % (fn["name"], (", " + args) if args else ""))
        if fn["ret"] is not None:
            w("    const %s ret_ = %s_body(%s);\n" % (fn["ret"].cpp(), cname(fn["name"]), args))
            w("    WGSL_HOOK_RETURN(FN_%s, ret_);\n    return ret_;\n}\n\n" % fn["name"])
        else:
            w("    %s_body(%s);\n    WGSL_HOOK_LEAVE(FN_%s);\n}\n\n" % (cname(fn["name"]), args, fn["name"]))

    # resource table: the driver binds by (group, binding) like src/core/ray_tracer.rs:19-130
    w("#define WGSL_RESOURCES(X) \\\n")
    for group, binding, name, vtype, space in mod.resources:
        w("    X(%d, %d, %s, %s) \\\n" % (group, binding, cname(name), vtype.kind))
    w("\n")
    entry = [fn for fn in mod.functions if any(a[0] == "compute" for a in fn["attrs"])]
    if len(entry) != 1:
        raise SyntaxError("make_ref: expected exactly one @compute entry point")
    wg = dict(entry[0]["attrs"]).get("workgroup_size", "1")
    w("#define WGSL_ENTRY %s\n#define WGSL_WORKGROUP_SIZE %s\n" % (cname(entry[0]["name"]), wg))
    w("\n} }\n\n#include \"wgsl_ref_driver.inc\"\n")
    return "".join(o), mod


CXXFLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
            "-fno-unsafe-math-optimizations", "-fno-math-errno", "-mfma", "-mavx2", "-Wall", "-Wno-unused-function",
            "-Wno-unused-variable", "-Wno-unused-but-set-variable"]


def build(wgsl=DEFAULT_WGSL, out=None, force=False, quiet=True):
    """Generate + compile. Returns the output directory, or None when the reference is not present
    (the GPU box: the prebuilt libraries travel there with the snapshot)."""
    out = out or os.path.join(HERE, "_ref")
    libs = [os.path.join(out, "librt_wgsl_ref.so"), os.path.join(out, "librt_wgsl_ref_plain.so")]
    if not os.path.exists(wgsl):
        return out if all(os.path.exists(l) for l in libs) else None
    deps = [wgsl, os.path.abspath(__file__)] + [os.path.join(HERE, f) for f in
                                                ("wgsl_builtins.h", "wgsl_ref_hooks.h", "wgsl_ref_driver.inc", "rto_math.h")]
    newest = max(os.path.getmtime(d) for d in deps)
    if not force and all(os.path.exists(l) and os.path.getmtime(l) >= newest for l in libs):
        return out
    os.makedirs(out, exist_ok=True)
    with open(wgsl) as f:
        src = f.read()
    code, _ = generate(src, wgsl)
    gen = os.path.join(out, "ray_tracer_wgsl.gen.cpp")
    with open(gen, "w") as f:
        f.write(code)
    for lib, extra in zip(libs, (["-DWGSL_CONTRACT"], [])):
        cmd = ["/usr/bin/g++"] + CXXFLAGS + extra + ["-I", HERE, gen, "-o", lib]
        if not quiet:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--wgsl", default=DEFAULT_WGSL)
    ap.add_argument("--out", default=None)
    ap.add_argument("--force", action="store_true")
    a = ap.parse_args()
    r = build(a.wgsl, a.out, a.force, quiet=False)
    print("oracle/_ref:", r if r else "reference not present and no prebuilt libraries")
    sys.exit(0 if r else 1)
