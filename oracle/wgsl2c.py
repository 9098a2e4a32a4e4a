#!/usr/bin/env python3
"""
wgsl2c.py -- mechanical WGSL -> C11 transliterator (TEST INFRASTRUCTURE, part of the oracle).

Purpose: pin the hand-written oracle (trt_oracle_core.inc) to the reference's OWN source text.  The
reference's path tracer is one WGSL compute shader (/root/reference/src/shader.wgsl); it cannot be
run here (no Rust, no wgpu adapter), but it is plain scalar code, so this tool reads it at build
time and emits the same program as C: one C statement per WGSL statement, one C operator per WGSL
operator, every function kept (the dead helpers too), nothing reordered, nothing simplified.  The one
structural change: a call to a function with side effects (it writes var<private> rand_state or a storage
buffer, directly or through a callee) is hoisted into its own `const T _wN = f(..);` statement just ahead of
the statement that uses it, in source order -- WGSL evaluates operands left to right, C leaves the order
unspecified.  The
output (oracle/_ref/shader_wgsl.inc) is compiled by oracle/Makefile with -ffp-contract=off into
oracle/_ref/libref.so, "the reference compiled here".  No reference source is copied into the repo:
the generated file lives only under oracle/_ref/ (git-ignored).

What is NOT taken from the shader, because WGSL leaves it to the implementation (naga + driver):
the accuracy of sqrt / sin / cos / tan and division, the f32<->u32 conversion edge cases and the
dispatch loop.  Those live in the hand-written oracle/wgsl_rt.h + oracle/ref_driver.c, documented
there.

Subset handled (everything shader.wgsl uses, and a little more so that small upstream edits still
translate; anything else stops the build with an error rather than being guessed at):
  declarations: struct, fn, var<private>, @group/@binding var<uniform|storage>, entry-point attributes
  types:        f32 u32 i32 bool, structs, vec3<u32>, array<T, N> members, array<T> storage bindings
  statements:   let, var, assignment (= += -= *= /=), if / else if / else, for, while, loop-free
                break / continue / return, call statements
  expressions:  C-like operators, calls, struct constructors, member and index access, literals
  WGSL typing:  AbstractInt / AbstractFloat literals are constant-folded in Python (f64 / int, as
                the WGSL spec evaluates const-expressions) and rounded ONCE to the concrete type they
                meet (f32 literals are emitted as hex floats, exact); `let x = 1.0` concretises to f32,
                `let i = 0` to i32.
  builtins:     sqrt abs min max tan sin cos floor ceil f32() u32() i32()  ->  wgsl_* (wgsl_rt.h)

Usage: wgsl2c.py <shader.wgsl> <out.inc>
"""
from __future__ import annotations

import re
import struct
import sys

C_KEYWORDS = {
    "auto", "break", "case", "char", "const", "continue", "default", "do", "double", "else", "enum", "extern",
    "float", "for", "goto", "if", "inline", "int", "long", "register", "restrict", "return", "short", "signed",
    "sizeof", "static", "struct", "switch", "typedef", "union", "unsigned", "void", "volatile", "while", "bool",
    "true", "false", "main", "index", "abs", "min", "max", "y0", "y1", "j0", "j1", "jn", "yn", "gamma", "exp", "log",
}
SCALARS = {"f32": "float", "u32": "uint32_t", "i32": "int32_t", "bool": "bool"}

TOKEN_RE = re.compile(r"""
    (?P<ws>\s+|//[^\n]*|/\*.*?\*/)
  | (?P<float>(?:\d+\.\d*(?:[eE][+-]?\d+)?|\.\d+(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)[fh]?|\d+[fh](?![A-Za-z0-9_]))
  | (?P<int>0[xX][0-9a-fA-F]+[iu]?|\d+[iu]?)
  | (?P<id>[A-Za-z_][A-Za-z0-9_]*)
  | (?P<op>->|<<|>>|<=|>=|==|!=|&&|\|\||\+=|-=|\*=|/=|[-+*/%<>=!&|^~(){}\[\],;:.@])
""", re.X | re.S)


class WgslError(Exception):
    pass


def tokenize(src: str):
    toks, pos, line = [], 0, 1
    while pos < len(src):
        m = TOKEN_RE.match(src, pos)
        if not m:
            raise WgslError(f"line {line}: cannot tokenise {src[pos:pos + 20]!r}")
        kind = m.lastgroup
        text = m.group()
        if kind != "ws":
            toks.append((kind, text, line))
        line += text.count("\n")
        pos = m.end()
    toks.append(("eof", "", line))
    return toks


def f32_round(v: float) -> float:
    """Round a Python float (f64) to the nearest binary32, ties to even; overflow is an error in WGSL."""
    try:
        return struct.unpack("<f", struct.pack("<f", v))[0]
    except OverflowError:
        raise WgslError(f"constant {v!r} does not fit f32")


def f32_literal(v: float) -> str:
    r = f32_round(v)
    if r != r or r in (float("inf"), float("-inf")):
        raise WgslError("non-finite constant")
    return f"{r.hex()}f/*{r!r}*/"


class E:
    """A typed expression: C text, WGSL type, and the Python value while it is still abstract."""
    __slots__ = ("c", "t", "v", "lvalue", "g")

    def __init__(self, c, t, v=None, lvalue=False, g=False):
        self.c, self.t, self.v, self.lvalue, self.g = c, t, v, lvalue, g


def is_abstract(t):
    return t in ("AI", "AF")


class Translator:
    def __init__(self, src: str, origin: str, impure=None):
        self.toks = tokenize(src)
        self.impure = impure            # None: analysis pass (nothing hoisted); else the set of functions with effects
        self.writes: dict[str, bool] = {}
        self.calls: dict[str, set[str]] = {}
        self.cur_fn = None
        self.pre: list[str] = []        # hoisted calls waiting for the statement they belong to
        self.no_hoist = 0
        self.ntmp = 0
        self.p = 0
        self.origin = origin
        self.structs: dict[str, list[tuple[str, object]]] = {}
        self.struct_order: list[str] = []
        self.funcs: dict[str, tuple[list[tuple[str, object]], object]] = {}
        self.globals: dict[str, tuple[object, str]] = {}      # name -> (type, C name)
        self.global_decls: list[str] = []
        self.bindings: list[tuple[int, str, str, object]] = []  # (binding, space, name, type)
        self.entry: list[str] = []
        self.scopes: list[dict[str, tuple[object, str]]] = []
        self.out: list[str] = []
        self.cur_ret = None

    # ---- token helpers -------------------------------------------------------------------------
    def peek(self, k=0):
        return self.toks[self.p + k]

    def at(self, text):
        return self.toks[self.p][1] == text and self.toks[self.p][0] in ("op", "id")

    def accept(self, text):
        if self.at(text):
            self.p += 1
            return True
        return False

    def expect(self, text):
        if not self.accept(text):
            k, t, ln = self.peek()
            raise WgslError(f"line {ln}: expected {text!r}, found {t!r}")

    def ident(self):
        k, t, ln = self.peek()
        if k != "id":
            raise WgslError(f"line {ln}: expected identifier, found {t!r}")
        self.p += 1
        return t

    def err(self, msg):
        raise WgslError(f"line {self.peek()[2]}: {msg}")

    # ---- types ---------------------------------------------------------------------------------
    def parse_type(self):
        name = self.ident()
        if name in ("vec2", "vec3", "vec4"):
            self.expect("<")
            elem = self.parse_type()
            self.expect(">")
            return ("vec", int(name[3]), elem)
        if name == "array":
            self.expect("<")
            elem = self.parse_type()
            n = None
            if self.accept(","):
                k, t, ln = self.peek()
                if k != "int":
                    self.err("array size must be an integer literal")
                self.p += 1
                n = int(t.rstrip("iu"), 0)
            self.expect(">")
            return ("array", elem, n)
        if name in SCALARS or name in self.structs:
            return name
        self.err(f"unknown type {name!r}")

    def ctype(self, t) -> str:
        if isinstance(t, str):
            if t in SCALARS:
                return SCALARS[t]
            if t in self.structs:
                return f"W_{t}"
        if isinstance(t, tuple) and t[0] == "vec" and t[2] in ("u32", "f32", "i32"):
            return f"wgsl_vec{t[1]}{t[2][0]}"
        raise WgslError(f"no C type for {t!r}")

    def cdecl(self, t, name) -> str:
        if isinstance(t, tuple) and t[0] == "array":
            if t[2] is None:
                raise WgslError("runtime-sized array outside a storage binding")
            return f"{self.ctype(t[1])} {name}[{t[2]}]"
        return f"{self.ctype(t)} {name}"

    @staticmethod
    def cname(name: str) -> str:
        return name + "_" if name in C_KEYWORDS else name

    # ---- scopes --------------------------------------------------------------------------------
    def lookup(self, name):
        for sc in reversed(self.scopes):
            if name in sc:
                return sc[name]
        if name in self.globals:
            return self.globals[name]
        return None

    def declare(self, name, t):
        c = self.cname(name)
        # a WGSL block may shadow an outer name; C allows the same, so the name is kept
        self.scopes[-1][name] = (t, c)
        return c

    # ---- abstract values -----------------------------------------------------------------------
    def concretise(self, e: E, t) -> E:
        if e.t == t:
            return e
        if is_abstract(e.t):
            if t == "f32":
                return E(f32_literal(float(e.v)), "f32")
            if t in ("u32", "i32") and e.t == "AI":
                if t == "u32" and not (0 <= e.v <= 0xFFFFFFFF):
                    self.err(f"constant {e.v} does not fit u32")
                return E(f"{e.v}u" if t == "u32" else f"{e.v}", t)
        self.err(f"type mismatch: have {e.t!r}, need {t!r} (WGSL has no implicit conversion)")

    def default_concrete(self, e: E) -> E:
        if e.t == "AI":
            return self.concretise(e, "i32")
        if e.t == "AF":
            return self.concretise(e, "f32")
        return e

    # ---- expressions ---------------------------------------------------------------------------
    BIN_LEVELS = [["||"], ["&&"], ["|"], ["^"], ["&"], ["==", "!="], ["<", ">", "<=", ">="], ["<<", ">>"],
                  ["+", "-"], ["*", "/", "%"]]

    def parse_expr(self, level=0) -> E:
        if level == len(self.BIN_LEVELS):
            return self.parse_unary()
        lhs = self.parse_expr(level + 1)
        while self.peek()[0] == "op" and self.peek()[1] in self.BIN_LEVELS[level]:
            op = self.peek()[1]
            self.p += 1
            self.no_hoist += op in ("&&", "||")      # a short-circuited operand is evaluated conditionally
            rhs = self.parse_expr(level + 1)
            self.no_hoist -= op in ("&&", "||")
            lhs = self.binary(op, lhs, rhs)
        return lhs

    def binary(self, op, a: E, b: E) -> E:
        if op in ("&&", "||"):
            if a.t != "bool" or b.t != "bool":
                self.err(f"{op} needs bool operands")
            return E(f"({a.c} {op} {b.c})", "bool")
        if is_abstract(a.t) and is_abstract(b.t):
            return self.fold(op, a, b)
        if op in ("<<", ">>"):
            a = self.default_concrete(a)
            b = self.concretise(b, "u32")
            if a.t not in ("u32", "i32"):
                self.err("shift of a non-integer")
            return E(f"({a.c} {op} {b.c})", a.t)
        if is_abstract(a.t):
            a = self.concretise(a, b.t)
        elif is_abstract(b.t):
            b = self.concretise(b, a.t)
        if a.t != b.t:
            self.err(f"operands of {op!r} differ: {a.t!r} vs {b.t!r}")
        if op in ("==", "!=", "<", ">", "<=", ">="):
            if a.t not in SCALARS:
                self.err(f"comparison of {a.t!r}")
            return E(f"({a.c} {op} {b.c})", "bool")
        if op in ("&", "|", "^"):
            if a.t not in ("u32", "i32", "bool"):
                self.err(f"{op!r} on {a.t!r}")
            return E(f"({a.c} {op} {b.c})", a.t)
        if a.t not in ("f32", "u32", "i32"):
            self.err(f"arithmetic on {a.t!r}")
        if op == "%" and a.t == "f32":
            return E(f"wgsl_mod_f32({a.c}, {b.c})", "f32")
        if op in ("/", "%") and a.t in ("u32", "i32"):
            return E(f"wgsl_{'div' if op == '/' else 'mod'}_{a.t}({a.c}, {b.c})", a.t)
        return E(f"({a.c} {op} {b.c})", a.t)

    def fold(self, op, a: E, b: E) -> E:
        """Const-expression over abstract operands: AbstractInt is exact, AbstractFloat is f64."""
        t = "AF" if "AF" in (a.t, b.t) else "AI"
        x, y = a.v, b.v
        if t == "AF":
            x, y = float(x), float(y)
        if op == "+":
            v = x + y
        elif op == "-":
            v = x - y
        elif op == "*":
            v = x * y
        elif op == "/":
            if y == 0:
                self.err("division by zero in a const-expression")
            v = x / y if t == "AF" else int(abs(x) // abs(y)) * (1 if (x >= 0) == (y >= 0) else -1)
        elif op in ("==", "!=", "<", ">", "<=", ">="):
            r = {"==": x == y, "!=": x != y, "<": x < y, ">": x > y, "<=": x <= y, ">=": x >= y}[op]
            return E("true" if r else "false", "bool")
        else:
            self.err(f"const-expression operator {op!r} not supported")
        return E(None, t, v)

    def parse_unary(self) -> E:
        if self.accept("-"):
            e = self.parse_unary()
            if is_abstract(e.t):
                return E(None, e.t, -e.v)
            if e.t not in ("f32", "i32"):
                self.err(f"negation of {e.t!r}")
            return E(f"(-{e.c})", e.t)
        if self.accept("!"):
            e = self.parse_unary()
            if e.t != "bool":
                self.err("! needs a bool")
            return E(f"(!{e.c})", "bool")
        if self.accept("~"):
            e = self.default_concrete(self.parse_unary())
            return E(f"(~{e.c})", e.t)
        return self.parse_postfix()

    def parse_postfix(self) -> E:
        e = self.parse_primary()
        while True:
            if self.accept("."):
                f = self.ident()
                if isinstance(e.t, str) and e.t in self.structs:
                    ft = dict(self.structs[e.t]).get(f)
                    if ft is None:
                        self.err(f"struct {e.t} has no member {f!r}")
                    e = E(f"{e.c}.{self.cname(f)}", ft, lvalue=e.lvalue, g=e.g)
                elif isinstance(e.t, tuple) and e.t[0] == "vec" and f in "xyzw"[: e.t[1]] and len(f) == 1:
                    e = E(f"{e.c}.{f}", e.t[2], lvalue=e.lvalue, g=e.g)
                else:
                    self.err(f"member access .{f} on {e.t!r}")
            elif self.accept("["):
                i = self.parse_expr()
                self.expect("]")
                if is_abstract(i.t):
                    i = self.concretise(i, "u32")
                if i.t not in ("u32", "i32"):
                    self.err("array index must be an integer")
                if not (isinstance(e.t, tuple) and e.t[0] == "array"):
                    self.err(f"indexing {e.t!r}")
                e = E(f"{e.c}[{i.c}]", e.t[1], lvalue=e.lvalue, g=e.g)
            else:
                return e

    def parse_args(self):
        args = []
        self.expect("(")
        if not self.accept(")"):
            while True:
                args.append(self.parse_expr())
                if self.accept(")"):
                    break
                self.expect(",")
                if self.accept(")"):
                    break
        return args

    def parse_primary(self) -> E:
        k, t, ln = self.peek()
        if k == "float":
            self.p += 1
            if t[-1] == "h":
                self.err("f16 literals are not supported")
            if t[-1] == "f":
                return E(f32_literal(float(t[:-1])), "f32")
            return E(None, "AF", float(t))
        if k == "int":
            self.p += 1
            if t[-1] == "u":
                return E(f"{int(t[:-1], 0)}u", "u32")
            if t[-1] == "i":
                return E(f"{int(t[:-1], 0)}", "i32")
            return E(None, "AI", int(t, 0))
        if self.accept("("):
            e = self.parse_expr()
            self.expect(")")
            return e
        if k != "id":
            self.err(f"unexpected token {t!r}")
        name = self.ident()
        if name in ("true", "false"):
            return E(name, "bool")
        if self.at("("):
            return self.call(name, self.parse_args())
        if name in ("vec2", "vec3", "vec4") and self.at("<"):
            self.err("vector constructors are not supported")
        sym = self.lookup(name)
        if sym is None:
            self.err(f"unknown identifier {name!r}")
        is_global = not any(name in sc for sc in self.scopes)
        return E(sym[1], sym[0], lvalue=True, g=is_global)

    def call(self, name, args) -> E:
        if name in self.funcs:
            params, ret = self.funcs[name]
            if len(params) != len(args):
                self.err(f"{name}: {len(args)} arguments for {len(params)} parameters")
            cargs = [self.concretise(a, pt).c for a, (_, pt) in zip(args, params)]
            text = f"w_{name}({', '.join(cargs)})"
            self.calls.setdefault(self.cur_fn, set()).add(name)
            if self.impure is not None and name in self.impure and ret is not None:
                # C leaves the order of operand / argument evaluation unspecified, WGSL evaluates left to right:
                # a call with side effects becomes its own statement, in source order, ahead of its user.
                if self.no_hoist:
                    self.err(f"call of {name}() (has side effects) in a conditionally evaluated position")
                self.ntmp += 1
                tmp = f"_w{self.ntmp}"
                self.pre.append(f"const {self.ctype(ret)} {tmp} = {text};")
                return E(tmp, ret)
            return E(text, ret)
        if name in self.structs:
            fields = self.structs[name]
            if len(args) == 0:
                return E(f"(W_{name}){{0}}", name)
            flat = []
            for ft_name, ft in fields:
                if isinstance(ft, tuple) and ft[0] == "array":
                    self.err("constructor of a struct with an array member")
                flat.append(ft)
            if len(flat) != len(args):
                self.err(f"{name}(): {len(args)} values for {len(flat)} members")
            cargs = [self.concretise(a, ft).c for a, ft in zip(args, flat)]
            return E(f"(W_{name}){{{', '.join(cargs)}}}", name)
        if name in ("f32", "u32", "i32"):
            if len(args) != 1:
                self.err(f"{name}() takes one value")
            a = args[0]
            if is_abstract(a.t):
                if name != "f32" and a.t == "AF":
                    return E(f"{int(a.v)}u" if name == "u32" else f"{int(a.v)}", name)
                return self.concretise(a, name)
            if a.t == name:
                return a
            if a.t not in ("f32", "u32", "i32", "bool"):
                self.err(f"{name}({a.t!r})")
            return E(f"wgsl_{name}_from_{a.t}({a.c})", name)
        one = {"sqrt", "tan", "sin", "cos", "floor", "ceil", "abs"}
        two = {"min", "max"}
        if name in one or name in two:
            n = 1 if name in one else 2
            if len(args) != n:
                self.err(f"{name}() takes {n} value(s)")
            if all(is_abstract(a.t) for a in args):
                args = [self.default_concrete(a) for a in args] if name not in ("sqrt", "tan", "sin", "cos") \
                    else [self.concretise(a, "f32") for a in args]
            t = next(a.t for a in args if not is_abstract(a.t))
            args = [self.concretise(a, t) for a in args]
            if name in ("sqrt", "tan", "sin", "cos", "floor", "ceil") and t != "f32":
                self.err(f"{name}({t!r})")
            if t not in ("f32", "u32", "i32"):
                self.err(f"{name}({t!r})")
            return E(f"wgsl_{name}_{t}({', '.join(a.c for a in args)})", t)
        self.err(f"call of unknown function / unsupported builtin {name!r}")

    # ---- statements ----------------------------------------------------------------------------
    def emit(self, depth, text):
        for h in self.pre:
            self.out.append("    " * depth + h)
        self.pre = []
        self.out.append("    " * depth + text)

    def parse_block(self, depth):
        self.expect("{")
        self.scopes.append({})
        while not self.accept("}"):
            self.parse_statement(depth)
        self.scopes.pop()

    def simple_statement(self) -> str:
        """let / var / assignment / call, without the trailing ';' (shared with the for header)."""
        if self.at("let") or self.at("var"):
            kw = self.ident()
            name = self.ident()
            t = None
            if self.accept(":"):
                t = self.parse_type()
            init = None
            if self.accept("="):
                init = self.parse_expr()
                init = self.concretise(init, t) if t is not None else self.default_concrete(init)
                t = init.t
            elif kw == "let":
                self.err("let without an initialiser")
            c = self.declare(name, t)
            qual = "const " if kw == "let" else ""
            if init is None:
                return f"{self.cdecl(t, c)} = {{0}}"      # WGSL zero-initialises every var
            return f"{qual}{self.cdecl(t, c)} = {init.c}"
        lhs = self.parse_unary()
        for op in ("=", "+=", "-=", "*=", "/="):
            if self.accept(op):
                if not lhs.lvalue:
                    self.err("assignment to a non-reference")
                if lhs.g:
                    self.writes[self.cur_fn] = True
                rhs = self.concretise(self.parse_expr(), lhs.t)
                if op == "=":
                    return f"{lhs.c} = {rhs.c}"
                return f"{lhs.c} = {self.binary(op[0], lhs, rhs).c}"
        if lhs.c is None:
            self.err("expression statement without effect")
        return lhs.c

    def parse_statement(self, depth):
        if self.at("{"):
            self.emit(depth, "{")
            self.parse_block(depth + 1)
            self.emit(depth, "}")
        elif self.accept("if"):
            cond = self.parse_expr()
            if cond.t != "bool":
                self.err("if condition must be bool")
            self.emit(depth, f"if ({cond.c}) {{")
            self.parse_block(depth + 1)
            while self.accept("else"):
                if self.accept("if"):
                    self.no_hoist += 1
                    cond = self.parse_expr()
                    self.no_hoist -= 1
                    if cond.t != "bool":
                        self.err("if condition must be bool")
                    self.emit(depth, f"}} else if ({cond.c}) {{")
                    self.parse_block(depth + 1)
                else:
                    self.emit(depth, "} else {")
                    self.parse_block(depth + 1)
                    break
            self.emit(depth, "}")
        elif self.accept("for"):
            self.expect("(")
            self.scopes.append({})
            init = "" if self.at(";") else self.simple_statement()
            self.expect(";")
            self.no_hoist += 1
            cond = None if self.at(";") else self.parse_expr()
            if cond is not None and cond.t != "bool":
                self.err("for condition must be bool")
            self.expect(";")
            step = "" if self.at(")") else self.simple_statement()
            self.no_hoist -= 1
            self.expect(")")
            self.emit(depth, f"for ({init}; {cond.c if cond else ''}; {step}) {{")
            self.parse_block(depth + 1)
            self.scopes.pop()
            self.emit(depth, "}")
        elif self.accept("while"):
            self.no_hoist += 1
            cond = self.parse_expr()
            self.no_hoist -= 1
            if cond.t != "bool":
                self.err("while condition must be bool")
            self.emit(depth, f"while ({cond.c}) {{")
            self.parse_block(depth + 1)
            self.emit(depth, "}")
        elif self.accept("return"):
            if self.accept(";"):
                self.emit(depth, "return;")
                return
            e = self.concretise(self.parse_expr(), self.cur_ret)
            self.expect(";")
            self.emit(depth, f"return {e.c};")
        elif self.accept("break"):
            self.expect(";")
            self.emit(depth, "break;")
        elif self.accept("continue"):
            self.expect(";")
            self.emit(depth, "continue;")
        elif self.accept(";"):
            pass
        else:
            s = self.simple_statement()
            self.expect(";")
            self.emit(depth, s + ";")

    # ---- top level -----------------------------------------------------------------------------
    def skip_attributes(self):
        attrs = []
        while self.accept("@"):
            a = self.ident()
            vals = []
            if self.accept("("):
                d = 1
                while d:
                    k, t, ln = self.peek()
                    if k == "eof":
                        self.err("unterminated attribute")
                    self.p += 1
                    d += (t == "(") - (t == ")")
                    if d:
                        vals.append(t)
            attrs.append((a, "".join(vals)))
        return attrs

    def skip_braces(self):
        self.expect("{")
        d = 1
        while d:
            k, t, ln = self.peek()
            if k == "eof":
                self.err("unterminated block")
            self.p += 1
            d += (t == "{") - (t == "}")

    def parse_fn_header(self):
        name = self.ident()
        self.expect("(")
        params = []
        while not self.accept(")"):
            self.skip_attributes()
            pn = self.ident()
            self.expect(":")
            params.append((pn, self.parse_type()))
            if not self.at(")"):
                self.expect(",")
        ret = None
        if self.accept("->"):
            self.skip_attributes()
            ret = self.parse_type()
        return name, params, ret

    def pass_declarations(self):
        """Pass 1: structs, globals and function signatures (WGSL allows use before declaration)."""
        fn_bodies = []
        while self.peek()[0] != "eof":
            attrs = self.skip_attributes()
            if self.accept("struct"):
                name = self.ident()
                self.expect("{")
                fields = []
                self.structs[name] = fields
                while not self.accept("}"):
                    self.skip_attributes()
                    fn_ = self.ident()
                    self.expect(":")
                    fields.append((fn_, self.parse_type()))
                    if not self.at("}"):
                        self.expect(",")
                self.accept(";")
                self.struct_order.append(name)
            elif self.accept("var"):
                space = "private"
                if self.accept("<"):
                    space = self.ident()
                    while not self.accept(">"):
                        self.p += 1
                name = self.ident()
                self.expect(":")
                t = self.parse_type()
                if self.accept("="):
                    self.err("initialised module-scope var not supported")
                self.expect(";")
                binding = dict(attrs).get("binding")
                if space in ("uniform", "storage"):
                    if binding is None:
                        self.err(f"{name}: resource without @binding")
                    self.bindings.append((int(binding), space, name, t))
                    if isinstance(t, tuple) and t[0] == "array" and t[2] is None:
                        self.globals[name] = (t, f"g_{name}")
                        self.global_decls.append(f"static _Thread_local {self.ctype(t[1])} *g_{name};"
                                                 f"   /* @binding({binding}) {space} */")
                    else:
                        self.globals[name] = (t, f"g_{name}")
                        self.global_decls.append(f"static {self.cdecl(t, 'g_' + name)};"
                                                 f"   /* @binding({binding}) {space} */")
                elif space == "private":
                    self.globals[name] = (t, f"g_{name}")
                    self.global_decls.append(f"static _Thread_local {self.cdecl(t, 'g_' + name)};   /* var<private> */")
                else:
                    self.err(f"address space {space!r} not supported")
            elif self.accept("fn"):
                start = self.p
                name, params, ret = self.parse_fn_header()
                if name in self.funcs:
                    self.err(f"function {name!r} declared twice")
                self.funcs[name] = (params, ret)
                if any(a == "compute" for a, _ in attrs):
                    self.entry.append(name)
                fn_bodies.append((start, name))
                self.skip_braces()
            elif self.accept(";"):
                pass
            else:
                self.err(f"unsupported module-scope declaration {self.peek()[1]!r}")
        return fn_bodies

    def effects(self) -> set[str]:
        """Functions that write a module-scope variable, directly or through a callee (after translate())."""
        imp = {f for f, w in self.writes.items() if w}
        changed = True
        while changed:
            changed = False
            for f, cs in self.calls.items():
                if f not in imp and cs & imp:
                    imp.add(f)
                    changed = True
        return imp

    def proto(self, name):
        params, ret = self.funcs[name]
        ps = ", ".join(self.cdecl(t, self.cname(n)) for n, t in params) or "void"
        return f"static {self.ctype(ret) if ret is not None else 'void'} w_{name}({ps})"

    def translate(self) -> str:
        bodies = self.pass_declarations()
        head = [
            f"/* GENERATED by oracle/wgsl2c.py from {self.origin} -- do not edit, do not commit.",
            " * One C statement per WGSL statement; AbstractFloat constants folded in f64 and rounded once to f32",
            " * (hex-float literals); builtins are the wgsl_* functions of oracle/wgsl_rt.h. */",
            "",
        ]
        for s in self.struct_order:
            head.append(f"typedef struct W_{s} {{")
            for fn_, ft in self.structs[s]:
                head.append(f"    {self.cdecl(ft, self.cname(fn_))};")
            head.append(f"}} W_{s};")
        head.append("")
        head += self.global_decls
        head.append("")
        for _, name in bodies:
            head.append(self.proto(name) + ";")
        head.append("")
        for start, name in bodies:
            self.p = start
            _, params, ret = self.parse_fn_header()
            self.cur_ret = ret
            self.cur_fn = name
            self.scopes = [{}]
            for pn, pt in params:
                self.declare(pn, pt)
            self.emit(0, self.proto(name) + " {")
            self.parse_block(1)
            self.emit(0, "}")
            self.emit(0, "")
            self.scopes = []
        tail = ["/* ---- reflection for the hand-written driver (oracle/ref_driver.c) ---- */"]
        for b, space, name, t in sorted(self.bindings):
            tail.append(f"#define WGSL_BINDING_{b}_{name} 1")
        for e in self.entry:
            tail.append(f"#define WGSL_ENTRY_{e} 1")
        for s in self.struct_order:
            tail.append(f"#define WGSL_STRUCT_{s} 1")
        return "\n".join(head + self.out + tail) + "\n"


def main(argv):
    if len(argv) != 3:
        print(__doc__)
        return 2
    with open(argv[1], encoding="utf-8") as f:
        src = f.read()
    try:
        probe = Translator(src, argv[1])
        probe.translate()                                   # analysis pass: which functions have side effects
        text = Translator(src, argv[1], impure=probe.effects()).translate()
    except WgslError as e:
        print(f"wgsl2c: {argv[1]}: {e}", file=sys.stderr)
        return 1
    with open(argv[2], "w", encoding="utf-8") as f:
        f.write(text)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
