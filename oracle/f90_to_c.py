"""TEST INFRASTRUCTURE ONLY -- a mechanical translator from a small Fortran 90 subset to C.

This image has no Fortran compiler, so the reference's Fortran cannot be compiled here.  For a subroutine that stays inside
a small, regular subset of the language this script does the next best thing: it reads the REFERENCE'S OWN SOURCE where it
lies (never copied into the repo), translates one internal subroutine statement by statement into C, and writes the C under
oracle/_ref/ (git-ignored), where `make -C oracle ref` compiles it with `gcc -O2 -ffp-contract=off` (no FMA contraction, no
reassociation -- what gfortran does without -ffast-math).  Nothing is interpreted or re-derived: every Fortran assignment
becomes one C assignment with the same operands in the same order, so a transcription error in the hand-written oracle shows
up as a bit difference against this build.  Anything outside the subset is an error, never a guess.

    python oracle/f90_to_c.py SOURCE.F90 SUBROUTINE --runtime nslices,nz,nx --out oracle/_ref/x.c

Subset (enough for mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90, subroutine advect_scalar2D_cpu, lines 477-642):
  * host-associated `integer, parameter :: name = <integer expression>` of the enclosing program (the ones named in --runtime
    become arguments of the C function instead: the source marks them "User changeable"); `rp = selected_real_kind(13)` makes
    real(rp) a C double; host-associated real arrays the body reads become trailing pointer arguments
  * declarations real(rp) / integer / integer(8) / logical, with intent, explicit-shape arrays with optional lower bounds
    (column-major, as Fortran stores them); local arrays are malloc'ed and pre-filled with NaN (Fortran leaves them undefined)
  * statement functions  name(a,b,..) = expression   ->  static inline functions
  * do / enddo, if (..) then / else / endif, the one-statement `if (..) x = ..`, assignments to scalars, array elements and whole sections `a(:,:,:,k) = expr`
  * expressions: + - * / and unary minus, parentheses, integer and real literals, `_rp` kind suffix, .true. / .false.,
    relational operators (< <= > >= == /= and their .lt. forms) and .and. / .or. / .not. in conditions,
    array references, statement-function calls, the intrinsics max, min (any number of arguments), abs
  * call system_clock(..) and write(*,*) .. are timing / printing only and are dropped; !$acc lines are comments here
  * --check-bounds (what gfortran -fcheck=bounds would abort on): a statement one of whose subscripts leaves the declared
    bounds is skipped and counted, <subroutine>_violations() reports the count and the first one (line, array, dimension)

Mixed-mode arithmetic follows from C's own usual conversions, which coincide with Fortran's for integer / real(4) / real(8):
a default-real literal such as 1.e-10 or 0.5 is emitted as a C float literal (1.e-10f), so `eps = 1.e-10` assigns
float32(1e-10) to the double exactly as the Fortran does, and 2./(kc-kb) is a single-precision quotient in both languages.
max / min compare their arguments left to right with > / < (gfortran's expansion for finite data), so signed zeros and ties
resolve as they do there.
"""
import argparse
import re
import sys

INT, FLT, DBL, LOG = "int", "float", "double", "logical"
RANK = {INT: 0, FLT: 1, DBL: 2}
CTYPE = {INT: "int", FLT: "float", DBL: "double", LOG: "int"}


class Unsupported(Exception):
    pass


# ------------------------------------------------------------------------------------------------ source lines
def logical_lines(text):
    """Comment-free statements with continuation lines joined; yields (first line number, statement)."""
    out, cur, start = [], "", None
    for no, raw in enumerate(text.splitlines(), 1):
        line = raw
        # strip comments (no character strings with '!' inside occur in the subset except in write(), which is dropped)
        if "!" in line:
            q = None
            for idx, ch in enumerate(line):
                if ch in "'\"":
                    q = None if q == ch else (ch if q is None else q)
                elif ch == "!" and q is None:
                    line = line[:idx]
                    break
        line = line.strip()
        if not line:
            continue
        if cur:
            if line.startswith("&"):
                line = line[1:].lstrip()
            cur += " " + line
        else:
            cur, start = line, no
        if cur.endswith("&"):
            cur = cur[:-1].rstrip()
            continue
        out.append((start, cur))
        cur = ""
    if cur:
        out.append((start, cur))
    return out


# ------------------------------------------------------------------------------------------------ expressions
TOKEN = re.compile(
    r"\s*(?:(?P<lop>\.(?:and|or|not|lt|le|gt|ge|eq|ne)\.)|(?P<num>(?:\d+\.(?![A-Za-z]{2,}\.)\d*|\.\d+|\d+)(?:[eEdD][-+]?\d+)?(?:_[A-Za-z]\w*)?)|(?P<log>\.(?:true|false)\.)"
    r"|(?P<id>[A-Za-z]\w*)|(?P<op>\*\*|>=|<=|==|/=|[-+*/(),:=<>]))",
    re.I,
)


def tokenize(s):
    pos, toks = 0, []
    s = s.rstrip()
    while pos < len(s):
        m = TOKEN.match(s, pos)
        if not m:
            raise Unsupported(f"cannot tokenize {s[pos:]!r}")
        kind = m.lastgroup
        val = m.group(kind)
        if kind == "lop":  # .lt. style relational operators are the same tokens as their symbols
            val = {".lt.": "<", ".le.": "<=", ".gt.": ">", ".ge.": ">=", ".eq.": "==", ".ne.": "/="}.get(val.lower(), val.lower())
            kind = "op"
        toks.append((kind, val))
        pos = m.end()
    return toks


class Node:
    def __init__(self, kind, **kw):
        self.kind = kind
        self.__dict__.update(kw)


class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else (None, None)

    def take(self, val=None):
        k, v = self.peek()
        if val is not None and v != val:
            raise Unsupported(f"expected {val!r}, found {v!r}")
        self.i += 1
        return k, v

    # precedence, lowest first: .or.  .and.  .not.  relational  + -  * /
    def lexpr(self):
        node = self.land()
        while self.peek()[1] == ".or.":
            self.take()
            node = Node("lbin", op="||", a=node, b=self.land())
        return node

    def land(self):
        node = self.lnot()
        while self.peek()[1] == ".and.":
            self.take()
            node = Node("lbin", op="&&", a=node, b=self.lnot())
        return node

    def lnot(self):
        if self.peek()[1] == ".not.":
            self.take()
            return Node("lnot", a=self.lnot())
        return self.rel()

    def rel(self):
        node = self.expr()
        if self.peek()[1] in ("<", "<=", ">", ">=", "==", "/="):
            op = self.take()[1]
            node = Node("rel", op="!=" if op == "/=" else op, a=node, b=self.expr())
        return node

    def expr(self):
        k, v = self.peek()
        if v in ("+", "-"):
            self.take()
            node = self.term()
            if v == "-":
                node = Node("neg", a=node)
        else:
            node = self.term()
        while self.peek()[1] in ("+", "-"):
            op = self.take()[1]
            node = Node("bin", op=op, a=node, b=self.term())
        return node

    def term(self):
        node = self.factor()
        while self.peek()[1] in ("*", "/"):
            op = self.take()[1]
            node = Node("bin", op=op, a=node, b=self.factor())
        return node

    def factor(self):
        k, v = self.take()
        if self.peek()[1] == "**" or v == "**":
            raise Unsupported("the ** operator is outside the subset")
        if k == "num":
            return Node("num", text=v)
        if k == "log":
            return Node("log", val=v.lower() == ".true.")
        if v == "(":
            e = self.lexpr()
            self.take(")")
            return Node("paren", a=e)
        if k == "id":
            name = v.lower()
            if self.peek()[1] == "(":
                self.take("(")
                args = []
                if self.peek()[1] != ")":
                    while True:
                        if self.peek()[1] == ":":
                            self.take()
                            args.append(Node("colon"))
                        else:
                            args.append(self.expr())
                        if self.peek()[1] == ",":
                            self.take()
                            continue
                        break
                self.take(")")
                return Node("call", name=name, args=args)
            return Node("var", name=name)
        raise Unsupported(f"unexpected token {v!r}")


def parse_expr(s):
    p = Parser(tokenize(s))
    e = p.lexpr()
    if p.i != len(p.t):
        raise Unsupported(f"trailing tokens in expression {s!r}")
    return e


# ------------------------------------------------------------------------------------------------ translation unit
class Sym:
    def __init__(self, name, typ, dims=None, role="local"):
        self.name, self.typ, self.dims, self.role = name, typ, dims, role  # dims: list of (lo expr str | None, hi expr str)


class Translator:
    def __init__(self, text, sub, runtime, check_bounds=False):
        self.check = check_bounds
        self.cur_line = 0
        self.lines = logical_lines(text)
        self.sub = sub.lower()
        self.runtime = [r.lower() for r in runtime]
        self.host_params = []   # (name, expr string) in source order
        self.host_arrays = {}   # name -> Sym (real arrays of the enclosing program)
        self.syms = {}
        self.stmt_funcs = {}    # name -> (arg names, expr node)
        self.real_kind = {}     # kind parameter name -> C type
        self.body = []
        self.used_host_arrays = []
        self.tmp = 0

    # ---- declarations
    def parse_decl(self, stmt, role_default, table):
        m = re.match(r"(real|integer|logical)\s*(\(([^)]*)\))?\s*((?:,\s*[^:]+?)*)\s*::\s*(.*)$", stmt, re.I)
        if not m:
            return False
        base, kind, attrs, rest = m.group(1).lower(), (m.group(3) or "").strip().lower(), m.group(4).lower(), m.group(5)
        if base == "real":
            if kind == "":
                typ = FLT
            elif kind in self.real_kind:
                typ = self.real_kind[kind]
            elif kind in ("8", "kind=8"):
                typ = DBL
            else:
                raise Unsupported(f"real kind {kind!r}")
        elif base == "integer":
            typ = INT
        else:
            typ = LOG
        if "parameter" in attrs:
            for name, val in split_top(rest, with_init=True):
                if typ == INT and re.match(r"selected_real_kind\s*\(\s*(\d+)\s*\)", val or "", re.I):
                    digits = int(re.match(r"selected_real_kind\s*\(\s*(\d+)\s*\)", val, re.I).group(1))
                    self.real_kind[name.lower()] = DBL if digits > 6 else FLT
                    continue
                if typ != INT:
                    raise Unsupported(f"non-integer parameter {name}")
                self.host_params.append((name.lower(), val))
            return True
        role = "arg" if "intent" in attrs else role_default
        for name, dims in split_top(rest, with_init=False):
            table[name.lower()] = Sym(name.lower(), typ, dims, role)
        return True

    # ---- expression -> (C text, type)
    def cx(self, n, locals_=None):
        locals_ = locals_ or {}
        k = n.kind
        if k == "num":
            t = n.text
            m = re.match(r"^(.*?)(?:_([A-Za-z]\w*))?$", t)
            body, kind = m.group(1), (m.group(2) or "").lower()
            if re.match(r"^\d+$", body) and not kind:
                return body, INT
            if re.search(r"[dD]", body):
                return body.replace("d", "e").replace("D", "e"), DBL
            if kind:
                if kind not in self.real_kind:
                    raise Unsupported(f"kind suffix _{kind}")
                typ = self.real_kind[kind]
            else:
                typ = FLT
            if "." not in body and "e" not in body.lower():
                body += ".0"
            if body.endswith("."):
                body += "0"
            body = re.sub(r"\.([eE])", r".0\1", body)
            return (body + "f", FLT) if typ == FLT else (body, DBL)
        if k == "log":
            return ("1" if n.val else "0"), LOG
        if k == "paren":
            c, t = self.cx(n.a, locals_)
            return f"({c})", t
        if k == "neg":
            c, t = self.cx(n.a, locals_)
            return f"(-({c}))", t
        if k == "lbin":
            a, ta = self.cx(n.a, locals_)
            b, tb = self.cx(n.b, locals_)
            if (ta, tb) != (LOG, LOG):
                raise Unsupported(".and. / .or. on non-logicals")
            return f"({a}) {n.op} ({b})", LOG
        if k == "lnot":
            a, ta = self.cx(n.a, locals_)
            if ta != LOG:
                raise Unsupported(".not. on a non-logical")
            return f"!({a})", LOG
        if k == "rel":
            a, ta = self.cx(n.a, locals_)
            b, tb = self.cx(n.b, locals_)
            if LOG in (ta, tb):
                raise Unsupported("relational operator on logicals")
            return f"({a}) {n.op} ({b})", LOG
        if k == "bin":
            a, ta = self.cx(n.a, locals_)
            b, tb = self.cx(n.b, locals_)
            if LOG in (ta, tb):
                raise Unsupported("arithmetic on logicals")
            t = ta if RANK[ta] >= RANK[tb] else tb
            return f"{a} {n.op} {b}", t
        if k == "var":
            if n.name in locals_:
                return n.name, locals_[n.name]
            s = self.lookup(n.name)
            if s.dims:
                raise Unsupported(f"whole-array reference to {n.name} inside an expression")
            return n.name, s.typ
        if k == "call":
            name = n.name
            if name in ("max", "min"):
                parts = [self.cx(a, locals_) for a in n.args]
                types = {t for _, t in parts}
                if types == {INT}:
                    fn = f"f90_i{name}"
                    rt = INT
                elif types <= {INT, FLT}:
                    raise Unsupported(f"{name} of default reals only")  # would have to stay single precision
                else:
                    fn, rt = f"f90_d{name}", DBL
                acc = parts[0][0]
                for c, _ in parts[1:]:
                    acc = f"{fn}({acc}, {c})"
                return acc, rt
            if name == "abs":
                c, t = self.cx(n.args[0], locals_)
                if t == DBL:
                    return f"f90_dabs({c})", DBL
                raise Unsupported("abs of a non-double")
            if name in self.stmt_funcs:
                args = [self.cx(a, locals_)[0] for a in n.args]
                return f"sf_{name}({', '.join(args)})", DBL
            s = self.lookup(name)
            if not s.dims:
                raise Unsupported(f"{name} is not an array, a statement function or a supported intrinsic")
            if len(n.args) != len(s.dims):
                raise Unsupported(f"rank mismatch in reference to {name}")
            idx = []
            for a in n.args:
                if a.kind == "colon":
                    raise Unsupported("array section inside an expression")
                c, t = self.cx(a, locals_)
                if t != INT:
                    raise Unsupported("non-integer subscript")
                idx.append(c)
            return f"{name}[{self.linear(name, idx)}]", s.typ
        raise Unsupported(k)

    def linear(self, name, idx):
        s = self.lookup(name)
        e = ""
        for d in reversed(range(len(idx))):
            term = f"(({idx[d]}) - {name}_lo{d})"
            if self.check:  # like gfortran -fcheck=bounds, but the offending statement is skipped and counted instead of aborting
                term = f'f90_ck({term}, {name}_n{d}, {self.cur_line}, "{name}", {d + 1}, {name}_lo{d})'
            e = term if not e else f"{term} + (long){name}_n{d} * ({e})"
        return e

    def lookup(self, name):
        if name in self.syms:
            return self.syms[name]
        if name in self.host_arrays:
            if name not in self.used_host_arrays:
                self.used_host_arrays.append(name)
            return self.host_arrays[name]
        if name in dict(self.host_params):
            return Sym(name, INT)
        raise Unsupported(f"undeclared name {name!r}")

    # ---- statements
    def run(self):
        # pass 1: host scope (everything before `contains`) and the subroutine's own lines
        in_sub, sub_lines, host_lines = False, [], []
        seen_contains = False
        for no, st in self.lines:
            low = st.lower()
            if re.match(r"contains\b", low):
                seen_contains = True
                continue
            m = re.match(r"subroutine\s+(\w+)\s*\(([^)]*)\)", low)
            if m and m.group(1) == self.sub:
                in_sub = True
                self.args = [a.strip() for a in m.group(2).split(",") if a.strip()]
                self.first_line = no
                continue
            if in_sub and re.match(r"end\s*subroutine", low):
                in_sub = False
                self.last_line = no
                continue
            if in_sub:
                sub_lines.append((no, st))
            elif not seen_contains:
                host_lines.append((no, st))
        if not sub_lines:
            raise Unsupported(f"subroutine {self.sub} not found")
        for no, st in host_lines:
            if re.match(r"(program|implicit|call|end)\b", st, re.I):
                continue
            if not self.parse_decl(st, "host", self.host_arrays):
                raise Unsupported(f"line {no}: host-scope statement outside the subset: {st}")
        # pass 2: the subroutine
        executable = False
        out = []
        depth = 1
        for no, st in sub_lines:
            low = st.lower()
            self.cur_line = no
            try:
                if low.startswith("implicit"):
                    continue
                if not executable and self.parse_decl(st, "local", self.syms):
                    continue
                if re.match(r"call\s+system_clock\b", low) or re.match(r"write\s*\(", low):
                    continue  # timing and printing only
                m = re.match(r"do\s+(\w+)\s*=\s*(.*)$", st, re.I)
                if m:
                    executable = True
                    var = m.group(1).lower()
                    bounds = split_commas(m.group(2))
                    if len(bounds) != 2:
                        raise Unsupported("do loop with a stride")
                    lo, tl = self.cx(parse_expr(bounds[0]))
                    hi, th = self.cx(parse_expr(bounds[1]))
                    if (tl, th) != (INT, INT) or self.lookup(var).typ != INT:
                        raise Unsupported("non-integer do loop")
                    self.tmp += 1
                    out.append("  " * depth + f"{{ const int hi{self.tmp} = {hi}; for ({var} = {lo}; {var} <= hi{self.tmp}; ++{var}) {{  /* line {no} */")
                    depth += 1
                    continue
                if re.match(r"end\s*do$", low):
                    depth -= 1
                    out.append("  " * depth + "} }")
                    continue
                m = re.match(r"if\s*\((.*)\)\s*then$", st, re.I)
                if m:
                    executable = True
                    c, t = self.cx(parse_expr(m.group(1)))
                    if t != LOG:
                        raise Unsupported("if condition that is not logical")
                    out.append("  " * depth + f"if ({c}) {{  /* line {no} */")
                    depth += 1
                    continue
                if low == "else":
                    out.append("  " * (depth - 1) + "} else {")
                    continue
                guard = ""
                if re.match(r"if\s*\(", low):  # logical IF: if (cond) <one assignment>
                    close = matching_paren(st, st.index("("))
                    c, t = self.cx(parse_expr(st[st.index("(") + 1:close]))
                    if t != LOG:
                        raise Unsupported("if condition that is not logical")
                    guard, st = f"if ({c}) ", st[close + 1:].strip()
                    executable = True
                if re.match(r"end\s*if$", low):
                    depth -= 1
                    out.append("  " * depth + "}")
                    continue
                # assignment or statement function
                lhs, rhs = split_assignment(st)
                ln = parse_expr(lhs)
                if ln.kind == "call" and ln.name in self.syms and not self.syms[ln.name].dims and not executable:
                    # statement function: name(dummy, ...) = expr
                    args = [a.name for a in ln.args]
                    locs = {a: self.syms[a].typ for a in args}
                    if any(t != DBL for t in locs.values()) or self.syms[ln.name].typ != DBL:
                        raise Unsupported("statement function that is not double -> double")
                    self.stmt_funcs[ln.name] = (args, parse_expr(rhs), locs)
                    continue
                executable = True
                rc, rt = self.cx(parse_expr(rhs))
                if ln.kind == "var":
                    s = self.lookup(ln.name)
                    if s.dims:
                        raise Unsupported("whole-array assignment without a section")
                    if self.check:
                        out.append("  " * depth + f"{guard}{{ f90_bad = 0; const {CTYPE[s.typ]} v_ = {rc}; if (!f90_bad) {ln.name} = v_; }}  /* line {no} */")
                    else:
                        out.append("  " * depth + f"{guard}{ln.name} = {rc};  /* line {no} */")
                    continue
                s = self.lookup(ln.name)
                if not s.dims or len(ln.args) != len(s.dims):
                    raise Unsupported(f"assignment to {ln.name}")
                idx, opened = [], 0
                if guard:
                    out.append("  " * depth + guard.rstrip())
                    opened = 1
                for d, a in enumerate(ln.args):
                    if a.kind == "colon":
                        v = f"s{d}_"
                        out.append("  " * (depth + opened) + f"for (int {v} = {ln.name}_lo{d}; {v} < {ln.name}_lo{d} + {ln.name}_n{d}; ++{v})")
                        opened += 1
                        idx.append(v)
                    else:
                        c, t = self.cx(a)
                        if t != INT:
                            raise Unsupported("non-integer subscript")
                        idx.append(c)
                if self.check:
                    out.append("  " * (depth + opened) + f"{{ f90_bad = 0; const long ix_ = {self.linear(ln.name, idx)}; const {CTYPE[s.typ]} v_ = {rc}; "
                               f"if (!f90_bad) {ln.name}[ix_] = v_; }}  /* line {no} */")
                else:
                    out.append("  " * (depth + opened) + f"{ln.name}[{self.linear(ln.name, idx)}] = {rc};  /* line {no} */")
            except Unsupported as e:
                raise Unsupported(f"line {no}: {e}: {st}") from None
        if depth != 1:
            raise Unsupported("unbalanced do / if nesting")
        self.body = out
        return self.emit()

    def emit(self):
        sub = self.sub
        for a in self.args:
            if a.lower() not in self.syms:
                raise Unsupported(f"dummy argument {a} is not declared")
        c = []
        c.append(f"/* GENERATED by oracle/f90_to_c.py from the reference's subroutine {sub} (source lines {self.first_line}-{self.last_line}).")
        c.append("   Test infrastructure; lives under oracle/_ref/ (git-ignored), never committed.  Do not edit. */")
        c.append("#include <math.h>\n#include <stdlib.h>\n")
        c.append("static inline int f90_imax(int a, int b) { return b > a ? b : a; }")
        c.append("static inline int f90_imin(int a, int b) { return b < a ? b : a; }")
        c.append("static inline double f90_dmax(double a, double b) { return b > a ? b : a; }")
        c.append("static inline double f90_dmin(double a, double b) { return b < a ? b : a; }")
        c.append("static inline double f90_dabs(double a) { return fabs(a); }\n")
        if self.check:
            c.append("/* --check-bounds: a statement with a subscript outside its declared bounds is skipped; the first one is kept for the report */")
            c.append('static int f90_bad; static long f90_violations; static int f90_first[4]; static const char *f90_first_name = "";')
            c.append("static inline long f90_ck(long off, int n, int line, const char *name, int dim, int lo) {")
            c.append("  if (off >= 0 && off < n) return off;")
            c.append("  if (!f90_violations++) { f90_first[0] = line; f90_first[1] = dim; f90_first[2] = (int)off + lo; f90_first[3] = lo + n - 1; f90_first_name = name; }")
            c.append("  f90_bad = 1; return 0; }")
            c.append(f"long {sub}_violations(int *line, int *dim, int *subscript, int *upper, const char **name) {{")
            c.append("  *line = f90_first[0]; *dim = f90_first[1]; *subscript = f90_first[2]; *upper = f90_first[3]; *name = f90_first_name; return f90_violations; }\n")
        for name, (args, node, locs) in self.stmt_funcs.items():
            body, _ = self.cx(node, locs)
            c.append(f"static inline double sf_{name}({', '.join('double ' + a for a in args)}) {{ return {body}; }}")
        ctype = {INT: "int", FLT: "float", DBL: "double", LOG: "int"}
        params = [f"int {r}" for r in self.runtime]
        for a in self.args:
            s = self.syms[a.lower()]
            params.append(f"{ctype[s.typ]} *{s.name}" if s.dims else f"{ctype[s.typ]} *{s.name}_p")
        for h in self.used_host_arrays:
            params.append(f"const {ctype[self.host_arrays[h].typ]} *{h}")
        c.append(f"\nint {sub}({', '.join(params)})\n{{")
        if self.check:
            c.append("  f90_violations = 0;")
        names = dict(self.host_params)
        for r in self.runtime:
            if r not in names:
                raise Unsupported(f"--runtime {r} is not an integer parameter of the program")
        for name, val in self.host_params:
            if name in self.runtime:
                continue
            e, t = self.cx(parse_expr(val))
            c.append(f"  const int {name} = {e};")
        arrays = [s for s in self.syms.values() if s.dims] + [self.host_arrays[h] for h in self.used_host_arrays]
        for s in arrays:
            for d, (lo, hi) in enumerate(s.dims):
                loc = self.cx(parse_expr(lo))[0] if lo is not None else "1"
                hic = self.cx(parse_expr(hi))[0]
                c.append(f"  const int {s.name}_lo{d} = {loc}, {s.name}_n{d} = ({hic}) - ({loc}) + 1;")
        # names that are no variables of the C function: the statement functions and their dummies (unless the body uses one)
        skip = set(self.stmt_funcs)
        for args, _, _ in self.stmt_funcs.values():
            skip |= {a for a in args if not self.used_as_variable(a)}
        for s in self.syms.values():
            if s.role == "arg" or s.name in skip:
                continue
            if s.dims:
                total = " * ".join(f"(size_t){s.name}_n{d}" for d in range(len(s.dims)))
                c.append(f"  const size_t {s.name}_len = {total};")
                c.append(f"  {ctype[s.typ]} *{s.name} = ({ctype[s.typ]} *)malloc({s.name}_len * sizeof({ctype[s.typ]}));")
                c.append(f"  if (!{s.name}) return 1;")
                if s.typ in (DBL, FLT):
                    c.append(f"  for (size_t t_ = 0; t_ < {s.name}_len; ++t_) {s.name}[t_] = NAN;  /* undefined in Fortran */")
            else:
                c.append(f"  {ctype[s.typ]} {s.name};")
        c.append("")
        c.extend(self.body)
        c.append("")
        for s in self.syms.values():
            if s.role != "arg" and s.dims:
                c.append(f"  free({s.name});")
        # silence unused warnings for scalars the dropped timing statements used
        unused = [s.name for s in self.syms.values() if s.role != "arg" and not s.dims and s.name not in skip]
        c.append("  " + " ".join(f"(void){u};" for u in unused))
        c.append("  return 0;\n}")
        return "\n".join(c) + "\n"

    def used_as_variable(self, name):
        return any(re.search(rf"\b{name}\b", ln) for ln in self.body)


def matching_paren(s, start):
    depth = 0
    for i in range(start, len(s)):
        if s[i] == "(":
            depth += 1
        elif s[i] == ")":
            depth -= 1
            if depth == 0:
                return i
    raise Unsupported("unbalanced parentheses")


def split_commas(s):
    parts, depth, cur = [], 0, ""
    for ch in s:
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        parts.append(cur.strip())
    return parts


def split_top(rest, with_init):
    """Entities of a declaration: [(name, dims or init)]"""
    out = []
    for ent in split_commas(rest):
        if with_init:
            name, _, val = ent.partition("=")
            out.append((name.strip(), val.strip()))
            continue
        m = re.match(r"(\w+)\s*(?:\((.*)\))?$", ent.strip())
        if not m:
            raise Unsupported(f"declaration entity {ent!r}")
        dims = None
        if m.group(2) is not None:
            dims = []
            for d in split_commas(m.group(2)):
                if ":" in d:
                    lo, hi = d.split(":")
                    dims.append((lo.strip(), hi.strip()))
                else:
                    dims.append((None, d.strip()))
        out.append((m.group(1), dims))
    return out


def split_assignment(st):
    depth = 0
    for i, ch in enumerate(st):
        if ch == "(":
            depth += 1
        elif ch == ")":
            depth -= 1
        elif ch == "=" and depth == 0:
            return st[:i].strip(), st[i + 1:].strip()
    raise Unsupported("statement outside the subset")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("source")
    ap.add_argument("subroutine")
    ap.add_argument("--runtime", default="", help="integer parameters of the program that become C arguments")
    ap.add_argument("--check-bounds", action="store_true", help="skip and count statements whose subscripts leave the declared bounds")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    text = open(a.source, errors="replace").read()
    try:
        c = Translator(text, a.subroutine, [r for r in a.runtime.split(",") if r], a.check_bounds).run()
    except Unsupported as e:
        sys.exit(f"f90_to_c: {e}")
    open(a.out, "w").write(c)


if __name__ == "__main__":
    main()
