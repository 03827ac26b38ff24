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

Unit mode (--module M --uses U1,U2 ENTRY): every routine of module M becomes a static C function and ENTRY is exported;
enough for the three modules of atmosphere/biharmonic_wk_kernel.F90 (CPU lines 100-202, GPU1 209-362, GPU2 369-536):

    python oracle/f90_to_c.py F.F90 biharmonic_wk_scalar --module biharmonic_wk_scalar_cpu --uses kernel_utils \
           --runtime nlev,qsize --export atm_cpu_biharmonic_wk_scalar --out oracle/_ref/atm_f90_cpu.c

  * the specification parts of the used modules: integer and real parameters (a real parameter keeps the precision of its
    literal: `rrearth = 0.00000016666666666666` is float32-rounded, as in Fortran), initialised integer module variables,
    derived types (-> C structs whose array components keep the column-major order; components by `%`)
  * subroutines and functions with a `result(..)` that is an array (-> a trailing output pointer); `x = f(..)` evaluates into a
    temporary and then copies, as Fortran does, so `q(:,:,k) = f(q(:,:,k))` is safe; `call` of the module's own subroutines
  * actual arguments: whole arrays, whole components (`elem%Dinv`), contiguous leading sections (`qtens(:,:,k,q,ie)`), elements
    and whole sections of derived-type arrays (`elem(ie)`, `elem(:)`), scalars by value; assumed-shape dummies `elem(:)`
  * whole-array copies `a = b`, rank-1 section copies `a = b(i,j,:,k)`, `sum(<elementwise expression over rank-1 sections>)`
    (first to last from zero), declarations without `::`, parameters repeated inside routines (same value only)

State mode (--module M --uses U --state R1,R2 [--also OTHER.F90]): for routines that take no arguments and work on module
state -- nested_loops/nested.F90, internal subroutines run_original_cpu_directive (lines 495-564) and
run_gpu_optimized_directive (593-630) of `program nested`, on the allocatable arrays of `module nested_vars`:

    python oracle/f90_to_c.py nested.F90 run_original_cpu_directive,run_gpu_optimized_directive --also nested_vars.F90 \
           --module nested --uses nested_vars --state --export nested_f90 --out oracle/_ref/nested_f90.c

  * #ifdef / #ifndef / #else / #endif are evaluated with nothing defined (the plain CPU build); line numbers are kept
  * module scalars and allocatable arrays become file-scope C variables; array shapes come from the module's own
    `allocate(..)` statements; <export>_set_int / <export>_bind set the integers and hand over the arrays
  * `dimension(..)` attributes, routines without an argument list, the `sign` intrinsic; specification statements no routine
    needs (namelists, c_ptr globals, timers, the main program's executable part) are skipped in this mode only
  * also in the unit / state modes: `optional` dummies with present(), keyword and omitted optional arguments, `save`d
    initialised locals, functions with a scalar result, x**2, sqrt, mod, dble, product(shape(x)), whole arrays inside sum(..)
    -- what module kernel_utils of the atmosphere file needs (initialize_data + myrandom, save_qtens, compute_l2norm), so
    that the reference PROGRAM's whole sequence runs translated
This is where the translator itself is checked against COMPILED reference code: the reference has C++ forms of the same two
loops, g++ builds them, and tests/test_zz_nested.py requires the translated Fortran to equal them bit for bit.

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
def preprocess(text, defines=()):
    """#ifdef / #ifndef / #else / #endif / #define with the given set of defined names (none by default: the plain CPU build);
    dropped lines become empty so that line numbers stay the reference's."""
    defines, out, stack = set(defines), [], []  # stack of (taking this branch, some parent is off)
    for raw in text.splitlines():
        st = raw.strip()
        m = re.match(r"#\s*(ifdef|ifndef|else|endif|define|undef|include|if|elif)\b\s*(\w+)?", st)
        if not m:
            out.append(raw if all(t for t, _ in stack) else "")
            continue
        d, name = m.group(1), m.group(2)
        if d in ("ifdef", "ifndef"):
            on = (name in defines) == (d == "ifdef")
            stack.append((on, False))
        elif d == "else":
            if not stack:
                raise Unsupported("#else without #ifdef")
            on, _ = stack.pop()
            stack.append((not on, False))
        elif d == "endif":
            if not stack:
                raise Unsupported("#endif without #ifdef")
            stack.pop()
        elif d == "define":
            pass  # macros of the reference (C_WRAP) are used only inside blocks this build drops
        else:
            raise Unsupported(f"preprocessor directive #{d}")
        out.append("")
    if stack:
        raise Unsupported("unterminated #ifdef")
    return "\n".join(out)


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
    r"|(?P<id>[A-Za-z]\w*)|(?P<op>\*\*|>=|<=|==|/=|[-+*/(),:=<>%]))",
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
        node = self.primary()
        if self.peek()[1] == "**":  # binds tighter than * and unary minus, groups right to left
            self.take()
            return Node("pow", a=node, b=self.factor())
        return node

    def primary(self):
        k, v = self.take()
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
                return self.components(Node("call", name=name, args=self.arglist()))
            return self.components(Node("var", name=name))
        raise Unsupported(f"unexpected token {v!r}")

    def arglist(self):
        self.take("(")
        args = []
        if self.peek()[1] != ")":
            while True:
                if self.peek()[1] == ":":
                    self.take()
                    args.append(Node("colon"))
                elif self.peek()[0] == "id" and self.i + 1 < len(self.t) and self.t[self.i + 1][1] == "=":
                    name = self.take()[1].lower()  # keyword argument  name = value
                    self.take("=")
                    args.append(Node("kwarg", name=name, a=self.lexpr()))
                else:
                    args.append(self.expr())
                if self.peek()[1] == ",":
                    self.take()
                    continue
                break
        self.take(")")
        return args

    def components(self, node):
        while self.peek()[1] == "%":
            self.take()
            k, v = self.take()
            if k != "id":
                raise Unsupported("component name expected after %")
            args = self.arglist() if self.peek()[1] == "(" else None
            node = Node("comp", base=node, name=v.lower(), args=args)
        return node


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
        self.optional = False   # `optional` dummy: passed by address, absent = NULL
        self.save_init = None   # `integer, save :: old = 11`: a static local with that initial value


class Translator:
    def __init__(self, text, sub, runtime, check_bounds=False):
        self.check = check_bounds
        self.cur_line = 0
        # `text` may be a list of sources: the used modules' files first, the unit's own file last (its line numbers are the
        # ones the generated comments quote)
        texts = text if isinstance(text, (list, tuple)) else [text]
        self.lines = [ln for t in texts for ln in logical_lines(preprocess(t))]
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
        if k == "pow":
            a, ta = self.cx(n.a, locals_)
            if n.b.kind != "num" or n.b.text != "2" or ta != DBL:
                raise Unsupported("the ** operator with anything but a double base and the literal exponent 2")
            return f"({{ const double p_ = {a}; p_ * p_; }})", DBL  # x**2 is x*x
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
            if name == "sign" and len(n.args) == 2:
                (a, ta), (b, tb) = self.cx(n.args[0], locals_), self.cx(n.args[1], locals_)
                if (ta, tb) != (DBL, DBL):
                    raise Unsupported("sign of non-doubles")
                return f"copysign({a}, {b})", DBL  # |a| with the sign of b (the sign bit, as Fortran 95 and later)
            if name == "sqrt" and len(n.args) == 1:
                c, t = self.cx(n.args[0], locals_)
                if t != DBL:
                    raise Unsupported("sqrt of a non-double")
                return f"sqrt({c})", DBL  # correctly rounded in both languages
            if name == "mod" and len(n.args) == 2:
                (a, ta), (b, tb) = self.cx(n.args[0], locals_), self.cx(n.args[1], locals_)
                if (ta, tb) != (INT, INT):
                    raise Unsupported("mod of non-integers")
                return f"(({a}) % ({b}))", INT  # the sign follows the dividend in both languages
            if name == "dble" and len(n.args) == 1:
                c, t = self.cx(n.args[0], locals_)
                return f"((double)({c}))", DBL
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
        self.body = self.run_body(sub_lines)
        return self.emit()

    def run_body(self, sub_lines):
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
                if re.match(r"call\s+\w+\s*\(", low):
                    executable = True
                    out.append("  " * depth + self.call_statement(st[4:].strip()) + f"  /* line {no} */")
                    continue
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
                special = self.special_assignment(ln, parse_expr(rhs), guard)
                if special is not None:
                    out.append("  " * depth + special + f"  /* line {no} */")
                    continue
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
        return out

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

    # hooks of the unit (module) mode; the single-subroutine mode has neither calls nor array-valued assignments
    def call_statement(self, text):
        raise Unsupported("call statement")

    def special_assignment(self, lhs, rhs, guard):
        return None

    def used_as_variable(self, name):
        return any(re.search(rf"\b{name}\b", ln) for ln in self.body)


# ================================================================================================ unit (module) mode
class Routine:
    def __init__(self, name, kind, args, result, first):
        self.name, self.kind, self.args, self.result, self.first, self.last = name, kind, args, result, first, first
        self.lines, self.syms, self.body = [], {}, []


class UnitTranslator(Translator):
    """All routines of one module (which may `use` others) -> one C file: the routines become static functions, `entry`
    is exported.  On top of the single-subroutine subset: derived types (-> C structs, components by `%`), functions whose
    result is an array (-> a trailing output pointer; `x = f(..)` evaluates into a temporary and copies, as Fortran does),
    `call` of the module's own subroutines, actual arguments that are whole arrays, components, contiguous leading sections
    `a(:,:,k)` or elements of a derived-type array, whole-array copies `a = b`, real parameters, declarations without `::`."""

    def __init__(self, text, module, uses, entry, runtime, export=None):
        super().__init__(text, entry, runtime, False)
        self.module, self.uses, self.entry = module.lower(), [u.lower() for u in uses], entry.lower()
        self.export = export or f"{self.module}_{self.entry}"
        self.types = {}       # type name -> {component name -> Sym}
        self.routines = {}    # name -> Routine
        self.param_types = {}
        self.host_scalars = {}
        self.host_vars = []      # initialised integer module variables
        self.section_var = None  # inside sum(..) / an array assignment: the loop variable a `:` subscript stands for
        self.section_extents = []

    # ---- declarations of the unit mode
    def parse_decl(self, stmt, role_default, table):
        st = stmt
        m = re.match(r"type\s*\(\s*(\w+)\s*\)\s*((?:,\s*[^:]+?)*)\s*::\s*(.*)$", st, re.I)
        if m:
            tname, attrs = m.group(1).lower(), m.group(2).lower()
            if tname not in self.types:
                raise Unsupported(f"unknown derived type {tname}")
            role = "arg" if "intent" in attrs else role_default
            for name, dims in split_top(m.group(3), with_init=False):
                table[name.lower()] = Sym(name.lower(), ("type", tname), dims, role)
            return True
        # `integer i,j` (no double colon), `real(kind=k)`, `integer :: n = 16` (initialised module variable)
        m = re.match(r"(real|integer|logical)\s*(\([^)]*\))?\s+([A-Za-z]\w*(?:\s*,\s*[A-Za-z]\w*)*)$", st, re.I)
        if m and "::" not in st:
            st = f"{m.group(1)}{m.group(2) or ''} :: {m.group(3)}"
        st = re.sub(r"\(\s*kind\s*=\s*(\w+)\s*\)", r"(\1)", st, count=1, flags=re.I)
        m = re.match(r"(real|integer|logical)\s*(\(([^)]*)\))?\s*((?:,\s*[^:]+?)*)\s*::\s*(.*)$", st, re.I)
        if not m:
            return False
        base, kind, attrs, rest = m.group(1).lower(), (m.group(3) or "").strip().lower(), m.group(4).lower(), m.group(5)
        if "parameter" in attrs and base == "real":
            typ = self.real_kind.get(kind, FLT if kind == "" else None)
            if typ is None:
                raise Unsupported(f"real kind {kind!r}")
            for name, val in split_top(rest, with_init=True):
                self.host_params.append((name.lower(), val))
                self.param_types[name.lower()] = typ
            return True
        if role_default == "local" and "save" in attrs and "=" in rest:
            for ent in split_commas(rest):
                name, val = (x.strip() for x in ent.split("=", 1))
                if base != "integer":
                    raise Unsupported("saved local that is not an integer")
                table[name.lower()] = Sym(name.lower(), INT, None, "local")
                table[name.lower()].save_init = val
            return True
        if "optional" in attrs:
            typ = INT if base == "integer" else LOG if base == "logical" else self.real_kind.get(kind)
            if typ is None:
                raise Unsupported("optional dummy of this type")
            for name, dims in split_top(rest, with_init=False):
                if dims:
                    raise Unsupported("optional array dummy")
                table[name.lower()] = Sym(name.lower(), typ, None, "arg")
                table[name.lower()].optional = True
            return True
        if "parameter" not in attrs and "=" in rest and role_default == "host":
            for ent in split_commas(rest):  # initialised module variables (`integer :: nelemd = 16`): file-scope statics
                name, val = (x.strip() for x in ent.split("=", 1))
                if base != "integer":
                    raise Unsupported("initialised module variable that is not an integer")
                self.host_vars.append((name.lower(), val))
                self.param_types[name.lower()] = INT
            return True
        if "allocatable" in attrs or re.search(r"\(\s*:", rest):
            if role_default == "host":
                for name, dims in split_top(re.sub(r"\(([\s:,]*)\)", "", rest), with_init=False):
                    self.host_scalars[name.lower()] = True  # deferred-shape module arrays: shadowed by dummies in the routines
                return True
        before = len(self.host_params)
        ok = super().parse_decl(st, role_default, table)
        new, seen = self.host_params[before:], dict(self.host_params[:before])
        del self.host_params[before:]
        for name, val in new:  # a routine may repeat a parameter of the module (kchunk): once, and only with the same value
            if name in seen:
                if seen[name].replace(" ", "") != val.replace(" ", ""):
                    raise Unsupported(f"parameter {name} redefined with another value")
                continue
            self.host_params.append((name, val))
            self.param_types[name] = INT
        return ok

    def lookup(self, name):
        if name in self.syms:
            return self.syms[name]
        if name in self.param_types:
            return Sym(name, self.param_types[name])
        if name in self.host_scalars:
            raise Unsupported(f"module variable {name} used inside a routine")
        raise Unsupported(f"undeclared name {name!r}")

    # ---- expressions: components and whole-array operands
    def struct_base(self, node):
        """C lvalue of the derived-type object a component is selected from, and its type name."""
        if node.kind == "var":
            s = self.lookup(node.name)
            if not isinstance(s.typ, tuple):
                raise Unsupported(f"{node.name} is not of derived type")
            if s.dims:
                raise Unsupported("component of a whole derived-type array")
            return f"{node.name}[0]", s.typ[1]
        if node.kind == "call":
            s = self.lookup(node.name)
            if not isinstance(s.typ, tuple) or not s.dims or len(node.args) != 1 or node.args[0].kind == "colon":
                raise Unsupported(f"{node.name}(..) is not an element of a derived-type array")
            c, t = self.cx(node.args[0])
            return f"{node.name}[({c}) - {node.name}_lo0]", s.typ[1]
        raise Unsupported("nested derived types")

    def cx(self, n, locals_=None):
        if n.kind == "comp":
            base, tname = self.struct_base(n.base)
            comp = self.types[tname].get(n.name)
            if comp is None:
                raise Unsupported(f"type {tname} has no component {n.name}")
            if n.args is None:
                if comp.dims:
                    raise Unsupported(f"whole component array {n.name} inside an expression")
                return f"{base}.{n.name}", comp.typ
            if not comp.dims or len(n.args) != len(comp.dims):
                raise Unsupported(f"rank mismatch in component {n.name}")
            idx = self.subscripts(tname + "_" + n.name, n.args, locals_)
            return f"{base}.{n.name}[{self.linear(tname + '_' + n.name, idx)}]", comp.typ
        if n.kind == "call" and n.name in self.routines:
            raise Unsupported(f"call of {n.name} inside an expression (only `x = {n.name}(..)` is in the subset)")
        if n.kind == "call" and n.name == "present" and len(n.args) == 1 and n.args[0].kind == "var":
            sy = self.lookup(n.args[0].name)
            if not sy.optional:
                raise Unsupported("present() of a dummy that is not optional")
            return f"({sy.name} != 0)", LOG
        if n.kind == "call" and n.name == "product" and len(n.args) == 1 and n.args[0].kind == "call" and n.args[0].name == "shape":
            _, sy = self.pointer_to(n.args[0].args[0])  # product(shape(x)): the number of elements of x
            prefix = sy.name if not getattr(sy, "comp_of", None) else sy.comp_of + "_" + sy.name
            return "(" + " * ".join(f"{prefix}_n{d}" for d in range(len(sy.dims))) + ")", INT
        if n.kind == "var" and n.name in self.syms and self.syms[n.name].optional:
            return f"(*{n.name})", self.syms[n.name].typ
        if n.kind == "var" and self.section_var and (locals_ is None or n.name not in locals_):
            sy = self.lookup(n.name)
            if sy.dims and not isinstance(sy.typ, tuple):  # a whole array inside sum(..): element t_ in array-element order
                self.section_extents.append(self.total(n.name, len(sy.dims)))
                return f"{n.name}[{self.section_var}]", sy.typ
        if n.kind == "call" and n.name == "sum":
            if len(n.args) != 1 or self.section_var:
                raise Unsupported("sum with more than one argument, or nested")
            self.section_var, self.section_extents = "t_", []
            try:
                c, t = self.cx(n.args[0], locals_)
                ext = self.section_extents
            finally:
                self.section_var, self.section_extents = None, []
            if t != DBL or not ext:
                raise Unsupported("sum of something that is not a double array section")
            # first to last from zero, as gfortran expands it (for the two-element sums of the reference the order is immaterial)
            return f"({{ double s_ = 0.0; for (long t_ = 0; t_ < {ext[0]}; ++t_) s_ = s_ + ({c}); s_; }})", DBL
        if n.kind == "call" and n.name in self.syms and self.syms[n.name].dims and not isinstance(self.syms[n.name].typ, tuple) \
                and any(a.kind == "colon" for a in n.args):
            sy = self.syms[n.name]
            if len(n.args) != len(sy.dims):
                raise Unsupported(f"rank mismatch in reference to {n.name}")
            return f"{n.name}[{self.linear(n.name, self.subscripts(n.name, n.args, locals_))}]", sy.typ
        return super().cx(n, locals_)

    def subscripts(self, prefix, args, locals_=None):
        idx, colons = [], 0
        for d, a in enumerate(args):
            if a.kind == "colon":
                colons += 1
                if not self.section_var or colons > 1:
                    raise Unsupported("array section inside an expression")
                idx.append(f"{prefix}_lo{d} + {self.section_var}")
                self.section_extents.append(f"{prefix}_n{d}")
            else:
                c, t = self.cx(a, locals_)
                if t != INT:
                    raise Unsupported("non-integer subscript")
                idx.append(c)
        return idx

    def pointer_to(self, node):
        """C pointer to the first element of an array-valued operand (whole array, whole component, leading section)."""
        if node.kind == "var":
            s = self.lookup(node.name)
            if not s.dims:
                raise Unsupported(f"{node.name} is not an array")
            return node.name, s
        if node.kind == "comp" and node.args is None:
            base, tname = self.struct_base(node.base)
            comp = self.types[tname][node.name]
            if not comp.dims:
                raise Unsupported(f"component {node.name} is not an array")
            comp.comp_of = tname
            return f"{base}.{node.name}", comp
        if node.kind == "call":
            s = self.lookup(node.name)
            if not s.dims or len(node.args) != len(s.dims):
                raise Unsupported(f"{node.name}(..) is not an array section")
            idx, colons = [], 0
            for d, a in enumerate(node.args):
                if a.kind == "colon":
                    if d != colons:
                        raise Unsupported("a section whose colons are not the leading subscripts is not contiguous")
                    colons += 1
                    idx.append(f"{node.name}_lo{d}")
                else:
                    c, t = self.cx(a)
                    if t != INT:
                        raise Unsupported("non-integer subscript")
                    idx.append(c)
            if isinstance(s.typ, tuple):
                if colons == 1 and len(node.args) == 1:
                    return node.name, s  # elem(:)
                if colons:
                    raise Unsupported("section of a derived-type array")
                return f"&{node.name}[({idx[0]}) - {node.name}_lo0]", Sym(node.name, s.typ)
            if not colons:
                raise Unsupported("an array element passed where an array is expected")
            return f"&{node.name}[{self.linear(node.name, idx)}]", Sym(node.name, s.typ, s.dims[:colons])
        raise Unsupported("operand that is not an array")

    def actual(self, node, dummy):
        if dummy.dims or isinstance(dummy.typ, tuple):
            c, s = self.pointer_to(node) if (dummy.dims or node.kind == "call") else (None, None)
            if c is None:  # scalar of derived type passed on
                s = self.lookup(node.name) if node.kind == "var" else None
                if s is None or s.typ != dummy.typ or s.dims:
                    raise Unsupported("actual argument for a derived-type dummy")
                return node.name
            if s.typ != dummy.typ:
                raise Unsupported("type mismatch between actual and dummy argument")
            return c
        c, t = self.cx(node)
        if (dummy.typ == INT) != (t == INT):
            raise Unsupported("integer / real mismatch between actual and dummy argument")
        return c

    def actuals(self, r, args):
        given, pos = {}, 0
        for a in args:
            if a.kind == "kwarg":
                if a.name not in r.args:
                    raise Unsupported(f"{r.name} has no dummy {a.name}")
                given[a.name] = a.a
            else:
                if pos >= len(r.args):
                    raise Unsupported(f"{r.name} takes {len(r.args)} arguments")
                given[r.args[pos]] = a
                pos += 1
        out = []
        for d in r.args:
            sy = r.syms[d]
            if d not in given:
                if not sy.optional:
                    raise Unsupported(f"{r.name}: argument {d} is missing")
                out.append("0")  # absent optional argument
            elif sy.optional:
                c, t = self.cx(given[d])
                if (t == LOG) != (sy.typ == LOG):
                    raise Unsupported("type mismatch for an optional argument")
                out.append(f"&({CTYPE[sy.typ]}){{{c}}}")  # by address, as Fortran passes it
            else:
                out.append(self.actual(given[d], sy))
        return out

    def call_statement(self, text):
        n = parse_expr(text)
        if n.kind != "call" or n.name not in self.routines or self.routines[n.name].kind != "subroutine":
            raise Unsupported("call of something that is not a subroutine of this module")
        return f"r_{n.name}({', '.join(self.actuals(self.routines[n.name], n.args))});"

    def total(self, prefix, ndims):
        return " * ".join(f"(long){prefix}_n{d}" for d in range(ndims))

    def special_assignment(self, lhs, rhs, guard):
        if rhs.kind == "call" and rhs.name in self.routines and self.routines[rhs.name].kind == "function":
            r = self.routines[rhs.name]
            res = r.syms[r.result]
            dst, _ = self.pointer_to(lhs)
            saved, self.syms = self.syms, {}
            try:  # the result's extents are module constants
                size = " * ".join("((" + self.cx(parse_expr(hi))[0] + ") - (" + (self.cx(parse_expr(lo))[0] if lo else "1") + ") + 1)" for lo, hi in res.dims)
            finally:
                self.syms = saved
            args = ", ".join(self.actuals(r, rhs.args) + ["t_"])
            return (f"{guard}{{ double t_[{size}]; r_{r.name}({args}); double *d_ = {dst}; "
                    f"for (long q_ = 0; q_ < {size}; ++q_) d_[q_] = t_[q_]; }}")
        def has_colon(n):
            if n is None or not isinstance(n, Node):
                return False
            if n.kind in ("call", "comp") and n.args and any(a.kind == "colon" for a in n.args):
                return True
            return any(has_colon(getattr(n, f, None)) for f in ("a", "b", "base")) or \
                (n.kind == "call" and n.name != "sum" and any(has_colon(a) for a in n.args))
        if lhs.kind == "var" and rhs.kind != "var" and has_colon(rhs) and not (rhs.kind == "call" and rhs.name in self.routines):
            a = self.lookup(lhs.name)
            if not a.dims or len(a.dims) != 1 or isinstance(a.typ, tuple):
                raise Unsupported("section assigned to something that is not a rank-1 array")
            def mentions(n, name):
                if not isinstance(n, Node):
                    return False
                if getattr(n, "name", None) == name:
                    return True
                return any(mentions(getattr(n, f, None), name) for f in ("a", "b", "base")) or any(mentions(x, name) for x in (getattr(n, "args", None) or []))
            if mentions(rhs, lhs.name):
                raise Unsupported("array assignment whose right side reads the left side")
            self.section_var, self.section_extents = "t_", []
            try:
                c, t = self.cx(rhs)
                ext = self.section_extents
            finally:
                self.section_var, self.section_extents = None, []
            return f"{guard}for (int t_ = 0; t_ < {ext[0]}; ++t_) {lhs.name}[t_] = {c};"
        if lhs.kind == "var" and rhs.kind == "var":
            a, b = self.lookup(lhs.name), self.lookup(rhs.name)
            if a.dims and b.dims:
                if len(a.dims) != len(b.dims) or a.typ != b.typ:
                    raise Unsupported("whole-array copy between different shapes")
                return f"{guard}for (long q_ = 0; q_ < {self.total(lhs.name, len(a.dims))}; ++q_) {lhs.name}[q_] = {rhs.name}[q_];"
        return None

    # ---- the unit
    def run(self):
        # split the file into units; keep the specification parts of the used modules and of the target, and its routines
        unit, in_contains, spec, cur = None, False, [], None
        for no, st in self.lines:
            low = st.lower()
            m = re.match(r"(module|program)\s+(\w+)$", low)
            if m and unit is None:
                unit, in_contains = m.group(2), False
                continue
            if re.match(r"end\s*(module|program)\b", low):
                unit, cur = None, None
                continue
            if unit is None:
                continue
            wanted = unit in self.uses or unit == self.module
            if low == "contains":
                in_contains = True
                continue
            if not in_contains:
                if wanted:
                    spec.append((no, st))
                continue
            if unit != self.module:
                continue
            m = re.match(r"(subroutine|function)\s+(\w+)\s*(?:\(([^)]*)\))?\s*(?:result\s*\(\s*(\w+)\s*\))?$", low)
            if m and cur is None:
                cur = Routine(m.group(2), m.group(1), [a.strip() for a in (m.group(3) or "").split(",") if a.strip()], m.group(4), no)
                if cur.kind == "function" and not cur.result:
                    raise Unsupported(f"line {no}: function without a result clause")
                self.routines[cur.name] = cur
                continue
            if re.match(r"end\s*(subroutine|function)\b", low):
                cur.last = no
                cur = None
                continue
            if cur is None:
                raise Unsupported(f"line {no}: statement between routines: {st}")
            cur.lines.append((no, st))
        roots = getattr(self, "entries", [self.entry])
        for e in roots:
            if e not in self.routines:
                raise Unsupported(f"{e} is not a routine of module {self.module}")
        # only what the entries reach is translated (the module may hold routines outside the subset that nothing here calls)
        reach, todo = set(), list(roots)
        while todo:
            name = todo.pop()
            if name in reach:
                continue
            reach.add(name)
            words = set(re.findall(r"[a-z]\w*", " ".join(st.lower() for _, st in self.routines[name].lines)))
            todo.extend(w for w in words if w in self.routines and w not in reach)
        self.routines = {k: v for k, v in self.routines.items() if k in reach}
        # specification parts
        in_type = None
        for no, st in spec:
            low = st.lower()
            try:
                if re.match(r"(implicit|use|private|public)\b", low):
                    continue
                m = re.match(r"type\s*(?:::)?\s*(\w+)$", low)
                if m and not low.startswith("type("):
                    in_type = m.group(1)
                    self.types[in_type] = {}
                    continue
                if re.match(r"end\s*type\b", low):
                    in_type = None
                    continue
                if in_type:
                    if not self.parse_decl(st, "component", self.types[in_type]):
                        raise Unsupported("component declaration")
                    continue
                if not self.parse_decl(st, "host", self.host_arrays):
                    raise Unsupported("specification statement outside the subset")
            except Unsupported as e:
                if not self.spec_unknown(no, st, e):
                    raise Unsupported(f"line {no}: {e}: {st}") from None
        self.after_spec()
        # declarations of every routine first (calls need the callee's dummies), then the bodies
        for r in self.routines.values():
            decl_end = 0
            for i, (no, st) in enumerate(r.lines):
                if st.lower().startswith("implicit"):
                    decl_end = i + 1
                    continue
                try:
                    if self.parse_decl(st, "local", r.syms):
                        decl_end = i + 1
                        continue
                except Unsupported as e:
                    raise Unsupported(f"line {no}: {e}: {st}") from None
                break
            r.exec_lines = r.lines[decl_end:]
            for a in r.args + ([r.result] if r.result else []):
                if a not in r.syms:
                    raise Unsupported(f"{r.name}: dummy argument {a} is not declared")
                r.syms[a].role = "arg"
        for r in self.routines.values():
            self.syms, self.stmt_funcs = r.syms, {}
            r.body = self.translate_body(r.exec_lines)
        return self.emit_unit()

    def spec_unknown(self, no, st, err):
        """A specification statement outside the subset: an error (the state mode skips what no routine needs)."""
        return False

    def after_spec(self):
        pass

    def translate_body(self, lines):
        """The statement loop of the base class on a routine's executable part."""
        return self.run_body(lines)

    def dims_consts(self, prefix, s, c, indent="  ", const="const int"):
        for d, (lo, hi) in enumerate(s.dims):
            loc = self.cx(parse_expr(lo))[0] if lo else "1"
            if hi in (None, ""):
                c.append(f"{indent}{const} {prefix}_lo{d} = {loc};")
            else:
                c.append(f"{indent}{const} {prefix}_lo{d} = {loc}, {prefix}_n{d} = ({self.cx(parse_expr(hi))[0]}) - ({loc}) + 1;")

    def ctype_of(self, s):
        return s.typ[1] if isinstance(s.typ, tuple) else CTYPE[s.typ]

    def signature(self, r):
        ps = []
        scalar_result = bool(r.result) and not r.syms[r.result].dims
        for a in r.args + ([r.result] if r.result and not scalar_result else []):
            s = r.syms[a]
            ps.append(f"{self.ctype_of(s)} *{a}" if (s.dims or isinstance(s.typ, tuple) or s.optional) else f"{CTYPE[s.typ]} {a}")
        ret = CTYPE[r.syms[r.result].typ] if scalar_result else "void"
        return f"{ret} r_{r.name}({', '.join(ps) or 'void'})"

    def emit_unit(self):
        c = [f"/* GENERATED by oracle/f90_to_c.py from the reference's module {self.module} (routines "
             + ", ".join(f"{r.name} {r.first}-{r.last}" for r in self.routines.values()) + ").",
             "   Test infrastructure; lives under oracle/_ref/ (git-ignored), never committed.  Do not edit. */",
             "#include <math.h>\n#include <stdlib.h>\n",
             "static inline int f90_imax(int a, int b) { return b > a ? b : a; }",
             "static inline int f90_imin(int a, int b) { return b < a ? b : a; }",
             "static inline double f90_dmax(double a, double b) { return b > a ? b : a; }",
             "static inline double f90_dmin(double a, double b) { return b < a ? b : a; }",
             "static inline double f90_dabs(double a) { return fabs(a); }\n"]
        self.syms = {}
        for name, val in self.host_params:
            typ = self.param_types[name]
            if name in self.runtime:
                if typ != INT:
                    raise Unsupported(f"--runtime {name} is not an integer parameter")
                c.append(f"static int {name};  /* compile-time parameter of the reference, a run-time argument here */")
            elif typ == INT:
                c.append(f"enum {{ {name} = {self.cx(parse_expr(val))[0]} }};")
            else:
                c.append(f"static const {CTYPE[typ]} {name} = {self.cx(parse_expr(val))[0]};")
        for name, val in self.host_vars:
            c.append(f"static int {name} = {self.cx(parse_expr(val))[0]};  /* module variable of the reference" + (", set from an argument here */" if name in self.runtime else " */"))
        for r in self.runtime:
            if r not in self.param_types:
                raise Unsupported(f"--runtime {r} is not a parameter of the used modules")
        for tname, comps in self.types.items():
            fields = []
            for cname, s in comps.items():
                if s.dims:
                    self.dims_consts(f"{tname}_{cname}", s, c, indent="", const="static const int")
                    fields.append(f"{CTYPE[s.typ]} {cname}[{' * '.join('((' + self.cx(parse_expr(hi))[0] + ') - (' + (self.cx(parse_expr(lo))[0] if lo else '1') + ') + 1)' for lo, hi in s.dims)}];")
                else:
                    fields.append(f"{CTYPE[s.typ]} {cname};")
            c.append(f"typedef struct {{ {' '.join(fields)} }} {tname};")
        c.append("")
        self.emit_state(c)
        for r in self.routines.values():
            c.append(f"static {self.signature(r)};")
        for r in self.routines.values():
            self.syms = r.syms
            c.append(f"\n/* {r.kind} {r.name}, source lines {r.first}-{r.last} */\nstatic {self.signature(r)}\n{{")
            for s in r.syms.values():
                if s.dims:
                    self.dims_consts(s.name, s, c)
            for s in r.syms.values():
                if s.role == "arg" and not (r.result == s.name and not s.dims):
                    continue
                if s.save_init is not None:
                    c.append(f"  static {CTYPE[s.typ]} {s.name} = {self.cx(parse_expr(s.save_init))[0]};  /* save */")
                    continue
                if s.dims:
                    c.append(f"  {self.ctype_of(s)} {s.name}[{self.total(s.name, len(s.dims))}];")
                    if s.typ in (DBL, FLT):
                        c.append(f"  for (long t_ = 0; t_ < {self.total(s.name, len(s.dims))}; ++t_) {s.name}[t_] = NAN;  /* undefined in Fortran */")
                else:
                    c.append(f"  {self.ctype_of(s)} {s.name};")
            c.append("")
            c.extend(r.body)
            unused = [s.name for s in r.syms.values() if not (s.role == "arg" and (s.dims or isinstance(s.typ, tuple)))]
            c.append("  " + " ".join(f"(void){u};" for u in unused))
            dimvars = [f"(void){s.name}_lo{d};" + (f" (void){s.name}_n{d};" if hi not in (None, "") else "") for s in r.syms.values() if s.dims for d, (lo, hi) in enumerate(s.dims)]
            c.append("  " + " ".join(dimvars))
            if r.result and not r.syms[r.result].dims:
                c.append(f"  return {r.result};")
            c.append("}")
        self.emit_entry(c)
        return "\n".join(c) + "\n"

    def emit_state(self, c):
        pass

    def emit_entry(self, c):
        e = self.routines[self.entry]
        ps = [f"int {r}_" for r in self.runtime]
        call = []
        for a in e.args:
            s = e.syms[a]
            ps.append(f"{self.ctype_of(s)} *{a}" if (s.dims or isinstance(s.typ, tuple)) else f"{CTYPE[s.typ]} {a}")
            call.append(a)
        c.append(f"\nint {self.export}({', '.join(ps)})\n{{")
        for r in self.runtime:
            c.append(f"  {r} = {r}_;")
        c.append(f"  r_{e.name}({', '.join(call)});\n  return 0;\n}}")


class StateUnitTranslator(UnitTranslator):
    """Unit mode for routines that work on MODULE STATE instead of dummy arguments (nested_loops/nested.F90: the internal
    subroutines of `program nested` take no arguments and use the allocatable arrays and scalars of `module nested_vars`).
    The module variables become file-scope C variables, as they are file-scope Fortran variables; the shapes of the allocatable
    arrays are read from the module's own `allocate(..)` statements; the caller sets the integers with <export>_set_int and
    hands over the arrays with <export>_bind, then calls <export>_<routine>().  Specification statements no routine needs
    (namelists, c_ptr globals, timers, the main program's executable part) are skipped."""

    def __init__(self, text, module, uses, entries, export):
        super().__init__(text, module, uses, entries[0], [], export)
        self.entries = [e.lower() for e in entries]
        self.state_scalars = {}  # name -> type
        self.state_init = {}     # name -> initial value of `integer :: nelemd = 16`
        self.state_arrays = {}   # name -> Sym (dims None until an allocate statement gives them)

    def parse_decl(self, stmt, role_default, table):
        if role_default != "host":
            return super().parse_decl(stmt, role_default, table)
        m = re.match(r"type\s*\(\s*(\w+)\s*\)\s*((?:,\s*[^:]+?)*)\s*::\s*(.*)$", stmt, re.I)
        if m and m.group(1).lower() in self.types:  # a module variable of derived type: bound by the caller, like the arrays
            for name, dims in split_top(re.sub(r"\(([\s:,]*)\)", lambda g: "(" + g.group(1) + ")", m.group(3)), with_init=False):
                sy = Sym(name.lower(), ("type", m.group(1).lower()), None, "state")
                if dims:
                    sy.rank = len(dims)
                else:
                    sy.dims = []  # a scalar: nothing to allocate
                    sy.scalar_struct = True
                self.state_arrays[name.lower()] = sy
            return True
        st = re.sub(r"\(\s*kind\s*=\s*(\w+)\s*\)", r"(\1)", stmt, count=1, flags=re.I)
        m = re.match(r"(real|integer|logical)\s*(\(([^)]*)\))?\s*((?:,\s*(?:dimension\s*\([^)]*\)|[^:,]+))*)\s*::\s*(.*)$", st, re.I)
        if not m:
            return False
        base, kind, attrs, rest = m.group(1).lower(), (m.group(3) or "").strip().lower(), m.group(4).lower(), m.group(5)
        if "parameter" in attrs:
            return super().parse_decl(stmt, role_default, table)
        if base == "real":
            typ = self.real_kind.get(kind, FLT if kind == "" else None)
            if typ is None:
                raise Unsupported(f"real kind {kind!r}")
        else:
            typ = INT if base == "integer" else LOG
        dm = re.search(r"dimension\s*\(([^)]*)\)", attrs)
        for ent in split_commas(rest):
            em = re.match(r"(\w+)\s*(?:\((.*)\))?\s*(?:=\s*(.*))?$", ent.strip())
            if not em:
                raise Unsupported(f"declaration entity {ent!r}")
            name, edims, init = em.group(1).lower(), em.group(2), em.group(3)
            dims = edims if edims is not None else (dm.group(1) if dm else None)
            if init is not None:
                if dims is not None or typ != INT:
                    raise Unsupported("initialised module variable that is not an integer scalar")
                self.state_init[name] = init
            if dims is None:
                self.state_scalars[name] = typ
            elif ":" in dims:
                self.state_arrays[name] = Sym(name, typ, None, "state")
                self.state_arrays[name].rank = len(split_commas(dims))
            else:
                self.state_arrays[name] = Sym(name, typ, [(None, d.strip()) for d in split_commas(dims)], "state")
        return True

    def spec_unknown(self, no, st, err):
        return True  # namelist, type(c_ptr) globals, timers, the main program's executable statements

    def after_spec(self):
        for no, st in self.lines:  # the shapes: every allocate(..) of the used modules and of the unit
            m = re.match(r"allocate\s*\((.*)\)$", st, re.I)
            if not m:
                continue
            for ent in split_commas(m.group(1)):
                em = re.match(r"(\w+)\s*\((.*)\)$", ent.strip())
                if not em or em.group(1).lower() not in self.state_arrays:
                    continue
                a = self.state_arrays[em.group(1).lower()]
                dims = [(None, d.strip()) for d in split_commas(em.group(2))]
                if getattr(a, "rank", len(dims)) != len(dims):
                    raise Unsupported(f"line {no}: allocate({a.name}) with a rank other than declared")
                a.dims = dims

    def lookup(self, name):
        if name in self.syms:
            return self.syms[name]
        if name in self.param_types:
            return Sym(name, self.param_types[name])
        if name in self.state_scalars:
            return Sym(name, self.state_scalars[name])
        if name in self.state_arrays:
            a = self.state_arrays[name]
            if a.dims is None:
                raise Unsupported(f"no allocate statement gives the shape of {name}")
            if getattr(a, "scalar_struct", False):
                return Sym(name, a.typ)
            return a
        raise Unsupported(f"undeclared name {name!r}")

    def run(self):
        for e in self.entries:
            self.entry = e
        return super().run()

    def emit_state(self, c):
        c.append("/* module state of the reference: scalars and allocatable arrays, file scope there and here */")
        self.syms = {}
        for name, typ in self.state_scalars.items():
            init = f" = {self.cx(parse_expr(self.state_init[name]))[0]}" if name in self.state_init else ""
            c.append(f"static {CTYPE[typ]} {name}{init};")
        used = [a for a in self.state_arrays.values() if a.dims is not None]
        for a in used:
            c.append(f"static {self.ctype_of(a)} *{a.name};")
            if a.dims:
                c.append("static int " + ", ".join(f"{a.name}_lo{d} = 1, {a.name}_n{d}" for d in range(len(a.dims))) + ";")
        self.syms = {}
        c.append("static void f90_shapes(void)\n{  /* the extents the module's allocate statements give */")
        for a in used:
            for d, (_, hi) in enumerate(a.dims):
                c.append(f"  {a.name}_n{d} = {self.cx(parse_expr(hi))[0]};")
        c.append("}\n")
        c.append("#include <string.h>")
        c.append(f"int {self.export}_set_int(const char *name, int v)\n{{")
        for name, typ in self.state_scalars.items():
            if typ == INT:
                c.append(f'  if (!strcmp(name, "{name}")) {{ {name} = v; return 0; }}')
        c.append("  return 1;\n}")
        c.append(f"int {self.export}_bind(const char *name, void *p)\n{{")
        for a in used:
            c.append(f'  if (!strcmp(name, "{a.name}")) {{ {a.name} = ({self.ctype_of(a)} *)p; return 0; }}')
        c.append("  return 1;\n}\n")

    def emit_entry(self, c):
        for e in self.entries:
            if e not in self.routines or self.routines[e].args:
                raise Unsupported(f"{e} is not an argument-less routine of {self.module}")
            r = self.routines[e]
            if r.result:  # a scalar function: its value is the entry's value
                if r.syms[r.result].dims or r.syms[r.result].typ != DBL:
                    raise Unsupported("an entry function whose result is not a double scalar")
                c.append(f"\ndouble {self.export}_{e}(void)\n{{\n  f90_shapes();\n  return r_{e}();\n}}")
            else:
                c.append(f"\nint {self.export}_{e}(void)\n{{\n  f90_shapes();\n  r_{e}();\n  return 0;\n}}")


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
    ap.add_argument("subroutine", help="the subroutine to translate (single-subroutine mode) or the entry routine (--module)")
    ap.add_argument("--module", help="unit mode: translate every routine of this module, export the entry")
    ap.add_argument("--uses", default="", help="unit mode: modules whose specification part the module uses")
    ap.add_argument("--export", help="unit mode: name of the exported C function")
    ap.add_argument("--also", default="", help="unit mode: further source files (comma list) holding the used modules")
    ap.add_argument("--state", action="store_true", help="unit mode on module state: SUBROUTINE is a comma list of argument-less routines")
    ap.add_argument("--runtime", default="", help="integer parameters of the program that become C arguments")
    ap.add_argument("--check-bounds", action="store_true", help="skip and count statements whose subscripts leave the declared bounds")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    text = open(a.source, errors="replace").read()
    if a.also:
        text = [open(f, errors="replace").read() for f in a.also.split(",") if f] + [text]
    try:
        runtime = [r for r in a.runtime.split(",") if r]
        if a.module and a.state:
            c = StateUnitTranslator(text, a.module, [u for u in a.uses.split(",") if u], a.subroutine.split(","), a.export or a.module).run()
        elif a.module:
            c = UnitTranslator(text, a.module, [u for u in a.uses.split(",") if u], a.subroutine, runtime, a.export).run()
        else:
            c = Translator(text, a.subroutine, runtime, a.check_bounds).run()
    except Unsupported as e:
        sys.exit(f"f90_to_c: {e}")
    open(a.out, "w").write(c)


if __name__ == "__main__":
    main()
