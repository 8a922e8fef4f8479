"""A small interpreter for the subset of R that TCLB model descriptions are written in.

The reference generates every model with R (``tools/RT`` -> rtemplate, ``src/conf.R`` sources the model's
``Dynamics.R``); there is no R on the target machines, so the front-end evaluates ``Dynamics.R`` (and the R chunks of
``Dynamics.c.Rt``, see :mod:`tclb_b200.gen.rtemplate`) itself.  Supported: numbers/strings/logicals/NULL/NA, vectors with
recycling and 1-based/negative/logical/name indexing, matrices, lists, data frames (``$``, ``[i,j]``, ``rbind``),
``function`` closures, ``if/else``, ``for``, ``while``, ``repeat``/``break``/``next``, the arithmetic/comparison/logical
operators, ``%in% %*% %% %/% %o%``, ``:`` ranges, formulas are NOT supported.  Builtins cover what the 61 shipped models
use (c, paste, rep, seq, expand.grid, matrix, solve, chol, diag, data.frame, merge-free helpers ...).
"""
from __future__ import annotations

import math
import re
from collections import OrderedDict
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np


class RError(Exception):
    pass


class _Null:
    def __repr__(self):
        return "NULL"

    def __bool__(self):
        return False

    def __len__(self):
        return 0


NULL = _Null()
NA = None


# ------------------------------------------------------------------------------------------ values
class RList(OrderedDict):
    """R list: ordered, optionally named (unnamed entries get keys '\x00<i>')."""

    def names(self):
        return [k if not k.startswith("\x00") else "" for k in self.keys()]

    @staticmethod
    def make(values, names=None):
        r = RList()
        for i, v in enumerate(values):
            n = names[i] if names and names[i] else f"\x00{i}"
            r[n] = v
        return r

    def values_list(self):
        return list(self.values())


class DataFrame:
    def __init__(self, cols: "OrderedDict[str, list]" = None):
        self.cols: "OrderedDict[str, list]" = OrderedDict()
        if cols:
            n = max((len(v) if isinstance(v, (list, np.ndarray)) else 1) for v in cols.values()) if cols else 0
            for k, v in cols.items():
                v = list(v) if isinstance(v, (list, tuple, np.ndarray)) else [v]
                if len(v) != n and len(v) > 0:
                    v = [v[i % len(v)] for i in range(n)]
                self.cols[k] = v

    def nrow(self):
        return len(next(iter(self.cols.values()))) if self.cols else 0

    def ncol(self):
        return len(self.cols)

    def row(self, i) -> RList:
        r = RList()
        for k, v in self.cols.items():
            c = v[i]
            # a cell of a list column (I(list(...)), closures, symbolic values) is handed out as it is
            r[k] = c if isinstance(c, (np.ndarray, RList, DataFrame)) or callable(c) or hasattr(c, "__rcell__") or (c is not None and not isinstance(c, (str, int, float, bool, np.generic))) else vec([c])
        return r

    def subset_rows(self, idx: List[int]) -> "DataFrame":
        d = DataFrame()
        for k, v in self.cols.items():
            d.cols[k] = [v[i] for i in idx]
        rn = getattr(self, "rownames", None)
        if rn:
            d.rownames = [rn[i] for i in idx]
        return d

    def rbind(self, other: "DataFrame") -> "DataFrame":
        if not self.cols:
            return other
        d = DataFrame()
        for k in self.cols:
            d.cols[k] = list(self.cols[k]) + list(other.cols.get(k, [NA] * other.nrow()))
        for k in other.cols:
            if k not in d.cols:
                d.cols[k] = [NA] * self.nrow() + list(other.cols[k])
        return d

    def __repr__(self):
        return f"DataFrame({dict(self.cols)})"


def vec(x) -> np.ndarray:
    """coerce to an R atomic vector (1-d numpy array; strings/NA -> object dtype)"""
    if isinstance(x, np.ndarray):
        return x
    if x is NULL:
        return np.array([], dtype=float)
    if isinstance(x, (list, tuple)):
        if any(isinstance(v, str) for v in x) or any(v is None for v in x):
            return np.array(list(x), dtype=object)
        if all(isinstance(v, (bool, np.bool_)) for v in x) and len(x) > 0:
            return np.array(x, dtype=bool)
        return np.array(x, dtype=float) if any(isinstance(v, float) for v in x) or not x else np.array(x, dtype=np.int64)
    if isinstance(x, str) or x is None:
        return np.array([x], dtype=object)
    if isinstance(x, (bool, np.bool_)):
        return np.array([bool(x)], dtype=bool)
    if isinstance(x, (int, np.integer)):
        return np.array([int(x)], dtype=np.int64)
    if isinstance(x, (float, np.floating)):
        return np.array([float(x)], dtype=float)
    return x


def is_vec(x):
    return isinstance(x, np.ndarray)


def is_str_vec(x):
    return isinstance(x, np.ndarray) and x.dtype == object


def truthy(x) -> bool:
    if isinstance(x, np.ndarray):
        if x.size == 0:
            raise RError("argument is of length zero")
        v = x.flat[0]
        if v is None:
            raise RError("missing value where TRUE/FALSE needed")
        return bool(v)
    return bool(x)


def scalar(x):
    if isinstance(x, np.ndarray):
        if x.size == 0:
            return NULL
        v = x.flat[0]
        return v.item() if hasattr(v, "item") else v
    return x


def to_str(v) -> str:
    if v is None:
        return "NA"
    if isinstance(v, (bool, np.bool_)):
        return "TRUE" if v else "FALSE"
    if isinstance(v, (float, np.floating)):
        if float(v).is_integer() and abs(v) < 1e15:
            return str(int(v))
        return repr(float(v)) if len(repr(float(v))) < 17 else ("%.15g" % v)
    return str(v)


# ------------------------------------------------------------------------------------------ lexer
_TOKEN = re.compile(r"""
    (?P<ws>[ \t\r]+) | (?P<comment>\#[^\n]*) | (?P<nl>\n) |
    (?P<num>(?:\d+\.\d*|\.\d+|\d+)(?:[eE][+-]?\d+)?L?|0x[0-9a-fA-F]+L?) |
    (?P<str>"(?:[^"\\]|\\.)*"|'(?:[^'\\]|\\.)*') |
    (?P<op>%[^%\n]*%|<<-|<-|->|<=|>=|==|!=|&&|\|\||\*\*|\[\[|::|[-+*/^<>=!&|~?:$@\[\](){},;]) |
    (?P<id>(?:[A-Za-z.][A-Za-z0-9._]*)|`[^`]+`)
""", re.X)


def tokenize(src: str):
    toks = []
    pos = 0
    line = 1
    while pos < len(src):
        m = _TOKEN.match(src, pos)
        if not m:
            raise RError(f"R lexer: unexpected character {src[pos]!r} at line {line}")
        pos = m.end()
        k = m.lastgroup
        if k in ("ws", "comment"):
            continue
        if k == "nl":
            toks.append(("nl", "\n", line))
            line += 1
            continue
        toks.append((k, m.group(k), line))
    toks.append(("eof", "", line))
    return toks


def _unescape(s: str) -> str:
    body = s[1:-1]
    return body.encode("utf-8").decode("unicode_escape") if "\\" in body else body


# ------------------------------------------------------------------------------------------ parser (Pratt)
class Parser:
    def __init__(self, src: str):
        self.t = tokenize(src)
        self.i = 0
        self.depth = 0     # inside () or [] newlines are insignificant

    def peek(self, skip_nl=False):
        j = self.i
        if skip_nl or self.depth > 0:
            while self.t[j][0] == "nl":
                j += 1
        return self.t[j]

    def next(self, skip_nl=False):
        if skip_nl or self.depth > 0:
            while self.t[self.i][0] == "nl":
                self.i += 1
        tok = self.t[self.i]
        self.i += 1
        return tok

    def skip_newlines(self):
        while self.t[self.i][0] == "nl" or (self.t[self.i][0] == "op" and self.t[self.i][1] == ";"):
            self.i += 1

    def expect(self, val):
        tok = self.next(skip_nl=True)
        if tok[1] != val:
            raise RError(f"R parser: expected {val!r} but found {tok[1]!r} at line {tok[2]}")
        return tok

    def parse_program(self):
        out = []
        self.skip_newlines()
        while self.peek()[0] != "eof":
            out.append(self.parse_expr(0))
            self.skip_newlines()
        return ("block", out)

    BINOPS = {
        "?": (1, "L"), "=": (3, "R"), "<-": (5, "R"), "<<-": (5, "R"), "->": (6, "L"), "~": (8, "L"),
        "||": (10, "L"), "|": (10, "L"), "&&": (12, "L"), "&": (12, "L"),
        "==": (16, "L"), "!=": (16, "L"), "<": (16, "L"), ">": (16, "L"), "<=": (16, "L"), ">=": (16, "L"),
        "+": (20, "L"), "-": (20, "L"), "*": (22, "L"), "/": (22, "L"), ":": (26, "L"), "^": (30, "R"), "**": (30, "R"),
    }

    def binop_info(self, tok):
        if tok[0] != "op":
            return None
        v = tok[1]
        if v.startswith("%") and v.endswith("%") and len(v) >= 2:
            return (24, "L")
        return self.BINOPS.get(v)

    def parse_expr(self, minbp):
        left = self.parse_unary()
        while True:
            tok = self.peek()
            info = self.binop_info(tok)
            if info is None:
                break
            bp, assoc = info
            if bp < minbp:
                break
            self.next()
            # a newline after a binary operator continues the expression
            while self.t[self.i][0] == "nl":
                self.i += 1
            right = self.parse_expr(bp + (0 if assoc == "R" else 1))
            op = tok[1]
            if op in ("=", "<-", "<<-"):
                left = ("assign", left, right, op == "<<-")
            elif op == "->":
                left = ("assign", right, left, False)
            else:
                left = ("bin", "^" if op == "**" else op, left, right)
        return left

    def parse_unary(self):
        tok = self.peek()
        if tok[0] == "op" and tok[1] in ("-", "+"):
            self.next()
            e = self.parse_expr(28)       # binds tighter than * but looser than ^
            return ("un", tok[1], e)
        if tok[0] == "op" and tok[1] == "!":
            self.next()
            return ("un", "!", self.parse_expr(14))
        return self.parse_postfix(self.parse_primary())

    def parse_postfix(self, e):
        while True:
            tok = self.peek()
            if tok[0] != "op":
                break
            v = tok[1]
            if v == "(":
                self.next()
                args = self.parse_args(")")
                e = ("call", e, args)
            elif v == "[[":
                self.next()
                self.depth += 1
                idx = self.parse_expr(0)
                self.depth -= 1
                self.expect("]")
                self.expect("]")
                e = ("index2", e, idx)
            elif v == "[":
                self.next()
                args = self.parse_args("]", allow_empty=True)
                e = ("index", e, args)
            elif v in ("$", "@"):
                self.next()
                nm = self.next()
                name = nm[1].strip("`") if nm[0] == "id" else _unescape(nm[1]) if nm[0] == "str" else nm[1]
                e = ("dollar", e, name)
            else:
                break
        return e

    def parse_args(self, close, allow_empty=False):
        self.depth += 1
        args = []
        while True:
            tok = self.peek()
            if tok[1] == close and tok[0] == "op":
                self.next()
                break
            if tok[0] == "op" and tok[1] == ",":
                self.next()
                if allow_empty:
                    args.append((None, ("missing",)))
                continue
            # named argument?
            name = None
            if tok[0] in ("id", "str"):
                j = self.i
                while self.t[j][0] == "nl":
                    j += 1
                k = j + 1
                while self.t[k][0] == "nl":
                    k += 1
                if self.t[k][0] == "op" and self.t[k][1] == "=" and not (self.t[k + 1][0] == "op" and self.t[k + 1][1] == "="):
                    name = tok[1].strip("`") if tok[0] == "id" else _unescape(tok[1])
                    self.i = k + 1
            nt = self.peek()
            if name is not None and nt[0] == "op" and nt[1] in (",", close):
                args.append((name, ("missing",)))
            else:
                args.append((name, self.parse_expr(4 if name is None else 4)))
            tok = self.peek()
            if tok[0] == "op" and tok[1] == ",":
                self.next()
                nt = self.peek()
                if allow_empty and nt[0] == "op" and nt[1] == close:
                    args.append((None, ("missing",)))
                continue
        self.depth -= 1
        return args

    def parse_block_or_expr(self):
        self.skip_only_newlines()
        return self.parse_expr(0)

    def skip_only_newlines(self):
        while self.t[self.i][0] == "nl":
            self.i += 1

    def parse_primary(self):
        tok = self.next(skip_nl=False) if self.depth == 0 else self.next(skip_nl=True)
        k, v, line = tok
        if k == "num":
            if v.startswith("0x"):
                return ("const", vec(int(v.rstrip("L"), 16)))
            if v.endswith("L"):
                return ("const", vec(int(float(v[:-1]))))
            return ("const", vec(float(v)))
        if k == "str":
            return ("const", vec(_unescape(v)))
        if k == "id":
            v = v.strip("`")
            if v == "TRUE":
                return ("const", vec(True))
            if v == "FALSE":
                return ("const", vec(False))
            if v == "NULL":
                return ("const", NULL)
            if v in ("NA", "NA_integer_", "NA_real_", "NA_character_"):
                return ("const", vec(None))
            if v == "Inf":
                return ("const", vec(float("inf")))
            if v == "function":
                self.expect("(")
                params = []
                self.depth += 1
                while True:
                    t2 = self.next(skip_nl=True)
                    if t2[1] == ")":
                        break
                    if t2[1] == ",":
                        continue
                    pname = t2[1]
                    default = None
                    if self.peek()[1] == "=":
                        self.next()
                        default = self.parse_expr(4)
                    params.append((pname, default))
                self.depth -= 1
                saved = self.depth
                self.depth = 0
                body = self.parse_block_or_expr()
                self.depth = saved
                return ("function", params, body)
            if v == "if":
                self.expect("(")
                self.depth += 1
                cond = self.parse_expr(0)
                self.depth -= 1
                self.expect(")")
                saved = self.depth
                self.depth = 0
                yes = self.parse_block_or_expr()
                # else may follow after newlines only inside a block; we accept it always
                j = self.i
                while self.t[j][0] == "nl":
                    j += 1
                no = None
                if self.t[j][0] == "id" and self.t[j][1] == "else":
                    self.i = j + 1
                    no = self.parse_block_or_expr()
                self.depth = saved
                return ("if", cond, yes, no)
            if v == "for":
                self.expect("(")
                self.depth += 1
                var = self.next(skip_nl=True)[1]
                t2 = self.next(skip_nl=True)
                if t2[1] != "in":
                    raise RError(f"R parser: expected 'in' at line {line}")
                seq = self.parse_expr(0)
                self.depth -= 1
                self.expect(")")
                saved = self.depth
                self.depth = 0
                body = self.parse_block_or_expr()
                self.depth = saved
                return ("for", var, seq, body)
            if v == "while":
                self.expect("(")
                self.depth += 1
                cond = self.parse_expr(0)
                self.depth -= 1
                self.expect(")")
                saved = self.depth
                self.depth = 0
                body = self.parse_block_or_expr()
                self.depth = saved
                return ("while", cond, body)
            if v == "repeat":
                saved = self.depth
                self.depth = 0
                body = self.parse_block_or_expr()
                self.depth = saved
                return ("while", ("const", vec(True)), body)
            if v == "break":
                return ("break",)
            if v == "next":
                return ("next",)
            return ("name", v)
        if k == "op" and v == "(":
            self.depth += 1
            e = self.parse_expr(0)
            self.depth -= 1
            self.expect(")")
            return ("paren", e)
        if k == "op" and v == "{":
            saved = self.depth
            self.depth = 0
            body = []
            self.skip_newlines()
            while not (self.peek()[0] == "op" and self.peek()[1] == "}"):
                if self.peek()[0] == "eof":
                    raise RError("R parser: unterminated '{'")
                body.append(self.parse_expr(0))
                self.skip_newlines()
            self.next()
            self.depth = saved
            return ("block", body)
        if k == "nl":
            return self.parse_primary()
        raise RError(f"R parser: unexpected token {v!r} at line {line}")


def parse(src: str):
    return Parser(src).parse_program()
