"""Template engine for ``*.Rt`` model sources (the reference pipes them through the external ``rtemplate`` R package:
``tools/RT``, ``src/makefile.main.Rt:120-160``).

A template is turned into ONE program of the R subset: literal text becomes an output statement, ``<?R code ?>`` is
spliced in verbatim (so control structures may open in one chunk and close in a later one), ``<?%s expr ?>`` /
``<?%d ?>`` / ``<?%f ?>`` print ``sprintf(fmt, expr)`` and ``<?RT file ?>`` renders another template in place.  The
program runs in the interpreter that already evaluated ``Dynamics.R``, so the model's symbolic objects are visible.
"""
from __future__ import annotations

import os
import re
from typing import List, Optional

from .rinterp import Env, Interp
from .rlang import NULL, RError, scalar, vec

_TAG = re.compile(r"<\?(RT\b|R\b|%[-0-9.]*[a-zA-Z]|python\b)?(.*?)\?>", re.S)


class _Lit(str):
    """literal chunk of a template that remembers where it ends (file, line): error messages point there"""
    where = ("", 0)


def template_to_program(text: str, texts: List[str], path: str = "") -> str:
    """R-subset source for a template; literal chunks are appended to `texts` and referenced by index."""
    out: List[str] = []
    pos = 0
    for m in _TAG.finditer(text):
        lit = text[pos:m.start()]
        if lit:
            lit = _Lit(lit)
            lit.where = (path, text.count("\n", 0, m.start()) + 1)
            texts.append(lit)
            out.append(f".rt_text({len(texts) - 1})")
        kind, body = m.group(1) or "R", m.group(2)
        if kind == "R":
            out.append(body)
        elif kind == "RT":
            out.append(f".rt_include({body.strip()!r})" if not body.strip().startswith(("'", '"')) else f".rt_include({body.strip()})")
        elif kind.startswith("%"):
            out.append(f".rt_fmt({kind!r}, {body.strip()})")
        elif kind == "python":
            texts.append(body)
            out.append(f".rt_python({len(texts) - 1})")
        else:
            raise RError(f"template tag <?{kind} is not supported")
        pos = m.end()
    lit = text[pos:]
    if lit:
        texts.append(lit)
        out.append(f".rt_text({len(texts) - 1})")
    return "\n".join(out) + "\n"


_PY_TABLES = ("Density", "DensityAll", "DensityAD", "Fields", "Settings", "ZoneSettings", "Globals", "Quantities", "NodeTypes",
              "Stages", "Actions")
_PY2_PRINT = re.compile(r"^(\s*)print[ \t]+(?!\()(.+?)[ \t]*$", re.M)


def run_python_chunk(I: Interp, env: Env, body: str) -> str:
    """``<?python ... ?>``: Python embedded in a model template.  Running it means executing code that ships with the
    model directory, so it is OFF unless the user opts in with ``TCLB_ALLOW_PYTHON_CHUNKS=1`` (only do that for model
    files you trust).  When enabled, the chunks of one template share ONE namespace (a preamble chunk defines what later
    chunks print), the model tables are visible as dicts of columns (what rPython hands over) and stdout goes into the
    generated file.  The chunks in existing models were written for Python 2 + sympy (print statements, list-returning
    zip/map, cStringIO), so those are provided."""
    import builtins
    import contextlib
    import io
    import sys
    import textwrap
    import types
    from .rlang import DataFrame
    if os.environ.get("TCLB_ALLOW_PYTHON_CHUNKS") != "1":
        raise RError("template tag <?python is not supported unless TCLB_ALLOW_PYTHON_CHUNKS=1 (it executes Python code "
                     "from the model directory)")
    ns = getattr(I, "_py_namespace", None)
    if ns is None:
        if "cStringIO" not in sys.modules:
            mod = types.ModuleType("cStringIO")
            mod.StringIO = io.StringIO
            sys.modules["cStringIO"] = mod
        ns = {"__name__": "__tclb_template__", "zip": lambda *a: list(builtins.zip(*a)), "map": lambda *a: list(builtins.map(*a)),
              "filter": lambda *a: list(builtins.filter(*a)), "xrange": range}
        I._py_namespace = ns
    for name in _PY_TABLES:
        try:
            v = env.get(name)
        except RError:
            continue
        if isinstance(v, DataFrame):
            ns[name] = {c: [x.item() if hasattr(x, "item") else x for x in col] for c, col in v.cols.items()}
    code = _PY2_PRINT.sub(lambda m: f"{m.group(1)}print({m.group(2)})", textwrap.dedent(body.expandtabs(8)))
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            exec(compile(code, "<?python chunk>", "exec"), ns)
    except Exception as e:
        raise RError(f"<?python chunk failed: {type(e).__name__}: {e}")
    return buf.getvalue()


def _read(path: str) -> str:
    raw = open(path, "rb").read()
    try:
        return raw.decode("utf-8")
    except UnicodeDecodeError:
        return raw.decode("latin-1")      # a few model sources carry Latin-1 author names in comments


def render(I: Interp, env: Env, path: str, search: Optional[List[str]] = None) -> str:
    """Render the template file `path` inside interpreter I / environment env and return the produced text."""
    search = list(search or []) + [os.path.dirname(os.path.abspath(path))]
    texts: List[str] = []

    def rt_text(I_, a, k):
        t = texts[int(scalar(a[0]))]
        I_.out.append(t)
        I_.rt_where = getattr(t, "where", None)
        return NULL

    def rt_fmt(I_, a, k):
        fmt = scalar(a[0])
        for v in vec(a[1]).ravel(order="F").tolist() if not isinstance(a[1], str) else [a[1]]:
            if fmt.endswith("d"):
                I_.out.append(fmt % int(v))
            elif fmt[-1] in "feEgG":
                I_.out.append(fmt % float(v))
            else:
                from .rbuiltins import to_str
                I_.out.append(fmt % to_str(v))
        return NULL

    def rt_include(I_, a, k):
        name = scalar(a[0])
        for d in search:
            p = os.path.join(d, name)
            if os.path.exists(p):
                prog = template_to_program(_read(p), texts, p)
                I_.run(prog, env)
                return NULL
        raise RError(f"template include not found: {name}")

    def rt_python(I_, a, k):
        I_.out.append(run_python_chunk(I_, env, texts[int(scalar(a[0]))]))
        return NULL

    env.set(".rt_python", rt_python)
    env.set(".rt_text", rt_text)
    env.set(".rt_fmt", rt_fmt)
    env.set(".rt_include", rt_include)
    saved = I.out
    I.out = []
    try:
        try:
            I.run(template_to_program(_read(path), texts, path), env)
        except Exception as e:
            if not getattr(e, "rt_where", None):
                e.rt_where = getattr(I, "rt_where", None)
                e.args = (f"{e.args[0] if e.args else ''} (after the literal text ending at {e.rt_where})",) + e.args[1:]
            raise
        return "".join(I.out)
    finally:
        I.out = saved
