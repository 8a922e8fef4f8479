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


def template_to_program(text: str, texts: List[str]) -> str:
    """R-subset source for a template; literal chunks are appended to `texts` and referenced by index."""
    out: List[str] = []
    pos = 0
    for m in _TAG.finditer(text):
        lit = text[pos:m.start()]
        if lit:
            texts.append(lit)
            out.append(f".rt_text({len(texts) - 1})")
        kind, body = m.group(1) or "R", m.group(2)
        if kind == "R":
            out.append(body)
        elif kind == "RT":
            out.append(f".rt_include({body.strip()!r})" if not body.strip().startswith(("'", '"')) else f".rt_include({body.strip()})")
        elif kind.startswith("%"):
            out.append(f".rt_fmt({kind!r}, {body.strip()})")
        else:
            raise RError(f"template tag <?{kind} is not supported")
        pos = m.end()
    lit = text[pos:]
    if lit:
        texts.append(lit)
        out.append(f".rt_text({len(texts) - 1})")
    return "\n".join(out) + "\n"


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
        I_.out.append(texts[int(scalar(a[0]))])
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
                prog = template_to_program(_read(p), texts)
                I_.run(prog, env)
                return NULL
        raise RError(f"template include not found: {name}")

    env.set(".rt_text", rt_text)
    env.set(".rt_fmt", rt_fmt)
    env.set(".rt_include", rt_include)
    saved = I.out
    I.out = []
    try:
        I.run(template_to_program(_read(path), texts), env)
        return "".join(I.out)
    finally:
        I.out = saved
