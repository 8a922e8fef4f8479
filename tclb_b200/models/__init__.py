"""Model registry.  A model *variant* is a base model plus a set of enabled options and is
named ``<model>[_<opt>...]`` exactly like the reference's build targets (``make d2q9_bc``,
variant expansion in ``src/models.R:20-89``)."""
from __future__ import annotations

import importlib
from typing import Dict, List, Tuple

# base model name -> module
_MODULES: Dict[str, str] = {
    "d2q9": "flow_d2q9",
    "d3q19": "flow_d3q19",
    "d3q27_cumulant": "flow_d3q27_cumulant",
}


def register(name: str, module: str) -> None:
    _MODULES[name] = module


def base_models() -> List[str]:
    return sorted(_MODULES)


def _module(base: str):
    return importlib.import_module(f"tclb_b200.models.{_MODULES[base]}")


def split_variant(variant: str) -> Tuple[str, Dict[str, bool]]:
    """'d3q27_cumulant_AVG_SMAG' -> ('d3q27_cumulant', {'AVG': True, 'SMAG': True})."""
    best = None
    for b in _MODULES:
        if variant == b or variant.startswith(b + "_"):
            if best is None or len(b) > len(best):
                best = b
    if best is None:
        raise KeyError(f"unknown model '{variant}' (known: {', '.join(base_models())})")
    rest = variant[len(best):].strip("_")
    opts = {}
    if rest:
        known = getattr(_module(best), "OPTIONS", [])
        for o in rest.split("_"):
            if o not in known:
                raise KeyError(f"model {best} has no option '{o}' (options: {known})")
            opts[o] = True
    return best, opts


def variants(base: str) -> List[str]:
    mod = _module(base)
    out = []
    for v in getattr(mod, "VARIANTS", [dict()]):
        on = [o for o in getattr(mod, "OPTIONS", []) if v.get(o)]
        out.append("_".join([base] + on))
    return out


def build_model(variant: str):
    base, opts = split_variant(variant)
    m = _module(base).build(opts)
    m.finalize()
    return m
