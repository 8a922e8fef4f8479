"""Model registry.  A model *variant* is a base model plus a set of enabled options and is
named ``<model>[_<opt>...]`` exactly like the reference's build targets (``make d2q9_bc``,
variant expansion in ``src/models.R:20-89``).

A model module either describes ONE model (module attributes NAME, OPTIONS, VARIANTS, build(options))
or several (module attribute MODELS = {name: ModelSpec})."""
from __future__ import annotations

import importlib
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple


@dataclass
class ModelSpec:
    name: str
    build: Callable
    options: List[str] = field(default_factory=list)
    variants: List[dict] = field(default_factory=lambda: [dict()])
    reference: str = ""        # directory of the counterpart in the reference tree (models/...)


_MODULES = ["flow_d2q9", "flow_d3q19", "flow_d3q27_cumulant", "pde", "flow_simple", "heat", "multiphase_simple", "electrokinetic"]
_SPECS: Optional[Dict[str, ModelSpec]] = None


def specs() -> Dict[str, ModelSpec]:
    global _SPECS
    if _SPECS is None:
        out: Dict[str, ModelSpec] = {}
        for mod in _MODULES:
            try:
                m = importlib.import_module(f"tclb_b200.models.{mod}")
            except ModuleNotFoundError as e:
                if e.name and e.name.endswith(mod):
                    continue
                raise
            if hasattr(m, "MODELS"):
                out.update(m.MODELS)
            else:
                out[m.NAME] = ModelSpec(m.NAME, m.build, list(getattr(m, "OPTIONS", [])), list(getattr(m, "VARIANTS", [dict()])),
                                        getattr(m, "REFERENCE", ""))
        _SPECS = out
    return _SPECS


def register(spec: ModelSpec) -> None:
    specs()[spec.name] = spec


def base_models() -> List[str]:
    return sorted(specs())


def split_variant(variant: str) -> Tuple[str, Dict[str, bool]]:
    """'d3q27_cumulant_AVG_SMAG' -> ('d3q27_cumulant', {'AVG': True, 'SMAG': True})."""
    best = None
    for b in specs():
        if variant == b or variant.startswith(b + "_"):
            rest = variant[len(b):].strip("_")
            known = specs()[b].options
            if rest and not all(o in known for o in rest.split("_")):
                continue
            if best is None or len(b) > len(best):
                best = b
    if best is None:
        raise KeyError(f"unknown model '{variant}' (known: {', '.join(base_models())})")
    rest = variant[len(best):].strip("_")
    return best, {o: True for o in rest.split("_")} if rest else {}


def variants(base: str) -> List[str]:
    sp = specs()[base]
    out = []
    for v in sp.variants:
        on = [o for o in sp.options if v.get(o)]
        out.append("_".join([base] + on))
    return out


def build_model(variant: str):
    base, opts = split_variant(variant)
    from ..gen import poly
    poly._tmp_counter[0] = 0          # temporaries of generated code are numbered per model: same source every time
    m = specs()[base].build(opts)
    m.name = base
    m.finalize()
    return m
