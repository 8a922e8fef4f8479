"""Documentation generators (reference: src/Model.md.Rt, src/Models.md.Rt, src/schema.xsd.Rt, src/catalog.xml.Rt,
tools/gen.doc.R + doc/elements.yaml):

    python -m tclb_b200.docs --out docs/generated [models...]

writes one ``<model>.md`` per model (settings with defaults/derivations, zonal settings, Globals, Quantities, node
types per group, densities with their streaming vectors, fields with their stencils, stages and actions), the index
``Models.md``, ``elements.md`` (the XML elements the case runner understands, with their attributes) and ``schema.xsd``.
The model pages are produced from the generator IR (ModelDef), so they also work for model directories of a TCLB tree
(``--dir``)."""
from __future__ import annotations

import argparse
import os
import sys
from typing import List

from .gen.modeldef import ModelDef

# element -> (kind, attributes, description); kinds follow the reference's handler classes
ELEMENTS = {
    "CLBConfig": ("container", "version output permissive oversubscribe_gpu", "root element; writes <out>_config_P00_00000000.xml and runs its children in order"),
    "Units": ("container", "", "children <Param name value gauge> define the unit gauge"),
    "Geometry": ("action", "nx ny nz px py pz predef model save", "paints node types, setting zones and wall cuts; children are node-type brushes around primitives"),
    "Model": ("container", "", "runs its <Param> children, then the Init action"),
    "Init": ("action", "", "runs the Init action again"),
    "Param": ("action", "name value zone gauge", "sets a setting or a zonal setting (unit-aware value)"),
    "Solve": ("action", "Iterations output", "main loop; periodic children fire inside it"),
    "RunAction": ("action", "name Iterations", "main loop over a named action"),
    "Repeat": ("container", "Times", "repeats its children"),
    "Container": ("container", "", "grouping"),
    "EvalIf": ("container", "opt_present opt_missing", "children run only if the model has / lacks an option"),
    "VTK": ("callback", "Iterations name what dx dy dz nx ny nz", "ImageData pieces + .pvti index"),
    "HDF5": ("callback", "Iterations name what compress write_xdmf point_data precision dx dy dz nx ny nz", "one .h5 for the output region + .xmf (3DCoRectMesh) index; compress=true|1..9: chunked datasets with deflate; one process per GPU: one piece per rank + an XDMF spatial collection"),
    "TXT": ("callback", "Iterations name what gzip", "one text file per quantity"),
    "Image": ("callback", "Iterations name plane scale", "the picture of the reference's interactive view (colour kernel of one z-plane) as a binary PPM"),
    "BIN": ("callback", "Iterations name", "raw binary per quantity"),
    "Log": ("callback", "Iterations name", "CSV log of settings, zone settings and Globals"),
    "Stop": ("callback", "Iterations Times *Change *PercentChange *Above *Below", "stops the loop on a Global criterion"),
    "Failcheck": ("callback", "Iterations what dx dy dz nx ny nz", "NaN scan on the device; runs children and stops on failure"),
    "Average": ("callback", "Iterations", "resets the running averages"),
    "Sample": ("callback", "Iterations what", "point probes (<Point dx dy dz/>) to CSV"),
    "Keep": ("callback", "Iterations What Above Below Equal Force", "penalty weight on a Global"),
    "PID": ("callback", "Iterations integral control zone scale DerivativeTime IntegrationTime", "PID controller: Global -> zonal setting"),
    "DumpSettings": ("callback", "Iterations", "zone-settings time series CSV"),
    "SaveCheckpoint": ("callback", "Iterations keep", "per-rank .pri images + restart XML"),
    "SaveBinary": ("callback", "Iterations file filename comp", "whole solution or one component"),
    "SaveMemoryDump": ("callback", "Iterations file filename", "whole solution"),
    "LoadBinary": ("action", "file comp", "load what SaveBinary wrote"),
    "LoadMemoryDump": ("action", "file", "load what SaveMemoryDump wrote"),
    "Control": ("action", "Iterations", "time-dependent zonal settings from <CSV file Time> + <Param value=expr>"),
    "SyntheticTurbulence": ("action", "Modes Spread Spectrum *WaveLength *WaveNumber *WaveFrequency", "random Fourier modes for turbulent inlets"),
    "RunPython": ("callback", "Iterations file", "Python code with the Solver object (python -m tclb_b200.run)"),
    "Adjoint": ("container", "type Iterations tape", "type=unsteady: children run with every iteration on the tape, then the reverse sweep; type=steady: Iterations adjoint iterations around the state the children left"),
    "Objective": ("action", "<objective>=weight ...", "weighted model objectives (every Global, and AddObjective expressions of the Globals) linearised at the current Globals into the <Global>InObj zonal settings"),
    "FDTest": ("container", "Samples Epsilon parameters", "adjoint gradient against central differences for some (parameters=\"a:b\") or all design parameters (children re-run from the same state)"),
    "OptSolve": ("action", "Iterations Descent direction", "one-shot optimisation: primal iteration + adjoint iteration + descent step of the parameter fields"),
    "Optimize": ("container", "MaxEvaluations method Material XAbsTolerance RelTolerance AbsTolerance StopAtValue Step", "maximises the objective accumulated by the children over the design parameters; method = MMA, LBFGS, NELDERMEAD (= COBYLA), GD (built in)"),
    "OptimalControl": ("design", "what lower upper", "time series of a zonal setting (what=\"Setting-Zone\") as design parameters"),
    "OptimalControlSecond": ("design", "what lower upper", "every second time level of a zonal setting, the levels in between interpolated"),
    "Fourier": ("design", "modes lower upper", "wraps one design child: its parameters as a truncated Fourier series"),
    "BSpline": ("design", "nodes order periodic lower upper", "wraps one design child: its parameters as a B-spline"),
    "RepeatControl": ("design", "length flip lower upper", "wraps one design child: a pattern of `length` parameters repeated (optionally mirrored)"),
    "InternalTopology": ("design", "", "the parameter fields on the DesignSpace nodes as design parameters"),
    "FieldParameter": ("design", "field where zone", "one field on the nodes of a node type / zone as design parameters"),
    "ControlParameter": ("design", "", "parameters of a <Control> element as design parameters"),
    "Extrude": ("design", "direction theta margin", "wraps one design child: parameters constant along one direction (x, y, z or t)"),
    "Threshold": ("container", "Levels", "runs the children on thresholded copies of the design, level i/Levels"),
    "ThresholdNow": ("action", "Level", "thresholds the parameter fields to 0/1"),
    "Andersen": ("container", "Directions Times Eps", "Anderson mixing of the fixed-point iteration state -> children(state)"),
    "RunR": ("callback", "Iterations file", "R code with the Solver list, run in the R-subset interpreter (python -m tclb_b200.run)"),
    "RemoteForceInterface": ("action", "integrator omega torque spawn connect", "particle coupling: built-in SIMPLEPART (child <SimplePart>), or an external integrator program started (spawn=) or listening (connect=)"),
    "ESYSParticle": ("action", "", "recognised; points to RemoteForceInterface spawn=/connect= (ESyS-Particle is not installed)"),
    "Catalyst": ("callback", "", "recognised; unavailable on this box"),
    "Run": ("marker", "", "ignored"),
}
PRIMITIVES = "Box Sphere HalfSphere OffgridSphere OffgridPipe XPipe YPipe ZPipe XAnnulus Pipe PipeY PipeZ Cylinder Wedge STL Sweep Text Zone Inlet Outlet Channel Tunnel"


def model_page(m: ModelDef) -> str:
    m.finalize()
    L: List[str] = [f"# Model `{m.name}`", ""]
    desc = getattr(m, "description", None)
    if isinstance(desc, dict):
        desc = desc.get("long") or desc.get("short")
    if desc and desc != m.name:
        L += [str(desc), ""]
    opts = [k for k, v in m.options.items()]
    if opts:
        L += ["Options: " + ", ".join(f"`{k}`={'on' if m.options[k] else 'off'}" for k in opts), ""]
    L += ["## Settings", "", "| name | default | derives | unit | comment |", "|---|---|---|---|---|"]
    for s in m.settings:
        L.append(f"| `{s.name}` | {s.default} | {('`' + s.derived + '` = ' + str(s.equation)) if s.derived else ''} | {s.unit} | {s.comment} |")
    L += ["", "## Zonal settings (per setting zone, may be time series)", "", "| name | default | unit | comment |", "|---|---|---|---|"]
    for s in m.zonesettings:
        L.append(f"| `{s.name}` | {s.default} | {s.unit} | {s.comment} |")
    L += ["", "## Globals", "", "| name | reduction | unit | comment |", "|---|---|---|---|"]
    for g in m.globals:
        L.append(f"| `{g.name}` | {g.op} | {g.unit} | {g.comment} |")
    L += ["", "## Quantities (exported by VTK/TXT/BIN/HDF5, sampled, scanned by Failcheck)", "", "| name | type | unit | comment |", "|---|---|---|---|"]
    for q in m.quantities:
        L.append(f"| `{q.name}` | {'vector' if q.vector else 'scalar'} | {q.unit} | {q.comment} |")
    L += ["", "## Node types", "", "| group | mask | types (value) |", "|---|---|---|"]
    for g in m.nodetypegroups:
        types = ", ".join(f"`{n.name}` ({n.value})" for n in m.nodetypes if n.group == g.name)
        L.append(f"| `{g.name}` | {g.mask} | {types} |")
    L += ["", f"Flag word: {m.flag_bits} bits, setting zone in the bits from {m.zone_shift} up (max zone {m.zone_max}).", ""]
    L += ["## Densities", "", "| name | field | streaming vector | group |", "|---|---|---|---|"]
    for d in m.densities:
        L.append(f"| `{d.name}` | `{d.field}` | ({d.dx},{d.dy},{d.dz}) | {d.group} |")
    L += ["", "## Fields", "", "| name | x range | y range | z range | group |", "|---|---|---|---|---|"]
    for f in m.fields:
        L.append(f"| `{f.name}` | {f.minx}..{f.maxx} | {f.miny}..{f.maxy} | {f.minz}..{f.maxz} | {f.group} |")
    L += ["", "## Stages and actions", "", "| stage | function |", "|---|---|"]
    for s in m.stages:
        L.append(f"| `{s.name}` | `{s.main}()` |")
    L += ["", "| action | stages |", "|---|---|"]
    for a in m.actions:
        L.append(f"| `{a.name}` | {' -> '.join(a.stages)} |")
    return "\n".join(L) + "\n"


def elements_page() -> str:
    L = ["# XML elements", "", "| element | kind | attributes | behaviour |", "|---|---|---|---|"]
    for n, (kind, attrs, doc) in ELEMENTS.items():
        L.append(f"| `{n}` | {kind} | {' '.join('`' + a + '`' for a in attrs.split())} | {doc} |")
    L += ["", "Geometry primitives: " + ", ".join(f"`{p}`" for p in PRIMITIVES.split()) + ".",
          "Every callback takes `Iterations=` (unit aware); without it the callback fires once, where it stands.", ""]
    return "\n".join(L)


def schema_xsd(models: List[ModelDef]) -> str:
    nts = sorted({n.name for m in models for n in m.nodetypes})
    out = ['<?xml version="1.0"?>', '<xs:schema xmlns:xs="http://www.w3.org/2001/XMLSchema">']
    for n, (kind, attrs, doc) in ELEMENTS.items():
        out.append(f'  <xs:element name="{n}"><xs:annotation><xs:documentation>{doc}</xs:documentation></xs:annotation><xs:complexType mixed="true">'
                   '<xs:sequence><xs:any minOccurs="0" maxOccurs="unbounded" processContents="lax"/></xs:sequence>')
        for a in attrs.split():
            if "*" not in a:
                out.append(f'    <xs:attribute name="{a}" type="xs:string"/>')
        out.append('    <xs:anyAttribute processContents="lax"/>\n  </xs:complexType></xs:element>')
    for n in nts + PRIMITIVES.split():
        out.append(f'  <xs:element name="{n}"><xs:complexType><xs:sequence><xs:any minOccurs="0" maxOccurs="unbounded" processContents="lax"/>'
                   '</xs:sequence><xs:anyAttribute processContents="lax"/></xs:complexType></xs:element>')
    out.append("</xs:schema>")
    return "\n".join(out) + "\n"


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m tclb_b200.docs")
    ap.add_argument("models", nargs="*")
    ap.add_argument("--dir", action="append", default=[], help="model directory of a TCLB tree (repeatable)")
    ap.add_argument("--out", default="docs/generated")
    a = ap.parse_args(argv)
    from .build import all_variants
    from .models import build_model
    os.makedirs(a.out, exist_ok=True)
    defs: List[ModelDef] = [build_model(v) for v in (a.models or all_variants())]
    if a.dir:
        from .gen import rfrontend
        defs += [rfrontend.load_model(d) for d in a.dir]
    index = ["# Models", "", "| model | options | fields | settings | globals | stages per iteration |", "|---|---|---|---|---|---|"]
    seen = set()
    for m in defs:
        on = [k for k, v in m.options.items() if v]
        name = "_".join([m.name] + on)
        if name in seen:
            continue
        seen.add(name)
        open(os.path.join(a.out, name + ".md"), "w").write(model_page(m))
        it = [x for x in m.actions if x.name == "Iteration"]
        index.append(f"| [{name}]({name}.md) | {' '.join(on)} | {len(m.fields)} | {len(m.settings) + len(m.zonesettings)} | {len(m.globals)} | {len(it[0].stages) if it else 0} |")
    open(os.path.join(a.out, "Models.md"), "w").write("\n".join(index) + "\n")
    open(os.path.join(a.out, "elements.md"), "w").write(elements_page())
    open(os.path.join(a.out, "schema.xsd"), "w").write(schema_xsd(defs))
    print(f"wrote {len(seen)} model pages, Models.md, elements.md, schema.xsd to {a.out}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
