"""<?python ?> chunks in model templates: off by default, opt-in with TCLB_ALLOW_PYTHON_CHUNKS=1.
The template here is this repository's own (nothing from a model tree is executed)."""
import pytest

from tclb_b200.gen import rtemplate
from tclb_b200.gen.rinterp import Interp
from tclb_b200.gen.rlang import RError

TEMPLATE = """// head
<?R n = 3 ?>
<?python
from sympy import var
from sympy.printing import print_ccode
f = var(['f[0]', 'f[1]', 'f[2]'])
pairs = zip([0, 1], [1, 2])
print "// pairs:", len(pairs)
?>
real_t s;
<?python
for i, j in pairs:
    print_ccode(f[i] + 2*f[j], assign_to=f[i])
?>
// n = <?%d n ?>
"""


def _render(tmp_path):
    p = tmp_path / "x.c.Rt"
    p.write_text(TEMPLATE)
    I = Interp()
    return rtemplate.render(I, I.globals, str(p))


def test_python_chunks_are_off_by_default(tmp_path, monkeypatch):
    monkeypatch.delenv("TCLB_ALLOW_PYTHON_CHUNKS", raising=False)
    with pytest.raises(RError, match="TCLB_ALLOW_PYTHON_CHUNKS"):
        _render(tmp_path)


def test_python_chunks_opt_in(tmp_path, monkeypatch):
    pytest.importorskip("sympy")
    monkeypatch.setenv("TCLB_ALLOW_PYTHON_CHUNKS", "1")
    out = _render(tmp_path)
    assert "// pairs: 2" in out                      # Python 2 print statement, list-returning zip
    assert "f[0] = f[0] + 2*f[1];" in out             # the namespace of the first chunk is visible in the second
    assert "f[1] = f[1] + 2*f[2];" in out
    assert "// n = 3" in out and out.startswith("// head")
