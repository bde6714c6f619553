"""``baseline/install_reference.py``: the reference arm's installer never raises, reports how the package got there, and
falls back from pip to a plain copy (the reference is a pure-Python package)."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mod():
    spec = importlib.util.spec_from_file_location("install_reference", os.path.join(ROOT, "baseline", "install_reference.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _fake_reference(tmp_path):
    src = tmp_path / "reference"
    (src / "scaletorch" / "utils").mkdir(parents=True)
    (src / "scaletorch" / "__init__.py").write_text("__version__ = 'fake'\n")
    (src / "scaletorch" / "utils" / "__init__.py").write_text("")
    (src / "docs").mkdir()
    (src / "docs" / "big.md").write_text("x" * 10)
    return src


def test_unavailable_source_is_reported_not_raised(tmp_path, monkeypatch):
    m = _mod()
    monkeypatch.setattr(m, "REF_DIR", tmp_path / "_ref")
    monkeypatch.setattr(m, "REF_SRC", tmp_path / "nowhere")
    assert not m.installed()
    ok, how = m.ensure()
    assert not ok and "not visible" in how


def test_present_installation_is_left_alone(tmp_path, monkeypatch):
    m = _mod()
    ref = tmp_path / "_ref"
    (ref / "scaletorch").mkdir(parents=True)
    (ref / "scaletorch" / "__init__.py").write_text("")
    monkeypatch.setattr(m, "REF_DIR", ref)
    monkeypatch.setattr(m, "REF_SRC", tmp_path / "nowhere")
    assert m.installed() and m.ensure() == (True, "present")


def test_copy_fallback_when_pip_cannot_build(tmp_path, monkeypatch):
    m = _mod()
    src = _fake_reference(tmp_path)
    monkeypatch.setattr(m, "REF_DIR", tmp_path / "_ref")
    monkeypatch.setattr(m, "REF_SRC", src)
    calls = []

    def no_pip(work, no_deps):
        calls.append(no_deps)
        assert (work / "scaletorch" / "__init__.py").is_file() and not (work / "docs").exists()   # built from a scratch copy
        return False, "simulated failure"

    monkeypatch.setattr(m, "_pip", no_pip)
    ok, how = m.ensure(verbose=True)
    assert ok and how == "copytree" and calls == [False, True]        # pip, then pip --no-deps, then the copy
    assert (tmp_path / "_ref" / "scaletorch" / "utils" / "__init__.py").is_file()
    assert (src / "scaletorch" / "__init__.py").read_text() == "__version__ = 'fake'\n"       # the source tree is untouched
