"""Documentation that is checked mechanically: the generated API reference is current, every public top-level symbol of
the package has a docstring, and the paths the documents name exist."""
import importlib.util
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_api_reference_is_current_and_complete():
    gen = _load("tools/gen_api_reference.py", "gen_api_reference")
    text = gen.generate()
    with open(os.path.join(ROOT, "docs", "api_reference.md")) as f:
        assert f.read() == text, "docs/api_reference.md is stale: run python tools/gen_api_reference.py"
    assert "(no docstring)" not in text, [l for l in text.splitlines() if "(no docstring)" in l][:10]
    for must in ("`scaletorch_b200.dist.collective_ops`", "`scaletorch_b200.trainer.graphed`", "class `ProcessGroupManager`",
                 "class `PeerShardedAdamW", "`flash_attention("):
        assert must in text, must


def test_paths_named_in_the_documents_exist():
    """Back-quoted repository paths in the Markdown documents must point at files that are there."""
    docs = [os.path.join(ROOT, "docs", f) for f in sorted(os.listdir(os.path.join(ROOT, "docs"))) if f.endswith(".md")]
    docs += [os.path.join(ROOT, f) for f in ("README.md", "DESIGN.md", "INTEGRATION.md")]
    roots = ("scaletorch_b200/", "csrc/", "tools/", "scripts/", "tests/", "docs/", "profiles/", "examples/", "include/", "baseline/", "configs/")
    # names of the REFERENCE's documents, a pair written as one token, and a file whose removal is being reported
    allowed = {"docs/device_mesh.md", "docs/FSDP2-fully-shard-guide.md", "scripts/benchmark_comprehensive.py/.sh",
               "tests/test_experimental_gpu.py"}
    missing = []
    for d in docs:
        text = open(d).read()
        for m in re.finditer(r"`([A-Za-z0-9_./\-]+\.(?:py|cu|cuh|cpp|h|md|sh|json|jsonl|txt|yaml|toml))(?:::[^`]*)?`", text):
            path = m.group(1)
            if path.startswith("../"):
                path = path[3:]
            if not path.startswith(roots):
                continue
            if "*" in path or "<" in path or "…" in path:
                continue
            if path.startswith("baseline/_ref") or path in allowed:
                continue
            if not os.path.exists(os.path.join(ROOT, path)):
                missing.append((os.path.relpath(d, ROOT), path))
    assert not missing, missing
