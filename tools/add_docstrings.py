#!/usr/bin/env python
"""Insert one-line docstrings for named top-level functions / classes that have none.

    python tools/add_docstrings.py spec.json      # {"path/to/file.py": {"symbol": "docstring text", ...}, ...}

A maintenance helper for ``tools/gen_api_reference.py``: it only touches symbols WITHOUT a docstring, keeps every other
byte of the file, and refuses definitions whose body starts on the ``def`` line.  ``Class.method`` names address methods.
"""
import ast
import json
import sys


def _find(tree, dotted):
    node = tree
    for part in dotted.split("."):
        for child in node.body:
            if isinstance(child, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)) and child.name == part:
                node = child
                break
        else:
            return None
    return node


def apply(path, entries):
    src = open(path).read()
    lines = src.split("\n")
    tree = ast.parse(src)
    inserts = []
    for name, doc in entries.items():
        node = _find(tree, name)
        if node is None:
            print(f"  {path}: {name} not found")
            continue
        if ast.get_docstring(node) is not None:
            print(f"  {path}: {name} already has a docstring")
            continue
        first = node.body[0]
        if first.lineno == node.lineno:
            print(f"  {path}: {name} has its body on the def line; edit by hand")
            continue
        indent = " " * first.col_offset
        text = doc.replace('"""', "'\'\'\'")
        pad = " " if text.endswith('"') else ""
        start = min([first.lineno] + [d.lineno for d in getattr(first, "decorator_list", [])])
        inserts.append((start - 1, f'{indent}"""{text}{pad}"""'))
    for at, line in sorted(inserts, reverse=True):
        lines.insert(at, line)
    open(path, "w").write("\n".join(lines))
    ast.parse(open(path).read())
    return len(inserts)


if __name__ == "__main__":
    spec = json.load(open(sys.argv[1]))
    total = sum(apply(p, e) for p, e in spec.items())
    print(f"inserted {total} docstrings")
