#!/usr/bin/env python
"""Static gate for the Python tree (no third-party linter is available offline): zero findings or exit 1.

Counterpart of the reference's checkstyle + findbugs gate (build.gradle:112-113,290-304).  Checks, per file:
syntax (ast.parse), unused imports, names imported twice, bare ``except:``, mutable default arguments,
``== None`` / ``!= None`` comparisons, tabs and trailing whitespace, and a 130-column limit.  ``# noqa`` on a line silences that line.

    python tools/lint.py [paths...]        (default: tony_b200 examples tests tools bench.py __graft_entry__.py)
"""
import ast
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT = ["tony_b200", "examples", "tests", "tools", "bench.py", "__graft_entry__.py"]
MAX_COL = 130


def py_files(paths):
    for p in paths:
        p = os.path.join(ROOT, p) if not os.path.isabs(p) else p
        if os.path.isfile(p) and p.endswith(".py"):
            yield p
        for d, _, files in os.walk(p) if os.path.isdir(p) else ():
            if "__pycache__" in d or "/lib/" in d + "/":
                continue
            for f in sorted(files):
                if f.endswith(".py"):
                    yield os.path.join(d, f)


class Visitor(ast.NodeVisitor):
    def __init__(self):
        self.imported = {}      # name -> line
        self.used = set()
        self.findings = []

    def visit_Import(self, node):
        for a in node.names:
            name = (a.asname or a.name).split(".")[0]
            self._imp(name, node.lineno)

    def visit_ImportFrom(self, node):
        if node.module == "__future__":
            return
        for a in node.names:
            if a.name != "*":
                self._imp(a.asname or a.name, node.lineno)

    def _imp(self, name, line):
        self.imported.setdefault(name, line)

    def visit_Name(self, node):
        self.used.add(node.id)

    def visit_Attribute(self, node):
        self.generic_visit(node)

    def visit_ExceptHandler(self, node):
        if node.type is None:
            self.findings.append((node.lineno, "bare except"))
        self.generic_visit(node)

    def _defaults(self, node):
        for d in list(node.args.defaults) + [d for d in node.args.kw_defaults if d is not None]:
            if isinstance(d, (ast.List, ast.Dict, ast.Set)):
                self.findings.append((d.lineno, "mutable default argument"))
        self.generic_visit(node)

    visit_FunctionDef = _defaults
    visit_AsyncFunctionDef = _defaults

    def visit_Compare(self, node):
        for op, c in zip(node.ops, node.comparators):
            if isinstance(op, (ast.Eq, ast.NotEq)) and isinstance(c, ast.Constant) and c.value is None:
                self.findings.append((node.lineno, "comparison to None with == / !="))
        self.generic_visit(node)



def lint(path):
    src = open(path, encoding="utf-8").read()
    lines = src.splitlines()
    out = []
    try:
        tree = ast.parse(src, path)
    except SyntaxError as e:
        return [(e.lineno or 0, f"syntax error: {e.msg}")]
    v = Visitor()
    v.visit(tree)
    exported = set()
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id == "__all__" for t in node.targets):
            if isinstance(node.value, (ast.List, ast.Tuple)):
                exported |= {e.value for e in node.value.elts if isinstance(e, ast.Constant)}
    # names used inside strings of type comments / docstrings do not count; attribute roots do
    for node in ast.walk(tree):
        if isinstance(node, ast.Attribute):
            r = node
            while isinstance(r, ast.Attribute):
                r = r.value
            if isinstance(r, ast.Name):
                v.used.add(r.id)
    is_init = os.path.basename(path) == "__init__.py"
    for name, line in v.imported.items():
        if name not in v.used and name not in exported and not is_init and not name.startswith("_"):
            out.append((line, f"unused import '{name}'"))
    out += v.findings
    for i, l in enumerate(lines, 1):
        if "\t" in l:
            out.append((i, "tab character"))
        if l != l.rstrip():
            out.append((i, "trailing whitespace"))
        if len(l) > MAX_COL:
            out.append((i, f"line longer than {MAX_COL} columns"))
    return [(ln, msg) for ln, msg in out if not (0 < ln <= len(lines) and "noqa" in lines[ln - 1])]


def main():
    paths = sys.argv[1:] or DEFAULT
    total = 0
    for f in py_files(paths):
        for ln, msg in sorted(lint(f)):
            print(f"{os.path.relpath(f, ROOT)}:{ln}: {msg}")
            total += 1
    print(f"lint: {total} finding(s)")
    return 1 if total else 0


if __name__ == "__main__":
    sys.exit(main())
