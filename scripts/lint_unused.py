"""Tiny static check used in lieu of pyflakes (not installed offline): undefined names are caught by the test suite;
this flags imported-but-unused names and duplicate function definitions per module."""
import ast
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
bad = 0
for base in ("dcnn_b200", "examples", "benchmarks", "scripts", "tests", "baseline"):
    for dp, _, fs in os.walk(os.path.join(ROOT, base)):
        if "_ref" in dp:
            continue
        for f in fs:
            if not f.endswith(".py"):
                continue
            path = os.path.join(dp, f)
            src = open(path).read()
            tree = ast.parse(src)
            imported = {}
            for node in ast.walk(tree):
                if isinstance(node, ast.Import):
                    for a in node.names:
                        imported[(a.asname or a.name).split(".")[0]] = node.lineno
                elif isinstance(node, ast.ImportFrom):
                    for a in node.names:
                        if a.name != "*":
                            imported[a.asname or a.name] = node.lineno
            used = {n.id for n in ast.walk(tree) if isinstance(n, ast.Name)} | \
                   {n.value.id for n in ast.walk(tree) if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name)}
            exported = set()
            for node in tree.body:
                if isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id == "__all__" for t in node.targets):
                    exported = {e.value for e in getattr(node.value, "elts", []) if isinstance(e, ast.Constant)}
            lines = src.split("\n")
            for name, ln in sorted(imported.items(), key=lambda kv: kv[1]):
                if name in used or name in exported or f == "__init__.py" or "noqa" in lines[ln - 1]:
                    continue
                if name in ("annotations",):
                    continue
                print(f"{os.path.relpath(path, ROOT)}:{ln}: '{name}' imported but unused")
                bad += 1
            seen = {}
            for node in tree.body:
                if isinstance(node, (ast.FunctionDef, ast.ClassDef)):
                    if node.name in seen:
                        print(f"{os.path.relpath(path, ROOT)}:{node.lineno}: '{node.name}' redefined (first at {seen[node.name]})")
                        bad += 1
                    seen[node.name] = node.lineno
sys.exit(1 if bad else 0)
