"""Rough undefined-name check (pyflakes is not available offline): for every function, names that are read but bound
nowhere -- not a parameter, local, enclosing-function local, module-level name or builtin.  Catches typos in code paths
the CPU test-suite cannot execute (CUDA-only branches, GPU scripts)."""
import ast
import builtins
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILTINS = set(dir(builtins)) | {"__file__", "__name__", "__doc__"}


def bound_names(node):
    """Names bound directly in this scope (not descending into nested function/class scopes' bodies)."""
    out = set()

    def visit(n, top=True):
        if isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)) and not top:
            out.add(n.name)
            return
        if isinstance(n, ast.Lambda) and not top:
            return
        if isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
            a = n.args
            for arg in a.posonlyargs + a.args + a.kwonlyargs:
                out.add(arg.arg)
            if a.vararg:
                out.add(a.vararg.arg)
            if a.kwarg:
                out.add(a.kwarg.arg)
        if isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
            out.add(n.id)
        elif isinstance(n, (ast.Import, ast.ImportFrom)):
            for al in n.names:
                out.add((al.asname or al.name).split(".")[0])
        elif isinstance(n, ast.ExceptHandler) and n.name:
            out.add(n.name)
        elif isinstance(n, (ast.Global, ast.Nonlocal)):
            out.update(n.names)
        elif isinstance(n, ast.NamedExpr):
            out.add(n.target.id)
        for c in ast.iter_child_nodes(n):
            visit(c, False)
    visit(node)
    return out


def check(path):
    tree = ast.parse(open(path).read())
    module_names = bound_names(tree) | BUILTINS
    problems = []

    def walk(fn, enclosing):
        scope = enclosing | bound_names(fn)

        def visit(n):
            for c in ast.iter_child_nodes(n):
                if isinstance(c, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                    walk(c, scope)
                    continue
                if isinstance(c, ast.ClassDef):
                    walk(c, scope)
                    continue
                if isinstance(c, (ast.ListComp, ast.SetComp, ast.DictComp, ast.GeneratorExp)):
                    pass                                  # targets are found by bound_names (treated as local: fine)
                if isinstance(c, ast.Name) and isinstance(c.ctx, ast.Load) and c.id not in scope:
                    problems.append((c.lineno, c.id))
                visit(c)
        visit(fn)
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            walk(node, module_names)
    # module-level statements (lambdas in them get their own scope)
    class _Top(ast.AST):
        _fields = ("body",)
    top = _Top()
    top.body = [n for n in tree.body if not isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef))]
    walk(top, module_names)
    return sorted(set(problems))


bad = 0
for base in ("dcnn_b200", "examples", "benchmarks", "scripts", "baseline", "bench.py", "__graft_entry__.py", "tests"):
    p = os.path.join(ROOT, base)
    files = [p] if p.endswith(".py") else [os.path.join(dp, f) for dp, _, fs in os.walk(p) for f in fs
                                           if f.endswith(".py") and "_ref" not in dp]
    for f in files:
        for ln, name in check(f):
            print(f"{os.path.relpath(f, ROOT)}:{ln}: undefined name '{name}'")
            bad += 1
sys.exit(1 if bad else 0)
