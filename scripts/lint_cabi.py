"""ctypes does not check argument counts.  This compares every ``<lib>.dcnn_*(...)`` call in the Python tree with the
prototype in include/dcnn_b200.h (device library) or the DCNN_HOST_API definitions in csrc/host/*.cpp (host library):
the number of arguments must match, and every called symbol must exist."""
import ast
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def param_list(s):
    s = s.strip()
    if s in ("", "void"):
        return []
    depth, cur, out = 0, "", []
    for ch in s:
        if ch in "([":
            depth += 1
        elif ch in ")]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    out.append(cur.strip())
    return out


def split_params(s):
    return len(param_list(s))


def kind(param):
    """Expected ctypes class of a C parameter: 'ptr', 'int', 'll', 'ull', 'float', 'double' or None (unknown)."""
    p = re.sub(r"\s+", " ", param)
    if "*" in p or "[" in p or "cudaStream_t" in p:
        return "ptr"
    if re.search(r"\bunsigned long long\b|\buint64_t\b|\bsize_t\b", p):
        return "ull"
    if re.search(r"\blong long\b|\bint64_t\b", p):
        return "ll"
    if re.search(r"\bdouble\b", p):
        return "double"
    if re.search(r"\bfloat\b", p):
        return "float"
    if re.search(r"\b(int|unsigned|uint32_t|int32_t)\b", p):
        return "int"
    return None


WRAPPERS = {"I": "int", "LL": "ll", "ULL": "ull", "F": "float", "D": "double", "P": "ptr", "_p": "ptr", "_ptr": "ptr",
            "_stream": "ptr", "c_int": "int", "c_longlong": "ll", "c_float": "float", "c_double": "double",
            "c_void_p": "ptr", "byref": "ptr", "c_ulonglong": "ull"}
ptypes = {}


protos = {}
hdr = open(os.path.join(ROOT, "include", "dcnn_b200.h")).read()
hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
hdr = re.sub(r"//[^\n]*", "", hdr)
for m in re.finditer(r"\b(dcnn_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
    protos[m.group(1)] = split_params(m.group(2))
    ptypes[m.group(1)] = [kind(x) for x in param_list(m.group(2))]
hostdir = os.path.join(ROOT, "csrc", "host")
for f in os.listdir(hostdir):
    if f.endswith(".cpp"):
        src = re.sub(r"//[^\n]*", "", open(os.path.join(hostdir, f)).read())
        for m in re.finditer(r"DCNN_HOST_API\s+[\w\s\*]+?\b(dcnn_\w+)\s*\(([^{;]*?)\)\s*\{", src, flags=re.S):
            protos[m.group(1)] = split_params(m.group(2))
            ptypes[m.group(1)] = [kind(x) for x in param_list(m.group(2))]

bad = 0
for base in ("dcnn_b200", "examples", "benchmarks", "scripts", "tests", "bench.py", "__graft_entry__.py"):
    p = os.path.join(ROOT, base)
    files = [p] if p.endswith(".py") else [os.path.join(dp, f) for dp, _, fs in os.walk(p) for f in fs if f.endswith(".py")]
    for f in files:
        tree = ast.parse(open(f).read())
        for n in ast.walk(tree):
            if isinstance(n, ast.Call) and isinstance(n.func, ast.Attribute) and n.func.attr.startswith("dcnn_"):
                name = n.func.attr
                if name not in protos:
                    print(f"{os.path.relpath(f, ROOT)}:{n.lineno}: {name} is not declared in the C ABI")
                    bad += 1
                    continue
                if any(isinstance(a, ast.Starred) for a in n.args) or n.keywords:
                    continue
                if len(n.args) != protos[name]:
                    print(f"{os.path.relpath(f, ROOT)}:{n.lineno}: {name} called with {len(n.args)} arguments, "
                          f"prototype has {protos[name]}")
                    bad += 1
                    continue
                for i, (a, want) in enumerate(zip(n.args, ptypes.get(name, []))):
                    if want is None or not isinstance(a, ast.Call):
                        continue
                    fn = a.func.attr if isinstance(a.func, ast.Attribute) else getattr(a.func, "id", None)
                    got = WRAPPERS.get(fn)
                    if got is not None and got != want and not (got == "ll" and want == "ull") \
                            and not (got == "ull" and want == "ll"):
                        print(f"{os.path.relpath(f, ROOT)}:{n.lineno}: {name} argument {i + 1} is passed as {got}, "
                              f"the prototype wants {want}")
                        bad += 1
# every DCNN_API definition must be declared in the header with the same parameter count, and vice versa
defs = {}
for f in os.listdir(os.path.join(ROOT, "csrc")):
    if f.endswith(".cu"):
        src = re.sub(r"//[^\n]*", "", open(os.path.join(ROOT, "csrc", f)).read())
        for m in re.finditer(r"DCNN_API\s+[\w\s\*]+?\b(dcnn_\w+)\s*\(([^{;]*?)\)\s*\{", src, flags=re.S):
            defs[m.group(1)] = (split_params(m.group(2)), f)
declared = {k for k in protos if k in set(re.findall(r"\b(dcnn_\w+)\s*\(", hdr))}
for name, (n, f) in sorted(defs.items()):
    if name not in declared:
        print(f"csrc/{f}: {name} is defined but not declared in include/dcnn_b200.h")
        bad += 1
    elif protos[name] != n:
        print(f"csrc/{f}: {name} has {n} parameters, the header declares {protos[name]}")
        bad += 1
for name in sorted(declared - set(defs)):
    print(f"include/dcnn_b200.h: {name} is declared but never defined")
    bad += 1
print(f"{len(protos)} prototypes, {len(defs)} device-library definitions checked")
sys.exit(1 if bad else 0)
