"""Static hygiene: no unused imports / duplicate definitions / undefined names anywhere in the Python tree -- the
undefined-name pass matters for CUDA-only branches that the CPU suite cannot execute."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_no_unused_imports_or_redefinitions():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "lint_unused.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout


def test_no_undefined_names():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "lint_undefined.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout


def test_c_abi_call_sites_match_prototypes():
    """ctypes checks nothing: argument counts of every dcnn_* call vs include/dcnn_b200.h / csrc/host, and header vs
    definitions."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "lint_cabi.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout


def test_host_header_is_in_sync_with_the_definitions():
    """include/dcnn_host.h is generated from csrc/host/*.cpp (scripts/gen_host_header.py); it must be current, and it
    must compile as C99 and as C++."""
    import shutil
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "gen_host_header.py"), "--check"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    if shutil.which("gcc"):
        for std, flag in (("c", "-std=c99"), ("c++", "-std=c++17")):
            r = subprocess.run(["gcc", "-x", std, flag, "-Wall", "-Werror", "-fsyntax-only", "-I",
                                os.path.join(root, "include"), "-"], input='#include "dcnn_host.h"\nint main(void) { return 0; }\n',
                               capture_output=True, text=True)
            assert r.returncode == 0, r.stderr
