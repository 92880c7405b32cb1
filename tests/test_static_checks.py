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
