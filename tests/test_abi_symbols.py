"""The C-ABI library must load without a GPU and export every symbol the headers declare."""
import ctypes as C
import os
import re

from conftest import ROOT
from linearham_b200 import _abi


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lhh?_[a-z0-9_]+)\s*\(", src)))


def test_device_header_symbols_exported():
    lib = C.CDLL(_abi.LIB_PATH)
    names = _declared("linearham_b200.h")
    assert set(names) == set(_abi.DEVICE_SYMBOLS)
    for n in names:
        assert hasattr(lib, n), n


def test_host_header_symbols_exported():
    lib = C.CDLL(_abi.LIB_PATH)
    names = _declared("linearham_b200_host.h")
    assert names
    for n in names:
        assert hasattr(lib, n), n


def test_no_cpu_fallback_without_device():
    """Without a CUDA device the product must fail loudly, not compute on the CPU."""
    import numpy as np
    from conftest import load_golden
    from linearham_b200 import device
    if device.device_count() > 0:
        return
    fam, _, _, _ = load_golden("catch_a")
    try:
        device.Family(fam)
    except device.DeviceError as e:
        assert "no CUDA device" in str(e) or "CUDA" in str(e)
    else:
        raise AssertionError("Family() succeeded without a GPU")


def test_product_does_not_reference_oracle():
    """linearham_b200/ must never import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "linearham_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cpp", ".cu", ".hpp", ".h", ".cuh")) or fn == "Makefile":
                text = open(os.path.join(dirpath, fn), errors="ignore").read()
                for pat in (r"(from|import)\s+oracle", r"oracle/", r"pyoracle", r"libport", r"libref_linearham"):
                    assert not re.search(pat, text), (pat, os.path.join(dirpath, fn))


def test_command_line_is_built_and_rejects_bad_usage():
    """linearham_b200/linearham: the reference's subcommands and flag names (src/linearham.cpp:268-455)."""
    import subprocess
    from linearham_b200 import host
    cli = os.path.join(os.path.dirname(os.path.abspath(host.__file__)), "linearham")
    assert os.access(cli, os.X_OK), "run __graft_entry__.build()"
    r = subprocess.run([cli, "--help"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0
    for sub in ("--compute-logl", "--sample", "--pipeline"):
        assert sub in r.stdout
        h = subprocess.run([cli, sub, "--help"], capture_output=True, text=True, timeout=60)
        assert h.returncode == 0 and "--yaml-path" in h.stdout and "--hmm-param-dir" in h.stdout
    r = subprocess.run([cli, "--frobnicate"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "ERROR: '--frobnicate' is not a valid subcommand." in r.stderr
    r = subprocess.run([cli, "--pipeline", "--yaml-path", "x"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and r.stderr.startswith("ERROR: ")
    # a missing input file is the constructor's error, not a crash
    r = subprocess.run([cli, "--pipeline", "--yaml-path", "/nonexistent.yaml", "--cluster-ind", "0", "--hmm-param-dir", "/nonexistent",
                        "--input-path", "a", "--output-path", "b"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and r.stderr.startswith("ERROR: ")


def test_repertoire_fails_loudly_without_device():
    from linearham_b200 import device
    if device.device_count() > 0:
        return
    try:
        device.Repertoire(n_slots=2)
    except device.DeviceError as e:
        assert "CUDA" in str(e)
    else:
        raise AssertionError("Repertoire() succeeded without a GPU")
