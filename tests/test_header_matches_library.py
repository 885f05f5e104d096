"""Every entry point declared in include/picd_b200.h is exported by the built library and vice versa, and every ctypes signature the Python
side binds (ops/native.py::_SIG) names an exported symbol (CPU-only: reads the ELF symbol table with nm)."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch-imageclassification-distributed_b200", "lib", "libpicd_b200.so")


def _exported():
    if not os.path.exists(LIB):
        import build
        build.build(force=False, verbose=False)
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if " T " in line}


@pytest.mark.skipif(shutil.which("nm") is None or shutil.which("nvcc") is None, reason="needs nm and nvcc")
def test_header_and_bindings_match_library():
    exported = _exported()
    header = open(os.path.join(ROOT, "include", "picd_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(picd_[a-z0-9_]+)\s*\(", header))
    assert len(declared) > 60
    missing = sorted(d for d in declared if d not in exported)
    assert not missing, f"declared in picd_b200.h but not exported: {missing}"
    undeclared = sorted(e for e in exported if e.startswith("picd_") and e not in declared)
    assert not undeclared, f"exported by the library but missing from picd_b200.h: {undeclared}"
    import sys
    sys.path.insert(0, ROOT)
    from picd_b200.ops import native
    unbound = sorted(n for n in native._SIG if n not in exported)
    assert not unbound, f"bound in ops/native.py but not exported: {unbound}"


@pytest.mark.skipif(shutil.which("gcc") is None or not os.path.exists("/usr/local/cuda/include/cuda_runtime.h"),
                    reason="needs gcc and the CUDA headers")
def test_public_header_is_plain_c_and_cpp():
    """Non-Python hosts link the same library: the header must parse as C and as C++."""
    hdr = os.path.join(ROOT, "include", "picd_b200.h")
    for lang, cc in (("c", "gcc"), ("c++", "g++")):
        r = subprocess.run([cc, "-I/usr/local/cuda/include", "-x", lang, "-fsyntax-only", hdr], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]


def _prototypes():
    header = open(os.path.join(ROOT, "include", "picd_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    header = re.sub(r"//[^\n]*", "", header)
    protos = {}
    for m in re.finditer(r"\b(picd_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        params = [p.strip() for p in m.group(2).replace("\n", " ").split(",")]
        protos[m.group(1)] = [] if params in ([""], ["void"]) else params
    return protos


def _kind(param: str) -> str:
    if "*" in param or "cudaStream_t" in param:
        return "p"
    t = param.rsplit(" ", 1)[0] if " " in param else param
    t = t.replace("const", "").strip()
    return {"int": "i", "float": "f", "long long": "l", "unsigned": "u", "unsigned int": "u", "size_t": "z", "uint32_t": "u"}.get(t, "?")


def test_ctypes_signatures_match_the_c_prototypes():
    """ops/native.py binds the library through hand-written signature strings: each must agree with the prototype in the public
    header in arity and in the class of every argument (pointer / int / float / long long / unsigned / size_t)."""
    import sys
    sys.path.insert(0, ROOT)
    from picd_b200.ops import native
    protos = _prototypes()
    bad = []
    for name, sig in native._SIG.items():
        if name not in protos:
            bad.append(f"{name}: no prototype in picd_b200.h")
            continue
        kinds = "".join(_kind(p) for p in protos[name])
        if kinds != sig:
            bad.append(f"{name}: header {kinds} != binding {sig}")
    assert not bad, "\n".join(bad)
