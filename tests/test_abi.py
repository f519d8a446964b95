"""The C-ABI boundary: the shared library loads, exports every symbol include/lsmc_b200.h declares, and the ctypes
mirror (what a P/Invoke host would declare) has the same struct layout as the C header.  No compute calls: CPU only."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

from storage_b200 import _abi, _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "lsmc_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?(?:int|void|char\s*\*|const char\s*\*)\s*\*?\s*(lsmc_[a-z_0-9]+)\s*\(", text, flags=re.M)
    return sorted(set(n for n in names if n != "lsmc_progress_cb"))


@pytest.fixture(scope="module")
def built():
    if not os.path.exists(_native.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "storage_b200", "csrc")])
    return _native.lib()


def test_header_declares_the_documented_entry_points():
    names = declared_functions()
    assert len(names) >= 20
    assert set(names) == set(_native.EXPORTS)


def test_library_loads_and_exports_every_declared_symbol(built):
    for name in declared_functions():
        assert hasattr(built, name), name
    assert b"sm_100a" in built.lsmc_version()


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", str(tmp_path / "absent.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _native.lib()


def test_create_without_a_gpu_reports_cuda_error(built):
    import ctypes
    try:
        C.CDLL("libcuda.so.1")
        pytest.skip("a CUDA driver is present; covered by the gpu tests")
    except OSError:
        pass
    with pytest.raises(_native.LsmcError) as e:
        _native.Context(0)
    assert e.value.status == _abi.LSMC_ECUDA


STRUCTS = {"lsmc_problem": _abi.LsmcProblem, "lsmc_rng": _abi.LsmcRng, "lsmc_layout": _abi.LsmcLayout,
           "lsmc_results": _abi.LsmcResults, "lsmc_timings": _abi.LsmcTimings}


def test_ctypes_mirror_matches_c_layout(tmp_path):
    """Compile a probe against the real header with gcc and compare sizeof/offsetof with the ctypes structs."""
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "lsmc_b200.h"', 'int main(void) {']
    for cname, cls in STRUCTS.items():
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for field, _ in cls._fields_:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, field, cname, field))
    lines += ['printf("LSMC_MAX_FACTORS %d\\n", LSMC_MAX_FACTORS);', 'printf("LSMC_MAX_BASIS %d\\n", LSMC_MAX_BASIS);',
              'printf("LSMC_NUM_TRIGGER_VOLUMES %d\\n", LSMC_NUM_TRIGGER_VOLUMES);',
              'printf("LSMC_ECUDA %d\\n", (int)LSMC_ECUDA);', 'printf("LSMC_RNG_PANELS %d\\n", (int)LSMC_RNG_PANELS);',
              'printf("LSMC_CONSTRAINT_STEP %d\\n", (int)LSMC_CONSTRAINT_STEP);', 'return 0; }']
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for cname, cls in STRUCTS.items():
        assert int(got[cname]) == C.sizeof(cls), cname
        for field, _ in cls._fields_:
            assert int(got["%s.%s" % (cname, field)]) == getattr(cls, field).offset, (cname, field)
    assert int(got["LSMC_MAX_FACTORS"]) == _abi.MAX_FACTORS and int(got["LSMC_MAX_BASIS"]) == _abi.MAX_BASIS
    assert int(got["LSMC_NUM_TRIGGER_VOLUMES"]) == _abi.NUM_TRIGGER_VOLUMES
    assert int(got["LSMC_ECUDA"]) == _abi.LSMC_ECUDA and int(got["LSMC_RNG_PANELS"]) == _abi.RNG_PANELS
    assert int(got["LSMC_CONSTRAINT_STEP"]) == _abi.CONSTRAINT_STEP


def test_product_package_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under storage_b200/ may import, link or execute it."""
    pkg = os.path.join(ROOT, "storage_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liblsmc_oracle" not in text and "from oracle" not in text and "import oracle" not in text, f
                assert "lsmc_oracle.cpp" not in text, f
