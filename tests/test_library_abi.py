"""CPU: the C-ABI library loads without a GPU, exports every symbol include/rthc.h declares,
and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "rthc.h")).read()
    return sorted(set(re.findall(r"\b(rthc_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound(hc):
    lib = hc.load_library()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(hc.EXPORTED_SYMBOLS)
    assert lib.rthc_version() >= 100


def test_create_fails_loudly_without_gpu(hc):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    with pytest.raises(hc.HCBackendError) as e:
        hc.HCContext()
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "reportree_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("oracle/", "").lower() or f == "partitioning_HC.py" and False, f


def test_cgmlst_dists_binary_is_built_and_refuses_to_run_without_a_gpu(tmp_path):
    """Seam B2 executable: present after build(); no CPU fallback (exit 1 + message) on a GPU-less host."""
    import subprocess
    import torch
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "reportree_b200", "bin", "cgmlst-dists")
    assert os.path.exists(exe)
    p = tmp_path / "m.tsv"
    p.write_text("FILE\ta\tb\ns1\t1\t2\ns2\t1\t3\n")
    r = subprocess.run([exe, str(p)], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0 and r.stdout.split("\n")[1] == "s1\t0\t1"
    else:
        assert r.returncode == 1 and "no CPU fallback" in r.stderr and r.stdout == ""
