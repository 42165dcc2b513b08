"""CPU tests of the small decision logic: algorithm autoselect, header/binding consistency, build entry."""
import json
import os
import re

import pytest

from b200dp.parallel import autoselect as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True)
def _fresh_tables(monkeypatch):
    monkeypatch.delenv("B200DP_ALLREDUCE_ALGO", raising=False)
    A.reset()
    yield
    A.reset()


def test_autoselect_uses_measured_table(tmp_path, monkeypatch):
    t = tmp_path / "table.json"
    t.write_text(json.dumps({"oneshot_max_bytes": {"4": 1000, "8": 500}, "nvls_min_bytes": {"8": 1 << 20}}))
    monkeypatch.setenv("B200DP_ALLREDUCE_TABLE", str(t))
    A.reset()
    assert A.choose(400, 8) == A.ONESHOT and A.choose(600, 8) == A.TWOSHOT and A.choose(600, 4) == A.ONESHOT
    assert A.choose(2 << 20, 8, nvls_ok=True) == A.NVLS and A.choose(2 << 20, 8, nvls_ok=False) == A.TWOSHOT
    assert A.choose(2 << 20, 4, nvls_ok=True) == A.TWOSHOT                 # no NVLS entry for 4 ranks
    assert A.choose(10 ** 9, 1) == A.ONESHOT                               # single rank: nothing to choose


def test_autoselect_forced_and_env(monkeypatch):
    assert A.choose(100, 8, "twoshot") == A.TWOSHOT and A.choose(10 ** 9, 8, "oneshot") == A.ONESHOT
    assert A.choose(10 ** 7, 8, "nvls", nvls_ok=True) == A.NVLS
    assert A.choose(10 ** 7, 8, "nvls", nvls_ok=False) == A.TWOSHOT        # hardware says no: peer two-shot instead
    assert A.choose(1000, 8, "nvls", nvls_ok=True) == A.ONESHOT            # tiny ranges never take the switch path
    monkeypatch.setenv("B200DP_ALLREDUCE_ALGO", "twoshot")
    assert A.choose(8, 2) == A.TWOSHOT


def test_repo_table_is_well_formed():
    d = json.load(open(os.path.join(ROOT, "profiles", "allreduce_table.json")))
    assert {"2", "4", "8"} <= set(d["oneshot_max_bytes"]) and all(int(v) > 0 for v in d["oneshot_max_bytes"].values())


def test_every_header_entry_point_has_a_ctypes_prototype():
    from b200dp.ops import _sigs
    hdr = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    bound = set(_sigs._SIGS)
    assert declared - bound == set(), f"declared in b200dp.h but not bound: {sorted(declared - bound)}"
    assert bound - declared == set(), f"bound but not declared: {sorted(bound - declared)}"


def test_built_library_exports_every_declared_symbol():
    import subprocess
    so = os.path.join(ROOT, "b200dp", "lib", "libb200dp.so")
    if not os.path.exists(so):
        pytest.skip("library not built")
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (b200_[a-z0-9_]+)", out))
    hdr = open(os.path.join(ROOT, "include", "b200dp.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared - exported == set(), f"declared but not exported: {sorted(declared - exported)}"


def test_sass_contains_blackwell_tensor_and_tma_instructions():
    """The build itself proves the tcgen05/TMA path: SASS of the library carries UTCHMMA / UTMALDG / LDTM."""
    import shutil
    import subprocess
    so = os.path.join(ROOT, "b200dp", "lib", "libb200dp.so")
    if not os.path.exists(so) or not shutil.which("cuobjdump"):
        pytest.skip("library or cuobjdump missing")
    sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    for mnem in ("UTCHMMA", "UTMALDG", "LDTM", "UBLKRED", "LDGMC"):
        assert mnem in sass, mnem
    assert "HMMA.16816" not in sass            # no legacy mma.sync tensor path anywhere
