"""CPU-side checks of the drop-in boundary: libcmbpo_b200.so loads without a GPU, exports every symbol
include/cmbpo_b200.h declares, the ctypes mirrors have the C structs' layout, and the product path
fails loudly (no CPU fallback) when no sm_100 device is present."""
import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "cmbpo_b200.h")


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    if not os.path.exists(ge.LIB):
        ge.build_library()
    from constrainedmbpo_b200 import _lib
    return _lib.load()


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cmbpo_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from constrainedmbpo_b200 import _lib
    names = _declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cmbpo_b200.h but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes prototype in _lib.SYMBOLS"
    assert sorted(_lib.SYMBOLS) == names


def test_ctypes_struct_layout_matches_header():
    from constrainedmbpo_b200 import _lib
    structs = {"cmbpo_cfg": _lib.cmbpo_cfg, "cmbpo_predict_io": _lib.cmbpo_predict_io,
               "cmbpo_step_io": _lib.cmbpo_step_io, "cmbpo_rollout_io": _lib.cmbpo_rollout_io}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void){"]
    for name, st in structs.items():
        lines.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for f, _ in st._fields_:
            lines.append(f'printf("{name}.{f} %zu\\n", offsetof({name}, {f}));')
    lines.append('printf("cmbpo_rollout_state %zu\\n", sizeof(cmbpo_rollout_state));')
    lines.append("return 0;}")
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, "l.c"), os.path.join(d, "l")
        open(src, "w").write("\n".join(lines))
        subprocess.run(["gcc", "-o", exe, src], check=True)
        out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    got = dict(l.split() for l in out.strip().splitlines())
    for name, st in structs.items():
        assert int(got[name]) == C.sizeof(st), name
        for f, _ in st._fields_:
            assert int(got[f"{name}.{f}"]) == getattr(st, f).offset, f"{name}.{f}"
    assert int(got["cmbpo_rollout_state"]) == 48   # 12 int32 words (rollout.py state tensor)


def test_no_gpu_means_loud_failure_not_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import constrainedmbpo_b200 as cm
    assert lib.cmbpo_device_ok(0) == 0
    with pytest.raises(cm.CmbpoError, match="no CPU fallback"):
        cm.construct_model(in_dim=5, out_dim=2, hidden_dims=(8,), num_networks=2, num_elites=1)
    cfg = cm._lib.cmbpo_cfg()
    cfg.num_members, cfg.num_layers, cfg.in_dim, cfg.out_dim, cfg.probabilistic = 2, 2, 5, 2, 1
    cfg.layer_width[0], cfg.layer_width[1] = 8, 4
    cfg.layer_act[0] = 1
    h = C.c_void_p()
    rc = lib.cmbpo_create(C.byref(cfg), 0, C.byref(h))
    assert rc == -3 and b"fallback" in lib.cmbpo_last_error()


def test_create_argument_validation(lib):
    from constrainedmbpo_b200 import _lib
    cfg = _lib.cmbpo_cfg()
    cfg.num_members, cfg.num_layers, cfg.in_dim, cfg.out_dim, cfg.probabilistic = 7, 2, 5, 2, 1
    cfg.layer_width[0], cfg.layer_width[1] = 8, 3     # head must be 2*out_dim
    h = C.c_void_p()
    assert lib.cmbpo_create(C.byref(cfg), 0, C.byref(h)) == -1
    assert b"bnn.py:190-191" in lib.cmbpo_last_error()
    cfg.layer_width[1] = 4
    cfg.num_members = 0
    assert lib.cmbpo_create(C.byref(cfg), 0, C.byref(h)) == -2
    assert lib.cmbpo_create(None, 0, C.byref(h)) == -1


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under constrainedmbpo_b200/ may import or execute it."""
    pkg = os.path.join(ROOT, "constrainedmbpo_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle", txt, flags=re.M), f
                assert "cmbpo_oracle" not in txt, f
    code = ("import sys; sys.path.insert(0, %r); import constrainedmbpo_b200; "
            "assert not any(m.startswith('oracle') for m in sys.modules), 'oracle imported by the product'" % ROOT)
    subprocess.run([sys.executable, "-c", code], check=True)
