"""CPU-side checks of the drop-in boundary: libshipsim.so loads, exports every symbol include/shipsim.h
declares, the ctypes mirrors have the C sizes, and the product fails loudly without a GPU (no CPU
fallback).  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "shipsim.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sit_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from sac_maritime_ast_b200 import _lib
    L = _lib.lib()
    declared = _declared_symbols()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/shipsim.h but not exported by libshipsim.so"
    assert set(declared) == set(_lib.EXPORTS), "python binding list and header disagree"
    assert L.sit_abi_version() == _lib.ABI_VERSION


def test_ctypes_struct_sizes_match_the_header(tmp_path):
    from sac_maritime_ast_b200 import _lib
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include "shipsim.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    "sizeof(SitShipParams),sizeof(SitMap),sizeof(SitConfig),sizeof(SitField),sizeof(SitFleetIO),"
                    "sizeof(SitEnvIO),sizeof(SitPolicy),sizeof(SitRollout),sizeof(SitSacConfig));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(prog)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(t) for t in (_lib.SitShipParams, _lib.SitMap, _lib.SitConfig, _lib.SitField, _lib.SitFleetIO,
                                  _lib.SitEnvIO, _lib.SitPolicy, _lib.SitRollout, _lib.SitSacConfig)]
    assert got == want


def test_status_strings_are_the_reference_strings():
    from sac_maritime_ast_b200 import _lib
    from oracle import cport
    assert _lib.status_strings() == cport.STATUS_STRINGS


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import sac_maritime_ast_b200 as sit
    sm, tc, ap = sit.default_ship([(0, 0), (3000, 3000)], 0.7)
    with pytest.raises(sit.ShipSimError):
        sit.ShipFleet(sm, tc, ap, 8.0, n_ships=4)
    # and straight through the C ABI
    from sac_maritime_ast_b200 import _lib
    cfg = _lib.SitConfig()
    cfg.abi_version, cfg.kind, cfg.precision, cfg.max_wp = _lib.ABI_VERSION, _lib.KIND_FLEET, _lib.F64, 4
    cfg.ship[0] = sit.ship_params(sm, tc, ap)
    h = C.c_void_p()
    rc = _lib.lib().sit_create(C.byref(cfg), 4, 0, C.byref(h))
    assert rc == -3 and b"no CPU fallback" in _lib.lib().sit_last_error()


def test_config_validation_errors():
    from sac_maritime_ast_b200 import _lib
    cfg = _lib.SitConfig()
    h = C.c_void_p()
    assert _lib.lib().sit_create(C.byref(cfg), 4, 0, C.byref(h)) == -1      # abi_version 0
    cfg.abi_version, cfg.kind, cfg.precision, cfg.max_wp = _lib.ABI_VERSION, 7, 0, 4
    assert _lib.lib().sit_create(C.byref(cfg), 4, 0, C.byref(h)) == -1 and b"kind" in _lib.lib().sit_last_error()


def test_decode_status_matches_reference_layout():
    import sac_maritime_ast_b200 as sit
    d = sit.MultiShipRLEnv.decode_status
    assert d(0) == " |Test ship not in terminal state| |Obstacle ship not in terminal state| "
    assert d(1 << 8) == " |Test ship not in terminal state| |Obstacle ship collides with the terrain| "
    assert d(1 << 6) == " |Test ship not in terminal state| |Obstacle ship reaches endpoint||Obstacle ship not in terminal state| "
    assert d((1 << 2) | (1 << 11)).endswith(" |Ship collision|")
