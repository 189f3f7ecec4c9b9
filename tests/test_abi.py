"""The drop-in boundary without a GPU: libsmartsb200.so builds for sm_100a, loads, exports every entry
point include/smarts_b200.h declares (and nothing the loader does not know), the ctypes mirrors in
smarts_b200/_abi.py have the C layouts, and the product refuses to run without CUDA (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

from conftest import ROOT, load_map, make_cfg
from smarts_b200 import _abi, build, engine

HEADER = os.path.join(ROOT, "include", "smarts_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sb_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    so = build.build()
    lib = C.CDLL(so)
    declared = _declared_functions()
    assert len(declared) >= 17
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in smarts_b200.h but not exported"
    assert sorted(engine.EXPORTS) == declared, "engine.EXPORTS and the header disagree"
    assert lib.sb_abi_version() == _abi.SB_ABI_VERSION
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert {e for e in exported if e.startswith("sb_")} == set(declared)
    assert not any(e.startswith("sbx_") for e in exported)  # the host widening pool stays internal


def test_ctypes_struct_layouts_match_the_header(tmp_path):
    names = ["SbConfig", "SbMap", "SbStatePtrs", "SbObsPtrs", "SbResetState"]
    probe = tmp_path / "probe.c"
    fields = {"SbConfig": ["dt", "nbr_radius", "lidar_wall_height", "dagm_resolution", "events_enable", "not_moving_distance"],
              "SbMap": ["lanepoint_spacing", "lane_length", "road_edges"], "SbObsPtrs": ["ogm", "ev_events", "ev_steps"],
              "SbResetState": ["route", "goal_radius"], "SbStatePtrs": ["alive", "status"]}
    body = "".join(f'printf("{n} %zu\\n", sizeof({n}));\n' for n in names)
    body += "".join(f'printf("{n}.{f} %zu\\n", offsetof({n}, {f}));\n' for n, fs in fields.items() for f in fs)
    probe.write_text(f'#include <stdio.h>\n#include <stddef.h>\n#include "{HEADER}"\nint main(void) {{\n{body}return 0; }}\n')
    exe = tmp_path / "probe"
    subprocess.run(["gcc", str(probe), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.splitlines())
    for n in names:
        assert int(got[n]) == C.sizeof(getattr(_abi, n)), n
    for n, fs in fields.items():
        for f in fs:
            assert int(got[f"{n}.{f}"]) == getattr(getattr(_abi, n), f).offset, f"{n}.{f}"
    for const in ("SB_ABI_VERSION", "SB_MAX_PATHS", "SB_MAX_LOOKAHEAD", "SB_MAX_TRAJ", "SB_CTRL_STATE", "SB_EV_REACHED_GOAL",
                  "SB_EV_AGENTS_ALIVE_DONE", "SB_DRIVEN_PATH_MAX", "SB_ACTION_TRAJECTORY"):
        m = re.search(rf"#define {const} (\d+)", open(HEADER).read())
        assert m and int(m.group(1)) == getattr(_abi, const), const


def test_error_paths_without_a_device():
    lib = engine.load_library()
    h = C.c_void_p()
    cfg = make_cfg(E=1, A=1)
    cfg.abi_version = 99
    assert lib.sb_create(C.byref(cfg), C.byref(h)) == -1 and b"abi_version" in lib.sb_last_error(h)
    lib.sb_destroy(h)
    cfg = make_cfg(E=0, A=1)
    assert lib.sb_create(C.byref(cfg), C.byref(h)) == -1
    lib.sb_destroy(h)
    cfg = make_cfg(E=1, A=1, ogm=(1025, 64, 0.5))  # width 1 .. 1024 (any width in that range is accepted)
    assert lib.sb_create(C.byref(cfg), C.byref(h)) == -1 and b"ogm" in lib.sb_last_error(h)
    lib.sb_destroy(h)
    lib.sb_destroy(None)
    assert lib.sb_step(None, None, None) == -1 and lib.sb_launch_count(None) == 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(engine.SmartsB200Error, match="no CPU fallback"):
        engine.Engine(make_cfg(E=1, A=1), load_map("4lane"))
    with pytest.raises(engine.SmartsB200Error):
        engine.load_library(os.path.join(ROOT, "smarts_b200", "does_not_exist.so"))
    # nothing in the product imports, includes or loads the oracle
    pat = re.compile(r"^\s*(from\s+oracle|import\s+oracle|#\s*include.*oracle)|oracle_py|liboracle|oracle/_", re.M)
    for dirpath, _, files in os.walk(os.path.join(ROOT, "smarts_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp")):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), f
