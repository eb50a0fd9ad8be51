"""CPU suite: libsumob200.so loads without a GPU and exports every symbol include/sumob200.h declares; without a CUDA
device every computing entry point fails loudly (there is no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "sumob200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sumob200_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from sumo_b200 import build
    from sumo_b200.engine import load_library
    build.build()
    lib = load_library()
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "include/sumob200.h declares %s but libsumob200.so does not export it" % n


def test_no_cpu_fallback():
    import shutil
    if shutil.which("nvidia-smi") and os.system("nvidia-smi -L >/dev/null 2>&1") == 0:
        pytest.skip("a GPU is present")
    from sumo_b200.engine import Engine, SumoB200Error
    flat = os.path.join(ROOT, "tests", "golden", "bench_grid10_tls.flat")
    with pytest.raises(SumoB200Error) as ei:
        Engine(flat)
    assert "no CUDA device" in str(ei.value) or "CUDA" in str(ei.value)


def test_bad_arguments_are_rejected():
    from sumo_b200.engine import load_library
    lib = load_library()
    out = ctypes.c_void_p()
    assert lib.sumob200_create_from_file(b"/nonexistent.flat", 0, ctypes.byref(out)) < 0
    assert lib.sumob200_step(None, 1, None) < 0
    assert lib.sumob200_num_vehicles(None) == 0


def test_unsupported_scenarios_are_refused_at_load(tmp_path):
    """a scenario whose junction logic draws random numbers inside the step (all-way stop, right-before-left) must be refused by
    sumob200_create, not simulated approximately; checked on the host side of the library (Model::build runs before any CUDA call)"""
    import numpy as np
    from sumo_b200.engine import Engine, SumoB200Error
    from sumo_b200.flatfile import load_flat
    base = load_flat(os.path.join(ROOT, "tests", "golden", "grid3_small.golden.flat"))
    for state, word in ((ord("w"), "allway_stop"), (ord("="), "right_before_left"), (ord("Z"), "zipper")):
        sc = dict(base)
        ls = base["link_state"].copy()
        ls[0] = state
        sc["link_state"] = ls
        with pytest.raises(SumoB200Error) as ei:
            Engine(sc)
        assert word in str(ei.value), str(ei.value)
    sc = dict(base)
    sc["lane_flags"] = base["lane_flags"] | 8   # LF_LANECHANGER: an edge with several lanes
    with pytest.raises(SumoB200Error):
        Engine(sc)


def test_malformed_arrays_are_einval(tmp_path):
    """the C ABI range-checks every index array and every CSR offset array it is handed (Model::validate): a malformed scenario
    is SUMOB200_EINVAL with a message naming the array, before any device memory is touched"""
    from sumo_b200.engine import Engine, SumoB200Error
    from sumo_b200.flatfile import load_flat
    base = load_flat(os.path.join(ROOT, "tests", "golden", "grid3_small.golden.flat"))
    bad = []
    for name, edit in (("link_to", lambda a: a.__setitem__(3, len(base["lane_length"]) + 5)),
                       ("route_edges", lambda a: a.__setitem__(0, -7)),
                       ("veh_type", lambda a: a.__setitem__(0, 99)),
                       ("lane_link_begin", lambda a: a.__setitem__(2, a[3] + 1)),
                       ("foelane", lambda a: a.__setitem__(0, 1 << 30)),
                       ("tllink_link", lambda a: a.__setitem__(0, -1) if len(a) else None),
                       ("veh_route", lambda a: a.__setitem__(1, len(base["route_begin"])))):
        sc = dict(base)
        arr = base[name].copy()
        if len(arr) == 0:
            continue
        edit(arr)
        sc[name] = arr
        with pytest.raises(SumoB200Error) as ei:
            Engine(sc)
        assert "(-1)" in str(ei.value) and name in str(ei.value), str(ei.value)
        bad.append(name)
    assert len(bad) >= 5
