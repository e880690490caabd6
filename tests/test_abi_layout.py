"""The ctypes mirrors in signalizer_b200/_abi.py against the C structs of include/sgz.h: a small C program prints
sizeof and the offset of every field; any drift between the header and the Python binding fails here, on the CPU."""
import ctypes as C
import os
import subprocess

from signalizer_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAIRS = {
    "sgz_spectrum_config": _abi.SpectrumConfig, "sgz_spectrum_info": _abi.SpectrumInfo, "sgz_tracker_result": _abi.TrackerResult,
    "sgz_tracker_state": _abi.TrackerState, "sgz_vscope_config": _abi.VScopeConfig, "sgz_vscope_filters": _abi.VScopeFilters,
    "sgz_scope_resample_config": _abi.ScopeResampleConfig, "sgz_scope_trigger_state": _abi.ScopeTriggerState,
    "sgz_scope_bin_record": _abi.ScopeBinRecord, "sgz_scope_period_state": _abi.ScopePeriodState,
    "sgz_scope_period_params": _abi.ScopePeriodParams, "sgz_scope_period_result": _abi.ScopePeriodResult,
    "sgz_scope_colour_config": _abi.ScopeColourConfig, "sgz_scope_colour_state": _abi.ScopeColourState,
}


def test_struct_layouts_match_the_header(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{ROOT}/include/sgz.h"', "int main(void) {"]
    for cname, cls in PAIRS.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(src)], check=True)   # a field the header does not have fails to compile
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    seen = 0
    for line in out.splitlines():
        cname, what, value = line.split()
        cls = PAIRS[cname]
        want = C.sizeof(cls) if what == "size" else getattr(cls, what).offset
        assert int(value) == want, (cname, what, value, want)
        seen += 1
    assert seen == sum(1 + len(c._fields_) for c in PAIRS.values())


def test_field_counts_match_the_header():
    # every member the header declares is mirrored (a member missing from the ctypes class would shift nothing above)
    import re
    text = open(os.path.join(ROOT, "include", "sgz.h")).read()
    for cname, cls in PAIRS.items():
        body = re.search(r"typedef struct " + cname + r"\s*\{(.*?)\}\s*" + cname + r"\s*;", text, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        members = 0
        for decl in body.split(";"):
            decl = decl.strip()
            if decl:
                members += len(decl.split(","))
        assert members == len(cls._fields_), (cname, members, len(cls._fields_))


def test_argument_counts_match_the_header():
    """every prototype in the header against the argtypes the ctypes loader declares for it"""
    import re
    from signalizer_b200 import _lib
    lib = _lib.load()
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "sgz.h")).read(), flags=re.S)
    protos = re.findall(r"\b(sgz_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, re.S)
    assert len(protos) >= len(_lib.EXPORTS)
    checked = 0
    for name, params in protos:
        fn = getattr(lib, name)
        if fn.argtypes is None:
            continue
        params = params.strip()
        n = 0 if params in ("", "void") else len(params.split(","))
        assert n == len(fn.argtypes), (name, n, len(fn.argtypes))
        checked += 1
    assert checked >= 40
