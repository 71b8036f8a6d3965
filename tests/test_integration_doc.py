"""INTEGRATION.md is the only evidence for the Julia side of the boundary (no Julia on either box), so its `ccall`
signatures are checked mechanically against the ctypes table the tests exercise (quantumsavory.jl_b200/_lib.py) and
against include/qsb200.h: same symbol, same argument count, same scalar widths, pointers where pointers are."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _split_top(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "({[":
            depth += 1
        if ch in ")}]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip()); cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def _ccalls():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    calls = []
    for m in re.finditer(r"ccall\(\(:(\w+), LIB\),\s*(\w+),\s*\(", text):
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(text[i], 0)
            i += 1
        calls.append((m.group(1), m.group(2), _split_top(text[m.end(): i - 1])))
    return calls


SCALARS = {"Cint": C.c_int, "Int32": C.c_int32, "Int64": C.c_int64, "UInt64": C.c_uint64, "Cdouble": C.c_double, "Csize_t": C.c_size_t}


def _is_pointer(t):
    return t in (C.c_void_p, C.c_char_p) or hasattr(t, "contents") or (isinstance(t, type) and issubclass(t, C._Pointer))


def test_every_ccall_of_the_glue_matches_the_ctypes_binding():
    from qsb200 import _lib
    calls = _ccalls()
    assert len(calls) >= 25
    table = None
    try:
        _lib.lib(); table = _lib.EXPORTS
    except OSError:
        pytest.skip("libqsb200.so not built")
    seen = set()
    for name, ret, args in calls:
        assert name in table, f"INTEGRATION.md calls {name}, which _lib.py does not bind"
        res, argtypes = table[name]
        seen.add(name)
        want_ret = {"Cint": C.c_int, "Cstring": C.c_char_p, "Int64": C.c_int64}[ret]
        assert res is want_ret, (name, ret, res)
        assert len(args) == len(argtypes), f"{name}: the glue passes {len(args)} arguments, the C function takes {len(argtypes)}: {args}"
        for j, (a, t) in enumerate(zip(args, argtypes)):
            if a.startswith(("Ptr{", "Ref{")) or a == "Cstring":
                assert _is_pointer(t), (name, j, a, t)
            else:
                assert a in SCALARS, (name, j, a)
                assert t is SCALARS[a] or (C.sizeof(t) == C.sizeof(SCALARS[a]) and not _is_pointer(t)), (name, j, a, t)
    # the generics of the boundary and the cohort / program entry points are all shown
    for must in ("qsb_ctx_create", "qsb_state_new", "qsb_state_from_host", "qsb_tab_from_host", "qsb_state_select", "qsb_state_copy", "qsb_tensor",
                 "qsb_apply", "qsb_apply_clifford", "qsb_channel", "qsb_measure", "qsb_ptrace", "qsb_expect", "qsb_state_kill", "qsb_mask_logic",
                 "qsb_record_begin", "qsb_record_end", "qsb_program_run", "qsb_ctx_retire", "qsb_wave_run", "qsb_program_run_lanes",
                 "qsb_cohort_count", "qsb_cohort_scatter"):
        assert must in seen, must


def test_glue_text_has_no_raw_pointer_masks_and_names_the_mask_aware_call_sites():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    assert "is2::Ptr" not in text                                   # round-1 defect: `outcome == 2` returned a raw pointer
    assert re.search(r"Base\.:\(==\)\(o::BatchOutcome, j::Integer\) = .*o\.is2", text) and "struct BatchMask" in text
    for site in ("src/CircuitZoo/CircuitZoo.jl:22-33", "src/CircuitZoo/CircuitZoo.jl:144-161", "src/ProtocolZoo/ProtocolZoo.jl:569-573",
                 "_wrap_state_for_slots", "consistent_representation", "qsb_state_select", "OBS_PROJ_STAB", "OBS_PAULI", "qsb_state_kill"):
        assert site in text, site
