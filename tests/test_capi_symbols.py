"""The C-ABI library loads and exports every symbol include/gsmfret_c.h declares; on a box
without a GPU the compute entry points refuse loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "gsmfret_c.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gsm_[A-Za-z0-9_]+)\s*\(", txt)))


def test_header_declares_the_mc_surface():
    syms = header_symbols()
    for must in ("gsm_create", "gsm_destroy", "gsm_set_device", "gsm_init_data", "gsm_set_nstates",
                 "gsm_set_params", "gsm_set_burst_bd", "gsm_init_randstate", "gsm_set_params_buff",
                 "gsm_run", "gsm_query", "gsm_result_mcE", "gsm_result_hist", "gsm_result_chi2",
                 "gsm_run_batch", "gsm_comm_init", "gsm_get_max_conversion_capacity"):
        assert must in syms


def test_library_exports_every_declared_symbol(gsm):
    lib = gsm.load()
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing
    import re
    declared = int(re.search(r"#define\s+GSM_ABI_VERSION\s+(\d+)", open(os.path.join(ROOT, "include", "gsmfret_c.h")).read()).group(1))
    assert lib.gsm_abi_version() == declared >= 3     # 2: exchange stamps + shared-device flag, 3: gsm_run_coalesced


def test_python_binding_types_every_symbol(gsm):
    assert sorted(gsm.SIGNATURES) == header_symbols()


def test_no_cpu_fallback(gsm):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = gsm.load()
    assert lib.gsm_device_count() == 0
    h = C.c_void_p()
    rc = lib.gsm_create(0, 2, 0, C.byref(h))
    assert rc == -3 and not h.value                       # GSM_ENODEVICE
    assert b"no CPU path" in lib.gsm_last_error()
    with pytest.raises(gsm.GsmError):
        gsm.Mc(0, 2)


def test_product_never_touches_the_oracle():
    """Nothing under gsmfretda_b200/ may import, link or call oracle/."""
    pkg = os.path.join(ROOT, "gsmfretda_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_py" not in src and "pda_oracle" not in src and "libref_cpu" not in src, f
