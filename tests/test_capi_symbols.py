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
    assert lib.gsm_abi_version() == declared >= 4     # 2: exchange stamps + shared-device flag, 3: gsm_run_coalesced, 4: gsm_launch_shape, 5: gsm_observed_hist_static, gsm_profiler


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


def test_static_entry_point_has_no_cpu_fallback_either(gsm):
    """gsm_observed_hist_static on a box without a GPU: argument checks first (GSM_ERANGE beyond the built-in
    scratch, whatever the box), then a loud GSM_ENODEVICE — the pointers are never touched on the host."""
    import torch
    lib = gsm.load()
    fake_in, fake_out = C.c_void_p(0x1000), C.c_void_p(0x2000)
    assert lib.gsm_observed_hist_static(fake_in, 10, 4097, fake_out, None) == -6      # GSM_ERANGE
    assert b"GSM_STATIC_MAX_BINS" in lib.gsm_last_error()
    assert lib.gsm_observed_hist_static(fake_in, 10, 200, None, None) == -1           # GSM_EINVAL
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert lib.gsm_observed_hist_static(fake_in, 10, 200, fake_out, None) == -3       # GSM_ENODEVICE
    assert b"no CPU path" in lib.gsm_last_error()


def test_product_never_touches_the_oracle():
    """Nothing under gsmfretda_b200/ may import, link or call oracle/."""
    pkg = os.path.join(ROOT, "gsmfretda_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_py" not in src and "pda_oracle" not in src and "libref_cpu" not in src, f


def _shape(lib, N, R, P, blocks, guided):
    out = (C.c_int64 * 6)()
    assert lib.gsm_launch_shape(N, R, P, blocks, guided, out) == 0
    return dict(zip(("chunk", "span", "tail_log2", "big", "per_p", "blocks"), [int(v) for v in out]))


@pytest.mark.parametrize("N,R,P,blocks,guided", [
    (7733, 4, 1, 592, 0), (100_000, 100, 1, 592, 0), (100_000, 100, 1, 888, 1), (1_000_000, 16, 1, 888, 1),
    (1_250_000, 4, 1, 888, 1), (10_000_000, 4, 1, 888, 1), (100_000, 16, 1024, 592, 0), (100_000, 16, 128, 592, 0),
    (1, 1, 1, 592, 1), (0, 7, 3, 592, 1), (333, 3, 9, 592, 1), (2049, 1, 1, 592, 1), (65_537, 33, 2, 888, 1),
])
def test_launch_shape_covers_every_sample_once(gsm, N, R, P, blocks, guided):
    """The ticket arithmetic of the persistent grid (pda_eval.cuh, shape_launch), on the host: the chunks of a
    vector — big ones, then the guided tail of small ones — tile [0, N*R) exactly, spans stay within
    bounds, and no more blocks are launched than there are chunks to hand out."""
    s = _shape(gsm.load(), N, R, P, blocks, guided)
    per_p = N * R
    assert s["chunk"] in (32, 64, 128, 256) and 1 <= s["span"] <= 32
    assert 0 <= s["big"] <= s["per_p"]
    tail = 1 << s["tail_log2"]
    assert tail <= s["chunk"]
    covered = 0
    for c in range(s["per_p"]):                      # the kernel's decode of chunk c of a vector
        t0 = c * s["chunk"] if c < s["big"] else s["big"] * s["chunk"] + (c - s["big"]) * tail
        length = s["chunk"] if c < s["big"] else tail
        assert t0 == covered
        cnt = min(per_p - t0, length)
        assert cnt > 0
        covered += cnt
    assert covered == per_p
    if not (guided and P == 1):
        assert s["big"] == s["per_p"]                # no tail without the guided schedule or for populations
    assert 1 <= s["blocks"] <= blocks
    assert s["blocks"] <= max(1, -(-s["per_p"] * P // 8))
    if per_p * P >= 8 * blocks * 8 * 32:
        assert s["blocks"] == blocks                 # big jobs fill the persistent grid
