"""The C-ABI library without a GPU: it loads, exports every symbol include/beholder_gpu.h declares,
fails loudly when there is no device (no CPU fallback), validates configurations, and its host-side
helpers (window division, synthetic stream generator) behave."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import beholder_b200 as bh
from beholder_b200 import _lib as L
from beholder_b200.synth import SynthSpec, generate_host, zipf_cdf
from conftest import HAS_GPU, ROOT


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "beholder_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bh_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    declared = _declared_functions()
    assert len(declared) >= 25
    out = subprocess.check_output(["nm", "-D", "--defined-only", L.LIB_PATH], text=True)
    exported = set(re.findall(r"\bT (bh_[a-z0-9_]+)", out))
    assert set(declared) <= exported, sorted(set(declared) - exported)
    assert set(declared) == set(L.PROTOTYPES), (sorted(set(declared) ^ set(L.PROTOTYPES)))


def test_struct_sizes_match_the_header():
    # compile a tiny C program against the header and compare sizeof
    import tempfile
    code = '#include "beholder_gpu.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu", sizeof(bh_config), sizeof(bh_record), sizeof(bh_counters), sizeof(bh_synth_config), sizeof(bh_field));return 0;}'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(code)
        subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(d, "t")], text=True).split()]
    assert sizes == [C.sizeof(L.BhConfig), C.sizeof(L.BhRecord), C.sizeof(L.BhCounters), C.sizeof(L.BhSynthConfig), C.sizeof(L.BhField)]


def test_js_wrapper_config_offsets_match_the_header():
    """addon/beholder-gpu.js packs a bh_config by hand: its OFF table against offsetof() of the C struct"""
    import tempfile
    js = open(os.path.join(ROOT, "addon", "beholder-gpu.js")).read()
    table = re.search(r"const OFF = \{([^}]*)\}", js).group(1)
    off = {k: int(v) for k, v in re.findall(r"(\w+):\s*(\d+)", table)}
    names = [k for k in off if k not in ("SIZE", "FIELD_SIZE")]
    prints = "".join(f'printf("{n} %zu\\n", offsetof(bh_config, {n}));' for n in names)
    code = ('#include "beholder_gpu.h"\n#include <stdio.h>\n#include <stddef.h>\nint main(){' + prints +
            'printf("SIZE %zu\\nFIELD_SIZE %zu\\n", sizeof(bh_config), sizeof(bh_field));return 0;}')
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(code)
        subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")])
        want = {k: int(v) for k, v in (line.split() for line in subprocess.check_output([os.path.join(d, "t")], text=True).splitlines())}
    assert off == want


@pytest.mark.skipif(HAS_GPU, reason="a GPU is visible")
def test_no_device_is_a_loud_error_not_a_fallback():
    assert L.lib.bh_device_count() == 0
    with pytest.raises(bh.BeholderError) as e:
        bh.Aggregator(mode=L.MODE_REFERENCE)
    assert e.value.code == L.BH_ERR_NO_DEVICE
    assert "no CPU path" in str(e.value)


def test_config_validation_happens_before_touching_the_device():
    def create(**kw):
        cfg = L.BhConfig(); cfg.struct_size = C.sizeof(cfg)
        for k, v in kw.items():
            setattr(cfg, k, v)
        ctx = C.c_void_p()
        return L.lib.bh_create(C.byref(cfg), C.byref(ctx)), (L.lib.bh_last_error(None) or b"").decode()
    rc, msg = create(struct_size=12)
    assert rc == L.BH_ERR_INVALID and "struct_size" in msg
    rc, msg = create(mode=L.MODE_REFERENCE)
    assert rc == L.BH_ERR_INVALID and "BH_R_STATUS" in msg
    rc, msg = create(mode=7)
    assert rc == L.BH_ERR_INVALID
    cfg = L.BhConfig(); cfg.struct_size = C.sizeof(cfg); cfg.mode = L.MODE_WINDOWED; cfg.n_fields = 1
    cfg.fields[0].number = 99; cfg.fields[0].kind = L.K_STRING
    ctx = C.c_void_p()
    assert L.lib.bh_create(C.byref(cfg), C.byref(ctx)) == L.BH_ERR_UNSUPPORTED
    cfg.fields[0].number = 3; cfg.fields[0].kind = 42
    assert L.lib.bh_create(C.byref(cfg), C.byref(ctx)) == L.BH_ERR_UNSUPPORTED
    cfg.fields[0].kind = L.K_DOUBLE; cfg.fields[0].role = L.R_KEY
    assert L.lib.bh_create(C.byref(cfg), C.byref(ctx)) == L.BH_ERR_UNSUPPORTED


def test_window_division_is_exact():
    rng = np.random.default_rng(11)
    ds = [1, 2, 3, 7, 10, 1000, 10_000, 4096, 60_000, 86_400_000, (1 << 32) - 1, (1 << 32) + 1, (1 << 63) - 25, (1 << 64) - 1]
    ns = [0, 1, 9999, 10_000, 10_001, 1_700_000_000_000, (1 << 53) + 1, (1 << 63), (1 << 64) - 1]
    ns += [int(x) for x in rng.integers(0, 1 << 63, 300)] + [int(x) * 2 + 1 for x in rng.integers(0, 1 << 63, 300)]
    ds += [int(x) for x in rng.integers(1, 1 << 40, 200)]
    for d in ds:
        for n in ns:
            assert L.lib.bh_debug_div64(n, d) == n // d, (n, d)


def test_synth_generator_is_deterministic_and_well_formed():
    spec = SynthSpec(n_keys=100, rate_per_ms=4, jitter_ms=100, payload_bytes=64)
    a = generate_host(spec, 50, 400)
    b = generate_host(spec, 50, 400)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    slab, off, ts = a
    assert (np.diff(off) == 64).all()
    # sub-ranges are consistent with the whole (events are a pure function of their global index)
    s2, o2, t2 = generate_host(spec, 150, 100)
    assert np.array_equal(t2, ts[100:200])
    assert np.array_equal(s2[: o2[-1]], slab[off[100]: off[200]])
    # time stamps follow t0 + g / rate +- jitter
    g = np.arange(50, 450, dtype=np.int64)
    assert (np.abs(ts.astype(np.int64) - (spec.t0_ms + g // 4)) <= 100).all()
    c = zipf_cdf(1000)
    assert c[-1] == np.uint64(0xFFFFFFFFFFFFFFFF) and (np.diff(c.astype(np.float64)) >= 0).all()
    z = generate_host(SynthSpec(n_keys=1000, zipf=True, payload_bytes=0), 0, 300, c)
    assert len(z[1]) == 301
