"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/ggrm.h declares, and fails loudly
(no CPU fallback) when no GPU is present.  No compute calls are made without a GPU."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

import gelex_b200 as g
from gelex_b200 import _lib
import plink_io

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "ggrm.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ggrm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ggrm.h but not exported by libggrm_b200.so"
    assert set(_lib.EXPORTS) == set(syms)


def test_header_is_plain_c_and_stage_calls_fail_loudly_without_gpu(tmp_path):
    # the boundary is a C header: a C99 compiler takes it as is (what a cgo / ctypes / JNI binding would include)
    import shutil
    import subprocess

    if shutil.which("gcc") is None:
        pytest.skip("no C compiler")
    src = tmp_path / "use.c"
    src.write_text('#include "ggrm.h"\nint main(void) { ggrm_config_t c; (void)c; return sizeof(ggrm_config_t) == 120 ? 0 : 1; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{ROOT / 'include'}", str(src), "-o", str(tmp_path / "use")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert subprocess.run([str(tmp_path / "use")]).returncode == 0
    if g.device_count() == 0:  # staging slots live in device memory: no device, no stage (and no host-memory stand-in)
        lib = _lib.load()
        st = ctypes.c_void_p()
        assert lib.ggrm_stage_create(0, 1 << 20, 2, ctypes.byref(st)) == -6 and not st.value
        assert b"no CPU fallback" in lib.ggrm_last_error(None)
        assert lib.ggrm_stage_create(0, 0, 2, ctypes.byref(st)) == -1  # GGRM_ERR_INVALID_ARGUMENT before any device work


def test_config_struct_layout_matches_header():
    # 8-byte aligned C layout of ggrm_config_t: 2*4 + 2*8 + 8 + 8 + 4*4 + 4(+4 pad) + 2*8 + 5*4(+4 pad) + 8 + 2*4
    assert ctypes.sizeof(_lib.Config) == 120


def test_no_gpu_fails_loudly():
    if g.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(g.GrmError) as e:
        g.Accumulator(100, 10)
    assert e.value.status == -6 and "no CPU fallback" in e.value.message
    with pytest.raises(g.GrmError):
        g.decode_chunk(np.zeros((2, 25), dtype=np.uint8), 100)


def test_synth_host_is_deterministic_and_hwe():
    a = g.synth_bed_host(1003, 5, 40, 20261104, 0.01)
    b = g.synth_bed_host(1003, 0, 50, 20261104, 0.01)
    assert np.array_equal(a, b[5:45])  # counter-based: any SNP range regenerates identical bytes
    assert a.shape == (40, 251)
    assert np.all((a[:, -1] >> 6) == 0)  # padding bits of the last byte are 0 (1003 = 4*250 + 3)
    geno = plink_io.decode_payload(a, 1003)
    miss = np.isnan(geno).mean()
    assert 0.003 < miss < 0.03
    p = np.nanmean(geno, axis=0) / 2
    assert np.all(p > 0.02) and np.all(p < 0.56)


def test_plink_helpers_follow_reference_conventions():
    # sorted-ID row order and the later-duplicate-wins projection (sample_manager.cpp:37-40)
    fam = [("fam2", "b"), ("fam1", "z"), ("fam1", "a"), ("fam2", "b")]
    ids, r2o = plink_io.sample_order(fam)
    assert ids == ["fam1\x1fa", "fam1\x1fz", "fam2\x1fb"]
    assert r2o.tolist() == [2, 1, 0, 2]
    # chromosome runs: a chromosome that re-appears yields a second group (grm_work_plan.cpp:54-81)
    assert plink_io.chrom_runs(["1", "1", "2", "1"]) == [("1", 0, 2), ("2", 2, 3), ("1", 3, 4)]
    # BED encoding (tests/bed_fixture.cpp:72-95)
    geno = np.array([[0.0], [1.0], [2.0], [np.nan], [2.0]])
    assert plink_io.encode_matrix(geno).tolist() == [[0b01001011, 0b00000000]]


def test_built_library_holds_blackwell_tensor_and_tma_instructions():
    # what proves a Blackwell-native kernel is in the SASS, not in the PTX names (profiles/r01_sass_mnemonics.md): the SYRK that
    # bench.py runs must hold UTCHMMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM (tcgen05.ld) and UTMALDG (TMA loads), the FP64
    # mode DMMA, and nothing may fall back to the legacy HMMA path
    import shutil
    import subprocess

    if shutil.which("cuobjdump") is None:
        pytest.skip("no cuobjdump")
    sass = subprocess.run(["cuobjdump", "-sass", str(_lib.LIB_PATH)], capture_output=True, text=True, check=True).stdout
    assert "arch = sm_100a" in sass
    funcs = {f.split("\n", 1)[0].strip(): f for f in re.split(r"\n\s*Function : ", sass)[1:]}
    tc2 = [body for name, body in funcs.items() if "syrk_tc_kernelILi2" in name]
    assert len(tc2) == 1
    for mnemonic in ("UTCHMMA.2CTA", "UTCBAR", "LDTM", "UTMALDG", "SYNCS"):
        assert mnemonic in tc2[0], mnemonic
    f64 = [body for name, body in funcs.items() if "syrk_f64_kernel" in name]
    assert len(f64) == 1 and "DMMA" in f64[0]
    assert not any(re.search(r"\bHMMA\b", body) for body in funcs.values())


def test_numpy_synth_matches_library():
    # bench.py's reference arm builds its PLINK triple with tests/synth_np.py so that it never maps libggrm_b200.so;
    # the bytes must be the ones the GPU arm's generator (csrc/synth.h) produces
    import synth_np

    for n, m, s0, seed, miss in ((1001, 300, 0, 7, 0.01), (4000, 130, 1234, 20261106, 0.03), (257, 70, 5, 1, 0.0), (64, 9, 0, 3, 1.0)):
        assert np.array_equal(synth_np.synth_bed(n, s0, m, seed, miss, threads=2, block=64), g.synth_bed_host(n, s0, m, seed, miss))
