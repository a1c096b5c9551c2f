"""CPU tests of the product's host side: libssn.so loads, exports every symbol include/ssn.h declares,
builds the reference's model tables bit-exactly (no device needed), and refuses to run without a GPU."""
import ctypes as C
import hashlib
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
TABLE_SETS = json.load(open(os.path.join(GOLD, "tables.json")))


@pytest.fixture(scope="module")
def api(built):
    from somatic_sniper_b200 import api
    return api


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_header_symbols_are_exported(api):
    hdr = open(os.path.join(ROOT, "include", "ssn.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ssn_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    lib = C.CDLL(api.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, "declared in include/ssn.h but not exported: %s" % missing
    assert declared == set(api.SIGNATURES), "python binding and header disagree: %s" % (declared ^ set(api.SIGNATURES))
    assert api.load_library().ssn_abi_version() == int(re.search(r"#define SSN_ABI_VERSION (\d+)", hdr).group(1))


def test_struct_layouts(api):
    """ssn_site / ssn_sample sizes the C side compiled with (216 / 104 bytes) match the numpy mirror."""
    assert api.SAMPLE_DTYPE.itemsize == 104 and api.SITE_DTYPE.itemsize == 216
    assert C.sizeof(api.Params) == 48


def _params(api, kw):
    m = dict(theta="theta", n_hap="n_hap", het_rate="het_rate", use_priors="use_priors", use_joint="use_joint_priors",
             somatic_rate="somatic_mutation_rate")
    return api.default_params(**{m[k]: v for k, v in kw.items()})


@pytest.mark.parametrize("name", sorted(TABLE_SETS))
def test_host_tables_bit_exact_with_reference(api, name):
    """sniper_cal_coef / sniper_cal_het / makeSoloPrior / make_joint_prior / qAddTableInit, built by the product's
    own host code (csrc/ssn_tables.c), against the byte hashes of the reference's tables."""
    meta = TABLE_SETS[name]
    t = api.host_tables(_params(api, meta["params"]))
    assert _sha(t["coef"]) == meta["coef_sha256"]
    assert _sha(t["lhet"]) == meta["lhet_sha256"]
    assert _sha(t["fk"]) == meta["fk_sha256"]
    assert float(t["q_r"]) == meta["q_r"]
    small = np.load(os.path.join(GOLD, "tables_small.npz"))
    for k in ("prior", "jointprior", "qadd", "fk"):
        assert np.array_equal(t[k], small["%s__%s" % (name, k)]), k


def test_defaults_follow_reference_main(api):
    p = api.default_params()          # main.c:69-78, sniper_maqcns.c:102-113
    assert (round(p.theta, 6), round(p.het_rate, 6), round(p.eta, 6)) == (0.85, 0.001, 0.03)
    assert (p.n_hap, p.cap_mapq, p.min_somatic_qual, p.include_loh, p.include_gor, p.use_priors, p.use_joint_priors) == (
        2, 60, 15, 1, 1, 1, 0)
    assert p.somatic_mutation_rate == 0.01


def test_nt16_table(api):
    lut = api.nt16_lut()
    for ch, code in zip("=ACMGRSVTWYHKDBN", range(16)):
        assert lut[ord(ch)] == code
        if ch.isalpha():
            assert lut[ord(ch.lower())] == code
    assert [lut[ord(c)] for c in "0123"] == [1, 2, 4, 8]
    assert lut[ord("x")] == 15 and lut[0] == 15 and lut[ord("*")] == 15


def test_no_cpu_fallback(api):
    """Without a CUDA device the product must fail loudly, not score on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.SsnError) as e:
        api.Scorer(api.default_params(), n_slots=1, max_columns=16, max_entries=256)
    assert "status -2" in str(e.value)      # SSN_ERR_CUDA


def test_bad_arguments(api):
    lib = api.load_library()
    assert lib.ssn_params_default(None) == -1
    assert b"null" in lib.ssn_last_error()
    ctx = C.c_void_p()
    assert lib.ssn_create(C.byref(api.default_params()), 0, 0, 16, 16, C.byref(ctx)) == -1
    bad = api.default_params(n_hap=1)
    t, h = api.Tables(), C.c_void_p()
    assert lib.ssn_host_tables_create(C.byref(bad), C.byref(t), C.byref(h)) == -1


def test_shim_binary_links_and_refuses_to_run_without_a_gpu(tmp_path):
    """tools/shim/ssn_shim.c built against the unmodified reference (oracle/build_ref.sh): the binary exists, resolves
    libssn.so, and - like everything else here - has no CPU scoring path."""
    import shutil, subprocess
    shim = os.path.join(ROOT, "oracle", "_ref", "bam-somaticsniper-shim")
    if not os.path.exists(shim):
        pytest.skip("oracle/_ref/bam-somaticsniper-shim is not built (no /root/reference here)")
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present; tests/test_cli_gpu.py runs the shim for real")
    except ImportError:
        pass
    integ = os.path.join(ROOT, "tests", "golden", "integration")
    fa = str(tmp_path / "small.fa")
    shutil.copy(os.path.join(integ, "small.fa"), fa)
    p = subprocess.run([shim, "-f", fa, os.path.join(integ, "t-small.bam"), os.path.join(integ, "n-small.bam"), str(tmp_path / "o")],
                       stderr=subprocess.PIPE, timeout=120)
    assert p.returncode == 1 and b"[ssn_shim] ssn_create" in p.stderr
