"""`danbing-tk-pred`: bias correction downstream of OUT.trkmc.ar.

CPU: the float32 restatement (tests/pred_common.py) is pinned bit for bit against the compiled reference
(oracle/_ref/danbing-tk-pred, built from /root/reference/src/pred.cpp by oracle/Makefile).
GPU: dbt_pred through the C ABI against the restatement, and the drop-in CLI against the live reference binary file for file.
Bit-exact: every float operation is one IEEE add or divide in the reference's order (no tolerance needed)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle_abi import PKG_DIR, abi, ref_bin
from pred_common import make_pred_case, pred_model, read_bias_tsv, read_matrix, eigen_sum_f32

REF_PRED = ref_bin("danbing-tk-pred")
CLI_PRED = os.path.join(PKG_DIR, "bin", "danbing-tk-pred")
needs_ref = pytest.mark.skipif(REF_PRED is None, reason="oracle/_ref/danbing-tk-pred not built")

CASES = [(12, 40, 1), (17, 25, 2), (10, 12, 3), (37, 30, 4), (300, 14, 5)]


def run_ref(c, d, tag="ref"):
    r = subprocess.run([REF_PRED, c["meta"], c["ikfn"], f"{d}/{tag}.raw", f"{d}/{tag}.cor", f"{d}/{tag}.bias.tsv"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-800:]
    return f"{d}/{tag}.raw", f"{d}/{tag}.cor", f"{d}/{tag}.bias.tsv"


@needs_ref
@pytest.mark.parametrize("ns,n_tr,seed", CASES)
def test_model_is_pinned_to_the_compiled_reference(ns, n_tr, seed, tmp_path):
    c = make_pred_case(str(tmp_path), ns, n_tr, seed)
    raw_fn, cor_fn, bias_fn = run_ref(c, str(tmp_path))
    raw, cor, bias = pred_model(c)
    assert read_matrix(raw_fn).tobytes() == raw.tobytes()
    assert read_matrix(cor_fn).tobytes() == cor.tobytes()
    got = read_bias_tsv(bias_fn)                      # printed with 6 significant digits
    assert got.shape == bias.shape and np.allclose(got, bias, rtol=1e-5, atol=0)
    assert float(np.abs(cor - raw).max()) > 1e-3      # the correction does something


def test_eigen_sum_order_differs_from_a_plain_loop_somewhere():
    """the reduction order matters at float32: the model must not degenerate into a left-to-right sum"""
    rng = np.random.default_rng(0)
    diff = 0
    for n in (9, 12, 17, 37, 300):
        a = rng.uniform(0.5, 1.5, n).astype(np.float32)
        seq = np.float32(0)
        for x in a:
            seq = np.float32(seq + x)
        diff += eigen_sum_f32(a) != seq
        assert abs(float(eigen_sum_f32(a)) - float(a.astype(np.float64).sum())) < 1e-4 * n
    assert diff > 0


def call_dbt_pred(c, bias_init=0.0):
    L = abi.load_library()
    ns, nk, n_tr, nik = c["ns"], c["nk"], c["n_tr"], c["nik"]
    inp = abi.PredInput()
    inp.ns, inp.nk, inp.n_tr, inp.n_ik = ns, nk, n_tr, nik
    arr = (abi.u64p * ns)(*[abi.ptr(v, abi.u64p) for v in c["kmc"]])
    inp.kmc = C.cast(arr, C.POINTER(abi.u64p))
    fp = C.POINTER(C.c_float)
    inp.rd = c["rd"].ctypes.data_as(fp)
    inp.nk_cum, inp.nik_cum = abi.ptr(c["nk_cum"], abi.u32p), abi.ptr(c["nik_cum"], abi.u32p)
    inp.iki, inp.ikmc = abi.ptr(c["iki"], abi.u32p), abi.ptr(c["ikmc"], abi.u8p)
    raw = np.zeros((nk, ns), np.float32)              # column-major (ns x nk)
    cor = np.zeros((nk, ns), np.float32)
    bias = np.full((n_tr, ns), bias_init, np.float32)
    ms = np.zeros(2, np.float32)
    rc = L.dbt_pred(0, C.byref(inp), raw.ctypes.data_as(fp), cor.ctypes.data_as(fp), bias.ctypes.data_as(fp), ms.ctypes.data_as(fp))
    return rc, raw.T, cor.T, bias.T, L


@pytest.mark.gpu
@pytest.mark.parametrize("ns,n_tr,seed", CASES)
def test_dbt_pred_equals_the_model_bit_for_bit(ns, n_tr, seed, tmp_path):
    c = make_pred_case(str(tmp_path), ns, n_tr, seed, skip_some=(seed % 2 == 0))
    rc, raw, cor, bias, L = call_dbt_pred(c, bias_init=-7.0)
    assert rc == 0, L.dbt_last_error()
    eraw, ecor, ebias = pred_model(c)
    assert np.ascontiguousarray(raw).tobytes() == eraw.tobytes()
    assert np.ascontiguousarray(cor).tobytes() == ecor.tobytes()
    skipped = np.isnan(ebias[0])
    assert np.ascontiguousarray(bias[:, ~skipped]).tobytes() == np.ascontiguousarray(ebias[:, ~skipped]).tobytes()
    assert (bias[:, skipped] == -7.0).all()           # skipped loci: the caller's values stay
    if seed % 2 == 0:
        assert skipped.any()


@pytest.mark.gpu
def test_dbt_pred_larger_case_and_bad_tables(tmp_path):
    c = make_pred_case(str(tmp_path), ns=70, n_tr=600, seed=9, kmax=400)
    rc, raw, cor, bias, L = call_dbt_pred(c)
    assert rc == 0, L.dbt_last_error()
    eraw, ecor, ebias = pred_model(c)
    assert np.ascontiguousarray(cor).tobytes() == ecor.tobytes() and np.ascontiguousarray(bias).tobytes() == ebias.tobytes()
    # an invariant k-mer that points outside its locus is refused (the loci are corrected independently)
    bad = dict(c)
    bad["iki"] = c["iki"].copy()
    bad["iki"][-1] = 0
    rc, *_ = call_dbt_pred(bad)
    assert rc == -6 and b"outside the locus" in L.dbt_last_error()
    bad = dict(c)
    bad["nk_cum"] = c["nk_cum"].copy()
    bad["nk_cum"][-1] += 1
    rc, *_ = call_dbt_pred(bad)
    assert rc == -6


@needs_ref
@pytest.mark.gpu
@pytest.mark.parametrize("ns,n_tr,seed", CASES[:4])
def test_pred_cli_vs_reference_live(ns, n_tr, seed, tmp_path):
    """the drop-in binary against the reference binary on the same files: the three outputs byte for byte"""
    d = str(tmp_path)
    c = make_pred_case(d, ns, n_tr, seed)
    ref = run_ref(c, d)
    r = subprocess.run([CLI_PRED, c["meta"], c["ikfn"], f"{d}/gpu.raw", f"{d}/gpu.cor", f"{d}/gpu.bias.tsv"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-800:]
    for a, b in zip(ref, (f"{d}/gpu.raw", f"{d}/gpu.cor", f"{d}/gpu.bias.tsv")):
        assert open(a, "rb").read() == open(b, "rb").read(), b
    # a sample file of another RPGG (nk differs) is refused like the reference refuses it
    import struct
    with open(f"{d}/other.trkmc.ar", "wb") as f:
        f.write(struct.pack("<Q", c["nk"] + 1) + bytes(8 * (c["nk"] + 1)))
    with open(c["meta"], "a") as f:
        f.write(f"{d}/other.trkmc.ar\t30\n")
    r = subprocess.run([CLI_PRED, c["meta"], c["ikfn"], f"{d}/x.raw", f"{d}/x.cor", f"{d}/x.bias.tsv"], capture_output=True, text=True)
    assert r.returncode != 0 and "nk" in r.stderr
