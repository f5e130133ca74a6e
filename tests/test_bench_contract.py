"""bench.py contract pieces that can be checked without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_and_exits_zero_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("only meaningful on a box without a GPU")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--grid", "64",
                        "--steps", "10", "--warmup", "4"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d


def test_reference_sources_are_staged_verbatim():
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import hashlib
    import ref_build as rb
    if not rb.staged():
        import pytest
        pytest.skip("reference not staged")
    up = "/root/reference"
    for f in rb.FILES:
        staged = (rb.SRC / f).read_bytes()
        if os.path.exists(os.path.join(up, f)):
            assert staged == open(os.path.join(up, f), "rb").read(), f"{f} differs from the upstream file"
        sums = dict(l.split()[::-1] for l in (rb.REF / "SHA256SUMS").read_text().splitlines())
        assert sums[f] == hashlib.sha256(staged).hexdigest()


def test_configure_rewrites_only_the_requested_defines():
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_build as rb
    if not rb.staged():
        import pytest
        pytest.skip("reference not staged")
    out = rb.configure("grad1612_cuda_heat.cu", {"NXPROB": 123, "NYPROB": 456, "STEPS": 7}, "test_cfg")
    a = (rb.SRC / "grad1612_cuda_heat.cu").read_text().splitlines()
    b = out.read_text().splitlines()
    assert len(a) == len(b)
    changed = [(x, y) for x, y in zip(a, b) if x != y]
    assert len(changed) == 3 and all(x.startswith("#define") and y.startswith("#define") for x, y in changed)
    assert any("123" in y for _, y in changed) and any("456" in y for _, y in changed)


def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("h2d_bench", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_both_arms_print_the_same_metric_and_an_e2e_block():
    """The driver divides the repo arm by the reference arm only when metric, unit and
    direction agree, and computes the end-to-end ratio only when both arms carry `e2e`
    (round 1 lost both ratios to a metric string that differed by three words)."""
    import inspect
    b = _bench_module()
    ref_src = inspect.getsource(b.reference_arm)
    main_src = inspect.getsource(b.main)
    assert '"metric": METRIC' in ref_src and '"metric": METRIC' in main_src
    assert ref_src.count('"metric"') == 1 and '"unit": "GLUP/s"' in ref_src and '"unit": "GLUP/s"' in main_src
    assert '"higher_is_better": True' in ref_src and '"higher_is_better": True' in main_src
    assert '"e2e"' in ref_src and '"h2d_bytes_per_step"' in ref_src and '"d2h_bytes_per_step"' in ref_src
    assert 'out["e2e"] = e2e' in main_src
    # the like-for-like number and the verification flag are part of the record
    assert '"same_arith"' in main_src and '"verified"' in main_src


def test_headline_arithmetic_is_the_reference_arithmetic():
    b = _bench_module()
    a = b.parse_args([])
    assert a.arith == "strict-fma" and a.gpus == 1
    assert b.parse_args(["--arith", "fp32"]).arith == "fp32"
