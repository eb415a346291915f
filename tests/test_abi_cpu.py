"""CPU-side checks of the product library: it loads, exports every symbol the header declares,
its host logic (process grid, decomposition, parameter defaults/validation) matches the reference,
and it fails loudly -- never falls back -- when there is no GPU."""
import os
import re

import pytest

from __graft_entry__ import ROOT, load_package

pb = load_package()


def test_exports_match_header():
    hdr = open(os.path.join(ROOT, "include", "poisson_b200.h")).read()
    declared = set(re.findall(r"\b(pb200_[A-Za-z0-9_]+)\s*\(", hdr))
    assert declared == set(pb.EXPORTS), declared ^ set(pb.EXPORTS)
    lib = pb.lib()
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.pb200_api_version() == 1


def test_default_params_are_the_reference_constants():
    p = pb.default_params()
    assert (p.M, p.N, p.delta, p.max_iter) == (40, 40, 1e-6, 0)          # S4:996-1007
    assert (p.A1, p.B1, p.A2, p.B2, p.f_val, p.ell_k) == (-1.0, 1.0, -0.6, 0.6, 1.0, 2.0)  # S4:11-13, 33
    assert p.norm_mode == pb.NORM_WEIGHTED and p.nranks == 1


def test_process_grid_and_decomposition(golden):
    for p, want in golden["process_grid"].items():
        assert list(pb.choose_process_grid(int(p))) == want
    for key, blocks in golden["decompose"].items():
        mn, p = key.split("@")
        m, n = map(int, mn.split("x"))
        px, py = pb.choose_process_grid(int(p))
        assert [list(pb.decompose_2d(m, n, px, py, r)) for r in range(int(p))] == blocks, key


def test_bad_arguments_are_rejected():
    with pytest.raises(pb.PB200Error):
        pb.decompose_2d(40, 40, 2, 2, 4)
    with pytest.raises(pb.PB200Error):
        pb.choose_process_grid(0)
    with pytest.raises(pb.PB200Error, match="M and N"):
        pb.run(pb.default_params(1, 40))


def test_no_gpu_means_error_not_fallback():
    if pb.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pb.PB200Error, match="no CPU path"):
        pb.run(pb.default_params(40, 40))
    with pytest.raises(pb.PB200Error, match="no CPU path"):
        pb.selftest_div(0.05)
