"""Multi-GPU runs of the `poisson_b200` driver (one host thread + one NCCL rank per GPU): the results must not
depend on the number of GPUs beyond reduction-order noise, iteration counts must equal the
reference's multi-rank counts, and the dumped solution must equal the single-GPU one. Skipped with < 2 GPUs."""
import json
import os
import subprocess

import numpy as np
import pytest

from __graft_entry__ import load_package

pb = load_package()
pytestmark = pytest.mark.gpu


def run_cli(tmp_path, M, N, gpus, extra=()):
    dump = tmp_path / f"w_{M}x{N}_{gpus}.npy"
    js = tmp_path / f"r_{M}x{N}_{gpus}.json"
    cmd = [pb.CLI_PATH, str(M), str(N), "--gpus", str(gpus), "--dump", str(dump), "--json", str(js), *extra]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    return out.stdout, json.load(open(js)), np.load(dump)


def need(gpus):
    if pb.device_count() < gpus:
        pytest.skip(f"needs {gpus} GPUs")


def test_cli_single_gpu_report_lines(tmp_path, golden):
    """The reference's stdout contract (S4:1002-1003, 901-903, 969-980, 1027-1035) on one GPU."""
    out, res, w = run_cli(tmp_path, 40, 40, 1)
    lines = out.splitlines()
    assert lines[0] == "MPI + CUDA 2D run with 1 processes; M=40, N=40"
    assert lines[1] == "Converged after 50 iterations (||w(k+1)-w(k)|| < 1e-06)."
    assert lines[2].startswith("   GPU compute time (Ap + D^{-1}r, max over ranks) ~ ")
    assert lines[3].startswith("   Host<->Device copy time (max over ranks)        ~ ")
    assert lines[4].startswith("   MPI halo exchange time (max over ranks)         ~ ")
    assert lines[5].startswith("   Preconditioner CPU part time (max over ranks)   ~ ")
    assert lines[6].startswith("   Dot products time (max over ranks)              ~ ")
    assert lines[7].startswith("M=40, N=40 | Iter=50 | Total Time=") and lines[7].endswith(" s")
    assert lines[8].startswith("   Init time (program)      ~ ")
    assert lines[9].startswith("   Solver time (MPI+CUDA)   ~ ")
    assert lines[10].startswith("   Finalization time        ~ ")
    assert res["iterations"] == golden["iters_stage3"]["40x40"] and w.shape == (41, 41)
    assert len(lines) == 11          # nothing but the reference's lines unless asked
    assert res["err_max_in_D"] == pytest.approx(6.418e-3, rel=1e-3) and 0 < res["err_l2_in_D"] < res["err_max_in_D"] * 2


def test_cli_report_error_and_geometry_flags(tmp_path):
    """--report-error appends one line after the reference's; --ellipse-ax / --rhs reach the solver."""
    out, res, _ = run_cli(tmp_path, 40, 40, 1, ("--report-error",))
    lines = out.splitlines()
    assert len(lines) == 12 and lines[11].startswith("   Error vs analytic solution in D: max = 6.41")
    _, res2, _ = run_cli(tmp_path, 40, 40, 1, ("--ellipse-ax", "0.8", "--rhs", "area"))
    assert res2["w_norm_E"] != res["w_norm_E"] and res2["iterations"] > 0


@pytest.mark.parametrize("M,N,gpus,extra", [(40, 40, 2, ()), (400, 600, 2, ()), (400, 600, 2, ("--grid", "2x1")),
                                            (37, 53, 2, ()), (400, 600, 4, ()), (400, 600, 8, ()), (37, 53, 6, ())])
def test_results_do_not_depend_on_gpu_count(tmp_path, golden, M, N, gpus, extra):
    need(gpus)
    _, r1, w1 = run_cli(tmp_path, M, N, 1)
    out, rp, wp = run_cli(tmp_path, M, N, gpus, extra)
    assert out.splitlines()[0] == f"MPI + CUDA 2D run with {gpus} processes; M={M}, N={N}"
    assert rp["iterations"] == r1["iterations"]
    key = f"{M}x{N}@{gpus}"
    if key in golden["iters_stage3_ranks"]:
        assert rp["iterations"] == golden["iters_stage3_ranks"][key]
    for f in ("w_norm_E", "w_max", "diff_norm", "zr"):
        assert abs(rp[f] - r1[f]) <= 1e-10 * abs(r1[f]), f      # north_star tolerance
    np.testing.assert_allclose(wp, w1, rtol=0, atol=1e-11)


@pytest.mark.parametrize("M,N,gpus,extra", [(400, 600, 2, ()), (400, 600, 2, ("--grid", "2x1")), (37, 53, 2, ()),
                                            (400, 600, 4, ()), (400, 600, 8, ()), (1016, 2030, 2, ())])
def test_peer_memory_and_nccl_modes_agree(tmp_path, monkeypatch, M, N, gpus, extra):
    """The default multi-GPU mode moves halos and partial sums through peer memory inside the kernels; --comm nccl uses
    NCCL calls between the kernels. With the two marching kernels in both (PB200_LOOP=0) it is the same arithmetic and the
    same rank-ordered sums: identical results. Blocks that fit in shared memory run the resident loop kernel in peer mode
    (one launch per GPU for the whole solve): other partial sums, so the bar between GPU counts applies."""
    need(gpus)
    _, b, wb = run_cli(tmp_path, M, N, gpus, extra + ("--comm", "nccl"))
    monkeypatch.setenv("PB200_LOOP", "0")
    _, a, wa = run_cli(tmp_path, M, N, gpus, extra + ("--comm", "peer"))
    assert a["iterations"] == b["iterations"] and a["kernel_launches"] > 2 * a["iterations"]
    assert abs(a["w_norm_E"] - b["w_norm_E"]) <= 1e-12 * abs(b["w_norm_E"])
    np.testing.assert_allclose(wa, wb, rtol=0, atol=1e-12)
    monkeypatch.delenv("PB200_LOOP")
    _, t, wt = run_cli(tmp_path, M, N, gpus, extra + ("--comm", "peer"))
    assert t["iterations"] == b["iterations"] and t["kernel_launches"] <= 6          # reset, init, its scalar tail, the loop, norms
    for f in ("w_norm_E", "w_max", "diff_norm", "zr", "err_max_in_D"):
        assert abs(t[f] - b[f]) <= 1e-10 * abs(b[f]), f
    np.testing.assert_allclose(wt, wb, rtol=0, atol=1e-11)


def test_multi_gpu_fixed_iterations_and_no_graph(tmp_path, monkeypatch):
    need(2)
    monkeypatch.setenv("PB200_LOOP", "0")          # the marching kernels: the graph only exists there
    _, a, wa = run_cli(tmp_path, 300, 200, 2, ("--fixed-iters", "37"))
    _, b, wb = run_cli(tmp_path, 300, 200, 2, ("--fixed-iters", "37", "--no-graph"))
    assert a["iterations"] == b["iterations"] == 37
    assert np.array_equal(wa, wb)            # same kernels, same order: bitwise
