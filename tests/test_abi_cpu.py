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
    assert lib.pb200_api_version() == 2


def test_default_params_are_the_reference_constants():
    p = pb.default_params()
    assert (p.M, p.N, p.delta, p.max_iter) == (40, 40, 1e-6, 0)          # S4:996-1007
    assert (p.A1, p.B1, p.A2, p.B2, p.f_val, p.ell_k) == (-1.0, 1.0, -0.6, 0.6, 1.0, 2.0)  # S4:11-13, 33
    assert p.norm_mode == pb.NORM_WEIGHTED and p.nranks == 1
    assert p.ell_ax == 1.0 and p.rhs_mode == pb.RHS_INDICATOR          # S4:32-34, 182-191
    import ctypes
    assert ctypes.sizeof(pb.Params) == 152 and ctypes.sizeof(pb.Result) == 168          # the v1 layout, reserved words reused


def test_process_grid_and_decomposition(golden):
    for p, want in golden["process_grid"].items():
        assert list(pb.choose_process_grid(int(p))) == want
    for key, blocks in golden["decompose"].items():
        mn, p = key.split("@")
        m, n = map(int, mn.split("x"))
        px, py = pb.choose_process_grid(int(p))
        assert [list(pb.decompose_2d(m, n, px, py, r)) for r in range(int(p))] == blocks, key


def bands_of(nx, plan):
    """Row ranges of the bands a launch plan describes (the arithmetic of make_part in pb200_kernels.cuh)."""
    nbA, nxA, nbB = plan[:3]
    out = [(1 + nxA * b // nbA, nxA * (b + 1) // nbA) for b in range(nbA)]
    out += [(1 + nxA + (nx - nxA) * k // nbB, nxA + (nx - nxA) * (k + 1) // nbB) for k in range(nbB)]
    return out


def test_launch_plan_covers_every_row_exactly_once():
    """The bands of the two iteration kernels must tile rows 1..nx without gap or overlap, whatever the block size:
    a row left out would silently keep its old values (the GPU parity tests only see the sizes they run)."""
    import random
    rnd = random.Random(7)
    cases = [(nx, ny) for nx in (1, 2, 3, 4, 5, 7, 8, 9, 11, 12, 13, 23, 24, 25, 39, 47, 48, 49, 399, 799, 2399, 4095,
                                 8191, 16383, 32767) for ny in (1, 39, 124, 125, 599, 3199, 8191, 32767)]
    cases += [(rnd.randint(1, 40000), rnd.randint(1, 40000)) for _ in range(300)]
    for resident in (592, 148, 1, 7):
        for nx, ny in cases:
            plan = pb.launch_plan(nx, ny, resident)
            nbA, nxA, nbB, ctas, wn, warps = plan
            assert wn == -(-ny // 124) and warps >= 1
            assert nbA >= 1 and 1 <= nxA <= nx and nbB >= 0 and (nbB > 0) == (nxA < nx), (nx, ny, plan)
            assert ctas == (nbA + nbB) * -(-wn // warps), (nx, ny, plan)
            bands = bands_of(nx, plan)
            assert bands[0][0] == 1 and bands[-1][1] == nx, (nx, ny, plan)
            for (lo, hi), (lo2, _) in zip(bands, bands[1:] + [(nx + 1, 0)]):
                assert lo <= hi and lo2 == hi + 1, (nx, ny, plan, lo, hi, lo2)
            if nbB:          # the tail bands are the lower ones and cover at most a fifth of the rows
                assert nx - nxA <= nx // 5 and (nx - nxA) / nbB <= nxA / nbA + 1, (nx, ny, plan)


def test_tile_plan_covers_the_block_and_fits():
    """The resident loop kernel's tiles: at most one per SM, every cell in exactly one tile, tile + halo + the 24 KB table of
    exact coefficients within the shared memory given, halo ring within four cells per thread; blocks beyond ~1.15 M
    cells are refused."""
    import random
    rnd = random.Random(11)
    cases = [(39, 39), (1, 1), (1, 5000), (5000, 1), (399, 599), (799, 1199), (1200, 800), (1199, 799), (2399, 1600),
             (1000, 1300), (3, 100000), (148, 148), (149, 1024), (1100, 1180)]
    cases += [(rnd.randint(1, 1500), rnd.randint(1, 1500)) for _ in range(300)]
    for sms, smem in ((148, 231000), (132, 231000), (148, 100000), (7, 231000)):
        for nx, ny in cases:
            plan = pb.tile_plan(nx, ny, sms, smem)
            if plan is None:
                assert nx * ny * 26 > sms * (smem - 24576) * 0.80, (nx, ny, sms, smem)          # only blocks that really do not fit
                continue
            R, C, tr, tc = plan
            assert 1 <= tr * tc <= sms and (tr - 1) * R < nx <= tr * R and (tc - 1) * C < ny <= tc * C, (nx, ny, plan)
            pc, cp = (C + 5) // 2 * 2, (C + 1) // 2 * 2          # pitches of the haloed / compact shared-memory arrays
            assert ((R + 2) * pc + 2 * R * cp) * 8 + (R + 2) * pc + R * cp + 24576 <= smem and 2 * (R + C) <= 2048 and C <= 1024, (nx, ny, plan)
    assert pb.tile_plan(1200, 800) is not None and pb.tile_plan(399, 599) is not None          # 2400x3200 on 8 GPUs; configs[1]
    assert pb.tile_plan(2399, 3199) is None and pb.tile_plan(8191, 8191) is None


def test_bad_arguments_are_rejected():
    with pytest.raises(pb.PB200Error):
        pb.decompose_2d(40, 40, 2, 2, 4)
    with pytest.raises(pb.PB200Error):
        pb.choose_process_grid(0)
    with pytest.raises(pb.PB200Error):
        pb.launch_plan(0, 40)
    with pytest.raises(pb.PB200Error, match="M and N"):
        pb.run(pb.default_params(1, 40))


def test_no_gpu_means_error_not_fallback():
    if pb.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pb.PB200Error, match="no CPU path"):
        pb.run(pb.default_params(40, 40))
    with pytest.raises(pb.PB200Error, match="no CPU path"):
        pb.selftest_div(0.05)
