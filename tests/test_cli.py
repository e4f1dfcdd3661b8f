"""The console driver (lettuce_b200/cli.py) -- reference/tests/test_cli.py:29-37 restated, plus the two
commands on the GPU: `benchmark` on stock flows and `convergence`, the reference's integration check
(cli.py:133-180: Taylor-Green 2-D in diffusive scaling, order(u) = 2 +- 0.1, order(p) = 1 +- 0.1)."""
import os

import pytest
import torch
from click.testing import CliRunner

from lettuce_b200 import cli


def test_command_line_interface():
    runner = CliRunner()
    result = runner.invoke(cli.main)
    assert result.exit_code == 0
    assert 'Boltzmann' in result.output
    help_result = runner.invoke(cli.main, ['--help'])
    assert help_result.exit_code == 0
    assert 'Show this message and exit.' in help_result.output
    for cmd in ("benchmark", "convergence"):
        r = runner.invoke(cli.main, [cmd, '--help'])
        assert r.exit_code == 0 and 'Show this message and exit.' in r.output


def test_cli_refuses_cpu_and_half():
    runner = CliRunner()
    r = runner.invoke(cli.main, ['--no-cuda', 'benchmark'])
    assert r.exit_code == 2 and 'no CPU path' in r.output
    r = runner.invoke(cli.main, ['-p', 'half', 'benchmark'])
    assert r.exit_code == 2 and 'single or double' in r.output
    if not torch.cuda.is_available():
        r = runner.invoke(cli.main, ['benchmark'])
        assert r.exit_code == 1 and 'CUDA not found.' in r.output


@pytest.mark.gpu
@pytest.mark.parametrize("flow,res", [("taylor2d", 64), ("taylor3d", 32), ("poiseuille2d", 32), ("shear2d", 64),
                                      ("decay2d", 32)])
def test_cli_benchmark(flow, res, tmp_path):
    runner = CliRunner()
    vtk = str(tmp_path / "vtk" / "out")
    r = runner.invoke(cli.main, ['-p', 'single', 'benchmark', '-s', '20', '-r', str(res), '-f', flow, '-v', vtk])
    assert r.exit_code == 0, r.output
    assert 'Finished 20 steps in 32 bit precision. MLUPS:' in r.output
    for i in (0, 10, 20):
        assert os.path.isfile(f"{vtk}_{i:08d}.vtr")


@pytest.mark.gpu
def test_cli_convergence():
    runner = CliRunner()
    r = runner.invoke(cli.main, ['-p', 'double', 'convergence'])
    assert r.exit_code == 0, r.output
    rows = [line.split() for line in r.output.splitlines() if line.split() and line.split()[0].isdigit()]
    assert [int(row[0]) for row in rows] == [16, 32, 64, 128, 256]
    # the CPU oracle gives 2.07 / 1.02 for the last pair of resolutions (fp64)
    assert float(rows[-1][2]) == pytest.approx(2.07, abs=0.02)
    assert float(rows[-1][4]) == pytest.approx(1.02, abs=0.02)
