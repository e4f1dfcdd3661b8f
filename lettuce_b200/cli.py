"""Console driver with the reference's commands and options (reference/lettuce/cli.py:23-180):

    python -m lettuce_b200 [--cuda/--no-cuda] [-i GPU] [-p single|double] benchmark [-s STEPS] [-r RES] [-f FLOW] [-v VTK]
    python -m lettuce_b200 [...] convergence

`benchmark` runs a stock flow for a few steps and prints MLUPS (`cli.py:83-125`); `convergence` is the
reference's integration check: Taylor-Green 2-D in diffusive scaling, resolutions 16..256, velocity error
order 2 and pressure error order 1 within 0.1 (`cli.py:133-180`).  Everything steps on the native CUDA
engine; `--no-cuda` and `-p half` are refused (no CPU path, no fp16 kernels)."""
import cProfile
import pstats
import sys

import click
import numpy as np
import torch

from . import __version__
from ._collision import BGKCollision
from ._context import Context
from ._flows import TaylorGreenVortex, flow_by_name
from ._force import Guo
from ._output import ErrorReporter, VTKReporter
from ._simulation import Simulation

__all__ = ['main', 'convergence_table']


@click.group(invoke_without_command=True)
@click.version_option(version=__version__)
@click.option("--cuda/--no-cuda", default=True, help="Use cuda (default=True).")
@click.option("-i", "--gpu-id", type=int, default=0, help="Device ID of the GPU (default=0).")
@click.option("-p", "--precision", type=click.Choice(["half", "single", "double"]), default="double",
              help="Numerical Precision; 16, 32, or 64 bit per float (default=double).")
@click.pass_context
def main(ctx, cuda, gpu_id, precision):
    """B200-native Lattice Boltzmann Solver behind the lettuce API
    """
    if ctx.invoked_subcommand is None:
        click.echo(ctx.get_help())
        return
    ctx.obj = {'cuda': cuda, 'gpu_id': gpu_id, 'precision': precision}


def _context(ctx) -> Context:
    """device / dtype of the group options (cli.py:42-55); resolved when a command runs, so that
    `--help` works on a machine without a GPU"""
    o = ctx.obj
    if not o['cuda']:
        raise click.UsageError("lettuce_b200 steps on the native CUDA engine only (no CPU path)")
    if o['precision'] == "half":
        raise click.UsageError("the native kernels compute in single or double precision")
    if not torch.cuda.is_available():
        print("CUDA not found.")
        raise click.Abort
    dtype = {"single": torch.float, "double": torch.double}[o['precision']]
    return Context(torch.device("cuda:{}".format(o['gpu_id'])), dtype, True)


@main.command()
@click.option("-s", "--steps", type=int, default=10, help="Number of time steps.")
@click.option("-r", "--resolution", type=int, default=1024, help="Grid Resolution")
@click.option("-o", "--profile-out", type=str, default="",
              help="File to write profiling information to (default=""; no profiling information gets written).")
@click.option("-f", "--flow", type=click.Choice(list(flow_by_name.keys())), default="taylor2d")
@click.option("-v", "--vtk-out", type=str, default="",
              help="VTK file basename to write the velocities and densities to (default=""; no info gets written).")
@click.pass_context
def benchmark(ctx, steps, resolution, profile_out, flow, vtk_out):
    """Run a short simulation and print performance in MLUPS.
    """
    if profile_out:
        profile = cProfile.Profile()
        profile.enable()
    flow_class, stencil = flow_by_name[flow]
    context = _context(ctx)
    flow = flow_class(context, resolution=[resolution] * stencil().d, reynolds_number=1, mach_number=0.05,
                      stencil=stencil)
    tau = flow.units.relaxation_parameter_lu
    force = Guo(flow, tau=tau, acceleration=flow.acceleration) if hasattr(flow, "acceleration") else None
    reporter = [VTKReporter(interval=10, filename_base=vtk_out)] if vtk_out else []
    simulation = Simulation(flow, BGKCollision(tau=tau, force=force), reporter)
    mlups = simulation(steps)
    if profile_out:
        profile.disable()
        stats = pstats.Stats(profile)
        stats.sort_stats('cumulative')
        stats.print_stats()
        profile.dump_stats(profile_out)
        click.echo(f"Saved profiling information to {profile_out}.")
    click.echo("Finished {} steps in {} bit precision. MLUPS: {:10.2f}".format(
        steps, str(context.dtype).replace("torch.float", ""), mlups))
    return 0


def convergence_table(context, exponents=range(4, 9)):
    """rows (resolution, error_u, error_p, MLUPS) of the diffusive-scaling series (cli.py:145-170)"""
    rows = []
    for i in exponents:
        resolution = 2 ** i
        flow = TaylorGreenVortex(context, [resolution] * 2, reynolds_number=10000, mach_number=8 / resolution)
        error_reporter = ErrorReporter(flow.analytic_solution, interval=1, out=None)
        simulation = Simulation(flow, BGKCollision(tau=flow.units.relaxation_parameter_lu), [error_reporter])
        mlups = simulation(10 * resolution)
        error_u, error_p = np.mean(np.abs(error_reporter.out), axis=0).tolist()
        rows.append((resolution, error_u, error_p, mlups))
    return rows


@main.command()
@click.pass_context
def convergence(ctx):
    """Use Taylor Green 2D for convergence test in diffusive scaling."""
    context = _context(ctx)
    print(("{:>15} " * 6).format("resolution", "error (u)", "order (u)", "error (p)", "order (p)", "MLUPS"))
    factor_u = factor_p = 0
    old = None
    for resolution, error_u, error_p, mlups in convergence_table(context):
        factor_u = 0 if old is None else old[0] / error_u
        factor_p = 0 if old is None else old[1] / error_p
        old = (error_u, error_p)
        print(f"{resolution:15} {error_u:15.2e} {factor_u / 2:15.2f} {error_p:15.2e} {factor_p / 2:15.2f} {mlups:15.2f}")
    tol = 1e-1
    if not (2 - tol) < factor_u / 2 < (2 + tol):
        print(f"FAILED: Velocity convergence order {factor_u / 2} is not in [1.9, 2.1]")
        sys.exit(1)
    if not (1 - tol) < factor_p / 2 < (1 + tol):
        print(f"FAILED: Pressure convergence order {factor_p / 2} is not in [0.9, 1.1].")
        sys.exit(1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
