"""The output side of the path: VTKReporter / write_vtk
(reference/lettuce/ext/_reporter/vtk_reporter.py:11-64), ErrorReporter (error_reporter.py:9-44) and
EnergySpectrum (observable_reporter.py:71-137).

What changes against the reference: the reporters never pull the populations to the host.  On a
CUDA flow ONE kernel (lb_output_fields) reads the populations once and writes p and u in physical
units, already in the file's point order (x fastest; tiles transposed through shared memory); the
1 + d fields cross PCIe once, into one pinned staging buffer; the bytes go to disk with a single
write per array.  The `.vtr` container is written here (the reference delegates to pyevtk 1.6
`gridToVTK`, absent from this image): VTK XML RectilinearGrid, version 1.0, little endian,
UInt64 block headers, raw appended data, coordinates first and then the point data in dict
order, every array in x-fastest (Fortran) point order -- pyevtk's published layout.  Parity of
the container is UNPINNED (no pyevtk here to diff against); `read_vtr` below reads the files
back and the tests pin the payload.

HDF5Reporter / LettuceDataset / write_image are out of scope: the names refuse."""
import os
import re
import sys
from typing import Dict, Optional

import numpy as np
import torch

from ._reporter import Observable, Reporter

__all__ = ['VTKReporter', 'write_vtk', 'read_vtr', 'ErrorReporter', 'EnergySpectrum',
           'HDF5Reporter', 'LettuceDataset', 'write_image']

_VTK_TYPE = {'float32': 'Float32', 'float64': 'Float64', 'int8': 'Int8', 'uint8': 'UInt8',
             'int16': 'Int16', 'uint16': 'UInt16', 'int32': 'Int32', 'uint32': 'UInt32',
             'int64': 'Int64', 'uint64': 'UInt64'}
_NP_TYPE = {v: k for k, v in _VTK_TYPE.items()}


def _point_order(a) -> np.ndarray:
    """[nx, ny, nz] array -> flat little-endian array in VTK point order (x fastest)"""
    a = np.asarray(a)
    if a.dtype.name not in _VTK_TYPE:
        raise TypeError(f"VTK cannot hold arrays of dtype {a.dtype}")
    flat = a.reshape(-1, order='F') if a.ndim > 1 else a
    return np.ascontiguousarray(flat, dtype=a.dtype.newbyteorder('<'))


def grid_to_vtr(path, x, y, z, point_data: Dict[str, np.ndarray], _preordered: bool = False) -> str:
    """Write `path + '.vtr'`: a rectilinear grid with 1-D coordinate arrays x, y, z and scalar point
    arrays of shape [len(x), len(y), len(z)] (what pyevtk.hl.gridToVTK(path, x, y, z, pointData=...)
    writes for 1-D coordinates).  `_preordered`: the arrays are already flat in point order."""
    x, y, z = (np.asarray(c) for c in (x, y, z))
    n = (x.size, y.size, z.size)
    ext = f"0 {n[0] - 1} 0 {n[1] - 1} 0 {n[2] - 1}"
    blocks, offset = [], 0

    def add(name, arr, flat):
        nonlocal offset
        data = arr if flat else _point_order(arr)
        line = (f'<DataArray Name="{name}" NumberOfComponents="1" type="{_VTK_TYPE[data.dtype.name]}" '
                f'format="appended" offset="{offset}"/>\n')
        blocks.append(data)
        offset += 8 + data.nbytes
        return line

    head = ['<?xml version="1.0"?>\n',
            '<VTKFile type="RectilinearGrid" version="1.0" byte_order="LittleEndian" header_type="UInt64">\n',
            f'<RectilinearGrid WholeExtent="{ext}">\n', f'<Piece Extent="{ext}">\n', '<Coordinates>\n']
    for name, c in (("x_coordinates", x), ("y_coordinates", y), ("z_coordinates", z)):
        head.append(add(name, c, False))
    head.append('</Coordinates>\n')
    first = next(iter(point_data), None)
    head.append(f'<PointData scalars="{first}">\n' if first is not None else '<PointData>\n')
    npoints = n[0] * n[1] * n[2]
    for name, arr in point_data.items():
        arr = np.asarray(arr)
        if arr.size != npoints or (not _preordered and tuple(arr.shape) != n):
            raise ValueError(f"point array '{name}' of shape {arr.shape} does not match the grid {n}")
        head.append(add(name, arr, _preordered))
    head += ['</PointData>\n', '</Piece>\n', '</RectilinearGrid>\n', '<AppendedData encoding="raw">\n_']
    filename = str(path) + ".vtr"
    header = "".join(head).encode("ascii")
    tail = b'\n</AppendedData>\n</VTKFile>\n'
    with open(filename, "wb") as fh:
        fh.write(header)
        for data in blocks:
            fh.write(np.uint64(data.nbytes).tobytes())
            fh.write(memoryview(data).cast('B'))
        fh.write(tail)
    return filename


def read_vtr(filename):
    """Read a file written by grid_to_vtr (or pyevtk's gridToVTK with 1-D coordinates) back:
    returns (x, y, z, {name: array [nx, ny, nz]})."""
    with open(filename, "rb") as fh:
        raw = fh.read()
    marker = raw.index(b'<AppendedData encoding="raw">')
    start = raw.index(b'_', marker) + 1
    header = raw[:marker].decode("ascii")
    arrays = {}
    for m in re.finditer(r'<DataArray Name="([^"]+)" NumberOfComponents="1" type="(\w+)" '
                         r'format="appended" offset="(\d+)"\s*/>', header):
        name, vtype, off = m.group(1), m.group(2), int(m.group(3))
        nbytes = int(np.frombuffer(raw, dtype='<u8', count=1, offset=start + off)[0])
        dt = np.dtype(_NP_TYPE[vtype]).newbyteorder('<')
        arrays[name] = np.frombuffer(raw, dtype=dt, count=nbytes // dt.itemsize, offset=start + off + 8)
    x, y, z = (arrays.pop(k) for k in ("x_coordinates", "y_coordinates", "z_coordinates"))
    shape = (x.size, y.size, z.size)
    return x, y, z, {k: v.reshape(shape, order='F') for k, v in arrays.items()}


def write_vtk(point_dict, id=0, filename_base="./data/output"):
    """vtk_reporter.py:11-16: integer node coordinates, one file per `id`"""
    shape = np.asarray(point_dict["p"]).shape
    return grid_to_vtr(f"{filename_base}_{id:08d}", np.arange(0, shape[0]), np.arange(0, shape[1]),
                       np.arange(0, shape[2]), point_dict)


class _ReportSlot:
    """one report in flight: device fields -> pinned host copy -> file"""

    def __init__(self, rows, nodes, dtype, device):
        import threading
        self.dev = torch.empty([rows, nodes], dtype=dtype, device=device)
        self.host = torch.empty([rows, nodes], dtype=dtype, pin_memory=True)
        self.fields_ready = torch.cuda.Event()
        self.copied = torch.cuda.Event()
        self.done = threading.Event()
        self.done.set()
        self.error = None


class VTKReporter(Reporter):
    """Pressure and velocity in physical units as `<filename_base>_<step:08d>.vtr` every `interval`
    steps (vtk_reporter.py:19-64).  `point_dict` holds the arrays of the last file, [nx, ny, nz]
    views (nz = 1 in 2-D) like the reference's.

    On a CUDA flow a report costs the time-stepping stream ONE kernel (lb_output_fields into a device buffer
    of the reporter's own); the device->host copy of the 1 + d fields runs on a second stream into
    double-buffered pinned staging and the file is written by a background thread, both overlapped with the
    following time steps (`asynchronous=False`: everything inside the call, as in the reference).
    `Simulation.__call__` calls `flush()` before it returns, so the files of a call exist when it is over;
    a report finds its slot busy only when reports arrive faster than PCIe + the disk take them."""

    def __init__(self, interval=50, filename_base="./data/output", asynchronous=True, buffers=2):
        super().__init__(interval)
        self.filename_base = filename_base
        directory = os.path.dirname(filename_base)
        if directory and not os.path.isdir(directory):
            os.makedirs(directory, exist_ok=True)     # the reference's os.mkdir (:25-26), nested paths included
        self.point_dict = dict()
        self.asynchronous = asynchronous
        self.buffers = max(1, int(buffers))
        self._stage = None     # synchronous path: pinned host staging buffer [1 + d, nodes]
        self._slots, self._next = [], 0
        self._copy_stream = None
        self._queue, self._writer = None, None

    def _staging(self, rows, nodes, dtype, pinned):
        st = self._stage
        if st is None or tuple(st.shape) != (rows, nodes) or st.dtype != dtype or st.is_pinned() != pinned:
            self._stage = st = torch.empty([rows, nodes], dtype=dtype, pin_memory=pinned)
        return st

    @staticmethod
    def _names_shape(flow):
        d = flow.stencil.d
        return ["p"] + [f"u{'xyz'[c]}" for c in range(d)], tuple(flow.resolution) + (1,) * (3 - d)

    def _write(self, flat, names, shape3, step):
        self.point_dict = {name: flat[k].reshape(shape3, order='F') for k, name in enumerate(names)}
        grid_to_vtr(f"{self.filename_base}_{step:08d}", *(np.arange(0, n) for n in shape3),
                    {name: flat[k] for k, name in enumerate(names)}, _preordered=True)

    def _writer_loop(self):
        while True:
            item = self._queue.get()
            if item is None:
                return
            slot, names, shape3, step = item
            try:
                slot.copied.synchronize()          # the D2H copy of this report has landed
                self._write(slot.host.numpy(), names, shape3, step)
            except Exception as e:                 # surfaced by flush()
                slot.error = e
            finally:
                slot.done.set()

    def __call__(self, simulation: 'Simulation'):
        flow = simulation.flow
        if flow.i % self.interval != 0:
            return
        names, shape3 = self._names_shape(flow)
        rows = len(names)
        if not (self.asynchronous and flow.f.is_cuda):
            ordered = flow.output_fields_pu()      # [1 + d, nodes] in file order; one kernel on a CUDA flow
            stage = self._staging(rows, ordered.shape[1], ordered.dtype, ordered.is_cuda)
            stage.copy_(ordered)                                     # synchronous for a pinned destination
            self._write(stage.numpy(), names, shape3, flow.i)
            return
        import queue
        import threading
        dev = flow.f.device
        nodes = flow.f[0].numel()
        if not self._slots or tuple(self._slots[0].dev.shape) != (rows, nodes) or self._slots[0].dev.dtype != flow.f.dtype:
            self.flush()
            with torch.cuda.device(dev):
                self._slots = [_ReportSlot(rows, nodes, flow.f.dtype, dev) for _ in range(self.buffers)]
                self._copy_stream = torch.cuda.Stream(device=dev)
        if self._writer is None:
            self._queue = queue.Queue()
            self._writer = threading.Thread(target=self._writer_loop, daemon=True)
            self._writer.start()
        slot = self._slots[self._next % self.buffers]
        self._next += 1
        slot.done.wait()                           # back-pressure: this slot's previous file is on disk
        if slot.error is not None:
            err, slot.error = slot.error, None
            raise err
        slot.done.clear()
        with torch.cuda.device(dev):
            flow.output_fields_pu(out=slot.dev)    # the ONE kernel on the time-stepping stream
            slot.fields_ready.record(torch.cuda.current_stream(dev))
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(slot.fields_ready)
                slot.host.copy_(slot.dev, non_blocking=True)
                slot.copied.record(self._copy_stream)
        self._queue.put((slot, names, shape3, flow.i))

    def flush(self):
        """wait until every report issued so far is on disk (and re-raise a writer error)"""
        for slot in self._slots:
            slot.done.wait()
            if slot.error is not None:
                err, slot.error = slot.error, None
                raise err

    def close(self):
        self.flush()
        if self._writer is not None:
            self._queue.put(None)
            self._writer.join()
            self._writer = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def output_mask(self, simulation: 'Simulation'):
        """no_collision_mask as `<filename_base>_mask.vtr` (vtk_reporter.py:49-64)"""
        flow = simulation.flow
        mask = flow.context.convert_to_ndarray(simulation.no_collision_mask).astype(int)
        if flow.stencil.d == 2:
            mask = mask[..., None]
        grid_to_vtr(str(self.filename_base) + "_mask", *(np.arange(0, n) for n in mask.shape), {"mask": mask})


class ErrorReporter(Reporter):
    """L2 errors of u and p (physical units) against `analytical_solution(t=...)`, normalised by
    resolution^(d/2) (error_reporter.py:9-44).  u and p come from the moments kernel; the norms are
    two device reductions."""

    def __init__(self, analytical_solution, interval=1, out=sys.stdout):
        Reporter.__init__(self, interval)
        self.analytical_solution = analytical_solution
        self.out = [] if out is None else out
        if not isinstance(self.out, list):
            print("#error_u         error_p", file=self.out)

    def __call__(self, simulation: 'Simulation'):
        flow = simulation.flow
        i = flow.i
        if i % self.interval != 0:
            return
        t = simulation.units.convert_time_to_pu(i)
        pref, uref = self.analytical_solution(t=t)
        pref = flow.context.convert_to_tensor(pref)
        uref = flow.context.convert_to_tensor(uref)
        rho, u = flow.rho_u()                                    # one pass over f (lb_moments)
        p = flow.units.convert_density_lu_to_pressure_pu(rho)
        u = flow.units.convert_velocity_to_pu(u)
        resolution = torch.pow(torch.prod(flow.context.convert_to_tensor(p.size())), 1 / flow.stencil.d)
        err_u = torch.norm(u - uref) / resolution ** (flow.stencil.d / 2)
        err_p = torch.norm(p - pref) / resolution ** (flow.stencil.d / 2)
        if isinstance(self.out, list):
            self.out.append([err_u.item(), err_p.item()])
        else:
            print(err_u.item(), err_p.item(), file=self.out)


class EnergySpectrum(Observable):
    """Kinetic energy per integer wavenumber shell (observable_reporter.py:71-137): u from the
    moments kernel, one FFT per component (cuFFT through torch.fft on a CUDA flow), and the shell
    sums as ONE matrix product with the [nodes, shells] membership mask (the reference multiplies
    the spectrum into a [*res, shells] temporary and sums it)."""

    def __init__(self, flow: 'Flow'):
        super().__init__(flow)
        self.dx = self.flow.units.convert_length_to_pu(1.0)
        self.dimensions = self.flow.resolution
        frequencies = [self.context.convert_to_tensor(np.fft.fftfreq(dim, d=1 / dim))
                       for dim in self.dimensions]
        wavenumbers = torch.stack(torch.meshgrid(*frequencies, indexing='ij'))
        wavenorms = torch.norm(wavenumbers, dim=0)
        if self.flow.stencil.d == 3:
            self.norm = self.dimensions[0] * np.sqrt(2 * np.pi) / self.dx ** 2
        else:
            self.norm = self.dimensions[0] / self.dx
        self.wavenumbers = torch.arange(int(torch.max(wavenorms)))
        k = self.wavenumbers.to(dtype=self.context.dtype, device=self.context.device)
        self.wavemask = (wavenorms[..., None] > k - 0.5) & (wavenorms[..., None] <= k + 0.5)
        self._shells = self.wavemask.reshape(-1, self.wavemask.shape[-1]).to(dtype=self.context.dtype)

    def __call__(self, f: Optional[torch.Tensor] = None):
        return self.spectrum_from_u(self.flow.u())

    def spectrum_from_u(self, u):
        u = self.flow.units.convert_velocity_to_pu(u)
        return self._ekin_spectrum(u).reshape(-1) @ self._shells

    def _ekin_spectrum(self, u):
        d = self.flow.stencil.d
        uh = torch.fft.fftn(u, dim=tuple(range(1, d + 1))) / self.norm
        return torch.sum(0.5 * (uh.imag ** 2 + uh.real ** 2), dim=0)


class _OutOfScope:
    """HDF5Reporter / LettuceDataset (util/datautils.py) and write_image (ext/_reporter/write_image.py) are
    OUT OF SCOPE of this package (HDF5 / image I/O, h5py and matplotlib are not in
    the image).  The names exist so that `from lettuce import *` code fails with a clear message instead of an
    AttributeError; use VTKReporter / Flow.dump / Flow.load."""
    _what = "this reporter"

    def __init__(self, *args, **kwargs):
        raise NotImplementedError(f"{self._what} is out of scope of lettuce_b200 (HDF5 / image output; "
                                  f"SURVEY.md section 2): use VTKReporter, Flow.dump / Flow.load")


class HDF5Reporter(_OutOfScope):
    _what = "HDF5Reporter"


class LettuceDataset(_OutOfScope):
    _what = "LettuceDataset"


def write_image(filename, array2d):
    raise NotImplementedError("write_image is out of scope of lettuce_b200 (matplotlib output; SURVEY.md section 2)")
