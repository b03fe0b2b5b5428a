"""VtkOutputter / CsvOutputter: the consumers on the other side of the hot path.

Host-side mirror of /root/reference/include/VtkOutputter.hpp:27-143, src/VtkOutputter.cpp:47-67 and
src/CsvOutputter.cpp:28-60 with the same method names.  The reference links VTK 7.1 for the XML writer; here the
.vtu (UnstructuredGrid, one scalar point array named after the field, every linear sub-element with its own points
exactly like insertAllCells, include/VtkOutputter.hpp:111-141) is written directly.

DeviceOutput is the B200 side of it: the derivation of eta, u, v and the interpolation to the split nodes run in
one kernel on the live device state (bdg_sw2d_output_begin), the copies to pinned host memory overlap the next steps,
and the files are written from the arrived arrays -- the coordinates and the connectivity are computed once.
"""
import base64
import os
import struct

import numpy as np

VTK_TRIANGLE, VTK_QUAD = 5, 9


def _data_array(fh, name, arr, ncomp, fmt):
    vtype = {"float64": "Float64", "int64": "Int64", "uint8": "UInt8"}[arr.dtype.name]
    comp = ' NumberOfComponents="%d"' % ncomp if ncomp > 1 else ""
    fh.write(('        <DataArray type="%s" Name="%s"%s format="%s">\n' % (vtype, name, comp, fmt)).encode())
    flat = np.ascontiguousarray(arr).reshape(-1)
    if fmt == "ascii":
        rep = "%.17g" if arr.dtype.kind == "f" else "%d"
        for i in range(0, flat.size, 6 * max(ncomp, 1)):
            fh.write(("          " + " ".join(rep % v for v in flat[i:i + 6 * max(ncomp, 1)]) + "\n").encode())
    else:  # inline base64, uncompressed: one UInt32 byte count followed by the raw little-endian data
        raw = flat.astype(flat.dtype.newbyteorder("<"), copy=False).tobytes()
        fh.write(b"          " + base64.b64encode(struct.pack("<I", len(raw)) + raw) + b"\n")
    fh.write(b"        </DataArray>\n")


def write_vtu(fileName, x, y, field, fieldName, fmt="binary"):
    """x, y, field: [corners][cells] (corners = 3 or 4), the layout splitElements returns.  Points are numbered cell by
    cell; quads are stored as (v1, v3, v4, v2), the node swap of insertAllCells (include/VtkOutputter.hpp:134-137)."""
    nc, ncell = field.shape
    if nc not in (3, 4):
        raise ValueError("write_vtu: %d corners per cell (3 or 4 expected)" % nc)
    if nc * ncell * 8 >= 2 ** 32 and fmt == "binary":
        raise ValueError("write_vtu: array too large for the UInt32 block header; split the output")
    pts = np.zeros((ncell * nc, 3))
    pts[:, 0] = x.T.reshape(-1)
    pts[:, 1] = y.T.reshape(-1)
    conn = np.arange(ncell * nc, dtype=np.int64).reshape(ncell, nc)
    if nc == 4:
        conn = conn[:, [0, 2, 3, 1]]
    offsets = np.arange(1, ncell + 1, dtype=np.int64) * nc
    types = np.full(ncell, VTK_TRIANGLE if nc == 3 else VTK_QUAD, dtype=np.uint8)
    with open(fileName, "wb") as fh:
        fh.write(b'<?xml version="1.0"?>\n<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian" header_type="UInt32">\n')
        fh.write(("  <UnstructuredGrid>\n    <Piece NumberOfPoints=\"%d\" NumberOfCells=\"%d\">\n" % (ncell * nc, ncell)).encode())
        fh.write(('      <PointData Scalars="%s">\n' % fieldName).encode())
        _data_array(fh, fieldName, np.ascontiguousarray(field.T, dtype=np.float64), 1, fmt)
        fh.write(b"      </PointData>\n      <CellData>\n      </CellData>\n      <Points>\n")
        _data_array(fh, "Points", pts, 3, fmt)
        fh.write(b"      </Points>\n      <Cells>\n")
        _data_array(fh, "connectivity", conn, 1, fmt)
        _data_array(fh, "offsets", offsets, 1, fmt)
        _data_array(fh, "types", types, 1, fmt)
        fh.write(b"      </Cells>\n    </Piece>\n  </UnstructuredGrid>\n</VTKFile>\n")


def read_vtu(fileName):
    """Minimal reader of what write_vtu writes (round-trip checks in tests/, post-processing without VTK)."""
    import xml.etree.ElementTree as ET

    root = ET.parse(fileName).getroot()
    piece = root.find("UnstructuredGrid/Piece")
    np_dt = {"Float64": "<f8", "Int64": "<i8", "UInt8": "u1"}
    out = {"npoints": int(piece.get("NumberOfPoints")), "ncells": int(piece.get("NumberOfCells")), "arrays": {}}
    for da in piece.iter("DataArray"):
        dt = np_dt[da.get("type")]
        if da.get("format") == "ascii":
            a = np.array(da.text.split(), dtype=float).astype(dt)
        else:
            raw = base64.b64decode(da.text.strip())
            n = struct.unpack("<I", raw[:4])[0]
            a = np.frombuffer(raw[4:4 + n], dtype=dt)
        nc = int(da.get("NumberOfComponents", "1"))
        out["arrays"][da.get("Name")] = a.reshape(-1, nc) if nc > 1 else a
    out["scalars"] = piece.find("PointData").get("Scalars")
    return out


class VtkOutputter:
    FileExtension = "vtu"

    def __init__(self, nodesProvisioner, directory=".", fmt="binary"):
        self.NodesProvisioner = nodesProvisioner
        self.directory = directory
        self.fmt = fmt

    def generateFileName(self, fieldName, fileNumber):  # src/VtkOutputter.cpp:47-51
        return "%s%07d.%s" % (fieldName, int(fileNumber), self.FileExtension)

    def writeFieldToFile(self, fileName, field, fieldName):  # include/VtkOutputter.hpp:48-99
        P = self.NodesProvisioner
        x, y = P.get_xGrid(), P.get_yGrid()
        field = np.asarray(field, dtype=np.float64)
        if P.NOrder > 1:
            x, y, field = P.splitElements(x, y, field)
        write_vtu(os.path.join(self.directory, fileName), x, y, field, fieldName, self.fmt)

    def writeFieldsToFiles(self, fields, tstep):  # src/VtkOutputter.cpp:53-61
        for fieldName, field in fields.items():
            self.writeFieldToFile(self.generateFileName(fieldName, tstep), field, fieldName)


class CsvOutputter:
    def __init__(self, directory="."):
        self.directory = directory

    @staticmethod
    def generateFileName(fieldName, fileNumber):  # src/CsvOutputter.cpp:46-50
        return "%s%07d.dat" % (fieldName, int(fileNumber))

    def writeFieldToFile(self, fileName, field, delimeter=" "):  # src/CsvOutputter.cpp:28-39 (ostream default: 6 significant digits)
        with open(os.path.join(self.directory, fileName), "w") as fh:
            for row in np.asarray(field):
                fh.write("".join("%g%s" % (v, delimeter) for v in row) + "\n")

    def writeFieldsToFiles(self, fields, tstep):  # src/CsvOutputter.cpp:52-60
        for fieldName, field in fields.items():
            self.writeFieldToFile(self.generateFileName(fieldName, tstep), field, " ")


class DeviceOutput:
    """Output of a running SW2dSolver without a full-state download: begin() after the steps of an output interval,
    keep stepping, write() when the files are wanted.  Two pinned buffer sets alternate, so a begin() may be issued
    while the previous snapshot is still being written."""

    def __init__(self, solver, nodesProvisioner, directory=".", fmt="binary"):
        import torch  # pinned host memory (plumbing only)

        self.sol, self.P, self.directory, self.fmt = solver, nodesProvisioner, directory, fmt
        rout, sout, cells = nodesProvisioner.splitTable()
        self.IM = nodesProvisioner.computeInterpMatrix(rout, sout)
        self.n_out, K = len(rout), nodesProvisioner.NumElements
        solver.set_output(self.IM)
        idx = cells.T
        self._corners = lambda a: np.ascontiguousarray(a[idx].transpose(0, 2, 1).reshape(idx.shape[0], -1))
        self.x = self._corners(self.IM @ nodesProvisioner.get_xGrid())  # static: computed once
        self.y = self._corners(self.IM @ nodesProvisioner.get_yGrid())
        self._pin = [[torch.empty((self.n_out, K), dtype=torch.float64).pin_memory() for _ in range(3)] for _ in range(2)]
        self._cur = 0
        self._pending = None
        self._writer = None

    def begin(self, tstep):
        if self._writer is not None:
            self._busy[self._cur].wait()  # the writer thread has finished with the buffer set about to be reused
        bufs = [t.numpy() for t in self._pin[self._cur]]
        self.sol.output_begin(*bufs)
        self._pending = (tstep, bufs)
        if self._writer is not None:
            self._busy[self._cur].clear()
            self._queue.put((self._cur, tstep, bufs))
        self._cur ^= 1

    # ---- optional writer thread: files are written while the time loop keeps running -------------------------------
    def start_writer(self):
        """After this, every begin() is followed -- on a background thread -- by the wait for its copies and the file
        writes; the caller only steps and calls begin().  finish() drains the queue and returns the file names."""
        import queue
        import threading

        self._queue = queue.Queue()
        self._busy = [threading.Event(), threading.Event()]
        for e in self._busy:
            e.set()
        self.written, self.writer_error = [], None

        def run():
            while True:
                item = self._queue.get()
                if item is None:
                    return
                slot, tstep, bufs = item
                try:
                    # waits for the copies of the LATEST begin(); copies complete in stream order, so this snapshot's
                    # have arrived too.  output_wait only synchronises on an event: safe beside the stepping thread.
                    self.sol.output_wait()
                    for name, a in zip(("eta", "u", "v"), bufs):
                        fn = os.path.join(self.directory, "%s%07d.vtu" % (name, int(tstep)))
                        write_vtu(fn, self.x, self.y, self._corners(a), name, self.fmt)
                        self.written.append(fn)
                except Exception as exc:  # surfaced by finish()
                    self.writer_error = exc
                finally:
                    self._busy[slot].set()

        self._writer = threading.Thread(target=run, daemon=True)
        self._writer.start()

    def finish(self):
        if self._writer is None:
            return []
        self._queue.put(None)
        self._writer.join()
        self._writer = None
        if self.writer_error is not None:
            raise self.writer_error
        return self.written

    def fields(self):
        """Waits for the snapshot of the last begin(); returns tstep, {"eta","u","v"} as [n_out][K] interpolated arrays."""
        tstep, bufs = self._pending
        self.sol.output_wait()
        return tstep, dict(zip(("eta", "u", "v"), bufs))

    def write(self):
        tstep, flds = self.fields()
        names = []
        for name, a in flds.items():
            fn = os.path.join(self.directory, "%s%07d.vtu" % (name, int(tstep)))
            write_vtu(fn, self.x, self.y, self._corners(a), name, self.fmt)
            names.append(fn)
        return names
