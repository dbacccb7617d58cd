"""A Gadget-1 snapshot writer for the file-level tests (format read by the reference's src/gadgetio.cxx:120-930 in its default
build: float32 positions and velocities, uint32 ids, 256-byte header, Fortran record markers; one file; dark matter = type 1
with the particle mass in the header's mass table, so there is no mass block)."""
import struct

import numpy as np


def write_snapshot(path, pos, vel, pid, box, mass, redshift=0.0, omega0=0.3, omega_lambda=0.7, hubble=1.0):
    n = len(pos)
    npart = [0, n, 0, 0, 0, 0]
    masses = [0.0, float(mass), 0.0, 0.0, 0.0, 0.0]
    hdr = struct.pack("<6I6ddd2i6I2i4d2i6Ii", *npart, *masses, 1.0 / (1.0 + redshift), redshift, 0, 0, *npart, 0, 1, float(box), omega0,
                      omega_lambda, hubble, 0, 0, *([0] * 6), 0)
    hdr = hdr + b"\0" * (256 - len(hdr))
    assert len(hdr) == 256

    def block(f, payload):
        f.write(struct.pack("<I", len(payload)))
        f.write(payload)
        f.write(struct.pack("<I", len(payload)))

    with open(path, "wb") as f:
        block(f, hdr)
        block(f, np.ascontiguousarray(pos, dtype="<f4").tobytes())
        block(f, np.ascontiguousarray(vel, dtype="<f4").tobytes())
        block(f, np.ascontiguousarray(pid, dtype="<u4").tobytes())
