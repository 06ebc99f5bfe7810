#!/usr/bin/env python
"""Write tests/golden/h5/corrupt_index.h5 and corrupt_index.bin: copies of earliest_contiguous.h5 /
cfg1_300.bin in which ONE burst's photon index points outside the photon arrays (istop of burst 3
= 2^40 in the HDF5 file; istart of burst 3 = -5 in the GSMB file).  The loader must refuse both
before indexing anything (tests/test_hdf5_loader.py)."""
import os
import struct
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(HERE)))
PROBE = os.path.join(ROOT, "gsmfretda_b200", "host", "h5_probe")


def main():
    src = os.path.join(HERE, "earliest_contiguous.h5")
    out = subprocess.run([PROBE, "read", src, "/sub_bursts_l/istop", "i64"], capture_output=True, text=True, check=True).stdout.split()
    vals = [int(v) for v in out[2:]]
    raw = open(src, "rb").read()
    needle = struct.pack("<%dq" % len(vals), *vals)      # contiguous layout: the values lie in the file as they are
    at = raw.find(needle)
    assert at > 0 and raw.find(needle, at + 1) < 0
    bad = bytearray(raw)
    bad[at + 8 * 3: at + 8 * 4] = struct.pack("<q", 1 << 40)
    open(os.path.join(HERE, "corrupt_index.h5"), "wb").write(bytes(bad))

    b = bytearray(open(os.path.join(HERE, "cfg1_300.bin"), "rb").read())
    B, = struct.unpack_from("<q", b, 8)
    istart_at = 40 + 16 * B                               # header, start[B], stop[B], then istart[B]
    struct.pack_into("<q", b, istart_at + 8 * 3, -5)
    open(os.path.join(HERE, "corrupt_index.bin"), "wb").write(bytes(b))


if __name__ == "__main__":
    main()
