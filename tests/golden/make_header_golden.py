"""Golden vectors for the .cool header writers (coolchic_b200/enc/bitstream/encode.py: gop_header_bytes,
frame_header_bytes): the GOP + frame header bytes of bitstreams SHIPPED with the reference
(/root/reference/results/image/*/bitstreams/*.cool), copied verbatim (89 bytes each).
Run in the build container (needs /root/reference):  python tests/golden/make_header_golden.py"""
import os

import numpy as np

REFERENCE_ROOT = os.environ.get("CC_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
FILES = [("kodak", "kodim01-lmbda-0001"), ("kodak", "kodim02-lmbda-002"), ("kodak", "kodim13-lmbda-00001"),
         ("clic20-pro-valid", "amy-zhang-15940-lmbda-0004"), ("clic20-pro-valid", "davide-ragusa-716-lmbda-0004")]

out = {}
for ds, name in FILES:
    with open(os.path.join(REFERENCE_ROOT, "results", "image", ds, "bitstreams", name + ".cool"), "rb") as f:
        data = f.read()
    n_gop = int.from_bytes(data[0:2], "big")
    n_frame = int.from_bytes(data[n_gop:n_gop + 2], "big")
    out[name] = np.frombuffer(data[:n_gop + n_frame], dtype=np.uint8).copy()
    print(name, len(data), "bytes; headers", n_gop, "+", n_frame)
np.savez(os.path.join(HERE, "cool_headers.npz"), **out)
