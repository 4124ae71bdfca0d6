"""Fixture readers -- mirror of crf/utils.py:46-109 (host side; no GPU work).

``read_image`` (:46-51), ``read_pfm`` (:59-90) and ``read_pgm`` (:92-109) with the reference's
return conventions: images as float64 in [0, 1] ``[h, w, 3]``; PFM rows flipped to top-down
order; PGM as the raw integer raster.  Differences, all on paths where the reference fails:
16-bit PGMs are read big-endian as the Netpbm format specifies (the reference names an undefined
``byteorder`` there, :104), colour PFMs come back as ``[h, w, 3]`` (the reference's reshape to
``(height, width)`` raises for them, :90), and a malformed header raises ``ValueError`` instead
of ``sys.exit`` / ``NameError``.
"""
import re

import numpy as np


def read_image(imgname):
    from PIL import Image
    img = Image.open(imgname).convert('RGB')
    return np.array(img).astype(float) / 255.


def read_pfm(file):
    with open(file, "rb") as f:
        kind = f.readline().decode('latin-1')
        if "PF" in kind:
            channels = 3
        elif "Pf" in kind:
            channels = 1
        else:
            raise ValueError(f"Not a PFM file: '{file}'")
        dims = re.findall(r'\d+', f.readline().decode('latin-1'))
        if len(dims) != 2:
            raise ValueError(f"Bad PFM header in '{file}'")
        width, height = int(dims[0]), int(dims[1])
        big_endian = "-" not in f.readline().decode('latin-1')   # negative scale = little endian
        samples = width * height * channels
        buffer = f.read(samples * 4)
    if len(buffer) != samples * 4:
        raise ValueError(f"Truncated PFM file: '{file}'")
    img = np.frombuffer(buffer, dtype=(">f4" if big_endian else "<f4")).astype(np.float64)
    shape = (height, width) if channels == 1 else (height, width, 3)
    return np.flip(img.reshape(shape), axis=0)


_PGM_HEADER = re.compile(
    rb"(^P5\s(?:\s*#.*[\r\n])*"
    rb"(\d+)\s(?:\s*#.*[\r\n])*"
    rb"(\d+)\s(?:\s*#.*[\r\n])*"
    rb"(\d+)\s(?:\s*#.*[\r\n]\s)*)")


def read_pgm(file):
    with open(file, 'rb') as f:
        buffer = f.read()
    m = _PGM_HEADER.search(buffer)
    if m is None:
        raise ValueError("Not a raw PGM file: '%s'" % file)
    header, width, height, maxval = m.groups()
    return np.frombuffer(buffer, dtype='u1' if int(maxval) < 256 else '>u2',
                         count=int(width) * int(height),
                         offset=len(header)).reshape((int(height), int(width)))
