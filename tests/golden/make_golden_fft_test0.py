"""Extracts the known-answer table of the reference's FFT example into a small
fixture. Run HERE (the container that has /root/reference); the GPU box only
reads the committed JSON.

Source: /root/reference/FFTXlib/examples/reference/fft_test0.out
  - header facts (grid, stick and G-vector counts) pin oracle/descriptor.py
  - the first 'function in Real space (i,j,k)' block pins the G->R sign (+i),
    the x-fastest box layout and the no-scaling convention of invfft.
"""
import json
import os
import re

SRC = "/root/reference/FFTXlib/examples/reference/fft_test0.out"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fft_test0.json")

lines = open(SRC).read().splitlines()
start = next(i for i, l in enumerate(lines) if "function in Real space" in l)
vals = []
i = start + 1
pat = re.compile(r"\(\s*(-?\d+\.\d+)\s+(-?\d+\.\d+)\)")
while len(vals) < 125:
    for re_, im_ in pat.findall(lines[i]):
        vals.append([float(re_), float(im_)])
    i += 1
hdr = "\n".join(lines[:30])
grid = [int(x) for x in re.search(r"nr1 =\s+(\d+)\s+nr2 =\s+(\d+)\s+nr3 =\s+(\d+)", hdr).groups()]
alat = float(re.search(r"alat\s+=\s+([\d.]+)", hdr).group(1))
ecutwfc = float(re.search(r"Ecutwfc =\s+([\d.]+)", hdr).group(1))
summ = [int(x) for x in re.search(r"Sum\s+(\d+)\s+(\d+)\s+(\d+)\s+(\d+)\s+(\d+)\s+(\d+)", hdr).groups()]
json.dump(
    dict(
        source="FFTXlib/examples/reference/fft_test0.out",
        alat=alat, ecutwfc=ecutwfc, grid=grid,
        sticks_dense=summ[0], sticks_smooth=summ[1], sticks_pw=summ[2],
        g_dense=summ[3], g_smooth=summ[4], g_pw=summ[5],
        # lattice rows as set in FFTXlib/tests/fft_test.f90:185-192 (alat_in=18.65)
        at_rows=[[0.5, 1.0, 0.0], [0.5, 0.0, 0.5], [0.0, 0.5, 1.5]], alat_in=18.65,
        coeff={"1,1,1": [1.0, 0.0], "2,1,1": [0.0, 0.5], "1,2,1": [0.0, 0.3], "1,1,2": [0.0, 0.7]},
        real_space_5x5x5_kji=vals,  # order: k (z) outer, j (y) rows, i (x) fastest
    ),
    open(OUT, "w"), indent=0,
)
print("wrote", OUT, len(vals))
