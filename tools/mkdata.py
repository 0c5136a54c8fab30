"""Write a synthetic corpus file: python tools/mkdata.py <kind> <path> <MiB per block> <copies>"""
import sys
import os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dmlc_core_b200 import datagen
kind, path, mb, reps = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
data, rows = datagen.generate(kind, 7, mb << 20)
with open(path, "wb") as f:
    for _ in range(reps):
        f.write(data.tobytes())
print(kind, path, rows * reps, "rows", len(data) * reps >> 20, "MB")
