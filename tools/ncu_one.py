"""Tiny driver for ncu captures: build the 16M Gray-Scott operator and launch one hot kernel a few times.
usage: python tools/ncu_one.py <nu> <kernel id 0..3> <reps> [real]"""
import sys, numpy as np
sys.path.insert(0, '.')
from fea_visualizer_b200 import feab, meshes
nu = int(sys.argv[1]); kid = int(sys.argv[2]); reps = int(sys.argv[3]); real = int(sys.argv[4]) if len(sys.argv) > 4 else 64
xyz, tri = meshes.icosphere(nu, radius=nu / 32.0); V = len(xyz); ob = np.zeros(V, np.uint8)
F = feab.Feab(real=real); F.set_mesh(xyz, tri, ob); F.set_equation(3)
ms, by = F.time_kernel(kid, reps)
print(f"kernel {kid}: {ms*1e3:.1f} us, {by/ms/1e6:.1f} GB/s algorithmic")
