import os, sys, time, numpy as np
sys.path.insert(0, '.')
from fea_visualizer_b200 import feab, meshes
nu = int(sys.argv[1]) if len(sys.argv) > 1 else 1265
variants = [int(v) for v in sys.argv[2].split(',')] if len(sys.argv) > 2 else [0, 1, 2, 3, 4, 5]
xyz, tri = meshes.icosphere(nu, radius=nu / 32.0); V = len(xyz); ob = np.zeros(V, np.uint8)
v0 = np.zeros(V, np.float32); v0[meshes.seed_nodes(V)] = 1.0
ref = None
for real in (64, 32):
  for var in variants:
    for pf in (0, 1):
        if var == 0 and pf: continue
        os.environ['FEAB_SPMV'] = str(var); os.environ['FEAB_PREFETCH_X'] = str(pf)
        F = feab.Feab(real=real); F.set_mesh(xyz, tri, ob)
        F.set_equation(0); F.set_values(v0); F.step(2)
        u = F.get_state(0); it = F.stats()['iterations']
        if var == 0: ref = u.copy()
        err = np.linalg.norm(u - ref) / np.linalg.norm(ref)
        F.set_equation(3)
        ms, by = F.time_kernel(0, 30)
        print(f"real={real} variant={var} prefetch_x={pf} spmv {ms*1e3:8.1f} us {by/ms/1e6:8.1f} GB/s  heat iters={it} rel_vs_v0={err:.2e}", flush=True)
        F.close()
