import sys, numpy as np
sys.path.insert(0, '.')
from fea_visualizer_b200 import feab, meshes
from oracle import oracle as O
def P(*a): print(*a, flush=True)
xyz, tri = meshes.icosphere(8); ob = meshes.boundary_flags(len(xyz), tri); n = len(xyz)
v0 = np.zeros(n, np.float32); v0[meshes.seed_nodes(n)] = 1.0
D = O.Oracle(xyz, tri, ob, 0, real=64); D.set_solver(mode=1); D.set_equation(2); D.set_values(v0); D.step(1)
P('oracle iters', D.last_stats())
for loop in (feab.LOOP_HOST_BATCHED, feab.LOOP_GRAPH_WHILE):
    P('loop mode', loop)
    F = feab.Feab(real=64)
    F.set_solver(solver=feab.SOLVER_BICGSTAB, loop_mode=loop, assembly=feab.ASSEMBLY_ORDERED)
    P('set_mesh'); F.set_mesh(xyz, tri, ob)
    F.set_equation(2); F.set_values(v0)
    P('step'); F.step(1)
    P('stats', F.stats())
    a, b = F.get_state(0), D.get_state(0)
    P('rel', np.linalg.norm(a-b)/np.linalg.norm(b))
