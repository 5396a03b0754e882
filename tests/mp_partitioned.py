"""One-process-per-GPU run of the partitioned solve, launched by torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mp_partitioned.py

Every rank owns one GPU and one partition of the same mesh; halos and all-reduces travel through cudaIpc-mapped peer
memory inside the solver kernels (csrc/comm.cuh).  Checked on every rank, for all four equations, with all three loop drivers (the persistent one runs Gray-Scott through pcg_dual):
  * the fields after each of 5 steps against the fp64 oracle (<= 1e-8 relative L2, the north star's per-step bar),
  * surface->values complete and identical on every rank after a synchronising step (the reference's data contract),
  * FEAB_SYNC_ROOT: values in and out through rank 0's host array only,
  * a multi-step device-resident call == the same steps one at a time.
Prints "MP_PARTITIONED_OK world=N" on rank 0 and exits 0; any failure exits non-zero on that rank.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from fea_visualizer_b200 import feab, meshes
    from fea_visualizer_b200.multigpu import connect_ranks
    from oracle import oracle as O
    from tests.util import rel_l2

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def gather_state(F, which, idx_map, row_node, n):
        """Global u / v assembled from the owners (all ranks get it)."""
        mine = torch.zeros(n, dtype=torch.float64, device="cuda")
        mine[torch.from_numpy(idx_map[row_node].astype(np.int64)).cuda()] = torch.from_numpy(F.get_state(which).astype(np.float64)).cuda()
        dist.all_reduce(mine)
        return mine.cpu().numpy()

    worst = 0.0
    for name, bc in (("ico20", 0), ("disk30", 0), ("disk30", 1)):
        if name == "ico20":
            xyz, tri = meshes.icosphere(20)
        else:
            xyz, tri = meshes.disk(30)
        ob = meshes.boundary_flags(len(xyz), tri)
        V = len(xyz)
        seeds = meshes.seed_nodes(V)
        D = O.Oracle(xyz, tri, ob, bc, real=64)
        for loop in (feab.LOOP_GRAPH_WHILE, feab.LOOP_HOST_BATCHED, feab.LOOP_PERSISTENT):
            F = feab.Feab(real=64, device=local)
            F.mg_config(rank, world)
            F.set_boundary_condition(bc)
            F.set_solver(loop_mode=loop)
            F.set_mesh(xyz, tri, ob)
            connect_ranks(F, dist)
            idx_map = F.idx_map()
            n = int((idx_map >= 0).sum())
            row_node, _ = F.mg_maps()
            for eq in (0, 1, 2, 3):
                for X in (F, D):
                    X.clear()
                    X.set_equation(eq)
                vals = np.zeros(V, np.float32)
                vals[seeds] = 1.0
                D.set_values(vals)
                host = vals.copy()
                for s in range(5):
                    if eq == 3 and 0 < s < 5:
                        host[seeds] = 1.0
                        dv = D.get_values()
                        dv[seeds] = 1.0
                        D.set_values(dv)
                    F.step(1, values=host, sync_values=True)   # the reference-facing call: values in, values out
                    D.step(1)
                    e = [rel_l2(host, D.get_values()), rel_l2(gather_state(F, 0, idx_map, row_node, n), D.get_state(0)),
                         rel_l2(gather_state(F, 1, idx_map, row_node, n), D.get_state(1))]
                    worst = max(worst, *e)
                    assert max(e) <= 1e-8, (name, bc, loop, eq, s, e)
                    # every rank holds the same complete values array
                    t = torch.from_numpy(host.copy()).cuda()
                    t0 = t.clone()
                    dist.broadcast(t0, 0)
                    assert bool((t == t0).all()), (name, eq, s, "values differ between ranks")
                # FEAB_SYNC_ROOT: only rank 0's host array is read and written; the peers' arrays are ignored and left alone
                host[seeds] = 0.75                      # the viewer painted between two steps: the peers must see it through rank 0 only
                root_in = host.copy() if rank == 0 else np.full(V, -7.0, np.float32)
                F.step(1, values=root_in, sync_values=feab.SYNC_ROOT)
                D.set_values(host)
                D.step(1)
                if rank == 0:
                    e = rel_l2(root_in, D.get_values())
                    assert e <= 1e-8, (name, bc, loop, eq, "root mode", e)
                else:
                    assert (root_in == -7.0).all(), "root mode wrote into a peer's host array"
                e = rel_l2(gather_state(F, 0, idx_map, row_node, n), D.get_state(0))
                assert e <= 1e-8, (name, bc, loop, eq, "root mode state", e)
                # K steps in one device-resident call == K single steps (bitwise on the owners' rows)
                F.step(3)
                D.step(3)
                e = rel_l2(gather_state(F, 0, idx_map, row_node, n), D.get_state(0))
                assert e <= 2e-8, (name, bc, loop, eq, "multi-step call", e)
            st = F.stats()
            assert st["kernel_launches"] > 0
            if loop == feab.LOOP_PERSISTENT:      # the last equation was Gray-Scott: both solves shared the dual kernel
                assert st["shared_iterations"] > 0, st
            dist.barrier()
            F.close()
    # ---- BASELINE config 4 on a partition: BiCGSTAB for the non-symmetric advection-diffusion operator (persistent kernel,
    #      y / z halos inside it) against the fp64 oracle in BiCGSTAB mode ----
    worst_bicg = 0.0
    for bc in (0, 1):
        xyz, tri = meshes.disk(30)
        ob = meshes.boundary_flags(len(xyz), tri)
        V = len(xyz)
        seeds = meshes.seed_nodes(V)
        D = O.Oracle(xyz, tri, ob, bc, real=64)
        D.set_solver(mode=1)
        F = feab.Feab(real=64, device=local)
        F.mg_config(rank, world)
        F.set_boundary_condition(bc)
        F.set_solver(solver=feab.SOLVER_BICGSTAB)
        F.set_mesh(xyz, tri, ob)
        connect_ranks(F, dist)
        idx_map = F.idx_map()
        n = int((idx_map >= 0).sum())
        row_node, _ = F.mg_maps()
        for X in (F, D):
            X.clear()
            X.set_equation(2)
        vals = np.zeros(V, np.float32)
        vals[seeds] = 1.0
        D.set_values(vals)
        host = vals.copy()
        for s in range(4):
            F.step(1, values=host, sync_values=True)
            D.step(1)
            e = [rel_l2(host, D.get_values()), rel_l2(gather_state(F, 0, idx_map, row_node, n), D.get_state(0))]
            worst_bicg = max(worst_bicg, *e)
            assert max(e) <= 1e-8, ("bicgstab", bc, s, e, F.stats()["iterations"], D.last_stats())
        dist.barrier()
        F.close()
    # ---- a mesh of the size the benchmark partitions (>= 1 M rows: many tiles per CTA, thousands of halo rows per
    #      neighbour), production loop driver only, against the fp64 oracle (MP_BIG_NU=0 skips it) ----
    big_nu = int(os.environ.get("MP_BIG_NU", "316"))
    worst_big = 0.0
    if big_nu > 0:
        xyz, tri = meshes.icosphere(big_nu, radius=big_nu / 32.0)
        V = len(xyz)
        ob = np.zeros(V, np.uint8)
        seeds = meshes.seed_nodes(V)
        D = O.Oracle(xyz, tri, ob, 0, real=64)
        F = feab.Feab(real=64, device=local)
        F.mg_config(rank, world)
        F.set_mesh(xyz, tri, ob)
        connect_ranks(F, dist)
        idx_map = F.idx_map()
        n = int((idx_map >= 0).sum())
        row_node, _ = F.mg_maps()
        for eq, nsteps in ((3, 3), (1, 2)):
            for X in (F, D):
                X.clear()
                X.set_equation(eq)
            vals = np.zeros(V, np.float32)
            vals[seeds] = 1.0
            D.set_values(vals)
            F.set_values(vals)
            for s in range(nsteps):
                if eq == 3 and s > 0:
                    F.brush(seeds, 1.0)
                    dv = D.get_values()
                    dv[seeds] = 1.0
                    D.set_values(dv)
                F.step(1)
                D.step(1)
                e = [rel_l2(gather_state(F, 0, idx_map, row_node, n), D.get_state(0)),
                     rel_l2(gather_state(F, 1, idx_map, row_node, n), D.get_state(1))]
                worst_big = max(worst_big, *e)
                assert max(e) <= 1e-8, ("big", big_nu, eq, s, e, F.stats()["iterations"], D.last_stats())
        dist.barrier()
        F.close()
    dist.barrier()
    if rank == 0:
        print(f"MP_PARTITIONED_OK world={world} worst_rel_l2={worst:.3e} bicgstab_worst_rel_l2={worst_bicg:.3e} big_nu={big_nu} worst_rel_l2_big={worst_big:.3e}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
