"""Host-side plumbing of the partitioned (multi-GPU) solve, SURVEY.md 8(e).

* ``connect_ranks``: one process per GPU (torchrun): exchanges the FeabMgExport blobs with torch.distributed
  (any backend; bytes only) and maps every peer's arena with cudaIpc.  torch is plumbing here, not the product.
* ``VirtualRanks``: P handles in ONE process on ONE device, advanced in lock-step by feab_mg_group_step.  Same
  kernels, same numbering, same mailboxes -- used by the GPU tests to check the partitioned data path against the
  single-GPU solver without P GPUs.
"""
from __future__ import annotations

import numpy as np

from . import feab as _feab


def gather_blobs(blob: bytes, dist, group=None):
    """all_gather of equally sized byte strings through torch.distributed (works with gloo on CPU tensors and nccl
    on CUDA tensors)."""
    import torch
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


def connect_ranks(F: "_feab.Feab", dist, group=None):
    """Call on every rank after set_mesh (and again after a boundary-condition change)."""
    rank = dist.get_rank(group)
    blobs = gather_blobs(F.mg_export(), dist, group)
    for q, b in enumerate(blobs):
        if q != rank:
            F.mg_connect(q, b, same_process=False)
    dist.barrier(group)
    return blobs


class VirtualRanks:
    """P partitioned handles on one device; mirrors the part of Feab's interface the tests drive."""

    def __init__(self, world, real=64, device=0, part=None, **solver):
        self.world = world
        self.real = np.float64 if real == 64 else np.float32
        self.hs = [_feab.Feab(real=real, device=device) for _ in range(world)]
        for r, h in enumerate(self.hs):
            h.mg_config(r, world)
            if solver:
                h.set_solver(**solver)
        self.part = part

    def set_boundary_condition(self, bc):
        for h in self.hs:
            h.set_boundary_condition(bc)
        if getattr(self, "V", 0):
            self._connect()

    def set_mesh(self, xyz, tri, on_boundary):
        self.V = len(xyz)
        for h in self.hs:
            if self.part is not None:
                h.mg_set_partition(self.part)
            h.set_mesh(xyz, tri, on_boundary)
        self._connect()

    def _connect(self):
        blobs = [h.mg_export() for h in self.hs]
        for r, h in enumerate(self.hs):
            for q, b in enumerate(blobs):
                if q != r:
                    h.mg_connect(q, b, same_process=True)
        self.rows = [h.mg_maps() for h in self.hs]        # (row_node, col_unknown) per rank
        self.idx_map = self.hs[0].idx_map()
        self.n = int((self.idx_map >= 0).sum())

    def _all(self, name, *a, **k):
        return [getattr(h, name)(*a, **k) for h in self.hs]

    def clear(self):
        self._all("clear")

    def set_equation(self, eq):
        self._all("set_equation", eq)

    def set_params(self, *a):
        self._all("set_params", *a)

    def set_values(self, v):
        self._all("set_values", v)

    def brush(self, nodes, strength=1.0):
        self._all("brush", nodes, strength)

    def step(self, nsteps=1):
        _feab.group_step(self.hs, nsteps)

    def get_values(self):
        """surface->values assembled from the owners (every rank's copy is only guaranteed on its own + ghost nodes)."""
        out = np.zeros(self.V, np.float32)
        for h, (row_node, _) in zip(self.hs, self.rows):
            out[row_node] = h.get_values()[row_node]
        return out

    def get_state(self, which):
        out = np.zeros(self.n, self.real)
        for h, (row_node, _) in zip(self.hs, self.rows):
            out[self.idx_map[row_node]] = h.get_state(which)
        return out

    def stats(self):
        return self.hs[0].stats()

    def unstable(self):
        return any(self._all("unstable"))

    def global_csr_dense(self, which):
        """The global matrix assembled from the ranks' local CSR pieces (small meshes only)."""
        A = np.zeros((self.n, self.n), self.real)
        P = np.zeros((self.n, self.n), bool)
        for h, (row_node, col_unknown) in zip(self.hs, self.rows):
            rowptr, col, val = h.csr(which)
            grow = self.idx_map[row_node]
            for r in range(len(row_node)):
                cs = col_unknown[col[rowptr[r]:rowptr[r + 1]]]
                A[grow[r], cs] = val[rowptr[r]:rowptr[r + 1]]
                P[grow[r], cs] = True
        return A, P

    def close(self):
        self._all("close")


class ThreadedRanks(VirtualRanks):
    """P partitioned handles on ONE device, each advanced by its own host thread through the ordinary feab_step: the
    PRODUCTION multi-GPU kernels (persistent pcg_dual / pcg_persistent / bicg_persistent with their in-kernel grid syncs,
    mailbox all-reduces, halo pushes and flags) run concurrently and talk through same-process peer pointers.  Their
    cooperative grids must be co-resident, so every handle's grid is capped at SMs / P CTAs (FEAB_PCG_MAX_GRID: half of what one
    handle alone may use, which keeps room for the short kernels of a rank that is one launch behind), and CUDA_DEVICE_MAX_CONNECTIONS must give
    every stream its own hardware queue (tests/conftest.py sets 32).
    This is what lets a ONE-GPU box check the kernels the N-GPU benchmark times."""

    def __init__(self, world, real=64, device=0, part=None, sm_count=148, **solver):
        import os
        old = os.environ.get("FEAB_PCG_MAX_GRID")
        os.environ["FEAB_PCG_MAX_GRID"] = str(max(1, sm_count // world))
        try:
            super().__init__(world, real=real, device=device, part=part, **solver)
        finally:
            if old is None:
                del os.environ["FEAB_PCG_MAX_GRID"]
            else:
                os.environ["FEAB_PCG_MAX_GRID"] = old

    def step(self, nsteps=1):
        import threading
        errs = []

        def run(r, h):
            try:
                h.step(nsteps)
            except Exception as e:  # noqa: BLE001 - re-raised below
                errs.append((r, e))

        ts = [threading.Thread(target=run, args=(r, h)) for r, h in enumerate(self.hs)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errs:
            if len(errs) == 1:
                raise errs[0][1]
            raise _feab.FeabError(getattr(errs[0][1], "code", -1), " || ".join(f"[rank {r}] {e}" for r, e in errs))
