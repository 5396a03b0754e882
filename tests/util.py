"""Shared test protocol: the same driving sequence for the oracle, the compiled reference and the CUDA path.

Mirrors Application::run (/root/reference/src/Application.cpp:506-565): the brush writes into surface->values,
then advance_time().  Seeds follow SURVEY.md 8(d).
"""
import os

import numpy as np

from fea_visualizer_b200 import meshes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


def run_steps(X, V, eq, nsteps, per_step=None):
    """X: any backend with clear/set_equation/set_values/get_values/step.  Returns surface->values."""
    seeds = meshes.seed_nodes(V)
    X.clear()
    X.set_equation(eq)
    v = np.zeros(V, np.float32)
    v[seeds] = 1.0
    X.set_values(v)
    for s in range(nsteps):
        if eq == 3 and 0 < s < 5:
            v = X.get_values()
            v[seeds] = 1.0
            X.set_values(v)
        X.step(1)
        if per_step is not None:
            per_step(s, X)
    return X.get_values()


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    nb = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / nb) if nb > 0 else float(np.linalg.norm(a))


def small_meshes():
    return {
        "ico4": meshes.icosphere(4),
        "grid": meshes.grid(12, 9),
        "disk": meshes.disk(6),
    }
