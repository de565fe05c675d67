"""Deterministic synthetic periodic cells of the BASELINE.json configs.

Geometry sources: the 8-atom cubic diamond cell of
examples/pbc/10-gamma_point_scf.py:14-34 (config 1), its 2x2x2 supercell built
with the rule of pyscf/pbc/tools/pbc.py:595-639 (config 2, mesh 35 -> 71), a
generated 32-molecule water box (config 3; the reference ships no such
geometry, so it is produced from a fixed seed here) and a 3x3x3 silicon
supercell (config 4).  Everything is gth-dzvp / gth-pade.
"""
import numpy as np
from .cell import Cell, super_cell

C8_FRAC = np.array([[0, 0, 0], [.25, .25, .25], [.5, .5, 0], [.75, .75, .25],
                    [.5, 0, .5], [.75, .25, .75], [0, .5, .5], [.25, .75, .75]])


def diamond_cubic(elem='C', a=3.5668, mesh=(35, 35, 35), basis='gth-dzvp'):
    atoms = [(elem, tuple(r)) for r in C8_FRAC * a]
    return Cell(atom=atoms, a=np.eye(3) * a, basis=basis, pseudo='gth-pade', mesh=list(mesh)).build()


def water_box(nside_mol=(4, 4, 2), L=9.8654, mesh=(96, 96, 96), seed=20240203):
    """32 rigid SPC/E-like molecules (r_OH = 1 Angstrom, HOH = 109.47 deg) on a
    jittered lattice, random orientations."""
    rng = np.random.default_rng(seed)
    n = np.asarray(nside_mol)
    atoms = []
    half = np.deg2rad(109.47) / 2
    for i in range(n[0]):
        for j in range(n[1]):
            for k in range(n[2]):
                o = (np.array([i, j, k]) + 0.5) / n * L + 0.3 * (rng.random(3) - .5)
                q, _ = np.linalg.qr(rng.standard_normal((3, 3)))
                h1 = o + q.dot([np.sin(half), 0.0, np.cos(half)])
                h2 = o + q.dot([-np.sin(half), 0.0, np.cos(half)])
                atoms += [('O', tuple(o)), ('H', tuple(h1)), ('H', tuple(h2))]
    return Cell(atom=atoms, a=np.eye(3) * L, basis='gth-dzvp', pseudo='gth-pade', mesh=list(mesh)).build()


WORKLOADS = {
    # name: (description, builder)
    'c8': ('diamond C8 cubic, gth-dzvp/gth-pade, mesh 35^3 (BASELINE configs[0])',
           lambda: diamond_cubic('C', 3.5668, (35, 35, 35))),
    'c64': ('diamond 2x2x2 supercell, 64 C, gth-dzvp/gth-pade, mesh 71^3 (BASELINE configs[1])',
            lambda: super_cell(diamond_cubic('C', 3.5668, (35, 35, 35)), [2, 2, 2])),
    'h2o32': ('32-molecule water box, gth-dzvp/gth-pade, mesh 96^3 (BASELINE configs[2])',
              lambda: water_box()),
    'si216': ('Si 3x3x3 supercell, 216 atoms, gth-dzvp/gth-pade, mesh 108^3 (BASELINE configs[3])',
              lambda: _with_mesh(super_cell(diamond_cubic('Si', 5.431, (36, 36, 36)), [3, 3, 3]), 108)),
    # small cells for smoke runs / CI
    'c8-small': ('diamond C8 cubic, gth-dzvp/gth-pade, mesh 20^3 (smoke)',
                 lambda: diamond_cubic('C', 3.5668, (20, 20, 20))),
}


def _with_mesh(cell, n):
    cell.mesh = [n, n, n]
    return cell


def make_cell(name):
    if name not in WORKLOADS:
        raise KeyError('unknown workload %s (have %s)' % (name, sorted(WORKLOADS)))
    return WORKLOADS[name][1]()


def describe(name):
    return WORKLOADS[name][0]


def synthetic_dm(nao, seed=2):
    """Random symmetric density matrix (the choice of
    pyscf/pbc/scf/test/test_hf.py:549-552)."""
    rng = np.random.default_rng(seed)
    d = rng.random((nao, nao))
    return d + d.T


def flop_model(nao, ng, npts, mesh, fit='implicit'):
    """Algorithmic FP64 flop / byte counts per stage."""
    ngh = int(mesh[0]) * int(mesh[1]) * (int(mesh[2]) // 2 + 1)
    f = {}
    f['select'] = 2.0 * ng * nao * npts + 1.0 * ng * npts ** 2
    if fit == 'implicit':
        f['fit'] = 2.0 * npts ** 3                     # W = U^-1 M U^-T: two triangular solves with Np right-hand sides
    elif fit == 'pivot_factor':
        f['fit'] = 1.0 * npts ** 2 * ng
    else:
        f['fit'] = 2.0 * npts * nao * ng + 2.0 * npts ** 2 * nao + npts ** 3 / 3.0 + 2.0 * npts ** 2 * ng
    f['coulomb_gemm'] = 2.0 * npts ** 2 * ngh           # upper triangle of Y Y^T, K = 2*ngh
    f['coulomb_fft_bytes'] = 8.0 * npts * ng + 16.0 * npts * ngh
    f['jk'] = 4.0 * npts ** 2 * nao + 6.0 * npts * nao ** 2 + 2.0 * npts ** 2
    f['jk_bytes'] = 8.0 * npts ** 2
    f['ao_bytes'] = 8.0 * ng * nao
    return f
