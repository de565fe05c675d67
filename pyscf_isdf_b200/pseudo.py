"""Local one-electron potentials on the FFT grid: nuclear attraction and the
local part of the GTH pseudopotential ("next" row 2 of SURVEY.md 8f).

What the reference does (pyscf/pbc/df/fft.py:39-84): build the potential in
reciprocal space from the per-atom form factor and the structure factor,
``vG = -sum_a SI[a,G] vloc[a,G]``, transform it once to real space and contract
it with the AO values on the grid, ``V = Phi diag(vR) Phi^T``.  Here the
reciprocal-space vector is assembled on the host (one vector of ``ng`` numbers,
once per geometry) and the contraction runs on the device over the AO values
the ISDF build already holds (``isdf_j_from_potential``).

Differences from a literal restatement, none of which change a number beyond
rounding: the form factor depends on ``|G|`` and the species only, so atoms are
grouped by species and the structure factor of a species is accumulated from the
separable products ``exp(-i G_x.r) exp(-i G_y.r) exp(-i G_z.r)`` of
pyscf/pbc/gto/cell.py:635-643 -- ``natm`` outer products instead of a dense
``natm x ng`` exponential.
"""
import numpy as np


def species_structure_factors(cell, mesh=None):
    """{symbol: sum over atoms of that symbol of SI[a, :]} as complex [ng]
    (structure factor of pyscf/pbc/gto/cell.py:613-646, separable branch)."""
    mesh = cell.mesh if mesh is None else mesh
    basex, basey, basez = cell.get_Gv_weights(mesh)[1]
    b = cell.reciprocal_vectors()
    rb = np.dot(cell.atom_coords(), b.T)
    out = {}
    for ia in range(cell.natm):
        sx = np.exp(-1j * rb[ia, 0] * basex)
        sy = np.exp(-1j * rb[ia, 1] * basey)
        sz = np.exp(-1j * rb[ia, 2] * basez)
        si = (sx[:, None, None] * sy[None, :, None]) * sz[None, None, :]
        sym = cell.atom_symbol(ia)
        if sym in out:
            out[sym] += si
        else:
            out[sym] = si
    return {s: v.reshape(-1) for s, v in out.items()}


def _coulG(G2):
    with np.errstate(divide='ignore'):
        c = 4 * np.pi / G2
    c[G2 == 0] = 0                      # pyscf/pbc/tools/pbc.py:352-356
    return c


def gth_form_factor(pp, Z, G2):
    """vloc(G) of one GTH species (positive-potential sign convention):
    pyscf/pbc/gto/pseudo/pp_int.py:51-71 (erf part, non-divergent G=0 term) and
    pyscf/pbc/gto/pseudo/pp.py:77-92 (C1..C4 polynomial part)."""
    rloc = pp['rloc']
    cexp = pp['cloc']
    nexp = len(cexp)
    g2r = G2 * rloc**2
    v = Z * _coulG(G2) * np.exp(-0.5 * g2r)
    v[G2 == 0] = -2 * np.pi * Z * rloc**2
    cfacs = 0
    if nexp >= 1:
        cfacs = cfacs + cexp[0]
    if nexp >= 2:
        cfacs = cfacs + cexp[1] * (3 - g2r)
    if nexp >= 3:
        cfacs = cfacs + cexp[2] * (15 - 10 * g2r + g2r**2)
    if nexp >= 4:
        cfacs = cfacs + cexp[3] * (105 - 105 * g2r + 21 * g2r**2 - g2r**3)
    v -= (2 * np.pi)**1.5 * rloc**3 * np.exp(-0.5 * g2r) * cfacs
    return v


def _ifft_real(vG, mesh):
    """tools.ifft(vG, mesh).real (pyscf/pbc/tools/pbc.py:177-211) for one vector."""
    return np.ascontiguousarray(np.fft.ifftn(vG.reshape(tuple(int(n) for n in mesh))).real.reshape(-1))


def pp_loc_on_grid(cell, mesh=None):
    """vpplocR[ng] of FFTDF.get_pp (pyscf/pbc/df/fft.py:70-77); the ifft's 1/ng makes it the
    potential times the grid weight vol/ng, so V = Phi diag(vpplocR) Phi^T needs no extra weight.
    Atoms without a pseudopotential contribute the bare -Z coulG (pp_int.py:65-66)."""
    mesh = cell.mesh if mesh is None else mesh
    Gv = cell.get_Gv(mesh)
    G2 = np.einsum('gx,gx->g', Gv, Gv)
    SI = species_structure_factors(cell, mesh)
    charges = cell.atom_charges()
    vG = np.zeros(len(G2), dtype=np.complex128)
    for sym, si in SI.items():
        ia = next(i for i in range(cell.natm) if cell.atom_symbol(i) == sym)
        Z = float(charges[ia])
        pp = cell._pseudo_for(sym)
        vG -= si * (gth_form_factor(pp, Z, G2) if pp is not None else Z * _coulG(G2))
    return _ifft_real(vG, mesh)


def nuc_on_grid(cell, mesh=None):
    """vneR[ng] of FFTDF.get_nuc (pyscf/pbc/df/fft.py:45-51): point charges -Z, coulG with G=0 -> 0."""
    mesh = cell.mesh if mesh is None else mesh
    Gv = cell.get_Gv(mesh)
    G2 = np.einsum('gx,gx->g', Gv, Gv)
    SI = species_structure_factors(cell, mesh)
    charges = cell.atom_charges()
    rhoG = np.zeros(len(G2), dtype=np.complex128)
    for sym, si in SI.items():
        ia = next(i for i in range(cell.natm) if cell.atom_symbol(i) == sym)
        rhoG -= float(charges[ia]) * si
    return _ifft_real(rhoG * _coulG(G2), mesh)


def has_nonlocal(cell):
    """True if any atom carries non-local GTH projectors (pyscf/pbc/df/fft.py:86-141 needs them)."""
    for ia in range(cell.natm):
        pp = cell._pseudo_for(cell.atom_symbol(ia))
        if pp is not None and any(nproj > 0 for _, nproj, _ in pp['nonlocal_']):
            return True
    return False
