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


# ---------------------------------------------------------------------------
# non-local projectors (pyscf/pbc/df/fft.py:86-141), host side, Gamma point
# ---------------------------------------------------------------------------
def solid_harmonics(l, q, qq=None):
    """Real solid harmonics r^l Y_lm at the points q[n, 3] -> [2l+1, n], in the normalisation and
    order of the AO evaluator (l = 1 as x, y, z; l = 2 as xy, yz, 2zz-xx-yy, xz, xx-yy)."""
    x, y, z = q[:, 0], q[:, 1], q[:, 2]
    if l == 0:
        return np.full((1, len(q)), 0.282094791773878143)
    if l == 1:
        return 0.488602511902919921 * np.stack([x, y, z])
    if l == 2:
        c1, c0, c2 = 1.092548430592079070, 0.315391565252520002, 0.546274215296039535
        return np.stack([c1 * x * y, c1 * y * z, c0 * (2 * z * z - x * x - y * y), c1 * x * z, c2 * (x * x - y * y)])
    raise NotImplementedError('solid harmonics for l = %d' % l)


def _qli(x2, l, i):
    """Polynomial factor of projector i (from 0) of channel l as a function of x^2 = (|G| r_l)^2
    (pyscf/pbc/gto/pseudo/pp.py:150-194)."""
    rt = np.sqrt
    if (l, i) == (0, 0):
        return 4 * rt(2.) + 0 * x2
    if (l, i) == (0, 1):
        return 8 * rt(2 / 15.) * (3 - x2)
    if (l, i) == (0, 2):
        return 16 / 3. * rt(2 / 105.) * (15 - 10 * x2 + x2 * x2)
    if (l, i) == (1, 0):
        return 8 * rt(1 / 3.) + 0 * x2
    if (l, i) == (1, 1):
        return 16 * rt(1 / 105.) * (5 - x2)
    if (l, i) == (1, 2):
        return 32 / 3. * rt(1 / 1155.) * (35 - 14 * x2 + x2 * x2)
    if (l, i) == (2, 0):
        return 8 * rt(2 / 15.) + 0 * x2
    if (l, i) == (2, 1):
        return 16 / 3. * rt(2 / 105.) * (7 - x2)
    if (l, i) == (2, 2):
        return 32 / 3. * rt(2 / 15015.) * (63 - 18 * x2 + x2 * x2)
    raise NotImplementedError('projector l = %d, i = %d' % (l, i))


def species_projectors(pp, q, qq):
    """Rows of the projector table of one species at the points q: for every channel l with
    projectors, i = 0..nl-1, m = -l..l (the reference's pYlm order, fft.py:118-124) -> real
    [nproj, n]: r_l^(l+3/2) pi^(5/4) S_lm(q) exp(-r_l^2 q^2 / 2) q_li(|q| r_l)."""
    rows = []
    for l, (rl, nl, _) in enumerate(pp['nonlocal_']):
        if nl == 0:
            continue
        base = solid_harmonics(l, q, qq) * (rl**(l + 1.5) * np.pi**1.25 * np.exp(-0.5 * rl * rl * qq))
        for i in range(nl):
            rows.append(base * _qli(qq * rl * rl, l, i))
    return np.concatenate(rows) if rows else np.zeros((0, len(q)))


def species_h_blocks(pp):
    """[(l, nl, h[nl, nl])] of the channels that carry projectors; h symmetric from the table's
    upper-triangular rows (pyscf/gto/basis/parse_cp2k_pp.py:86-89)."""
    out = []
    for l, (rl, nl, rows) in enumerate(pp['nonlocal_']):
        if nl == 0:
            continue
        h = np.zeros((nl, nl))
        for i, row in enumerate(rows):
            h[i, i:] = row
        out.append((l, nl, h + h.T - np.diag(h.diagonal())))
    return out


def species_ft_shells(cell, ia, q, qq):
    """Centred analytic Fourier transforms of the AOs of atom ``ia`` at the points q, without the
    atom's phase: complex [nao_atom, n].  FT[S_lm(r) exp(-a r^2)](q) = (-i)^l (pi/a)^(3/2) (2a)^(-l)
    S_lm(q) exp(-q^2/4a)  (the closed form libcint's ft_ao evaluates; pyscf/pbc/df/ft_ao.py)."""
    bas, env = cell._bas, cell._env
    rows = []
    for ib in np.where(bas[:, 0] == ia)[0]:
        l, npr, nc = int(bas[ib, 1]), int(bas[ib, 2]), int(bas[ib, 3])
        alpha = env[bas[ib, 5]:bas[ib, 5] + npr]
        coeff = env[bas[ib, 6]:bas[ib, 6] + npr * nc].reshape(nc, npr)
        prim = np.exp(-np.outer(0.25 / alpha, qq)) * ((np.pi / alpha)**1.5 * (2 * alpha)**(-float(l)))[:, None]
        rad = coeff.dot(prim)                                          # [nc, n]
        ang = solid_harmonics(l, q, qq)                                # [2l+1, n]
        rows.append(((-1j)**l * rad[:, None, :] * ang[None, :, :]).reshape(nc * (2 * l + 1), -1))
    return np.concatenate(rows)


def _pair_classes(mesh, g0, g1):
    """For the flat reciprocal-grid indices [g0, g1): which are representatives of a +-G pair (the lower flat
    index of the two) and which stand alone (G = 0, or a Nyquist component -n/2 on an even axis, whose
    negative is not on the grid).  The partner of a representative may lie outside [g0, g1)."""
    nx, ny, nz = (int(m) for m in mesh)
    flat = np.arange(g0, g1)
    ix, rem = np.divmod(flat, ny * nz)
    iy, iz = np.divmod(rem, nz)
    neg = (((-ix) % nx) * ny + ((-iy) % ny)) * nz + ((-iz) % nz)
    single = neg == flat                                         # G = 0 and pure-Nyquist corners
    for i, n in ((ix, nx), (iy, ny), (iz, nz)):
        if n % 2 == 0:
            single |= i == n // 2
    # a pair in which exactly one member has a Nyquist component cannot occur: negation keeps index n/2
    rep = (~single) & (flat < neg)
    return rep, single


def pp_nl_overlaps(cell, g0=0, g1=None, mesh=None, max_bytes=1 << 29):
    """Sum over the reciprocal vectors [g0, g1) of  conj(SI_a) P_a(G) x ft_ao(G)  for every projector
    of every atom: complex [nproj_total, nao] (the reference's SPG_lm_aoGs, fft.py:129-130, for all
    atoms at once and without the 1/sqrt(vol)).  The sum over G is a plain sum, so ranks can take
    disjoint ranges and add their results.

    At the Gamma point the term of -G is (-1)^l times the conjugate of the term of G (real AOs, real
    projectors of parity l), so a +-G pair contributes 2 Re (even l) or 2i Im (odd l) of one term: only one
    member of each pair is evaluated, with real matrix products (a quarter of the arithmetic of the
    complex product over all G).  Vectors without a partner on the grid are summed as they are."""
    mesh = cell.mesh if mesh is None else mesh
    Gv = cell.get_Gv(mesh)
    g1 = len(Gv) if g1 is None else g1
    coords = cell.atom_coords()
    nao = cell.nao_nr()
    pps = [cell._pseudo_for(cell.atom_symbol(ia)) for ia in range(cell.natm)]
    first = {}
    for ia in range(cell.natm):
        first.setdefault(cell.atom_symbol(ia), ia)
    nproj = [sum(nl * (2 * l + 1) for l, (_, nl, _) in enumerate(pp['nonlocal_'])) if pp is not None else 0 for pp in pps]
    Z = np.zeros((sum(nproj), nao), dtype=np.complex128)
    if Z.shape[0] == 0 or g1 <= g0:
        return Z
    odd = np.concatenate([np.concatenate([np.full(nl * (2 * l + 1), l % 2 == 1) for l, (_, nl, _) in enumerate(pp['nonlocal_'])] or [np.zeros(0, bool)])
                          for pp in pps if pp is not None and any(nl for _, nl, _ in pp['nonlocal_'])])
    rep, single = _pair_classes(mesh, g0, g1)

    def tables(q):
        qq = np.einsum('gx,gx->g', q, q)
        SI = np.exp(-1j * np.dot(coords, q.T))                         # [natm, n]
        ft = {s: species_ft_shells(cell, ia, q, qq) for s, ia in first.items()}
        pr = {s: species_projectors(pps[ia], q, qq) for s, ia in first.items() if nproj[ia]}
        aoG = np.concatenate([ft[cell.atom_symbol(ia)] * SI[ia] for ia in range(cell.natm)])       # [nao, n]
        spg = np.concatenate([pr[cell.atom_symbol(ia)] * SI[ia].conj() for ia in range(cell.natm) if nproj[ia]])
        return spg, aoG

    step = max(256, int(max_bytes // (16 * (nao + sum(nproj) + 1))))
    idx = g0 + np.nonzero(single)[0]
    for b0 in range(0, len(idx), step):                                # unpaired vectors: complex product
        spg, aoG = tables(Gv[idx[b0:b0 + step]])
        Z += spg.dot(aoG.T)
    idx = g0 + np.nonzero(rep)[0]
    Zr = np.zeros(Z.shape)
    for b0 in range(0, len(idx), step):                                # one of each pair: real products
        spg, aoG = tables(Gv[idx[b0:b0 + step]])
        A = np.where(odd[:, None], np.concatenate([spg.imag, spg.real], axis=1),
                     np.concatenate([spg.real, -spg.imag], axis=1))   # Im t (odd l) / Re t (even l)
        Zr += A.dot(np.concatenate([aoG.real, aoG.imag], axis=1).T)
    Z += 2 * np.where(odd[:, None], 1j * Zr, Zr)
    return Z


def pp_nl_overlaps_device(cell, h, g0=0, g1=None, mesh=None, batch=32768):
    """pp_nl_overlaps with the part that scales with the number of atoms on the device: per batch of G vectors the
    host prepares the per-SPECIES tables (closed forms above), `isdf_phase_expand` turns them into every atom's
    rows conj(SI_a) P(G) and SI_a ft_ao(G) (real and imaginary planes), and four `isdf_dgemm` calls accumulate
    Z = spg . aoG^T over the batches (the chain continues through Cin).  Returns two device tensors
    (Z_re, Z_im) of shape [nproj_total, nao], the sum over the reciprocal vectors [g0, g1)."""
    from . import _lib
    mesh = cell.mesh if mesh is None else mesh
    Gv = cell.get_Gv(mesh)
    g1 = len(Gv) if g1 is None else g1
    coords = cell.atom_coords()
    nao = cell.nao_nr()
    syms = [cell.atom_symbol(ia) for ia in range(cell.natm)]
    pps = [cell._pseudo_for(s) for s in syms]
    first = {}
    for ia, sym in enumerate(syms):
        first.setdefault(sym, ia)
    nproj = [sum(nl * (2 * l + 1) for l, (_, nl, _) in enumerate(pp['nonlocal_'])) if pp is not None else 0 for pp in pps]
    nao_atom = [int(sum((2 * cell._bas[ib, 1] + 1) * cell._bas[ib, 3] for ib in np.where(cell._bas[:, 0] == ia)[0]))
                for ia in range(cell.natm)]
    ao_off = np.concatenate([[0], np.cumsum(nao_atom)])
    p_off = np.concatenate([[0], np.cumsum(nproj)])
    ntot = int(p_off[-1])
    Zre, Zim = h.zeros((max(ntot, 1), nao)), h.zeros((max(ntot, 1), nao))
    if ntot == 0 or g1 <= g0:
        return Zre[:ntot], Zim[:ntot]
    ld = _lib.pad8(min(batch, g1 - g0))
    A_re, A_im = h.empty((nao, ld)), h.empty((nao, ld))
    S_re, S_im = h.empty((ntot, ld)), h.empty((ntot, ld))
    for b0 in range(g0, g1, batch):
        b1 = min(b0 + batch, g1)
        n = b1 - b0
        q = np.ascontiguousarray(Gv[b0:b1])
        qq = np.einsum('gx,gx->g', q, q)
        d_G = h.from_host(q)
        tabs = {}
        for sym, ia in first.items():
            ft = np.zeros((nao_atom[ia], ld), dtype=np.complex128)
            ft[:, :n] = species_ft_shells(cell, ia, q, qq)
            pr = None
            if nproj[ia]:
                pr = np.zeros((nproj[ia], ld))
                pr[:, :n] = species_projectors(pps[ia], q, qq)
            tabs[sym] = (h.from_host(np.ascontiguousarray(ft.real)), h.from_host(np.ascontiguousarray(ft.imag)),
                         h.from_host(pr) if pr is not None else None)
        for ia, sym in enumerate(syms):
            t_re, t_im, t_pr = tabs[sym]
            h.phase_expand(t_re, t_im, ld, nao_atom[ia], d_G, n, coords[ia], -1.0,
                           A_re[ao_off[ia]:], A_im[ao_off[ia]:], ld)                       # ft_ao * SI
            if nproj[ia]:
                h.phase_expand(t_pr, None, ld, nproj[ia], d_G, n, coords[ia], +1.0,
                               S_re[p_off[ia]:], S_im[p_off[ia]:], ld)                     # P * conj(SI)
        # Z += (S_re + i S_im) (A_re + i A_im)^T over this batch
        h.dgemm(ntot, nao, n, S_re, ld, 0, A_re, ld, 0, Zre, nao, Cin=Zre, ldcin=nao)
        h.dgemm(ntot, nao, n, S_im, ld, 0, A_im, ld, 0, Zre, nao, Cin=Zre, ldcin=nao, neg=1)
        h.dgemm(ntot, nao, n, S_re, ld, 0, A_im, ld, 0, Zim, nao, Cin=Zim, ldcin=nao)
        h.dgemm(ntot, nao, n, S_im, ld, 0, A_re, ld, 0, Zim, nao, Cin=Zim, ldcin=nao)
        h.synchronize()                               # the species tables of this batch are temporaries
    return Zre, Zim


def pp_nl_from_overlaps(cell, Z):
    """vppnl[nao, nao] = sum_atoms sum_l sum_m  Z^H h_l Z / vol^2  (fft.py:131-139; each ft_ao carries
    1/sqrt(vol) there and the result another 1/vol); real part at the Gamma point (fft.py:144-145)."""
    nao = cell.nao_nr()
    out = np.zeros((nao, nao), dtype=np.complex128)
    p = 0
    for ia in range(cell.natm):
        pp = cell._pseudo_for(cell.atom_symbol(ia))
        if pp is None:
            continue
        for l, nl, h in species_h_blocks(pp):
            n = nl * (2 * l + 1)
            z = Z[p:p + n].reshape(nl, 2 * l + 1, nao)
            out += np.einsum('imp,imq->pq', z.conj(), np.einsum('ij,jmp->imp', h, z))
            p += n
    return out.real / cell.vol**2


def pp_nl_matrix(cell, mesh=None):
    return pp_nl_from_overlaps(cell, pp_nl_overlaps(cell, mesh=mesh))


def has_nonlocal(cell):
    """True if any atom carries non-local GTH projectors (pyscf/pbc/df/fft.py:86-141 needs them)."""
    for ia in range(cell.natm):
        pp = cell._pseudo_for(cell.atom_symbol(ia))
        if pp is not None and any(nproj > 0 for _, nproj, _ in pp['nonlocal_']):
            return True
    return False
