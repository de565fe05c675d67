"""Minimal periodic-cell host model.

PySCF cannot be imported on the build/GPU image (no h5py, no libcint), so the
ISDF plugin carries its own ``Cell``.  It reproduces exactly those pieces of
``pyscf.pbc.gto.Cell`` that the ISDF hot path consumes:

* the libcint-layout integer/float tables ``_atm/_bas/_env`` that the AO
  kernel takes as input (slot layout: pyscf/gto/mole.py:59-89; coefficient
  normalisation: pyscf/gto/mole.py:976-1023),
* lattice / reciprocal vectors, uniform grid, G-vectors
  (pyscf/pbc/gto/cell.py:523-610, :874-897, :1545-1592),
* the precision-driven defaults ``rcut`` and ``mesh``
  (pyscf/pbc/gto/cell.py:389-461, :1418-1447; pyscf/pbc/tools/pbc.py:703-726),
* per-shell AO cutoff radii and the sorted lattice-image list used by the
  lattice-sum AO evaluation (pyscf/pbc/gto/eval_gto.py:132-137, :169-253),
* Ewald / Madelung for ``exxdiv='ewald'``
  (pyscf/pbc/gto/cell.py:666-813, pyscf/pbc/tools/pbc.py:483-493).

Only 3-D periodic, Gamma-point, GTH-pseudopotential cells are supported;
everything else raises.
"""
import math
import numpy as np
import scipy.special

from . import gth_data

BOHR = 0.52917721092          # pyscf/data/nist.py:24
PTR_ENV_START = 20            # pyscf/gto/mole.py:89
ATM_SLOTS, BAS_SLOTS = 6, 8
CHARGE_OF, PTR_COORD, NUC_MOD_OF, PTR_ZETA = 0, 1, 2, 3
ATOM_OF, ANG_OF, NPRIM_OF, NCTR_OF, KAPPA_OF, PTR_EXP, PTR_COEFF = 0, 1, 2, 3, 4, 5, 6


def cartesian_prod(arrays):
    """Row-major cartesian product, first array slowest
    (same ordering as pyscf.lib.cartesian_prod)."""
    grids = np.meshgrid(*arrays, indexing='ij')
    return np.stack([g.ravel() for g in grids], axis=1)


def gaussian_int(n, alpha):
    """int_0^inf x^n exp(-alpha x^2) dx"""
    n1 = (n + 1) * .5
    return scipy.special.gamma(n1) / (2. * alpha**n1)


def gto_norm(l, expnt):
    """1/sqrt(int r^2 (r^l e^{-a r^2})^2 dr)  (pyscf/gto/mole.py:121-151)."""
    return 1. / np.sqrt(gaussian_int(l * 2 + 2, 2 * np.asarray(expnt, dtype=float)))


def _normalize_contracted(l, es, cs):
    ee = gaussian_int(l * 2 + 2, es[:, None] + es[None, :])
    s1 = 1. / np.sqrt(np.einsum('pi,pq,qi->i', cs, ee, cs))
    return cs * s1


def _parse_atoms(atom, unit):
    if isinstance(atom, str):
        items = []
        for line in atom.replace(';', '\n').split('\n'):
            tok = line.replace(',', ' ').split()
            if not tok:
                continue
            items.append((tok[0], tuple(float(x) for x in tok[1:4])))
    else:
        items = []
        for a in atom:
            if len(a) == 2:
                items.append((a[0], tuple(float(x) for x in a[1])))
            else:
                items.append((a[0], tuple(float(x) for x in a[1:4])))
    scale = 1.0 if unit.upper().startswith(('B', 'AU')) else 1.0 / BOHR
    return [(s, tuple(c * scale for c in xyz)) for s, xyz in items]


def _pure_symbol(sym):
    return ''.join(ch for ch in sym if ch.isalpha()).capitalize()


def _shells_to_bas_env(shells, ptr):
    """Shell list -> (_bas rows without atom id, env chunk).  Mirrors the
    conventions of pyscf/gto/mole.py:976-1011: primitives sorted by
    descending exponent, coefficients multiplied by the primitive radial
    norm, then each contraction renormalised."""
    bas, env = [], []
    for sh in shells:
        l = int(sh[0])
        rows = np.array(sorted([tuple(r) for r in sh[1:]], reverse=True), dtype=float)
        es = rows[:, 0].copy()
        cs = rows[:, 1:] * gto_norm(l, es)[:, None]
        cs = _normalize_contracted(l, es, cs)
        nprim, nctr = cs.shape
        bas.append([0, l, nprim, nctr, 0, ptr, ptr + nprim, 0])
        env.append(es)
        env.append(cs.T.ravel())
        ptr += nprim + nprim * nctr
    return np.array(bas, dtype=np.int32).reshape(-1, BAS_SLOTS), np.concatenate(env), ptr


class Cell:
    """Subset of ``pyscf.pbc.gto.Cell`` (see module docstring)."""

    def __init__(self, atom=None, a=None, basis='gth-dzvp', pseudo='gth-pade',
                 mesh=None, precision=1e-8, unit='Angstrom', ke_cutoff=None,
                 charge=0, verbose=0):
        self.atom = atom
        self.a = a
        self.basis = basis
        self.pseudo = pseudo
        self.unit = unit
        self.precision = precision
        self.ke_cutoff = ke_cutoff
        self.charge = charge
        self.verbose = verbose
        self.dimension = 3
        self.omega = 0
        self.incore_anyway = False
        self.max_memory = 4000
        self._mesh = mesh
        self._rcut = None
        self._built = False

    # -- construction -----------------------------------------------------
    def build(self):
        if self.a is None or self.atom is None:
            raise ValueError('Cell needs .a and .atom')
        scale = 1.0 if self.unit.upper().startswith(('B', 'AU')) else 1.0 / BOHR
        self._a = np.asarray(self.a, dtype=float).reshape(3, 3) * scale
        self._atom = _parse_atoms(self.atom, self.unit)
        natm = len(self._atom)

        # env: 20 reserved slots, then (x,y,z,zeta) per atom, then basis data
        env = [np.zeros(PTR_ENV_START)]
        ptr = PTR_ENV_START
        atm = np.zeros((natm, ATM_SLOTS), dtype=np.int32)
        for ia, (sym, xyz) in enumerate(self._atom):
            elem = _pure_symbol(sym)
            atm[ia, CHARGE_OF] = self._atom_charge(sym, elem)
            atm[ia, PTR_COORD] = ptr
            atm[ia, NUC_MOD_OF] = 1
            atm[ia, PTR_ZETA] = ptr + 3
            env.append(np.array(list(xyz) + [0.0]))
            ptr += 4

        basdic = {}
        for sym, _ in self._atom:
            if sym in basdic:
                continue
            shells = self._shells_for(sym)
            b, e, ptr = _shells_to_bas_env(shells, ptr)
            basdic[sym] = b
            env.append(e)
        bas = []
        for ia, (sym, _) in enumerate(self._atom):
            b = basdic[sym].copy()
            b[:, ATOM_OF] = ia
            bas.append(b)
        self._atm = atm
        self._bas = np.concatenate(bas, axis=0).astype(np.int32) if bas else \
            np.zeros((0, BAS_SLOTS), np.int32)
        self._env = np.concatenate(env)
        self._built = True

        if self._rcut is None:
            self._rcut = self._estimate_rcut_overlap()
        if self._mesh is None:
            ke = self.ke_cutoff if self.ke_cutoff is not None else self._estimate_ke_cutoff()
            self._mesh = cutoff_to_mesh(self._a, ke)
        self._mesh = np.asarray(self._mesh, dtype=int)
        return self

    def _shells_for(self, sym):
        elem = _pure_symbol(sym)
        b = self.basis
        if isinstance(b, dict):
            b = b.get(sym, b.get(elem))
            if b is None:
                raise KeyError('no basis for %s' % sym)
        if isinstance(b, str):
            return gth_data.gth_basis(b, elem)
        return [list(sh) for sh in b]

    def _pseudo_for(self, sym):
        elem = _pure_symbol(sym)
        p = self.pseudo
        if isinstance(p, dict):
            p = p.get(sym, p.get(elem))
        if p is None:
            return None
        return gth_data.gth_pseudo(p, elem)

    def _atom_charge(self, sym, elem):
        pp = self._pseudo_for(sym)
        if pp is not None:
            return int(sum(pp['nelec']))
        return gth_data.ELEMENT_Z[elem]

    def copy(self):
        import copy
        return copy.copy(self)

    # -- basic queries ----------------------------------------------------
    @property
    def natm(self):
        return len(self._atm)

    @property
    def nbas(self):
        return len(self._bas)

    @property
    def mesh(self):
        return self._mesh

    @mesh.setter
    def mesh(self, x):
        self._mesh = None if x is None else np.asarray(x, dtype=int)

    @property
    def rcut(self):
        return self._rcut

    @rcut.setter
    def rcut(self, x):
        self._rcut = x

    @property
    def vol(self):
        return abs(np.linalg.det(self._a))

    @property
    def nelectron(self):
        return int(self._atm[:, CHARGE_OF].sum()) - self.charge

    def lattice_vectors(self):
        return self._a

    def reciprocal_vectors(self, norm_to=2 * np.pi):
        return norm_to * np.linalg.inv(self._a.T)

    def atom_coords(self):
        return np.array([self._env[p:p + 3] for p in self._atm[:, PTR_COORD]])

    def atom_charges(self):
        return self._atm[:, CHARGE_OF].astype(float)

    def atom_symbol(self, ia):
        return self._atom[ia][0]

    def bas_angular(self, ib):
        return int(self._bas[ib, ANG_OF])

    def bas_exp(self, ib):
        p, n = self._bas[ib, PTR_EXP], self._bas[ib, NPRIM_OF]
        return self._env[p:p + n]

    def _libcint_ctr_coeff(self, ib):
        n, c, p = self._bas[ib, NPRIM_OF], self._bas[ib, NCTR_OF], self._bas[ib, PTR_COEFF]
        return self._env[p:p + n * c].reshape(c, n).T

    def ao_loc_nr(self):
        dims = (self._bas[:, ANG_OF] * 2 + 1) * self._bas[:, NCTR_OF]
        return np.concatenate([[0], np.cumsum(dims)]).astype(np.int32)

    def nao_nr(self):
        return int(self.ao_loc_nr()[-1])

    @property
    def nao(self):
        return self.nao_nr()

    # -- precision-driven defaults -----------------------------------------
    def _extract_pgto(self, op):
        es, cs = [], []
        for ib in range(self.nbas):
            e = self.bas_exp(ib)
            c = self._libcint_ctr_coeff(ib)
            idx = e.argmin() if op == 'min' else e.argmax()
            es.append(e[idx])
            cs.append(abs(c[idx]).max())
        return np.array(es), np.array(cs)

    def _estimate_rcut_overlap(self):
        """Lattice-sum cutoff for the whole cell
        (pyscf/pbc/gto/cell.py:389-433)."""
        if self.nbas == 0:
            return 0.01
        es, cs = self._extract_pgto('min')
        ls = self._bas[:, ANG_OF]
        return float(_rcut_overlap(es, ls, cs, self.precision).max())

    def _estimate_ke_cutoff(self):
        """pyscf/pbc/gto/cell.py:435-461"""
        if self.nbas == 0:
            return 0.
        es, cs = self._extract_pgto('max')
        ls = self._bas[:, ANG_OF]
        norm_ang = (2 * ls + 1) / (4 * np.pi)
        fac = (32 * np.pi**2 * (2 * np.pi)**1.5 * cs**2 * norm_ang
               / (2 * es)**(2 * ls + .5) / self.precision)
        Ecut = 20.
        Ecut = np.log(fac * (Ecut * 2)**(ls - .5) + 1.) * 4 * es
        Ecut = np.log(fac * (Ecut * 2)**(ls - .5) + 1.) * 4 * es
        return float(Ecut.max())

    # -- grids --------------------------------------------------------------
    def get_uniform_grids(self, mesh=None, wrap_around=True):
        """pyscf/pbc/gto/cell.py:874-897"""
        mesh = self.mesh if mesh is None else mesh
        if wrap_around:
            qv = cartesian_prod([np.fft.fftfreq(int(x)) for x in mesh])
            return np.dot(qv, self._a)
        meshf = np.asarray(mesh, dtype=float)
        qv = cartesian_prod([np.arange(int(x)) for x in mesh])
        a_frac = np.einsum('i,ij->ij', 1. / meshf, self._a)
        return np.dot(qv, a_frac)

    def get_Gv_weights(self, mesh=None):
        """pyscf/pbc/gto/cell.py:537-596 (3-D branch)."""
        mesh = self.mesh if mesh is None else mesh
        base = tuple(np.fft.fftfreq(int(n), 1. / int(n)) for n in mesh)
        b = self.reciprocal_vectors()
        Gv = np.dot(cartesian_prod(base), b)
        weights = abs(np.linalg.det(b)) / (2 * np.pi)**3
        return Gv, base, weights

    def get_Gv(self, mesh=None):
        return self.get_Gv_weights(mesh)[0]

    def get_SI(self, Gv):
        return np.exp(-1j * np.dot(self.atom_coords(), Gv.T))

    def make_kpts(self, nks):
        """Gamma-centred uniform k-point mesh in absolute units, x slowest (pyscf/pbc/gto/cell.py
        Cell.make_kpts with its default options: scaled k = j / n, j = 0 .. n-1)."""
        ks = [np.arange(int(n), dtype=float) / int(n) for n in nks]
        return cartesian_prod(ks).dot(self.reciprocal_vectors())

    def get_scaled_atom_coords(self):
        return np.dot(self.atom_coords(), np.linalg.inv(self._a))

    # -- lattice sums ---------------------------------------------------------
    def get_lattice_Ls(self, rcut=None, discard=True):
        """Translation vectors for Gaussian lattice sums (used by ewald);
        pyscf/pbc/tools/pbc.py:528-592."""
        rcut = self.rcut if rcut is None else rcut
        if rcut <= 0 or self.natm == 0:
            return np.zeros((1, 3))
        a = self._a
        sc = self.get_scaled_atom_coords()
        bmax = np.minimum(sc.max(axis=0), 1)
        bmin = np.maximum(sc.min(axis=0), -1)
        penalty = bmax - bmin
        dR_basis = np.diag(penalty.dot(a))

        def bound(aa):
            r = np.linalg.qr(np.vstack([aa, dR_basis]).T)[1]
            return (rcut + abs(r[2, 3:]).sum()) / abs(r[2, 2])

        nb = np.ceil([bound(a[[1, 2, 0]]), bound(a[[2, 0, 1]]), bound(a)]).astype(int)
        Ts = cartesian_prod([np.arange(-n, n + 1) for n in nb])
        Ls = np.dot(Ts, a)
        if discard:
            penalty = penalty + 1e-200
            Ts_scaled = (Ts + 1e-200) / penalty
            fac = 1. / abs(Ts_scaled).min(axis=1)
            Ls = Ls[np.linalg.norm(Ls, axis=1) * (1 - fac) < rcut]
        return np.asarray(Ls, order='C')

    def ao_rcut(self):
        """Per-shell AO-value cutoff radius (pyscf/pbc/gto/eval_gto.py:169-186)."""
        precision = self.precision / max(self.vol, 1)
        out = []
        for ib in range(self.nbas):
            l = self.bas_angular(ib)
            es = self.bas_exp(ib)
            cs = abs(self._libcint_ctr_coeff(ib)).max(axis=1)
            norm_ang = ((2 * l + 1) / (4 * np.pi))**.5
            fac = 2 * np.pi / self.vol * cs * norm_ang / es / precision
            r = self.rcut
            r = (np.log(fac * r**(l + 1) + 1.) / es)**.5
            r = (np.log(fac * r**(l + 1) + 1.) / es)**.5
            out.append(r.max())
        return np.array(out)

    def ao_lattice_Ls(self, rcut):
        """Images whose atoms can reach the (wrap-around) grid box, sorted by
        |L| with a stable sort (pyscf/pbc/gto/eval_gto.py:135-136, :188-253)."""
        if rcut <= 0:
            return np.zeros((1, 3))
        a = self._a
        xyz = self.atom_coords()
        sc = np.linalg.solve(a.T, xyz.T).T
        bmax = np.minimum(sc.max(axis=0), 1)
        bmin = np.maximum(sc.min(axis=0), -1)
        bound1 = np.diag(bmax).dot(a)
        bound2 = np.diag(bmin).dot(a)

        def bound(aa):
            r = np.linalg.qr(np.vstack([aa, bound1, bound2]).T)[1]
            return (rcut + abs(r[2, 3:]).max()) / abs(r[2, 2])

        nb = np.ceil([bound(a[[1, 2, 0]]), bound(a[[2, 0, 1]]), bound(a)]).astype(int)
        Ts = cartesian_prod([np.arange(-n, n + 1) for n in nb])
        Ls = np.dot(Ts, a)
        edge = cartesian_prod([[-.5, 1.]] * 3).dot(a)
        lb, ub = edge.min(axis=0), edge.max(axis=0)
        g2a = Ls + xyz[:, None, :]
        below = g2a <= lb
        above = g2a >= ub
        dist = np.where(below, g2a - lb, np.where(above, g2a - ub, 0.))
        mask = (np.linalg.norm(dist, axis=2) < rcut).any(axis=0)
        Ls = Ls[mask]
        order = np.argsort(np.linalg.norm(Ls, axis=1), kind='stable')
        return np.ascontiguousarray(Ls[order])

    def ao_eval_params(self):
        """(rcut[nbas], Ls[nimg,3]) exactly as ``eval_gto`` derives them."""
        rcut = self.ao_rcut()
        return rcut, self.ao_lattice_Ls(rcut.max())

    # -- Ewald ------------------------------------------------------------------
    def get_ewald_params(self):
        """3-D branch of pyscf/pbc/gto/cell.py:648-690."""
        eta = 1. / self.vol**(1. / 6)
        cut = float(_rcut_overlap(eta**2, 0, 1., self.precision))
        return eta, cut

    def ewald(self, ew_eta=None, ew_cut=None):
        """Ewald energy of the point (valence) charges; 3-D branch of
        pyscf/pbc/gto/cell.py:692-813 (same optional splitting parameter and real-space cutoff)."""
        if self.natm == 0:
            return 0.
        q = self.atom_charges()
        eta, cut = self.get_ewald_params()
        eta = eta if ew_eta is None else float(ew_eta)
        cut = cut if ew_cut is None else float(ew_cut)
        logp = np.log(self.precision / (q.sum() * 16 * np.pi**2))
        mesh = cutoff_to_mesh(self._a, -2 * eta**2 * logp)
        xyz = self.atom_coords()
        Lall = self.get_lattice_Ls(rcut=cut)
        rLij = xyz[:, None, :] - xyz[None, :, :] + Lall[:, None, None, :]
        r = np.sqrt(np.einsum('Lijx,Lijx->Lij', rLij, rLij))
        r[r < 1e-16] = 1e200
        ewovrl = .5 * np.einsum('i,j,Lij->', q, q, scipy.special.erfc(eta * r) / r)
        ewself = -.5 * np.dot(q, q) * 2 * eta / np.sqrt(np.pi)
        ewself += -.5 * q.sum()**2 * np.pi / (eta**2 * self.vol)
        Gv, _, weights = self.get_Gv_weights(mesh)
        absG2 = np.einsum('gi,gi->g', Gv, Gv)
        absG2[absG2 == 0] = 1e200
        coulG = 4 * np.pi / absG2 * weights
        ZSI = np.einsum('i,ij->j', q, self.get_SI(Gv))
        ewg = .5 * np.einsum('i,i,i', ZSI.conj(), ZSI * np.exp(-absG2 / (4 * eta**2)), coulG).real
        return ewovrl + ewself + ewg

    energy_nuc = ewald


def _rcut_overlap(alpha, l, c, precision):
    """pyscf/pbc/gto/cell.py:389-406"""
    alpha = np.asarray(alpha, dtype=float)
    l = np.asarray(l)
    theta = alpha * .5
    a1 = (alpha * 2)**-.5
    norm_ang = (2 * l + 1) / (4 * np.pi)
    fac = 2 * np.pi * np.asarray(c)**2 * norm_ang / theta / precision
    fac = fac * 4 * alpha**2
    r0 = 20
    r0 = (np.log(fac * r0 * (r0 * .5 + a1)**(2 * l + 2) + 1.) / theta)**.5
    r0 = (np.log(fac * r0 * (r0 * .5 + a1)**(2 * l + 2) + 1.) / theta)**.5
    return r0


def cutoff_to_mesh(a, cutoff):
    """KE cutoff -> odd FFT mesh (pyscf/pbc/tools/pbc.py:703-726)."""
    b = 2 * np.pi * np.linalg.inv(np.asarray(a).T)
    rx = np.linalg.qr(b[[1, 2, 0]].T)[1][2, 2]
    ry = np.linalg.qr(b[[2, 0, 1]].T)[1][2, 2]
    rz = np.linalg.qr(b.T)[1][2, 2]
    Gmax = (2 * cutoff)**.5 / np.abs([rx, ry, rz])
    return np.ceil(Gmax).astype(int) * 2 + 1


def madelung(cell):
    """Gamma-point Madelung constant: -2 x Ewald energy of a unit probe
    charge in the same lattice (pyscf/pbc/tools/pbc.py:483-493)."""
    probe = Cell(atom=[('H', (0., 0., 0.))], a=cell.lattice_vectors(), unit='B',
                 basis={'H': [[0, (1.0, 1.0)]]}, pseudo=None, mesh=[3, 3, 3],
                 precision=cell.precision)
    probe._rcut = cell.rcut
    probe.build()
    probe._atm[:, CHARGE_OF] = 1
    return -2 * probe.ewald()


def super_cell(cell, ncopy, mesh=None):
    """ncopy[0] x ncopy[1] x ncopy[2] supercell, images stacked in +
    direction, odd mesh (pyscf/pbc/tools/pbc.py:595-639) unless ``mesh``
    is given (kmesh.py wants exactly ncopy * cell.mesh)."""
    a = cell.lattice_vectors()
    Ts = cartesian_prod([np.arange(n) for n in ncopy])
    shifts = np.dot(Ts, a)
    xyz = cell.atom_coords()
    atoms = []
    for L in shifts:
        for ia in range(cell.natm):
            atoms.append((cell.atom_symbol(ia), tuple(xyz[ia] + L)))
    if mesh is None:
        mesh = np.asarray(ncopy) * np.asarray(cell.mesh)
        mesh = (mesh // 2) * 2 + 1
    sup = Cell(atom=atoms, a=np.einsum('i,ij->ij', np.asarray(ncopy, float), a),
               basis=cell.basis, pseudo=cell.pseudo, unit='B',
               mesh=[int(m) for m in mesh], precision=cell.precision)
    return sup.build()


def M(**kw):
    return Cell(**kw).build()
