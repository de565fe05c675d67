"""k-point sampling on a Gamma-centred uniform mesh through the supercell equivalence.

The reference reaches its DF plugin at k-points through ``KSCF.get_jk -> with_df.get_jk(dm_kpts, kpts=...)``
(pyscf/pbc/scf/khf.py:519-532; FFTDF: pyscf/pbc/df/fft_jk.py:33-109, :177-302) and ships the transformation used
here as a tool (pyscf/pbc/tools/k2gamma.py): a calculation on the n1 x n2 x n3 mesh ``cell.make_kpts(kmesh)`` is
the Gamma-point calculation of the n1 x n2 x n3 supercell, with

    phi_k,mu(r) = sum_R exp(i k.R) chi_(R mu)(r)                      (chi: supercell AOs, R: the cells of the supercell)
    D_sc[(R mu), (R' nu)] = 1/Nk sum_k exp(i k.(R - R')) D_k[mu, nu]
    V_k[mu, nu] = 1/Nk sum_{R, R'} exp(i k.(R' - R)) V_sc[(R mu), (R' nu)]        (V = J or K of D_sc at Gamma)

when the supercell's FFT mesh is kmesh * cell.mesh (the same real-space points, the same plane waves; exact for
odd kmesh * mesh, Nyquist-plane differences otherwise).  Every flop runs in the Gamma-point ISDF of the supercell
(pyscf_isdf_b200/isdf.py) -- this module is host-side bookkeeping: mesh detection, the two Bloch transformations,
and the split of a complex supercell density matrix into its real symmetric and real antisymmetric parts (J and K
are linear, the device path is real).  It does NOT exploit translational symmetry: cost and memory are those of the
supercell.  The one-electron pieces (ISDF.get_nuc / get_pp / nr_rks) fold the same way (isdf.py: _folded).
Arbitrary k-point lists, kpts_band and twisted meshes raise NotImplementedError.
"""
import numpy as np

from .cell import super_cell


def make_kpts(cell, nks):
    """Gamma-centred Monkhorst-Pack mesh, absolute units, x slowest (pyscf/pbc/gto/cell.py make_kpts, default
    options: scaled k = j / n, j = 0 .. n-1)."""
    return cell.make_kpts(nks)


def detect_kmesh(cell, kpts, tol=1e-7):
    """(kmesh, index[nk, 3]) if ``kpts`` is a complete Gamma-centred uniform mesh (in any order), else None."""
    kpts = np.reshape(kpts, (-1, 3))
    scaled = kpts.dot(cell.lattice_vectors().T) / (2 * np.pi)      # k = s . b, a_i . b_j = 2 pi delta_ij
    scaled = scaled - np.floor(scaled + tol)
    kmesh = []
    for d in range(3):
        vals = np.unique(np.round(scaled[:, d], 7))
        n = len(vals)
        if abs(vals - np.arange(n) / n).max() > 1e-6:
            return None
        kmesh.append(n)
    if int(np.prod(kmesh)) != len(kpts):
        return None
    idx = np.rint(scaled * np.asarray(kmesh)).astype(int) % np.asarray(kmesh)
    if len({tuple(i) for i in idx}) != len(kpts):
        return None
    return tuple(kmesh), idx


class KMesh:
    """Bloch bookkeeping between the k-points of a uniform mesh and the supercell."""

    def __init__(self, cell, kpts):
        found = detect_kmesh(cell, kpts)
        if found is None:
            raise NotImplementedError('ISDF at k-points: only a complete Gamma-centred uniform mesh '
                                      '(cell.make_kpts(kmesh)) is supported, through its supercell')
        self.kmesh, self.kidx = found
        self.kpts = np.reshape(kpts, (-1, 3))
        self.nk = len(self.kpts)
        self.nao = cell.nao_nr()
        # cells of the supercell in the order super_cell() stacks them (x slowest), as lattice translations
        n1, n2, n3 = self.kmesh
        self.Ts = np.stack(np.meshgrid(np.arange(n1), np.arange(n2), np.arange(n3), indexing='ij'), axis=-1).reshape(-1, 3)
        self.R = self.Ts.dot(cell.lattice_vectors())
        self.phase = np.exp(1j * self.kpts.dot(self.R.T))          # [nk, nR]: exp(i k.R)
        self.supercell = super_cell(cell, self.kmesh, mesh=np.asarray(self.kmesh) * np.asarray(cell.mesh))

    def dm_to_supercell(self, dm_kpts):
        """[nk, nao, nao] -> complex [nR nao, nR nao]."""
        nk, nao = self.nk, self.nao
        dm = np.asarray(dm_kpts).reshape(nk, nao, nao)
        # D_sc[R, mu, R', nu] = 1/nk sum_k e^{ik.R} D_k e^{-ik.R'}
        out = np.einsum('kr,kmn,ks->rmsn', self.phase, dm, self.phase.conj(), optimize=True) / nk
        return out.reshape(nk * nao, nk * nao)

    def mat_to_kpts(self, m_sc):
        """[nR nao, nR nao] (real or complex) -> complex [nk, nao, nao]."""
        nk, nao = self.nk, self.nao
        m = np.asarray(m_sc).reshape(nk, nao, nk, nao)
        return np.einsum('kr,rmsn,ks->kmn', self.phase.conj(), m, self.phase, optimize=True) / nk


def get_jk_kmesh(mydf, dm_kpts, hermi=1, with_j=True, with_k=True, exxdiv=None):
    """J and K at the k-points of ``mydf._kmesh`` from the Gamma-point ISDF of the supercell (``mydf._sc``)."""
    km = mydf._kmesh
    nk, nao = km.nk, km.nao
    dm0 = np.asarray(dm_kpts)
    dms = dm0.reshape(-1, nk, nao, nao)
    nset = dms.shape[0]
    stack = []
    for i in range(nset):
        d = km.dm_to_supercell(dms[i])
        stack += [np.ascontiguousarray(d.real), np.ascontiguousarray(d.imag)]
    vj, vk = mydf._sc.get_jk(np.stack(stack), hermi=0, with_j=with_j, with_k=with_k, exxdiv=exxdiv)
    out = []
    for v, want in ((vj, with_j), (vk, with_k)):
        if not want:
            out.append(None)
            continue
        res = np.stack([km.mat_to_kpts(v[2 * i] + 1j * v[2 * i + 1]) for i in range(nset)])
        out.append(res.reshape(dm0.shape))
    return out[0], out[1]
