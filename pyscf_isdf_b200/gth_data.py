"""GTH (Goedecker-Teter-Hutter) basis-set and pseudopotential parameters.

These are published scientific constants (CP2K GTH_BASIS_SETS / GTH_POTENTIALS,
the same numbers the reference ships as data under
``pyscf/pbc/gto/basis/gth-dzvp.dat``, ``gth-szv.dat`` and
``pyscf/pbc/gto/pseudo/gth-pade.dat``).  They are stored here in this
package's own compact layout so the host model works on a box that has no
PySCF: only the elements the BASELINE.json configs need (H, He, C, O, Si).

Layout
------
``_VALENCE[elem] = (exps, {l: [col0, col1]})``: one primitive set shared by
the listed angular momenta; ``col0`` is the SZV contraction, ``col0``+``col1``
together are the two DZVP contractions (``col1`` selects the most diffuse
primitive).  ``_POLARISATION[elem] = (l, exponent)`` is the single DZVP
polarisation primitive.
"""

_VALENCE = {
    'H': ([8.3744350009, 1.8058681460, 0.4852528328, 0.1658236932],
          {0: [[-0.0283380461, -0.1333810052, -0.3995676063, -0.5531027541],
               [0.0, 0.0, 0.0, 1.0]]}),
    'He': ([13.1305278312, 4.1977275150, 1.3647725570, 0.4549715461, 0.1513197845],
           {0: [[-0.0500802904, -0.1474339352, -0.3245281495, -0.4365284986, -0.2164629317],
                [0.0, 0.0, 0.0, 0.0, 1.0]]}),
    'C': ([4.3362376436, 1.2881838513, 0.4037767149, 0.1187877657],
          {0: [[0.1490797872, -0.0292640031, -0.6882040510, -0.3964426906],
               [0.0, 0.0, 0.0, 1.0]],
           1: [[-0.0878123619, -0.2775560300, -0.4712295093, -0.4058039291],
               [0.0, 0.0, 0.0, 1.0]]}),
    'O': ([8.3043855492, 2.4579484191, 0.7597373434, 0.2136388632],
          {0: [[0.1510165999, -0.0393195364, -0.6971724029, -0.3841133622],
               [0.0, 0.0, 0.0, 1.0]],
           1: [[-0.0995679273, -0.3011422449, -0.4750857083, -0.3798777957],
               [0.0, 0.0, 0.0, 1.0]]}),
    'Si': ([1.2032403600, 0.4688385970, 0.1679853910, 0.0575616890],
           {0: [[0.3290356759, -0.2533162616, -0.7870936517, -0.1909870193],
                [0.0, 0.0, 0.0, 1.0]],
            1: [[0.0474536439, -0.2594495462, -0.5440932235, -0.3623984652],
                [0.0, 0.0, 0.0, 1.0]]}),
}

_POLARISATION = {
    'H': (1, 0.7270000000),
    'He': (1, 1.2750000000),
    'C': (2, 0.5500000000),
    'O': (2, 1.1850000000),
    'Si': (2, 0.2750000000),
}

# GTH-PADE pseudopotentials:
#   nelec per l-channel, rloc, local coefficients C1..C4,
#   nonlocal channels: list of (r_l, nproj, upper-triangular h matrix rows)
GTH_PADE = {
    'H': dict(nelec=[1], rloc=0.20000000, cloc=[-4.18023680, 0.72507482],
              nonlocal_=[]),
    'He': dict(nelec=[2], rloc=0.20000000, cloc=[-9.11202340, 1.69836797],
               nonlocal_=[]),
    'C': dict(nelec=[2, 2], rloc=0.34883045, cloc=[-8.51377110, 1.22843203],
              nonlocal_=[(0.30455321, 1, [[9.52284179]]),
                         (0.23267730, 0, [])]),
    'O': dict(nelec=[2, 4], rloc=0.24762086, cloc=[-16.58031797, 2.39570092],
              nonlocal_=[(0.22178614, 1, [[18.26691718]]),
                         (0.25682890, 0, [])]),
    'Si': dict(nelec=[2, 2], rloc=0.44000000, cloc=[-7.33610297],
               nonlocal_=[(0.42273813, 2, [[5.90692831, -1.26189397],
                                           [3.25819622]]),
                          (0.48427842, 1, [[2.72701346]])]),
}

ELEMENT_Z = {'H': 1, 'He': 2, 'Li': 3, 'Be': 4, 'B': 5, 'C': 6, 'N': 7,
             'O': 8, 'F': 9, 'Ne': 10, 'Na': 11, 'Mg': 12, 'Al': 13,
             'Si': 14, 'P': 15, 'S': 16, 'Cl': 17, 'Ar': 18}


def gth_basis(name, elem):
    """Return the basis of ``elem`` as a list of shells
    ``[l, (exp, c0, c1, ...), ...]`` ordered by angular momentum
    (same in-memory form the reference's basis loader produces,
    pyscf/gto/basis/parse_cp2k.py:70-118)."""
    key = name.lower().replace('_', '-')
    if key not in ('gth-szv', 'gth-dzvp'):
        raise KeyError('basis %s not bundled (have gth-szv, gth-dzvp)' % name)
    if elem not in _VALENCE:
        raise KeyError('element %s not bundled in %s' % (elem, name))
    exps, cols = _VALENCE[elem]
    ncontr = 1 if key == 'gth-szv' else 2
    shells = []
    for l in sorted(cols):
        shell = [l]
        for ip, e in enumerate(exps):
            shell.append(tuple([e] + [cols[l][ic][ip] for ic in range(ncontr)]))
        shells.append(shell)
    if key == 'gth-dzvp':
        l, e = _POLARISATION[elem]
        shells.append([l, (e, 1.0)])
    shells.sort(key=lambda s: s[0])
    return shells


def gth_pseudo(name, elem):
    key = name.lower().replace('_', '-')
    if key not in ('gth-pade', 'gth-lda'):
        raise KeyError('pseudo %s not bundled (have gth-pade)' % name)
    return GTH_PADE[elem]
