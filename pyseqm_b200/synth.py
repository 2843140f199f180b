"""
Deterministic synthetic geometries for the BASELINE.json configs.

numpy only (RandomState streams are stable across platforms), so the same inputs are produced
in the build container (where the reference generates golden vectors) and on the GPU box.
Every generator returns (species[int64 B,S], coordinates[float64 B,S,3]) with atoms sorted by
descending atomic number and zero padding at the end, which is the ordering the whole path
relies on (reference doc/documentation.md:152-163).
"""
import numpy as np

__all__ = ['benzene', 'c60', 'water_cluster', 'mixed_organics', 'formaldehyde_pair', 'small_molecules']


def _finish(species, coords):
    return np.ascontiguousarray(species, dtype=np.int64), np.ascontiguousarray(coords, dtype=np.float64)


def benzene(nmol, seed=0, sigma=0.05):
    """Ideal D6h benzene (r_CC ring 1.397 Å, H ring 2.481 Å) + N(0, sigma) per coordinate."""
    ang = np.arange(6) * (np.pi / 3.0)
    ring = np.stack([np.cos(ang), np.sin(ang), np.zeros(6)], axis=1)
    base = np.concatenate([1.397 * ring, 2.481 * ring], axis=0)
    rng = np.random.RandomState(seed)
    coords = base[None, :, :] + sigma * rng.standard_normal((nmol, 12, 3))
    species = np.tile(np.array([6] * 6 + [1] * 6), (nmol, 1))
    return _finish(species, coords)


def _c60_ideal():
    phi = (1.0 + np.sqrt(5.0)) / 2.0
    gens = [(0.0, 1.0, 3.0 * phi), (1.0, 2.0 + phi, 2.0 * phi), (phi, 2.0, 2.0 * phi + 1.0)]
    pts = set()
    for g in gens:
        for shift in range(3):                       # even permutations = cyclic shifts
            v = (g[(0 - shift) % 3], g[(1 - shift) % 3], g[(2 - shift) % 3])
            for sx in (1, -1):
                for sy in (1, -1):
                    for sz in (1, -1):
                        pts.add((round(sx * v[0], 10) + 0.0, round(sy * v[1], 10) + 0.0, round(sz * v[2], 10) + 0.0))
    xyz = np.array(sorted(pts))
    assert xyz.shape == (60, 3), xyz.shape
    # edge length of this construction is 2; scale to a mean bond of 1.42 Å
    return xyz * (1.42 / 2.0)


def c60(nmol, seed=0, sigma=0.02):
    """Truncated-icosahedron C60 (bond ≈1.42 Å) + N(0, sigma)."""
    base = _c60_ideal()
    rng = np.random.RandomState(seed)
    coords = base[None] + sigma * rng.standard_normal((nmol, 60, 3))
    species = np.full((nmol, 60), 6)
    return _finish(species, coords)


def _rand_rot(rng):
    q = rng.standard_normal(4)
    q /= np.linalg.norm(q)
    a, b, c, d = q
    return np.array([[a * a + b * b - c * c - d * d, 2 * (b * c - a * d), 2 * (b * d + a * c)],
                     [2 * (b * c + a * d), a * a - b * b + c * c - d * d, 2 * (c * d - a * b)],
                     [2 * (b * d - a * c), 2 * (c * d + a * b), a * a - b * b - c * c + d * d]])


def water_cluster(nwater, seed=0, spacing=3.1, jitter=0.2):
    """nwater waters on a jittered cubic lattice, random orientation; all O first, then all H.
    Returns a single 'molecule' (B=1, S=3*nwater)."""
    rng = np.random.RandomState(seed)
    side = int(np.ceil(nwater ** (1.0 / 3.0)))
    th = np.deg2rad(104.5) / 2.0
    h_local = 0.96 * np.array([[np.sin(th), 0.0, np.cos(th)], [-np.sin(th), 0.0, np.cos(th)]])
    O, H = [], []
    k = 0
    for ix in range(side):
        for iy in range(side):
            for iz in range(side):
                if k >= nwater:
                    break
                c = spacing * np.array([ix, iy, iz], dtype=np.float64) + jitter * rng.standard_normal(3)
                R = _rand_rot(rng)
                O.append(c)
                H.append(c + h_local[0] @ R.T)
                H.append(c + h_local[1] @ R.T)
                k += 1
    coords = np.concatenate([np.array(O), np.array(H)], axis=0)[None]
    species = np.array([[8] * nwater + [1] * (2 * nwater)])
    return _finish(species, coords)


_VALENCE = {6: 4, 7: 3, 8: 2}
_BOND = {(6, 6): 1.53, (6, 7): 1.47, (6, 8): 1.43, (7, 7): 1.45, (7, 8): 1.40, (8, 8): 1.47,
         (1, 6): 1.09, (1, 7): 1.01, (1, 8): 0.96}
_TETRA = np.array([[1.0, 1.0, 1.0], [1.0, -1.0, -1.0], [-1.0, 1.0, -1.0], [-1.0, -1.0, 1.0]]) / np.sqrt(3.0)


def _rot_to(v, target):
    """Rotation matrix taking unit vector v to unit vector target (Rodrigues)."""
    c = float(np.dot(v, target))
    if c > 1.0 - 1e-12:
        return np.eye(3)
    if c < -1.0 + 1e-12:
        a = np.array([1.0, 0.0, 0.0]) if abs(v[0]) < 0.9 else np.array([0.0, 1.0, 0.0])
        a = a - np.dot(a, v) * v
        a /= np.linalg.norm(a)
        return 2.0 * np.outer(a, a) - np.eye(3)
    k = np.cross(v, target)
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + K + K @ K / (1.0 + c)


def _grow_molecule(rng, natoms, attempts=50):
    """A saturated acyclic C/N/O skeleton with tetrahedral local geometry, every free valence
    capped by H: closed shell by construction and well-behaved in SCF.  Returns (Z, xyz) sorted by
    descending Z, with about `natoms` atoms."""
    nheavy = max(1, int(round((natoms - 2) / 3.0)))
    for _attempt in range(attempts):
        elems = rng.choice([6, 7, 8], size=nheavy, p=[0.75, 0.125, 0.125])
        elems[0] = 6
        pos = [np.zeros(3)]
        frames = [_rand_rot(rng)]
        free = [list(range(_VALENCE[int(elems[0])]))]       # unused tetrahedral slots per heavy atom
        ok = True
        for c in range(1, nheavy):
            placed = False
            for _try in range(40):
                cand = [p for p in range(len(pos)) if free[p]]
                if not cand:
                    break
                p = int(cand[rng.randint(len(cand))])
                s = int(free[p][rng.randint(len(free[p]))])
                d = frames[p] @ _TETRA[s]
                bl = _BOND[tuple(sorted((int(elems[p]), int(elems[c]))))]
                x = pos[p] + bl * d
                dist = np.linalg.norm(np.array(pos) - x, axis=1)
                dist[p] = 1e9
                if dist.min() < 2.2:
                    continue
                R = _rand_rot(rng)
                fr = _rot_to(R @ _TETRA[0], -d) @ R
                pos.append(x)
                frames.append(fr)
                free[p].remove(s)
                free.append(list(range(1, _VALENCE[int(elems[c])])))
                placed = True
                break
            if not placed:
                ok = False
                break
        if not ok:
            continue
        hpos = []
        for p in range(nheavy):
            for s in free[p]:
                hpos.append(pos[p] + _BOND[(1, int(elems[p]))] * (frames[p] @ _TETRA[s]))
        allpos = np.array(pos + hpos)
        Z = np.concatenate([elems, np.ones(len(hpos), dtype=np.int64)]).astype(np.int64)
        dm = np.linalg.norm(allpos[:, None, :] - allpos[None, :, :], axis=2) + 1e9 * np.eye(len(Z))
        if len(hpos) and dm[nheavy:, :].min() < 0.9:
            continue
        if len(hpos) > 1 and dm[nheavy:, nheavy:].min() < 1.45:
            continue
        order = np.argsort(-Z, kind='stable')
        return Z[order], allpos[order]
    raise RuntimeError("could not grow a molecule")


def mixed_organics(nmol, seed=0, min_atoms=5, max_atoms=30, attempts=50):
    """Random saturated C/N/O/H molecules of about min_atoms..max_atoms atoms, padded to max_atoms."""
    rng = np.random.RandomState(seed)
    S = max_atoms
    species = np.zeros((nmol, S), dtype=np.int64)
    coords = np.zeros((nmol, S, 3), dtype=np.float64)
    for b in range(nmol):
        n = int(rng.randint(min_atoms, max_atoms + 1))
        Z, xyz = _grow_molecule(rng, n, attempts)
        while len(Z) > S:
            Z, xyz = _grow_molecule(rng, n - 1, attempts)
        species[b, :len(Z)] = Z
        coords[b, :len(Z)] = xyz
    return _finish(species, coords)


def mixed_organics_sweep(nmol, seed=0, base=2048, sigma=0.02, max_atoms=30):
    """BASELINE.json configs[3] at full size without growing 65,536 molecules one by one: a library of `base`
    grown molecules (mixed_organics), each used nmol / base times with independent N(0, sigma) displacements of
    every real atom.  Molecule k is library entry k % base, so any contiguous shard holds the whole size mix."""
    sp0, xyz0 = mixed_organics(min(base, nmol), seed=seed, max_atoms=max_atoms)
    reps = (nmol + sp0.shape[0] - 1) // sp0.shape[0]
    species = np.tile(sp0, (reps, 1))[:nmol]
    coords = np.tile(xyz0, (reps, 1, 1))[:nmol].copy()
    rng = np.random.RandomState(seed + 1)
    coords += rng.normal(0.0, sigma, size=coords.shape) * (species > 0)[:, :, None]
    return species, coords


def formaldehyde_pair():
    """The reference's tests/test1.py molecules 0 and 2: H2CO [8,6,1,1] and the padded [8,8,6,0]."""
    species = np.array([[8, 6, 1, 1], [8, 8, 6, 0]])
    coords = np.array([[[0.0000, 0.0, 0.0], [1.22732374, 0.0, 0.0], [1.8194841064614802, 0.93941263319067747, 0.0],
                        [1.8193342232738994, -0.93951967178254525, 3.0565334533430606e-006]],
                       [[0.0000, 0.0, 0.0], [1.22732374, 0.0, 0.0], [1.8194841064614802, 0.93941263319067747, 0.0],
                        [0.0, 0.0, 0.0]]])
    return _finish(species, coords)


def small_molecules():
    """Config-1 class batch: H2, N2, CH4, H2O, NH3, HCN, CH3OH-like, padded to 6 atoms."""
    t = 1.09 / np.sqrt(3.0)
    th = np.deg2rad(104.5) / 2.0
    mols = [
        ([1, 1], [[0, 0, 0], [0.74, 0, 0]]),
        ([7, 7], [[0, 0, 0], [1.10, 0, 0]]),
        ([6, 1, 1, 1, 1], [[0, 0, 0], [t, t, t], [t, -t, -t], [-t, t, -t], [-t, -t, t]]),
        ([8, 1, 1], [[0, 0, 0], [0.96 * np.sin(th), 0, 0.96 * np.cos(th)], [-0.96 * np.sin(th), 0, 0.96 * np.cos(th)]]),
        ([7, 1, 1, 1], [[0, 0, 0.1], [0.94, 0, -0.27], [-0.47, 0.814, -0.27], [-0.47, -0.814, -0.27]]),
        ([7, 6, 1], [[0, 0, 1.156], [0, 0, 0], [0, 0, -1.064]]),
        ([8, 6, 1, 1, 1, 1], [[0.0, 0.0, 1.43], [0, 0, 0], [1.03, 0, -0.36], [-0.51, 0.89, -0.36],
                              [-0.51, -0.89, -0.36], [0.90, 0.0, 1.75]]),
    ]
    S = 6
    species = np.zeros((len(mols), S), dtype=np.int64)
    coords = np.zeros((len(mols), S, 3), dtype=np.float64)
    for b, (Z, xyz) in enumerate(mols):
        species[b, :len(Z)] = Z
        coords[b, :len(Z)] = np.array(xyz, dtype=np.float64)
    return _finish(species, coords)
