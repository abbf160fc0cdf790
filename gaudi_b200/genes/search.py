"""Search gene (gaudi/genes/search.py): allele = (translation 3x4, rotation 3x4); expression
sets ``xform = T . R . to_zero(origin)`` (search.py:201-214).  The GPU receives the product
``T . R`` (``Search.device_allele``); variation by xform interpolation stays on the host."""
from __future__ import annotations

import math
import random

import numpy as np

from .. import molio
from . import GeneProvider

IDENTITY = ((1.0, 0.0, 0.0, 0.0), (0.0, 1.0, 0.0, 0.0), (0.0, 0.0, 1.0, 0.0))


def enable(**kwargs):
    kwargs = Search.validate(kwargs)
    return Search(**kwargs)


def multiply_matrices(*mats):
    """Affine product of 3x4 matrices, right-most applied first (Chimera ``Matrix.multiply_matrices``)."""
    out = np.eye(4)
    for m in mats:
        a = np.eye(4)
        a[:3, :] = np.asarray(m, dtype=np.float64).reshape(3, 4)
        out = out @ a
    return out[:3, :]


def _quat_from_matrix(r):
    t = np.trace(r)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (r[2, 1] - r[1, 2]) / s, (r[0, 2] - r[2, 0]) / s, (r[1, 0] - r[0, 1]) / s])
    else:
        i = int(np.argmax(np.diag(r)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = math.sqrt(1.0 + r[i, i] - r[j, j] - r[k, k]) * 2
        q = np.zeros(4)
        q[0] = (r[k, j] - r[j, k]) / s
        q[1 + i] = 0.25 * s
        q[1 + j] = (r[j, i] + r[i, j]) / s
        q[1 + k] = (r[k, i] + r[i, k]) / s
    return q / np.linalg.norm(q)


def _matrix_from_quat(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def interpolate_xforms(xf1, centre, xf2, frac):
    """Restatement of Chimera ``Matrix.interpolate_xforms`` as used at search.py:227-246:
    rotate a fraction of the way from xf1 to xf2 about ``centre`` (slerp) while the image of
    ``centre`` moves linearly.  Host-side GA variation only — parity unpinned (no reference
    test covers it), off the kernel path."""
    a, b = np.asarray(xf1, float).reshape(3, 4), np.asarray(xf2, float).reshape(3, 4)
    c = np.asarray(centre, float)
    q1, q2 = _quat_from_matrix(a[:, :3]), _quat_from_matrix(b[:, :3])
    d = float(np.dot(q1, q2))
    if d < 0:
        q2, d = -q2, -d
    if d > 0.9995:
        q = q1 + frac * (q2 - q1)
    else:
        th = math.acos(d)
        q = (math.sin((1 - frac) * th) * q1 + math.sin(frac * th) * q2) / math.sin(th)
    r = _matrix_from_quat(q / np.linalg.norm(q))
    c1, c2 = a[:, :3] @ c + a[:, 3], b[:, :3] @ c + b[:, 3]
    cm = c1 + frac * (c2 - c1)
    return np.c_[r, cm - r @ c]


def random_rotation():
    """Uniform random rotation as a 3x4 with zero translation (``FitMap.search.random_rotation``)."""
    q = np.array([random.gauss(0, 1) for _ in range(4)])
    r = _matrix_from_quat(q / np.linalg.norm(q))
    return tuple(tuple(row) + (0.0,) for row in r.tolist())


def random_translation(center, r):
    """search.py:311-335."""
    r2 = r * r
    a, b, c = center
    while True:
        x, y, z = [random.uniform(-r, r) for _ in center]
        if x * x + y * y + z * z <= r2:
            break
    return ((1.0, 0.0, 0.0, a + x), (0.0, 1.0, 0.0, b + y), (0.0, 0.0, 1.0, c + z))


class Search(GeneProvider):
    _validate = {"target": "required", "center": None, "radius": None, "rotate": None, "precision": None,
                 "interpolation": None}

    def __init__(self, target=None, center=None, radius=None, rotate=True, precision=0, interpolation=0.5,
                 **kwargs):
        GeneProvider.__init__(self, **kwargs)
        self.radius = float(radius)
        self.rotate = bool(rotate)
        self.precision = int(precision)
        self._center = center
        self.target = target
        self.interpolation = float(interpolation)

    def __ready__(self):
        if isinstance(self.target, str):
            if "/" in self.target:
                mol, serial = self.target.split("/")
                self.target = (mol, int(serial))
            else:                       # search.py:176-180: nearest atom to the centre of mass
                gene = self.parent.find_molecule(self.target)
                self.target = (self.target, int(gene.mol.serial[molio.nearest_atom_to_centre(gene.mol)]))
        self.allele = self.random_transform()

    @property
    def molecule_gene(self):
        return self.parent.find_molecule(self.target[0])

    @property
    def origin_atom(self):
        return self.molecule_gene.find_atom(self.target[1])

    @property
    def origin(self):
        return tuple(float(v) for v in self.molecule_gene.mol.xyz[self.origin_atom])

    @property
    def center(self):
        if self._center is None:
            return self.origin
        if isinstance(self._center, (list, tuple)) and len(self._center) == 3:
            return tuple(float(v) for v in self._center)
        mol, serial = self._center if not isinstance(self._center, str) else self._center.split("/")
        gene = self.parent.find_molecule(mol)
        return tuple(float(v) for v in gene.mol.xyz[gene.find_atom(int(serial))])

    def device_allele(self):
        """(12,) row-major 3x4 product ``T . R`` handed to ``gmm_express`` (search.py:206-214;
        rounding to ``precision`` decimals first when precision > 0, search.py:207-210)."""
        t, r = self.allele
        if self.precision > 0:
            t, r = np.around(np.asarray(t), self.precision), np.around(np.asarray(r), self.precision)
        return multiply_matrices(t, r).reshape(12)

    def mate(self, mate):
        xf1 = multiply_matrices(*self.allele)
        xf2 = multiply_matrices(*mate.allele)
        interp = interpolate_xforms(xf1, (0, 0, 0), xf2, 0.5)
        rot = tuple(tuple(row[:3]) + (0.0,) for row in interp.tolist())
        tl = tuple(tuple(i[:3]) + (row[3],) for row, i in zip(interp.tolist(), IDENTITY))
        self.allele, mate.allele = (self.allele[0], rot), (tl, mate.allele[1])

    def mutate(self, indpb):
        if random.random() < self.indpb:
            xf1 = multiply_matrices(*self.allele)
            xf2 = multiply_matrices(*self.random_transform())
            interp = interpolate_xforms(xf1, (0, 0, 0), xf2, self.interpolation)
            rot = tuple(tuple(row[:3]) + (0.0,) for row in interp.tolist())
            tl = tuple(tuple(i[:3]) + (row[3],) for row, i in zip(interp.tolist(), IDENTITY))
            self.allele = tl, rot

    def random_transform(self):
        rotation = random_rotation() if self.rotate else IDENTITY
        translation = random_translation(self.center, self.radius)
        return translation, rotation
