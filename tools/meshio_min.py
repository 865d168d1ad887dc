"""Minimal mesh readers (binary/ascii STL, COLLADA float arrays, OBJ) used by the offline model compiler.

Only vertex positions are needed: Bullet turns every URDF <mesh> collision geometry into a convex hull of the
vertices (recalled: BulletUrdfImporter convertURDFToCollisionShape), so faces are irrelevant here.
"""
import re
import struct
import numpy as np


def load_stl(path):
    data = open(path, 'rb').read()
    ntri = struct.unpack('<I', data[80:84])[0]
    if len(data) == 84 + 50 * ntri:
        arr = np.frombuffer(data, dtype=np.uint8, count=50 * ntri, offset=84).reshape(ntri, 50)
        v = arr[:, 12:48].copy().view('<f4').reshape(ntri * 3, 3)
        return v.astype(np.float64)
    txt = data.decode('ascii', 'ignore')
    v = re.findall(r'vertex\s+(\S+)\s+(\S+)\s+(\S+)', txt)
    return np.array(v, dtype=np.float64)


def load_dae(path):
    """COLLADA: positions arrays only. The four gripper files have one geometry, one untransformed node, no <unit>
    element (so Bullet's unitMeterScaling stays 1) and Y_UP, which Bullet's collision path ignores (recalled)."""
    txt = open(path, 'r').read()
    out = []
    for m in re.finditer(r'<float_array[^>]*positions-array[^>]*>([^<]*)</float_array>', txt):
        out.append(np.array(m.group(1).split(), dtype=np.float64).reshape(-1, 3))
    assert out, path
    return np.concatenate(out, 0)


def load_obj(path):
    v = []
    for line in open(path):
        if line.startswith('v '):
            v.append([float(x) for x in line.split()[1:4]])
    return np.array(v, dtype=np.float64)


def load_mesh(path):
    p = path.lower()
    if p.endswith('.stl'):
        return load_stl(path)
    if p.endswith('.dae'):
        return load_dae(path)
    if p.endswith('.obj'):
        return load_obj(path)
    raise ValueError(path)
