#!/usr/bin/env python
"""Offline model compiler: URDF + collision meshes -> constant link table (C header + JSON).

Run in the build container (needs /root/reference, which does not exist on the GPU box):

    python tools/compile_model.py [--ref /root/reference]

Writes two identical generated headers, `oracle/model_data.h` (for the CPU restatement) and
`urgym_b200/csrc/model_data.h` (for the CUDA library), plus `urgym_b200/model/ur5_robotiq_85.json` for inspection.
Only numbers derived from the assets are emitted; no reference source text is copied.

What it follows (reference file:line):
  * link / joint table and Bullet's depth-first link numbering: urgym/urdf/ur5_robotiq_85.urdf:1-565, cross-checked
    with the env code's magic indices (urgym/base/robot.py:180 eef_id=7, env_cubes_push_v0.py:187 fingers 9..18).
  * mesh -> convex hull -> primitive: the north star sanctions "collision reduced to primitive"; every robot link
    becomes ONE oriented box (minimum-volume fit over single-axis rotations of the hull's vertex cloud).
  * inertia: pybullet without URDF_USE_INERTIA_FROM_FILE (robot.py:186-187 passes no such flag) discards the URDF
    inertia tensor and uses the box inertia of the collision shape's local AABB (recalled Bullet behaviour:
    btCompoundShape / btPolyhedralConvexShape ::calculateLocalInertia), in the URDF inertial frame.
  * a link without <inertial> (ee_link, urdf:200-212) gets mass 1 (recalled: UrdfParser warning "No inertial data for
    link, using mass=1").
"""
import argparse
import json
import math
import os
import sys
import xml.etree.ElementTree as ET

import numpy as np

sys.path.insert(0, os.path.dirname(__file__))
from meshio_min import load_mesh  # noqa: E402

HULL_MARGIN = 0.001  # gUrdfDefaultCollisionMargin (recalled)


def rpy_to_mat(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


def parse_urdf(path):
    root = ET.parse(path).getroot()
    links, joints = {}, []
    for l in root.findall('link'):
        links[l.get('name')] = l
    for j in root.findall('joint'):
        joints.append(j)
    children = {}
    for j in joints:  # file order, as Bullet's UrdfParser::initTreeAndRoot keeps it
        children.setdefault(j.find('parent').get('link'), []).append(j)
    childset = {j.find('child').get('link') for j in joints}
    roots = [n for n in links if n not in childset]
    assert roots == ['world'], roots
    order = []  # (link name, joint element, parent bullet index)

    def dfs(name, parent_idx):
        for j in children.get(name, []):
            c = j.find('child').get('link')
            idx = len(order)
            order.append((c, j, parent_idx))
            dfs(c, idx)

    dfs('world', -1)
    return links, order


def fvec(s, n=3):
    v = [float(x) for x in s.split()]
    assert len(v) == n
    return v


def min_volume_box(pts):
    """Oriented box fit: search rotations about each coordinate axis; return (center, R, half) in the input frame."""
    best = None
    for ax in range(3):
        for deg in np.arange(0.0, 90.0, 0.5):
            a = math.radians(deg)
            c, s = math.cos(a), math.sin(a)
            if ax == 0:
                R = np.array([[1, 0, 0], [0, c, -s], [0, s, c]])
            elif ax == 1:
                R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
            else:
                R = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]])
            q = pts @ R  # coordinates in the rotated box frame (columns of R are the box axes)
            lo, hi = q.min(0), q.max(0)
            vol = float(np.prod(hi - lo))
            if best is None or vol < best[0] * (1 - 1e-9):
                best = (vol, R, (lo + hi) / 2, (hi - lo) / 2)
    vol, R, cb, half = best
    return R @ cb, R, half


POLY_LINKS = (8, 9, 10, 11, 13, 14, 15, 16, 18)   # gripper links with mesh collision geometry (pads 12 / 17 are boxes)
POLY_MAXV, POLY_MAXF, POLY_MAXL, POLY_MAXE = 16, 32, 8, 12


def reduce_hull(pts, kmax=POLY_MAXV):
    """Inner approximation of the convex hull of `pts` by at most kmax of its own vertices: start from the six axis
    extremes, then keep adding the hull vertex farthest outside the current polytope. Being a subset of the hull it never
    protrudes beyond the true geometry (a pad must stay proud of the finger that carries it). Returns (vertices, max
    distance of a dropped hull vertex to the polytope)."""
    from scipy.spatial import ConvexHull
    H0 = ConvexHull(pts)
    V = pts[H0.vertices]
    sel = []
    for ax in range(3):
        for sg in (1, -1):
            i = int(np.argmax(sg * V[:, ax]))
            if i not in sel:
                sel.append(i)
    while True:
        H = ConvexHull(V[sel])
        d = (H.equations[:, :3] @ V.T + H.equations[:, 3:4]).max(0)
        d[sel] = -1
        i = int(np.argmax(d))
        if len(sel) >= kmax or d[i] < 1e-5:
            return V[sel], float(max(d[i], 0.0))
        sel.append(i)


def polytope_features(P, ang_tol_deg=2.0, off_tol=2e-5):
    """Faces (outward unit normal, offset, CCW vertex loop; coplanar hull triangles merged), and the unique edge directions."""
    from scipy.spatial import ConvexHull
    H = ConvexHull(P)
    n, d = H.equations[:, :3], -H.equations[:, 3]
    used = np.zeros(len(n), bool)
    ct = math.cos(math.radians(ang_tol_deg))
    faces = []
    for i in range(len(n)):
        if used[i]:
            continue
        grp = [j for j in range(len(n)) if not used[j] and n[i] @ n[j] > ct and abs(d[i] - d[j]) < off_tol + 1e-3 * (1 - n[i] @ n[j])]
        for j in grp:
            used[j] = True
        vs = sorted(set(int(v) for j in grp for v in H.simplices[j]))
        nn = n[grp].mean(0)
        nn /= np.linalg.norm(nn)
        dd = float(max(P[vs] @ nn))
        c = P[vs].mean(0)
        u = np.cross(nn, [1, 0, 0]) if abs(nn[0]) < 0.9 else np.cross(nn, [0, 1, 0])
        u /= np.linalg.norm(u)
        w = np.cross(nn, u)
        q = np.stack([(P[vs] - c) @ u, (P[vs] - c) @ w], 1)
        if len(vs) > 3:
            vs = [vs[k] for k in ConvexHull(q).vertices]        # CCW seen from outside
        else:
            ang = np.arctan2(q[:, 1], q[:, 0])
            vs = [vs[k] for k in np.argsort(ang)]
        faces.append((nn, dd, vs))
    # Edge directions for the edge-edge separating axes: the POLY_MAXE longest, merging directions within 8 degrees. The
    # vertex subset keeps many short edges around what were fillets of the mesh; their cross products with the partner's
    # edges are near-duplicates of these. (A coarser candidate set can only miss a separating axis, i.e. report a shallow
    # contact where the bodies are a hair apart.)
    edges = {}
    for nn, dd, vs in faces:
        for a in range(len(vs)):
            i, k = vs[a], vs[(a + 1) % len(vs)]
            edges[(min(i, k), max(i, k))] = P[k] - P[i]
    dirs = []
    for e in sorted(edges.values(), key=lambda v: -np.linalg.norm(v)):
        u = e / np.linalg.norm(e)
        if len(dirs) < POLY_MAXE and not any(abs(u @ x) > math.cos(math.radians(8.0)) for x in dirs):
            dirs.append(u)
    return faces, dirs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--ref', default='/root/reference')
    args = ap.parse_args()
    udir = os.path.join(args.ref, 'urgym', 'urdf')
    links_xml, order = parse_urdf(os.path.join(udir, 'ur5_robotiq_85.urdf'))
    table = []
    mesh_pts = {}
    for idx, (name, j, parent) in enumerate(order):
        l = links_xml[name]
        o = j.find('origin')
        jxyz = fvec(o.get('xyz')) if o is not None else [0, 0, 0]
        jrpy = fvec(o.get('rpy')) if o is not None else [0, 0, 0]
        jtype = j.get('type')
        axis = fvec(j.find('axis').get('xyz')) if j.find('axis') is not None else [1, 0, 0]
        lim = j.find('limit')
        lower = float(lim.get('lower')) if lim is not None else 0.0
        upper = float(lim.get('upper')) if lim is not None else 0.0
        effort = float(lim.get('effort')) if lim is not None else 0.0
        vel = float(lim.get('velocity')) if lim is not None else 0.0
        inert = l.find('inertial')
        if inert is not None:
            mass = float(inert.find('mass').get('value'))
            io = inert.find('origin')
            com = fvec(io.get('xyz')) if io is not None else [0, 0, 0]
            irpy = fvec(io.get('rpy')) if io is not None else [0, 0, 0]
            assert all(abs(x) < 1e-12 for x in irpy), 'inertial frames are axis-aligned in this URDF'
        else:
            mass, com = 1.0, [0.0, 0.0, 0.0]
        col = l.find('collision')
        geom = col.find('geometry')
        co = col.find('origin')
        cxyz = fvec(co.get('xyz')) if co is not None else [0, 0, 0]
        crpy = fvec(co.get('rpy')) if co is not None else [0, 0, 0]
        assert all(abs(x) < 1e-12 for x in crpy)
        single_identity_child = all(abs(c) < 1e-12 for c in com) and all(abs(c) < 1e-12 for c in cxyz)
        if geom.find('mesh') is not None:
            m = geom.find('mesh')
            sc = fvec(m.get('scale')) if m.get('scale') else [1, 1, 1]
            raw = load_mesh(os.path.normpath(os.path.join(udir, m.get('filename'))))
            if os.environ.get('URGYM_DAE_YUP') and m.get('filename').lower().endswith('.dae'):
                raw = np.stack([raw[:, 0], -raw[:, 2], raw[:, 1]], 1)   # experiment: Y_UP -> Z_UP (+90 deg about x)
            pts = raw * np.array(sc) + np.array(cxyz)
            lo, hi = pts.min(0), pts.max(0)
            aabb_c, aabb_h = (lo + hi) / 2, (hi - lo) / 2
            bc, bR, bh = min_volume_box(pts)
            inertia_h = aabb_h + 2 * HULL_MARGIN  # hull aabb incl. margin, plus the margin added once more (recalled)
            is_mesh = 1
            if idx in POLY_LINKS:
                mesh_pts[idx] = (m.get('filename'), pts)
        else:
            size = np.array(fvec(geom.find('box').get('size')))
            aabb_c, aabb_h = np.array(cxyz, dtype=float), size / 2
            bc, bR, bh = aabb_c.copy(), np.eye(3), size / 2
            inertia_h = aabb_h + (0.0 if single_identity_child else HULL_MARGIN)
            is_mesh = 0
        lx, ly, lz = 2 * inertia_h
        inertia = mass / 12.0 * np.array([ly * ly + lz * lz, lx * lx + lz * lz, lx * lx + ly * ly])
        table.append(dict(index=idx, name=name, joint=j.get('name'), parent=parent,
                          jtype=1 if jtype == 'revolute' else 0,
                          axis=axis, jxyz=jxyz, jrot=rpy_to_mat(*jrpy).reshape(-1).tolist(),
                          lower=lower, upper=upper, effort=effort, velocity=vel,
                          mass=mass, com=com, inertia=inertia.tolist(),
                          aabb_center=aabb_c.tolist(), aabb_half=aabb_h.tolist(),
                          box_center=bc.tolist(), box_rot=bR.reshape(-1).tolist(), box_half=bh.tolist(),
                          is_mesh=is_mesh))
    polys, poly_of_file, link_poly = [], {}, [-1] * len(table)
    for idx in POLY_LINKS:
        fname, pts = mesh_pts[idx]
        if fname not in poly_of_file:
            P, err = reduce_hull(pts)
            faces, dirs = polytope_features(P)
            assert len(P) <= POLY_MAXV and len(faces) <= POLY_MAXF and len(dirs) <= POLY_MAXE and max(len(f[2]) for f in faces) <= POLY_MAXL, (fname, len(P), len(faces), len(dirs))
            poly_of_file[fname] = len(polys)
            polys.append(dict(mesh=os.path.basename(fname), verts=P.tolist(), faces=[dict(n=f[0].tolist(), d=f[1], loop=f[2]) for f in faces],
                              edge_dirs=[e.tolist() for e in dirs], dropped_vertex_max_mm=err * 1000))
        link_poly[idx] = poly_of_file[fname]
    names = [t['name'] for t in table]
    expect = ['base_link', 'shoulder_link', 'upper_arm_link', 'forearm_link', 'wrist_1_link', 'wrist_2_link',
              'wrist_3_link', 'ee_link', 'robotiq_arg2f_base_link', 'left_outer_knuckle', 'left_outer_finger',
              'left_inner_finger', 'left_inner_finger_pad', 'left_inner_knuckle', 'right_outer_knuckle',
              'right_outer_finger', 'right_inner_finger', 'right_inner_finger_pad', 'right_inner_knuckle']
    assert names == expect, names
    pitch = load_mesh(os.path.join(args.ref, 'urgym', 'meshes', 'golf_pitch.obj'))
    plo, phi = pitch.min(0), pitch.max(0)
    model = dict(links=table, pitch_half=((phi - plo) / 2).tolist(), pitch_center=((phi + plo) / 2).tolist(), polytopes=polys, link_polytope=link_poly)

    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    os.makedirs(os.path.join(here, 'urgym_b200', 'model'), exist_ok=True)
    with open(os.path.join(here, 'urgym_b200', 'model', 'ur5_robotiq_85.json'), 'w') as f:
        json.dump(model, f, indent=1)

    def arr(vals):
        return '{' + ', '.join(repr(float(v)) for v in vals) + '}'

    H = []
    H.append('// GENERATED by tools/compile_model.py from urgym/urdf/ur5_robotiq_85.urdf and the collision meshes it names.')
    H.append('// Do not edit. Link order = Bullet depth-first numbering (0 base_link ... 18 right_inner_knuckle).')
    H.append('#pragma once')
    H.append('#define URG_NLINK 19')
    H.append('struct UrgLinkModel {')
    H.append('  int parent; int jtype;            // jtype: 0 fixed, 1 revolute')
    H.append('  double axis[3];                   // joint axis in the child link frame')
    H.append('  double jxyz[3]; double jrot[9];   // parent link frame -> child link frame at q=0 (row-major rotation)')
    H.append('  double lower, upper, effort, velocity;')
    H.append('  double mass; double com[3]; double inertia[3];  // inertial frame is axis-aligned with the link frame')
    H.append('  double aabb_center[3], aabb_half[3];            // collision geometry AABB in the link frame (no margin)')
    H.append('  double box_center[3], box_rot[9], box_half[3];  // fitted collision box in the link frame')
    H.append('};')
    H.append('static const UrgLinkModel URG_LINKS[URG_NLINK] = {')
    for t in table:
        H.append('  { %d, %d, %s, %s, %s, %r, %r, %r, %r, %r, %s, %s, %s, %s, %s, %s, %s },  // %d %s' % (
            t['parent'], t['jtype'], arr(t['axis']), arr(t['jxyz']), arr(t['jrot']),
            t['lower'], t['upper'], t['effort'], t['velocity'], t['mass'], arr(t['com']), arr(t['inertia']),
            arr(t['aabb_center']), arr(t['aabb_half']), arr(t['box_center']), arr(t['box_rot']), arr(t['box_half']),
            t['index'], t['name']))
    H.append('};')
    H.append('static const double URG_PITCH_HALF[3] = %s;' % arr(model['pitch_half']))
    H.append('// Convex polytopes of the gripper links whose collision geometry is a mesh (Bullet: convex hull of the mesh vertices):')
    H.append('// at most %d of the hull\'s own vertices (inner approximation), faces with outward normals and CCW loops, unique edge' % POLY_MAXV)
    H.append('// directions. Link frame, URDF scale and collision origin applied. Left / right links share a mesh.')
    H.append('#define URG_NPOLY %d' % len(polys))
    H.append('#define URG_POLY_MAXV %d' % POLY_MAXV)
    H.append('#define URG_POLY_MAXF %d' % POLY_MAXF)
    H.append('#define URG_POLY_MAXL %d' % POLY_MAXL)
    H.append('#define URG_POLY_MAXE %d' % POLY_MAXE)
    H.append('struct UrgPoly { int nv, nf, ne; double v[URG_POLY_MAXV][3]; double fn[URG_POLY_MAXF][4]; int fcount[URG_POLY_MAXF]; int floop[URG_POLY_MAXF][URG_POLY_MAXL]; double ed[URG_POLY_MAXE][3]; };')
    H.append('static const UrgPoly URG_POLYS[URG_NPOLY] = {')
    for pl in polys:
        vv = ', '.join(arr(v) for v in pl['verts'])
        ff = ', '.join(arr(f['n'] + [f['d']]) for f in pl['faces'])
        fc = ', '.join(str(len(f['loop'])) for f in pl['faces'])
        fl = ', '.join('{' + ', '.join(str(i) for i in f['loop']) + '}' for f in pl['faces'])
        ee = ', '.join(arr(e) for e in pl['edge_dirs'])
        H.append('  { %d, %d, %d, {%s}, {%s}, {%s}, {%s}, {%s} },  // %s (dropped hull vertices within %.2f mm)' % (
            len(pl['verts']), len(pl['faces']), len(pl['edge_dirs']), vv, ff, fc, fl, ee, pl['mesh'], pl['dropped_vertex_max_mm']))
    H.append('};')
    H.append('static const int URG_LINK_POLY[URG_NLINK] = {%s};' % ', '.join(str(x) for x in link_poly))
    H.append('')
    txt = '\n'.join(H)
    for rel in ('oracle/model_data.h', 'urgym_b200/csrc/model_data.h'):
        p = os.path.join(here, rel)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, 'w') as f:
            f.write(txt)
    for t in table:
        print('%2d %-26s par %2d m %.4f box_half %s vol_ratio %.2f' % (
            t['index'], t['name'], t['parent'], t['mass'], np.round(t['box_half'], 4),
            np.prod(t['box_half']) / max(np.prod(t['aabb_half']), 1e-12)))
    print('total mass', sum(t['mass'] for t in table))


if __name__ == '__main__':
    main()
