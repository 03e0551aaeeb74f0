#!/usr/bin/env python3
"""Generates the 256-case marching-cubes triangle table used by the mesh readback
(`tsdf_extract_mesh`, the replacement of fusion.TSDFVolume.get_mesh() that the reference's
plugin calls in SaveMesh, src/TSDFfusion.py.in:48-53).

The table is constructed, not transcribed: for every sign configuration of the 8 cell corners the
iso-surface is cut face by face (a segment joins the two crossed edges of a cube face; on a face
whose diagonal corners are inside, the two INSIDE corners are separated — a rule that depends on
the face's own four signs only, so both cells sharing a face cut it the same way and the mesh is
watertight), the segments close into loops, each loop is oriented so that its normal points from
the inside (tsdf < 0) corners to the outside ones, and fan-triangulated.

Conventions (shared by csrc/mesh.cu and oracle/tsdf_oracle.c):
  corner c = dx + 2*dy + 4*dz                       bit c of the case index = (tsdf[c] < 0)
  edge   e = 4*axis + k   axis 0 (x): k = dy + 2*dz, joins (0,dy,dz)-(1,dy,dz)
                          axis 1 (y): k = dx + 2*dz, joins (dx,0,dz)-(dx,1,dz)
                          axis 2 (z): k = dx + 2*dy, joins (dx,dy,0)-(dx,dy,1)
  row  = [n_tri, e0,e1,e2, e0,e1,e2, ...] padded with 255 to 1 + 3*MAX_TRI entries

usage: python tools/gen_mc_table.py            -> rewrites csrc/mc_table.inc and oracle/mc_table.inc
       python tools/gen_mc_table.py --print    -> to stdout
"""
import os
import sys

MAX_TRI = 5


def corner_pos(c):
    return (c & 1, (c >> 1) & 1, (c >> 2) & 1)


def corner_id(p):
    return p[0] + 2 * p[1] + 4 * p[2]


def edge_id(p, q):
    """Edge between two corners that differ in exactly one coordinate."""
    axis = [i for i in range(3) if p[i] != q[i]]
    assert len(axis) == 1
    a = axis[0]
    o = [p[i] for i in range(3) if i != a]  # the two fixed coordinates, in axis order
    return 4 * a + o[0] + 2 * o[1]


def edge_ends(e):
    a, k = divmod(e, 4)
    o = (k & 1, k >> 1)
    p, q = [0, 0, 0], [0, 0, 0]
    j = 0
    for i in range(3):
        if i == a:
            p[i], q[i] = 0, 1
        else:
            p[i] = q[i] = o[j]
            j += 1
    return tuple(p), tuple(q)


def face_cycles():
    """The 6 faces as cyclic corner lists."""
    faces = []
    for a in range(3):
        b, c = [i for i in range(3) if i != a]
        for s in (0, 1):
            cyc = []
            for (u, v) in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = [0, 0, 0]
                p[a], p[b], p[c] = s, u, v
                cyc.append(tuple(p))
            faces.append(cyc)
    return faces


def case_triangles(case):
    inside = [(case >> c) & 1 for c in range(8)]
    adj = {}
    for cyc in face_cycles():
        ins = [inside[corner_id(p)] for p in cyc]
        fe = [edge_id(cyc[i], cyc[(i + 1) % 4]) for i in range(4)]       # edge i joins corner i, i+1
        crossed = [i for i in range(4) if ins[i] != ins[(i + 1) % 4]]
        segs = []
        if len(crossed) == 2:
            segs.append((fe[crossed[0]], fe[crossed[1]]))
        elif len(crossed) == 4:
            for i in range(4):
                if ins[i]:  # separate the inside corners: join the two face edges that meet at corner i
                    segs.append((fe[(i - 1) % 4], fe[i]))
        else:
            assert len(crossed) == 0
        for (e0, e1) in segs:
            adj.setdefault(e0, []).append(e1)
            adj.setdefault(e1, []).append(e0)
    for e, nb in adj.items():
        assert len(nb) == 2, (case, e, nb)
    loops, seen = [], set()
    for start in sorted(adj):
        if start in seen:
            continue
        loop, prev, cur = [start], None, start
        seen.add(start)
        while True:
            nxt = [n for n in adj[cur] if n != prev]
            # a 2-cycle cannot occur (two faces never join the same pair of edges)
            n = nxt[0] if nxt[0] not in seen or (nxt[0] == start and len(loop) > 2) else nxt[-1]
            if n == start:
                break
            assert n not in seen, (case, loop, n)
            loop.append(n)
            seen.add(n)
            prev, cur = cur, n
        loops.append(loop)
    oriented = []
    for loop in loops:
        pts, grad = [], [0.0, 0.0, 0.0]
        for e in loop:
            p, q = edge_ends(e)
            pts.append([(p[i] + q[i]) / 2.0 for i in range(3)])
            pin, pout = (p, q) if inside[corner_id(p)] else (q, p)
            for i in range(3):
                grad[i] += pout[i] - pin[i]
        area = [0.0, 0.0, 0.0]
        n = len(pts)
        cen = [sum(p[i] for p in pts) / n for i in range(3)]
        for k in range(n):
            a = [pts[k][i] - cen[i] for i in range(3)]
            b = [pts[(k + 1) % n][i] - cen[i] for i in range(3)]
            area[0] += a[1] * b[2] - a[2] * b[1]
            area[1] += a[2] * b[0] - a[0] * b[2]
            area[2] += a[0] * b[1] - a[1] * b[0]
        d = sum(area[i] * grad[i] for i in range(3))
        assert abs(d) > 1e-9, (case, loop)
        if d < 0:
            loop = loop[::-1]
        # start at the smallest edge id: a canonical, rotation-free spelling
        k0 = loop.index(min(loop))
        oriented.append(loop[k0:] + loop[:k0])
    tris = []
    for loop in oriented:
        tris += best_triangulation(loop)
    return tris


def edge_faces(e):
    p, q = edge_ends(e)
    return {(a, p[a]) for a in range(3) if p[a] == q[a]}


def midpoint(e):
    p, q = edge_ends(e)
    return [(p[i] + q[i]) / 2.0 for i in range(3)]


def triangulations(poly):
    """All triangulations of a polygon (list of labels in order), orientation kept."""
    n = len(poly)
    if n == 3:
        return [[tuple(poly)]]
    out = []
    for k in range(1, n - 1):  # triangle (poly[0], poly[k], poly[n-1]) + the two sides
        left = triangulations(poly[:k + 1]) if k >= 2 else [[]]
        right = triangulations(poly[k:]) if n - 1 - k >= 2 else [[]]
        for l in left:
            for r in right:
                out.append(l + [(poly[0], poly[k], poly[n - 1])] + r)
    return out


def best_triangulation(loop):
    """A diagonal joining two vertices of the same cube face lies IN that face, where the
    neighbouring cell may own a segment or a diagonal between the same two vertices (a mesh edge
    shared by more than two triangles).  Pick the triangulation with the fewest such diagonals
    (none exists for every loop, see the assertion in table()), then the shortest diagonals, then
    the lexicographically first."""
    n = len(loop)
    sides = {frozenset((loop[k], loop[(k + 1) % n])) for k in range(n)}
    best = None
    for tri in triangulations(loop):
        diags = set()
        for t in tri:
            for k in range(3):
                e = frozenset((t[k], t[(k + 1) % 3]))
                if e not in sides:
                    diags.add(e)
        inface = sum(1 for d in diags if len(set.intersection(*[edge_faces(e) for e in d])) > 0)
        length = sum(sum((a - b) ** 2 for a, b in zip(*[midpoint(e) for e in d])) for d in diags)
        # rotate every triangle to start at its smallest label (orientation kept), sort triangles
        canon = sorted(t[t.index(min(t)):] + t[:t.index(min(t))] for t in tri)
        key = (inface, round(length, 9), canon)
        if best is None or key < best[0]:
            best = (key, canon)
    return best[1]


def inface_diagonals():
    """Number of in-face diagonals left per case (diagnostic)."""
    out = {}
    for case in range(256):
        tris = case_triangles(case)
        cnt = {}
        for t in tris:
            for k in range(3):
                e = frozenset((t[k], t[(k + 1) % 3]))
                cnt[e] = cnt.get(e, 0) + 1
        bad = [e for e, c in cnt.items() if c == 2 and len(set.intersection(*[edge_faces(x) for x in e])) > 0]
        if bad:
            out[case] = bad
    return out


def table():
    assert not inface_diagonals(), "a case is left with a diagonal lying in a cube face"
    rows = []
    for case in range(256):
        t = case_triangles(case)
        assert len(t) <= MAX_TRI, (case, len(t))
        row = [len(t)] + [e for tri in t for e in tri]
        row += [255] * (1 + 3 * MAX_TRI - len(row))
        rows.append(row)
    return rows


def render():
    rows = table()
    out = ["// GENERATED by tools/gen_mc_table.py — do not edit.  256 x 16: {n_tri, e0,e1,e2, ...}, 255 = unused.",
           "// corner c = dx + 2*dy + 4*dz; case bit c = (tsdf[c] < 0); edge e = 4*axis + k (see the generator)."]
    for case, row in enumerate(rows):
        out.append("{" + ",".join("%d" % v for v in row) + "}," + (" // %d" % case if case % 16 == 0 else ""))
    return "\n".join(out) + "\n"


if __name__ == "__main__":
    text = render()
    if "--print" in sys.argv:
        sys.stdout.write(text)
    else:
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        for rel in ("semantic-slam_b200/csrc/mc_table.inc", "oracle/mc_table.inc"):
            with open(os.path.join(root, rel), "w") as f:
                f.write(text)
        rows = table()
        print("max triangles per cell:", max(r[0] for r in rows), " total:", sum(r[0] for r in rows))
        print("cases with a diagonal lying in a cube face:", sorted(inface_diagonals()))
