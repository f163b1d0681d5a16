"""Generates arcnerf_b200/csrc/mc_table.cuh: the 256-case triangle table of the marching-cubes kernels (csrc/mcubes.cu).

The table is DERIVED, not transcribed: for every sign configuration of the 8 cube corners the iso-contour is traced on the six
faces (an edge whose end points differ carries one vertex; a face with two such edges carries one segment, a face with four -
the ambiguous face - carries two segments, each cutting off one INSIDE corner, a rule that depends on the face's four corner
states only, so the two cells sharing the face agree and the surface is watertight), the segments are chained into closed
loops and every loop is triangulated as a fan.  Segments are oriented with the inside region on their left as seen from outside
the cube, which orients all loops alike; the global sign is fixed on the one-corner case so that triangle normals point from
inside (value above the level, `grad_dir='descent'` in arcnerf/geometry/mesh.py:12-47) to outside.

Conventions shared with the kernel and the CPU oracle (oracle/marching_cubes.py imports this module):
  corner c = (x, y, z) bits (c & 1, (c >> 1) & 1, (c >> 2) & 1); case index bit c = 1 when corner c is inside;
  edge e = 4 * axis + t, along `axis`, anchored at the corner whose other two coordinates are the bits of t (lower axis first).
"""
import itertools
import os

CORNERS = [(c & 1, (c >> 1) & 1, (c >> 2) & 1) for c in range(8)]


def corner_id(p):
    return p[0] | (p[1] << 1) | (p[2] << 2)


def edge_table():
    """edge e -> (corner a, corner b, axis, anchor offset (dx,dy,dz))"""
    edges = []
    for axis in range(3):
        others = [a for a in range(3) if a != axis]
        for t in range(4):
            p = [0, 0, 0]
            p[others[0]], p[others[1]] = t & 1, (t >> 1) & 1
            q = list(p)
            q[axis] = 1
            edges.append((corner_id(p), corner_id(q), axis, tuple(p)))
    return edges


EDGES = edge_table()
EDGE_OF = {frozenset((a, b)): e for e, (a, b, _, _) in enumerate(EDGES)}


def faces():
    """six faces: (outward normal axis, side) -> the four corners counter-clockwise seen from outside"""
    out = []
    for axis in range(3):
        u, v = (axis + 1) % 3, (axis + 2) % 3            # u x v = +axis
        for side in (0, 1):
            quad = []
            for (a, b) in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = [0, 0, 0]
                p[axis] = side
                p[u], p[v] = a, b
                quad.append(corner_id(p))
            if side == 0:                                # outward normal is -axis: reverse to stay counter-clockwise from outside
                quad = quad[::-1]
            out.append(quad)
    return out


FACES = faces()


def case_triangles(case):
    inside = [(case >> c) & 1 for c in range(8)]
    nxt = {}                                             # oriented segments: edge vertex -> next edge vertex
    for quad in FACES:
        # walk the face counter-clockwise (seen from outside); crossing from an inside corner to an outside corner = the contour
        # LEAVES the inside corner's neighbourhood: with the inside on the left, a segment starts at an out->in crossing ... we
        # orient: start where the boundary walk goes inside -> outside, end where it goes outside -> inside (inside stays left)
        cross = []
        for i in range(4):
            a, b = quad[i], quad[(i + 1) % 4]
            if inside[a] != inside[b]:
                cross.append((i, EDGE_OF[frozenset((a, b))], inside[a]))        # inside[a] = 1: walk leaves the inside here
        if len(cross) == 2:
            (i0, e0, s0), (i1, e1, s1) = cross
            start, end = (e0, e1) if s0 == 1 else (e1, e0)
            # the walk inside->outside at `start`; inside region lies before `start` along the counter-clockwise walk, i.e. the
            # segment from `end` to `start` keeps it on the left
            nxt[end] = start
        elif len(cross) == 4:
            # ambiguous face: each segment cuts off one inside corner: it joins the two crossings adjacent to that corner
            for i in range(4):
                c = quad[i]
                if inside[c]:
                    e_prev = EDGE_OF[frozenset((quad[(i - 1) % 4], c))]          # walk enters the inside corner here (outside -> inside)
                    e_next = EDGE_OF[frozenset((c, quad[(i + 1) % 4]))]          # and leaves it here
                    nxt[e_prev] = e_next
        elif len(cross) != 0:
            raise AssertionError('odd number of crossings')
    tris, seen = [], set()
    for e0 in sorted(nxt):
        if e0 in seen:
            continue
        loop, e = [], e0
        while e not in seen:
            seen.add(e)
            loop.append(e)
            e = nxt[e]
        assert e == e0 and len(loop) >= 3, (case, loop)
        for k in range(1, len(loop) - 1):
            tris.append((loop[0], loop[k], loop[k + 1]))
    assert len(seen) == len(nxt)
    return tris


def build():
    table = [case_triangles(c) for c in range(256)]
    # global orientation: one inside corner at the origin -> the normal of its triangle must point away from the origin
    (a, b, c) = table[1][0]
    mid = lambda e: [0.5 * (CORNERS[EDGES[e][0]][k] + CORNERS[EDGES[e][1]][k]) for k in range(3)]
    pa, pb, pc = mid(a), mid(b), mid(c)
    u = [pb[k] - pa[k] for k in range(3)]
    v = [pc[k] - pa[k] for k in range(3)]
    nrm = [u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]]
    if sum(nrm[k] * pa[k] for k in range(3)) < 0:
        table = [[(t[0], t[2], t[1]) for t in tris] for tris in table]
    return table


TABLE = build()
MAX_TRIS = max(len(t) for t in TABLE)


def write_header(path):
    lines = ['// GENERATED by arcnerf_b200/tools/gen_mc_table.py - do not edit.  256 cases x up to %d triangles of cube-edge ids,' % MAX_TRIS,
             '// -1 terminated; edge e = 4 * axis + t (see the generator for the conventions).',
             '#pragma once', '#include <stdint.h>', 'constexpr int kMcMaxTris = %d;' % MAX_TRIS,
             '__device__ __constant__ int8_t kMcTriTable[256][%d] = {' % (3 * MAX_TRIS + 1)]
    for c, tris in enumerate(TABLE):
        flat = [e for t in tris for e in t] + [-1] * (3 * MAX_TRIS + 1 - 3 * len(tris))
        lines.append('    {' + ', '.join('%d' % e for e in flat) + '},')
    lines.append('};')
    lines.append('__device__ __constant__ uint8_t kMcNumTris[256] = {' + ', '.join(str(len(t)) for t in TABLE) + '};')
    lines.append('// edge e: corner a, corner b, axis, anchor offset x, y, z')
    lines.append('__device__ __constant__ uint8_t kMcEdges[12][6] = {' + ', '.join(
        '{%d, %d, %d, %d, %d, %d}' % (a, b, ax, off[0], off[1], off[2]) for (a, b, ax, off) in EDGES) + '};')
    with open(path, 'w') as f:
        f.write('\n'.join(lines) + '\n')


if __name__ == '__main__':
    here = os.path.dirname(os.path.abspath(__file__))
    write_header(os.path.join(here, '..', 'csrc', 'mc_table.cuh'))
    print('max triangles per cell', MAX_TRIS, 'total', sum(len(t) for t in TABLE))
