"""Writes include/na_mc_tables.h: the classic Lorensen-Cline marching-cubes case tables in the
vertex/edge numbering popularised by Paul Bourke's "Polygonising a scalar field" (the tables
PyMCubes' marchingcubes.cpp uses).  The 256-entry edge table is DERIVED (an edge is crossed iff
its two corner bits differ); the triangle table is the published one and is VALIDATED here:
for every case the triangles must (i) use exactly the crossed edges, (ii) form, on each cube
face, a legal marching-squares configuration of that face's corner bits, (iii) have every
non-face edge shared by exactly two triangles with opposite orientation, (iv) be consistently
oriented with respect to the inside corners.  Run: python scripts/gen_mc_tables.py
"""
import itertools
import os
import sys

CORNERS = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]
EDGES = [(0, 1), (1, 2), (2, 3), (3, 0), (4, 5), (5, 6), (6, 7), (7, 4), (0, 4), (1, 5), (2, 6), (3, 7)]

TRI = """
-
0 8 3
0 1 9
1 8 3 9 8 1
1 2 10
0 8 3 1 2 10
9 2 10 0 2 9
2 8 3 2 10 8 10 9 8
3 11 2
0 11 2 8 11 0
1 9 0 2 3 11
1 11 2 1 9 11 9 8 11
3 10 1 11 10 3
0 10 1 0 8 10 8 11 10
3 9 0 3 11 9 11 10 9
9 8 10 10 8 11
4 7 8
4 3 0 7 3 4
0 1 9 8 4 7
4 1 9 4 7 1 7 3 1
1 2 10 8 4 7
3 4 7 3 0 4 1 2 10
9 2 10 9 0 2 8 4 7
2 10 9 2 9 7 2 7 3 7 9 4
8 4 7 3 11 2
11 4 7 11 2 4 2 0 4
9 0 1 8 4 7 2 3 11
4 7 11 9 4 11 9 11 2 9 2 1
3 10 1 3 11 10 7 8 4
1 11 10 1 4 11 1 0 4 7 11 4
4 7 8 9 0 11 9 11 10 11 0 3
4 7 11 4 11 9 9 11 10
9 5 4
9 5 4 0 8 3
0 5 4 1 5 0
8 5 4 8 3 5 3 1 5
1 2 10 9 5 4
3 0 8 1 2 10 4 9 5
5 2 10 5 4 2 4 0 2
2 10 5 3 2 5 3 5 4 3 4 8
9 5 4 2 3 11
0 11 2 0 8 11 4 9 5
0 5 4 0 1 5 2 3 11
2 1 5 2 5 8 2 8 11 4 8 5
10 3 11 10 1 3 9 5 4
4 9 5 0 8 1 8 10 1 8 11 10
5 4 0 5 0 11 5 11 10 11 0 3
5 4 8 5 8 10 10 8 11
9 7 8 5 7 9
9 3 0 9 5 3 5 7 3
0 7 8 0 1 7 1 5 7
1 5 3 3 5 7
9 7 8 9 5 7 10 1 2
10 1 2 9 5 0 5 3 0 5 7 3
8 0 2 8 2 5 8 5 7 10 5 2
2 10 5 2 5 3 3 5 7
7 9 5 7 8 9 3 11 2
9 5 7 9 7 2 9 2 0 2 7 11
2 3 11 0 1 8 1 7 8 1 5 7
11 2 1 11 1 7 7 1 5
9 5 8 8 5 7 10 1 3 10 3 11
5 7 0 5 0 9 7 11 0 1 0 10 11 10 0
11 10 0 11 0 3 10 5 0 8 0 7 5 7 0
11 10 5 7 11 5
10 6 5
0 8 3 5 10 6
9 0 1 5 10 6
1 8 3 1 9 8 5 10 6
1 6 5 2 6 1
1 6 5 1 2 6 3 0 8
9 6 5 9 0 6 0 2 6
5 9 8 5 8 2 5 2 6 3 2 8
2 3 11 10 6 5
11 0 8 11 2 0 10 6 5
0 1 9 2 3 11 5 10 6
5 10 6 1 9 2 9 11 2 9 8 11
6 3 11 6 5 3 5 1 3
0 8 11 0 11 5 0 5 1 5 11 6
3 11 6 0 3 6 0 6 5 0 5 9
6 5 9 6 9 11 11 9 8
5 10 6 4 7 8
4 3 0 4 7 3 6 5 10
1 9 0 5 10 6 8 4 7
10 6 5 1 9 7 1 7 3 7 9 4
6 1 2 6 5 1 4 7 8
1 2 5 5 2 6 3 0 4 3 4 7
8 4 7 9 0 5 0 6 5 0 2 6
7 3 9 7 9 4 3 2 9 5 9 6 2 6 9
3 11 2 7 8 4 10 6 5
5 10 6 4 7 2 4 2 0 2 7 11
0 1 9 4 7 8 2 3 11 5 10 6
9 2 1 9 11 2 9 4 11 7 11 4 5 10 6
8 4 7 3 11 5 3 5 1 5 11 6
5 1 11 5 11 6 1 0 11 7 11 4 0 4 11
0 5 9 0 6 5 0 3 6 11 6 3 8 4 7
6 5 9 6 9 11 4 7 9 7 11 9
10 4 9 6 4 10
4 10 6 4 9 10 0 8 3
10 0 1 10 6 0 6 4 0
8 3 1 8 1 6 8 6 4 6 1 10
1 4 9 1 2 4 2 6 4
3 0 8 1 2 9 2 4 9 2 6 4
0 2 4 4 2 6
8 3 2 8 2 4 4 2 6
10 4 9 10 6 4 11 2 3
0 8 2 2 8 11 4 9 10 4 10 6
3 11 2 0 1 6 0 6 4 6 1 10
6 4 1 6 1 10 4 8 1 2 1 11 8 11 1
9 6 4 9 3 6 9 1 3 11 6 3
8 11 1 8 1 0 11 6 1 9 1 4 6 4 1
3 11 6 3 6 0 0 6 4
6 4 8 11 6 8
7 10 6 7 8 10 8 9 10
0 7 3 0 10 7 0 9 10 6 7 10
10 6 7 1 10 7 1 7 8 1 8 0
10 6 7 10 7 1 1 7 3
1 2 6 1 6 8 1 8 9 8 6 7
2 6 9 2 9 1 6 7 9 0 9 3 7 3 9
7 8 0 7 0 6 6 0 2
7 3 2 6 7 2
2 3 11 10 6 8 10 8 9 8 6 7
2 0 7 2 7 11 0 9 7 6 7 10 9 10 7
1 8 0 1 7 8 1 10 7 6 7 10 2 3 11
11 2 1 11 1 7 10 6 1 6 7 1
8 9 6 8 6 7 9 1 6 11 6 3 1 3 6
0 9 1 11 6 7
7 8 0 7 0 6 3 11 0 11 6 0
7 11 6
7 6 11
3 0 8 11 7 6
0 1 9 11 7 6
8 1 9 8 3 1 11 7 6
10 1 2 6 11 7
1 2 10 3 0 8 6 11 7
2 9 0 2 10 9 6 11 7
6 11 7 2 10 3 10 8 3 10 9 8
7 2 3 6 2 7
7 0 8 7 6 0 6 2 0
2 7 6 2 3 7 0 1 9
1 6 2 1 8 6 1 9 8 8 7 6
10 7 6 10 1 7 1 3 7
10 7 6 1 7 10 1 8 7 1 0 8
0 3 7 0 7 10 0 10 9 6 10 7
7 6 10 7 10 8 8 10 9
6 8 4 11 8 6
3 6 11 3 0 6 0 4 6
8 6 11 8 4 6 9 0 1
9 4 6 9 6 3 9 3 1 11 3 6
6 8 4 6 11 8 2 10 1
1 2 10 3 0 11 0 6 11 0 4 6
4 11 8 4 6 11 0 2 9 2 10 9
10 9 3 10 3 2 9 4 3 11 3 6 4 6 3
8 2 3 8 4 2 4 6 2
0 4 2 4 6 2
1 9 0 2 3 4 2 4 6 4 3 8
1 9 4 1 4 2 2 4 6
8 1 3 8 6 1 8 4 6 6 10 1
10 1 0 10 0 6 6 0 4
4 6 3 4 3 8 6 10 3 0 3 9 10 9 3
10 9 4 6 10 4
4 9 5 7 6 11
0 8 3 4 9 5 11 7 6
5 0 1 5 4 0 7 6 11
11 7 6 8 3 4 3 5 4 3 1 5
9 5 4 10 1 2 7 6 11
6 11 7 1 2 10 0 8 3 4 9 5
7 6 11 5 4 10 4 2 10 4 0 2
3 4 8 3 5 4 3 2 5 10 5 2 11 7 6
7 2 3 7 6 2 5 4 9
9 5 4 0 8 6 0 6 2 6 8 7
3 6 2 3 7 6 1 5 0 5 4 0
6 2 8 6 8 7 2 1 8 4 8 5 1 5 8
9 5 4 10 1 6 1 7 6 1 3 7
1 6 10 1 7 6 1 0 7 8 7 0 9 5 4
4 0 10 4 10 5 0 3 10 6 10 7 3 7 10
7 6 10 7 10 8 5 4 10 4 8 10
6 9 5 6 11 9 11 8 9
3 6 11 0 6 3 0 5 6 0 9 5
0 11 8 0 5 11 0 1 5 5 6 11
6 11 3 6 3 5 5 3 1
1 2 10 9 5 11 9 11 8 11 5 6
0 11 3 0 6 11 0 9 6 5 6 9 1 2 10
11 8 5 11 5 6 8 0 5 10 5 2 0 2 5
6 11 3 6 3 5 2 10 3 10 5 3
5 8 9 5 2 8 5 6 2 3 8 2
9 5 6 9 6 0 0 6 2
1 5 8 1 8 0 5 6 8 3 8 2 6 2 8
1 5 6 2 1 6
1 3 6 1 6 10 3 8 6 5 6 9 8 9 6
10 1 0 10 0 6 9 5 0 5 6 0
0 3 8 5 6 10
10 5 6
11 5 10 7 5 11
11 5 10 11 7 5 8 3 0
5 11 7 5 10 11 1 9 0
10 7 5 10 11 7 9 8 1 8 3 1
11 1 2 11 7 1 7 5 1
0 8 3 1 2 7 1 7 5 7 2 11
9 7 5 9 2 7 9 0 2 2 11 7
7 5 2 7 2 11 5 9 2 3 2 8 9 8 2
2 5 10 2 3 5 3 7 5
8 2 0 8 5 2 8 7 5 10 2 5
9 0 1 5 10 3 5 3 7 3 10 2
9 8 2 9 2 1 8 7 2 10 2 5 7 5 2
1 3 5 3 7 5
0 8 7 0 7 1 1 7 5
9 0 3 9 3 5 5 3 7
9 8 7 5 9 7
5 8 4 5 10 8 10 11 8
5 0 4 5 11 0 5 10 11 11 3 0
0 1 9 8 4 10 8 10 11 10 4 5
10 11 4 10 4 5 11 3 4 9 4 1 3 1 4
2 5 1 2 8 5 2 11 8 4 5 8
0 4 11 0 11 3 4 5 11 2 11 1 5 1 11
0 2 5 0 5 9 2 11 5 4 5 8 11 8 5
9 4 5 2 11 3
2 5 10 3 5 2 3 4 5 3 8 4
5 10 2 5 2 4 4 2 0
3 10 2 3 5 10 3 8 5 4 5 8 0 1 9
5 10 2 5 2 4 1 9 2 9 4 2
8 4 5 8 5 3 3 5 1
0 4 5 1 0 5
8 4 5 8 5 3 9 0 5 0 3 5
9 4 5
4 11 7 4 9 11 9 10 11
0 8 3 4 9 7 9 11 7 9 10 11
1 10 11 1 11 4 1 4 0 7 4 11
3 1 4 3 4 8 1 10 4 7 4 11 10 11 4
4 11 7 9 11 4 9 2 11 9 1 2
9 7 4 9 11 7 9 1 11 2 11 1 0 8 3
11 7 4 11 4 2 2 4 0
11 7 4 11 4 2 8 3 4 3 2 4
2 9 10 2 7 9 2 3 7 7 4 9
9 10 7 9 7 4 10 2 7 8 7 0 2 0 7
3 7 10 3 10 2 7 4 10 1 10 0 4 0 10
1 10 2 8 7 4
4 9 1 4 1 7 7 1 3
4 9 1 4 1 7 0 8 1 8 7 1
4 0 3 7 4 3
4 8 7
9 10 8 10 11 8
3 0 9 3 9 11 11 9 10
0 1 10 0 10 8 8 10 11
3 1 10 11 3 10
1 2 11 1 11 9 9 11 8
3 0 9 3 9 11 1 2 9 2 11 9
0 2 11 8 0 11
3 2 11
2 3 8 2 8 10 10 8 9
9 10 2 0 9 2
2 3 8 2 8 10 0 1 8 1 10 8
1 10 2
1 3 8 9 1 8
0 9 1
0 3 8
-
"""


def parse():
    rows = [ln.strip() for ln in TRI.strip().splitlines()]
    assert len(rows) == 256, len(rows)
    table = []
    for r in rows:
        vals = [] if r == "-" else [int(v) for v in r.split()]
        assert len(vals) % 3 == 0 and len(vals) <= 15
        table.append(vals)
    return table


def edge_mask(case):
    m = 0
    for e, (a, b) in enumerate(EDGES):
        if ((case >> a) & 1) != ((case >> b) & 1):
            m |= 1 << e
    return m


def edge_faces(e):
    """The two cube faces (axis, side) an edge lies on."""
    a, b = (CORNERS[i] for i in EDGES[e])
    return [(ax, a[ax]) for ax in range(3) if a[ax] == b[ax]]


def edge_mid(e):
    a, b = (CORNERS[i] for i in EDGES[e])
    return tuple((a[i] + b[i]) / 2 for i in range(3))


def validate(table):
    problems = []
    for case, tris in enumerate(table):
        mask = edge_mask(case)
        used = 0
        for e in tris:
            used |= 1 << e
        if used != mask:
            problems.append((case, "edges used != crossed edges"))
            continue
        tri_list = [tuple(tris[i:i + 3]) for i in range(0, len(tris), 3)]
        directed = {}
        for t in tri_list:
            if len(set(t)) != 3:
                problems.append((case, f"degenerate triangle {t}"))
            for i in range(3):
                key = (t[i], t[(i + 1) % 3])
                directed[key] = directed.get(key, 0) + 1
        boundary = []
        for (a, b), cnt in directed.items():
            if cnt != 1:
                problems.append((case, f"directed edge {(a, b)} used {cnt}x"))
            if (b, a) not in directed:
                boundary.append((a, b))
        # boundary segments must lie on a common cube face; each crossed edge is met by exactly two of them
        deg = {}
        per_face = {}
        for a, b in boundary:
            common = set(edge_faces(a)) & set(edge_faces(b))
            if len(common) != 1:
                problems.append((case, f"boundary segment {(a, b)} not on one cube face"))
                continue
            per_face.setdefault(next(iter(common)), []).append((a, b))
            deg[a] = deg.get(a, 0) + 1
            deg[b] = deg.get(b, 0) + 1
        for e in range(12):
            if (mask >> e) & 1 and deg.get(e, 0) != 2:
                problems.append((case, f"edge {e} has boundary degree {deg.get(e, 0)}"))
        for (ax, side), segs in per_face.items():
            bits = [(case >> i) & 1 for i, c in enumerate(CORNERS) if c[ax] == side]
            n_in = sum(bits)
            expect = {0: 0, 4: 0, 1: 1, 3: 1}.get(n_in)
            if n_in == 2:
                idx = [i for i, c in enumerate(CORNERS) if c[ax] == side]
                ins = [CORNERS[i] for i in idx if (case >> i) & 1]
                diag = sum(abs(ins[0][k] - ins[1][k]) for k in range(3)) == 2
                expect = 2 if diag else 1
            if len(segs) != expect:
                problems.append((case, f"face {(ax, side)} has {len(segs)} segments, expected {expect}"))
        # orientation: every triangle must see the set corners on the same side (checked through its 3 edges' inside ends)
        signs = set()
        for t in tri_list:
            p = [edge_mid(e) for e in t]
            u = [p[1][i] - p[0][i] for i in range(3)]
            v = [p[2][i] - p[0][i] for i in range(3)]
            n = (u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0])
            s = 0.0
            for e in t:
                a, b = EDGES[e]
                inside = CORNERS[a] if (case >> a) & 1 else CORNERS[b]
                m = edge_mid(e)
                s += sum(n[i] * (inside[i] - m[i]) for i in range(3))
            if abs(s) < 1e-9:
                problems.append((case, f"triangle {t} orientation undetermined"))
            signs.add(s > 0)
        if len(signs) > 1:
            problems.append((case, "mixed triangle orientation"))
    return problems


def emit(table, path):
    lines = ["/* GENERATED by scripts/gen_mc_tables.py -- do not edit.",
             " * Marching-cubes case tables (Lorensen & Cline 1987) in the corner/edge numbering of",
             " * P. Bourke, \"Polygonising a scalar field\": corners v0=(0,0,0) v1=(1,0,0) v2=(1,1,0) v3=(0,1,0),",
             " * v4..v7 the same at z+1; edges 0:v0v1 1:v1v2 2:v2v3 3:v3v0 4:v4v5 5:v5v6 6:v6v7 7:v7v4 8:v0v4",
             " * 9:v1v5 10:v2v6 11:v3v7.  Bit m of the case index is set when corner m is <= the iso value. */",
             "#ifndef NA_MC_TABLES_H", "#define NA_MC_TABLES_H", "",
             "#define NA_MC_EDGE_TABLE_INIT { \\"]
    em = [edge_mask(c) for c in range(256)]
    for i in range(0, 256, 8):
        lines.append("    " + ", ".join(f"0x{v:03x}" for v in em[i:i + 8]) + ", \\")
    lines[-1] = lines[-1][:-3] + " }"
    lines += ["", "#define NA_MC_TRI_COUNT_INIT { \\"]
    cnt = [len(t) // 3 for t in table]
    for i in range(0, 256, 32):
        lines.append("    " + ", ".join(str(v) for v in cnt[i:i + 32]) + ", \\")
    lines[-1] = lines[-1][:-3] + " }"
    lines += ["", "/* 16 entries per case, -1 terminated */", "#define NA_MC_TRI_TABLE_INIT { \\"]
    for t in table:
        row = t + [-1] * (16 - len(t))
        lines.append("    " + ", ".join(f"{v:2d}" for v in row) + ", \\")
    lines[-1] = lines[-1][:-3] + " }"
    lines += ["", "#endif /* NA_MC_TABLES_H */", ""]
    with open(path, "w") as fh:
        fh.write("\n".join(lines))


if __name__ == "__main__":
    tbl = parse()
    bad = validate(tbl)
    for b in bad[:40]:
        print("PROBLEM", b)
    print(len(bad), "problems;", sum(len(t) // 3 for t in tbl), "triangles in total")
    if bad:
        sys.exit(1)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    emit(tbl, os.path.join(root, "include", "na_mc_tables.h"))
    print("wrote include/na_mc_tables.h")
