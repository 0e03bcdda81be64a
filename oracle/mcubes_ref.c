/*
 * CPU restatement of PyMCubes' marching cubes -- TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * PARITY UNPINNED: the algorithm lives in a third-party dependency that is NOT vendored under the
 * reference tree: PyMCubes (pip, pinned version: none; reference requirements.txt:23).  Reference
 * call site: projects/neuralangelo/utils/mesh.py:121 `V, F = mcubes.marching_cubes(sdf, 0.)`.
 * The reference holds no test or golden vector at this boundary and PyMCubes cannot be installed
 * here, so this file restates PyMCubes' published algorithm (mcubes/src/marchingcubes.h:
 * `marching_cubes`, `mc_add_vertex`, `mc_isovalue_interpolation`) and is anchored on analytic
 * properties (tests/test_oracle_mcubes.py: closed 2-manifold, Euler characteristic, vertex radii).
 *
 * Algorithm restated:
 *   - cells visited x-outer, y, z-inner; corner m of the cell is "inside" when v[m] <= iso;
 *   - Lorensen-Cline tables in Bourke's numbering (include/na_mc_tables.h);
 *   - one vertex per crossed lattice edge, created by the first cell (in visiting order) that
 *     touches the edge: a cell always creates its edges 6, 5, 10 (the three meeting at corner 6)
 *     and, on the low boundaries, 0 (j==0&&k==0), 1 (k==0), 2 (k==0), 3 (i==0&&k==0), 4 (j==0),
 *     7 (i==0), 8 (i==0&&j==0), 9 (j==0), 11 (i==0) -- in exactly that order;
 *     CONVENTION CHOSEN HERE (mode 0, what csrc/mcubes.cu implements): the low-boundary conditions are
 *     conjunctions (`&&`), i.e. every crossed lattice edge owns exactly ONE vertex, also on the low
 *     faces of the volume -- the mesh of a block is a proper indexed manifold.  Upstream PyMCubes may
 *     instead re-create edges 0 / 3 / 8 whenever the cell lies on EITHER low face (`||`), which leaves
 *     duplicate vertices (same position, different index) along the low faces of every block; that cannot
 *     be checked here (PyMCubes absent).  mc_ref_run_mode(..., 1) produces that variant so that the tests
 *     can show what is and is not affected: vertex COUNT and vertex ids differ, the set of triangles as
 *     geometry (faces keyed by the lattice edge / position of their corners) is identical
 *     (mcubes_ref.canonical_faces, tests/test_oracle_mcubes.py, tests/test_mcubes_gpu.py).
 *   - vertex position by linear interpolation in double, x1 + (x2-x1)*(iso-f1)/(f2-f1), midpoint
 *     when f1 == f2, in index units; triangles in table order.
 * This file doubles as the timed single-thread CPU baseline for isosurface extraction.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "na_mc_tables.h"

static const uint16_t edge_table[256] = NA_MC_EDGE_TABLE_INIT;
static const int8_t tri_table[256][16] = NA_MC_TRI_TABLE_INIT;

typedef struct {
    double* v;
    size_t n, cap;
} dvec;
typedef struct {
    int64_t* v;
    size_t n, cap;
} ivec;

static void dpush3(dvec* a, double x, double y, double z) {
    if (a->n + 3 > a->cap) {
        a->cap = a->cap ? a->cap * 2 : 3072;
        a->v = (double*)realloc(a->v, a->cap * sizeof(double));
    }
    a->v[a->n++] = x;
    a->v[a->n++] = y;
    a->v[a->n++] = z;
}
static void ipush(ivec* a, int64_t x) {
    if (a->n + 1 > a->cap) {
        a->cap = a->cap ? a->cap * 2 : 3072;
        a->v = (int64_t*)realloc(a->v, a->cap * sizeof(int64_t));
    }
    a->v[a->n++] = x;
}

static double interp(double iso, double f1, double f2, double x1, double x2) {
    if (f2 == f1) return (x2 + x1) / 2;
    return (x2 - x1) * (iso - f1) / (f2 - f1) + x1;
}

/* axis: which coordinate is interpolated from its value in (x1,y1,z1) towards c2 */
static int64_t add_vertex(dvec* verts, double x1, double y1, double z1, double c2, int axis, double f1, double f2,
                          double iso) {
    int64_t index = (int64_t)(verts->n / 3);
    if (axis == 0) dpush3(verts, interp(iso, f1, f2, x1, c2), y1, z1);
    else if (axis == 1) dpush3(verts, x1, interp(iso, f1, f2, y1, c2), z1);
    else dpush3(verts, x1, y1, interp(iso, f1, f2, z1, c2));
    return index;
}

/* volume: float32 [nx][ny][nz] (C order).  Outputs are malloc'ed; release with mc_ref_free.
 * cases (optional, may be NULL): uint8 [(nx-1)*(ny-1)*(nz-1)] per-cell case index. */
int mc_ref_run_mode(const float* volume, int nx, int ny, int nz, double iso, double** out_verts, int64_t* out_nv,
                    int64_t** out_faces, int64_t* out_nf, uint8_t* cases, int dup_low_faces) {
#define LOW2(a, b) (dup_low_faces ? ((a) || (b)) : ((a) && (b)))
    const int cx = nx - 1, cy = ny - 1, cz = nz - 1;
    dvec verts = {0, 0, 0};
    ivec polys = {0, 0, 0};
    *out_verts = NULL, *out_faces = NULL, *out_nv = 0, *out_nf = 0;
    if (cx < 1 || cy < 1 || cz < 1) return 0;
    /* vertex ids of lattice edges, two x-slabs deep: [slab][y 0..cy][z 0..cz] per edge direction */
    const size_t slab = (size_t)(cy + 1) * (cz + 1);
    int64_t* sx = (int64_t*)malloc(2 * slab * sizeof(int64_t)); /* x-edges starting at lattice (i, Y, Z) */
    int64_t* sy = (int64_t*)malloc(2 * slab * sizeof(int64_t)); /* y-edges at lattice x-plane, (Y, Z)   */
    int64_t* sz = (int64_t*)malloc(2 * slab * sizeof(int64_t)); /* z-edges at lattice x-plane, (Y, Z)   */
#define VOL(a, b, c) ((double)volume[((size_t)(a) * ny + (b)) * nz + (c)])
#define SX(Y, Z) sx[(size_t)(Y) * (cz + 1) + (Z)]                        /* x-edge of the current column i */
#define SY(P, Y, Z) sy[(size_t)((P) & 1) * slab + (size_t)(Y) * (cz + 1) + (Z)] /* y-edge in lattice plane x=P */
#define SZ(P, Y, Z) sz[(size_t)((P) & 1) * slab + (size_t)(Y) * (cz + 1) + (Z)]
    for (int i = 0; i < cx; ++i) {
        const double x = i, x_dx = i + 1;
        for (int j = 0; j < cy; ++j) {
            const double y = j, y_dy = j + 1;
            for (int k = 0; k < cz; ++k) {
                const double z = k, z_dz = k + 1;
                double v[8];
                v[0] = VOL(i, j, k), v[1] = VOL(i + 1, j, k), v[2] = VOL(i + 1, j + 1, k), v[3] = VOL(i, j + 1, k);
                v[4] = VOL(i, j, k + 1), v[5] = VOL(i + 1, j, k + 1), v[6] = VOL(i + 1, j + 1, k + 1);
                v[7] = VOL(i, j + 1, k + 1);
                unsigned cube = 0;
                for (int m = 0; m < 8; ++m)
                    if (v[m] <= iso) cube |= 1u << m;
                if (cases) cases[((size_t)i * cy + j) * cz + k] = (uint8_t)cube;
                const unsigned edges = edge_table[cube];
                if (!edges) continue;
                int64_t idx[12];
                if (edges & 0x040) idx[6] = SX(j + 1, k + 1) = add_vertex(&verts, x_dx, y_dy, z_dz, x, 0, v[6], v[7], iso);
                if (edges & 0x020) idx[5] = SY(i + 1, j, k + 1) = add_vertex(&verts, x_dx, y, z_dz, y_dy, 1, v[5], v[6], iso);
                if (edges & 0x400) idx[10] = SZ(i + 1, j + 1, k) = add_vertex(&verts, x_dx, y_dy, z, z_dz, 2, v[2], v[6], iso);
                if (edges & 0x001) {
                    if (LOW2(j == 0, k == 0)) idx[0] = SX(j, k) = add_vertex(&verts, x, y, z, x_dx, 0, v[0], v[1], iso);
                    else idx[0] = SX(j, k);
                }
                if (edges & 0x002) {
                    if (k == 0) idx[1] = SY(i + 1, j, 0) = add_vertex(&verts, x_dx, y, z, y_dy, 1, v[1], v[2], iso);
                    else idx[1] = SY(i + 1, j, k);
                }
                if (edges & 0x004) {
                    if (k == 0) idx[2] = SX(j + 1, 0) = add_vertex(&verts, x_dx, y_dy, z, x, 0, v[2], v[3], iso);
                    else idx[2] = SX(j + 1, k);
                }
                if (edges & 0x008) {
                    if (LOW2(i == 0, k == 0)) idx[3] = SY(i, j, k) = add_vertex(&verts, x, y_dy, z, y, 1, v[3], v[0], iso);
                    else idx[3] = SY(i, j, k);
                }
                if (edges & 0x010) {
                    if (j == 0) idx[4] = SX(0, k + 1) = add_vertex(&verts, x, y, z_dz, x_dx, 0, v[4], v[5], iso);
                    else idx[4] = SX(j, k + 1);
                }
                if (edges & 0x080) {
                    if (i == 0) idx[7] = SY(0, j, k + 1) = add_vertex(&verts, x, y_dy, z_dz, y, 1, v[7], v[4], iso);
                    else idx[7] = SY(i, j, k + 1);
                }
                if (edges & 0x100) {
                    if (LOW2(i == 0, j == 0)) idx[8] = SZ(i, j, k) = add_vertex(&verts, x, y, z, z_dz, 2, v[0], v[4], iso);
                    else idx[8] = SZ(i, j, k);
                }
                if (edges & 0x200) {
                    if (j == 0) idx[9] = SZ(i + 1, 0, k) = add_vertex(&verts, x_dx, y, z, z_dz, 2, v[1], v[5], iso);
                    else idx[9] = SZ(i + 1, j, k);
                }
                if (edges & 0x800) {
                    if (i == 0) idx[11] = SZ(0, j + 1, k) = add_vertex(&verts, x, y_dy, z, z_dz, 2, v[3], v[7], iso);
                    else idx[11] = SZ(i, j + 1, k);
                }
                const int8_t* row = tri_table[cube];
                for (int m = 0; row[m] != -1; ++m) ipush(&polys, idx[row[m]]);
            }
        }
    }
#undef LOW2
#undef VOL
#undef SX
#undef SY
#undef SZ
    free(sx), free(sy), free(sz);
    *out_verts = verts.v, *out_nv = (int64_t)(verts.n / 3);
    *out_faces = polys.v, *out_nf = (int64_t)(polys.n / 3);
    return 0;
}

int mc_ref_run(const float* volume, int nx, int ny, int nz, double iso, double** out_verts, int64_t* out_nv,
               int64_t** out_faces, int64_t* out_nf, uint8_t* cases) {
    return mc_ref_run_mode(volume, nx, ny, nz, iso, out_verts, out_nv, out_faces, out_nf, cases, 0);
}

void mc_ref_free(void* p) { free(p); }
