// GPU marching cubes with stream-compacted, serially-ordered emission.
//
// Replaces PyMCubes' single-threaded `mcubes.marching_cubes(volume, 0.)` behind
// projects/neuralangelo/utils/mesh.py:121 (PyMCubes is an un-vendored pip dependency of the
// reference, requirements.txt:23; the algorithm restated here is SURVEY.md section 8c).
//
// The output is bit-identical to the serial algorithm, ordering included: vertex v is "created"
// by the first cell -- in x-outer / y / z-inner visiting order -- that touches its lattice edge,
// and within a cell in the fixed edge order 6,5,10,0,1,2,3,4,7,8,9,11.  Both facts are local: a
// cell knows from its own (i,j,k) which of its 12 edges it creates, so
//     vertex id  = exclusive_scan(#created-and-crossed edges per cell)[owner] + rank inside owner
//     face   id  = exclusive_scan(#triangles per cell)
// and a three-phase classify / scan / emit pipeline reproduces the serial numbering exactly.
// Scratch (caller-owned): 1 byte case index + 4 byte vertex offset per cell, 8 bytes per CTA.
// All kernels are HBM-streaming: 4 B/sample read in classify, 5 B/cell in emit, compacted writes.
#include "na_common.cuh"
#include "na_mc_tables.h"

namespace {

constexpr int kThreads = 256;

__constant__ uint16_t c_edge_table[256] = NA_MC_EDGE_TABLE_INIT;
__constant__ uint8_t c_tri_count[256] = NA_MC_TRI_COUNT_INIT;
__constant__ int8_t c_tri_table[256 * 16] = NA_MC_TRI_TABLE_INIT;

// creation order of a cell's edges and, per edge, the bit mask of the edges created before it
__constant__ uint8_t c_create_order[12] = {6, 5, 10, 0, 1, 2, 3, 4, 7, 8, 9, 11};

struct Dims {
    int nx, ny, nz;     // samples
    int cx, cy, cz;     // cells
    int64_t n_cells;
};

__host__ __device__ inline Dims make_dims(int nx, int ny, int nz) {
    Dims d{nx, ny, nz, nx - 1, ny - 1, nz - 1, 0};
    d.n_cells = (int64_t)d.cx * d.cy * d.cz;
    return d;
}

// Which of its 12 edges cell (i,j,k) creates (bit e set) -- see the header comment.
__device__ __forceinline__ uint32_t created_mask(int i, int j, int k) {
    uint32_t m = (1u << 6) | (1u << 5) | (1u << 10);
    if (j == 0 && k == 0) m |= 1u << 0;
    if (k == 0) m |= (1u << 1) | (1u << 2);
    if (i == 0 && k == 0) m |= 1u << 3;
    if (j == 0) m |= (1u << 4) | (1u << 9);
    if (i == 0) m |= (1u << 7) | (1u << 11);
    if (i == 0 && j == 0) m |= 1u << 8;
    return m;
}

// mask of edges that precede edge e in the creation order
__device__ __forceinline__ uint32_t before_mask(int e) {
    //            order: 6, 5, 10, 0, 1, 2, 3, 4, 7, 8, 9, 11
    switch (e) {
        case 6: return 0;
        case 5: return 1u << 6;
        case 10: return (1u << 6) | (1u << 5);
        case 0: return (1u << 6) | (1u << 5) | (1u << 10);
        case 1: return (1u << 6) | (1u << 5) | (1u << 10) | 0x001;
        case 2: return (1u << 6) | (1u << 5) | (1u << 10) | 0x003;
        case 3: return (1u << 6) | (1u << 5) | (1u << 10) | 0x007;
        case 4: return (1u << 6) | (1u << 5) | (1u << 10) | 0x00f;
        case 7: return (1u << 6) | (1u << 5) | (1u << 10) | 0x01f;
        case 8: return (1u << 6) | (1u << 5) | (1u << 10) | 0x09f;
        case 9: return (1u << 6) | (1u << 5) | (1u << 10) | 0x19f;
        default: return (1u << 6) | (1u << 5) | (1u << 10) | 0x39f;  // 11
    }
}

__device__ __forceinline__ void cell_coords(const Dims& d, int64_t c, int& i, int& j, int& k) {
    k = (int)(c % d.cz);
    const int64_t t = c / d.cz;
    j = (int)(t % d.cy);
    i = (int)(t / d.cy);
}

__device__ __forceinline__ float vol(const float* v, const Dims& d, int a, int b, int c) {
    return __ldg(v + ((int64_t)a * d.ny + b) * d.nz + c);
}

__device__ __forceinline__ void load_corners(const float* v, const Dims& d, int i, int j, int k, float f[8]) {
    f[0] = vol(v, d, i, j, k), f[1] = vol(v, d, i + 1, j, k), f[2] = vol(v, d, i + 1, j + 1, k);
    f[3] = vol(v, d, i, j + 1, k), f[4] = vol(v, d, i, j, k + 1), f[5] = vol(v, d, i + 1, j, k + 1);
    f[6] = vol(v, d, i + 1, j + 1, k + 1), f[7] = vol(v, d, i, j + 1, k + 1);
}

// CTA-wide exclusive scan of one uint per thread (256 threads); returns the exclusive prefix and the CTA total.
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* warp_sums, uint32_t& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        const uint32_t s = warp_sums[w];
        if (w < warp) base += s;
        tot += s;
    }
    __syncthreads();
    total = tot;
    return base + incl - v;
}

// ---- phase 1: per-cell case index + per-CTA (vertex, triangle) counts ---------------------------
__global__ void __launch_bounds__(kThreads)
mc_classify_kernel(const float* __restrict__ volume, Dims d, double iso, uint8_t* __restrict__ cases,
                   uint2* __restrict__ block_sums) {
    __shared__ uint32_t ws[kThreads / 32];
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    uint32_t nv = 0, nt = 0;
    if (c < d.n_cells) {
        int i, j, k;
        cell_coords(d, c, i, j, k);
        float f[8];
        load_corners(volume, d, i, j, k, f);
        uint32_t cube = 0;
#pragma unroll
        for (int m = 0; m < 8; ++m)
            if ((double)f[m] <= iso) cube |= 1u << m;
        cases[c] = (uint8_t)cube;
        nv = __popc((uint32_t)c_edge_table[cube] & created_mask(i, j, k));
        nt = c_tri_count[cube];
    }
    uint32_t tv, tt;
    block_exclusive_scan(nv, ws, tv);
    block_exclusive_scan(nt, ws, tt);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = make_uint2(tv, tt);
}

// ---- phase 2: exclusive scan of the per-CTA sums: ONE CTA of 1024 threads, one pass --------------
// Each thread owns a contiguous segment of the sums (8 of the 8 193 of a 129^3 block), the segment totals are scanned
// across the block with shuffles, then every thread rewrites its segment as exclusive offsets.  (The first version walked
// the array in chunks of one element per thread with two block scans per chunk: 25 us per block of extract_mesh.)
constexpr int kScanThreads = 1024;
__global__ void __launch_bounds__(kScanThreads)
mc_scan_blocks_kernel(uint2* __restrict__ block_sums, int64_t n_blocks, uint32_t* __restrict__ counts) {
    __shared__ uint32_t wv[kScanThreads / 32], wt[kScanThreads / 32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int64_t per = (n_blocks + kScanThreads - 1) / kScanThreads;
    const int64_t b0 = (int64_t)t * per, b1 = b0 + per < n_blocks ? b0 + per : n_blocks;
    uint32_t sv = 0, st = 0;
    for (int64_t b = b0; b < b1; ++b) {
        const uint2 s = block_sums[b];
        sv += s.x, st += s.y;
    }
    uint32_t iv = sv, it = st;   // inclusive scan inside the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t uv = __shfl_up_sync(0xffffffffu, iv, o), ut = __shfl_up_sync(0xffffffffu, it, o);
        if (lane >= o) iv += uv, it += ut;
    }
    if (lane == 31) wv[w] = iv, wt[w] = it;
    __syncthreads();
    if (w == 0) {                // exclusive scan of the 32 warp totals
        const uint32_t tv = wv[lane], tt = wt[lane];
        uint32_t xv = tv, xt = tt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t uv = __shfl_up_sync(0xffffffffu, xv, o), ut = __shfl_up_sync(0xffffffffu, xt, o);
            if (lane >= o) xv += uv, xt += ut;
        }
        wv[lane] = xv - tv, wt[lane] = xt - tt;
        if (lane == 31) counts[0] = xv, counts[1] = xt;
    }
    __syncthreads();
    uint32_t ev = wv[w] + iv - sv, et = wt[w] + it - st;
    for (int64_t b = b0; b < b1; ++b) {
        const uint2 s = block_sums[b];
        block_sums[b] = make_uint2(ev, et);
        ev += s.x, et += s.y;
    }
}

__device__ __forceinline__ double interp(double iso, double f1, double f2, double x1, double x2) {
    if (f2 == f1) return __ddiv_rn(__dadd_rn(x2, x1), 2.0);
    // (x2 - x1) * (iso - f1) / (f2 - f1) + x1 with every operation individually rounded (no FMA contraction)
    return __dadd_rn(__ddiv_rn(__dmul_rn(__dsub_rn(x2, x1), __dsub_rn(iso, f1)), __dsub_rn(f2, f1)), x1);
}

struct Xform {
    double scale;
    double origin[3];
};

// ---- phase 3: per-cell vertex offsets + vertex emission -------------------------------------------
__global__ void __launch_bounds__(kThreads)
mc_vertices_kernel(const float* __restrict__ volume, Dims d, double iso, const uint8_t* __restrict__ cases,
                   const uint2* __restrict__ block_offsets, uint32_t* __restrict__ vert_offset,
                   double* __restrict__ verts, int64_t cap_v, Xform xf) {
    __shared__ uint32_t ws[kThreads / 32];
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    uint32_t mask = 0;
    int i = 0, j = 0, k = 0;
    if (c < d.n_cells) {
        cell_coords(d, c, i, j, k);
        mask = (uint32_t)c_edge_table[cases[c]] & created_mask(i, j, k);
    }
    uint32_t total;
    const uint32_t off = block_offsets[blockIdx.x].x + block_exclusive_scan(__popc(mask), ws, total);
    if (c >= d.n_cells) return;
    vert_offset[c] = off;
    if (!mask) return;
    float f[8];
    load_corners(volume, d, i, j, k, f);
    const double x = i, y = j, z = k, x1 = i + 1, y1 = j + 1, z1 = k + 1;
    uint32_t out = off;
#pragma unroll
    for (int q = 0; q < 12; ++q) {
        const int e = c_create_order[q];
        if (!((mask >> e) & 1u)) continue;
        double px, py, pz;
        switch (e) {
            case 6: px = interp(iso, f[6], f[7], x1, x), py = y1, pz = z1; break;
            case 5: px = x1, py = interp(iso, f[5], f[6], y, y1), pz = z1; break;
            case 10: px = x1, py = y1, pz = interp(iso, f[2], f[6], z, z1); break;
            case 0: px = interp(iso, f[0], f[1], x, x1), py = y, pz = z; break;
            case 1: px = x1, py = interp(iso, f[1], f[2], y, y1), pz = z; break;
            case 2: px = interp(iso, f[2], f[3], x1, x), py = y1, pz = z; break;
            case 3: px = x, py = interp(iso, f[3], f[0], y1, y), pz = z; break;
            case 4: px = interp(iso, f[4], f[5], x, x1), py = y, pz = z1; break;
            case 7: px = x, py = interp(iso, f[7], f[4], y1, y), pz = z1; break;
            case 8: px = x, py = y, pz = interp(iso, f[0], f[4], z, z1); break;
            case 9: px = x1, py = y, pz = interp(iso, f[1], f[5], z, z1); break;
            default: px = x, py = y1, pz = interp(iso, f[3], f[7], z, z1); break;  // 11
        }
        if ((int64_t)out < cap_v) {
            double* o = verts + (int64_t)out * 3;
            o[0] = __dadd_rn(__dmul_rn(px, xf.scale), xf.origin[0]);
            o[1] = __dadd_rn(__dmul_rn(py, xf.scale), xf.origin[1]);
            o[2] = __dadd_rn(__dmul_rn(pz, xf.scale), xf.origin[2]);
        }
        ++out;
    }
}

// Vertex id of edge e of cell (i,j,k): find the creating cell and the edge's number there.
__device__ __forceinline__ uint32_t vertex_id(const Dims& d, const uint8_t* cases, const uint32_t* vert_offset, int i,
                                              int j, int k, int e) {
    int oi = i, oj = j, ok = k, oe = e;
    switch (e) {
        case 0:
            if (j > 0 && k > 0) oj = j - 1, ok = k - 1, oe = 6;
            else if (k > 0) ok = k - 1, oe = 4;
            else if (j > 0) oj = j - 1, oe = 2;
            break;
        case 1: if (k > 0) ok = k - 1, oe = 5; break;
        case 2: if (k > 0) ok = k - 1, oe = 6; break;
        case 3:
            if (i > 0 && k > 0) oi = i - 1, ok = k - 1, oe = 5;
            else if (i > 0) oi = i - 1, oe = 1;
            else if (k > 0) ok = k - 1, oe = 7;
            break;
        case 4: if (j > 0) oj = j - 1, oe = 6; break;
        case 7: if (i > 0) oi = i - 1, oe = 5; break;
        case 8:
            if (i > 0 && j > 0) oi = i - 1, oj = j - 1, oe = 10;
            else if (i > 0) oi = i - 1, oe = 9;
            else if (j > 0) oj = j - 1, oe = 11;
            break;
        case 9: if (j > 0) oj = j - 1, oe = 10; break;
        case 11: if (i > 0) oi = i - 1, oe = 10; break;
        default: break;  // 5, 6, 10: always created by the cell itself
    }
    const int64_t oc = ((int64_t)oi * d.cy + oj) * d.cz + ok;
    const uint32_t omask = (uint32_t)c_edge_table[cases[oc]] & created_mask(oi, oj, ok);
    return vert_offset[oc] + __popc(omask & before_mask(oe));
}

// ---- phase 4: triangle emission -------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
mc_faces_kernel(Dims d, const uint8_t* __restrict__ cases, const uint2* __restrict__ block_offsets,
                const uint32_t* __restrict__ vert_offset, int32_t* __restrict__ faces, int64_t cap_f) {
    __shared__ uint32_t ws[kThreads / 32];
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    uint32_t cube = 0;
    if (c < d.n_cells) cube = cases[c];
    const uint32_t nt = c_tri_count[cube];
    uint32_t total;
    const uint32_t off = block_offsets[blockIdx.x].y + block_exclusive_scan(nt, ws, total);
    if (c >= d.n_cells || nt == 0) return;
    int i, j, k;
    cell_coords(d, c, i, j, k);
    const uint32_t edges = c_edge_table[cube];
    uint32_t idx[12];
#pragma unroll
    for (int e = 0; e < 12; ++e) idx[e] = ((edges >> e) & 1u) ? vertex_id(d, cases, vert_offset, i, j, k, e) : 0u;
    for (uint32_t t = 0; t < nt; ++t) {
        if ((int64_t)(off + t) >= cap_f) break;
        int32_t* o = faces + (int64_t)(off + t) * 3;
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            const int e = c_tri_table[cube * 16 + t * 3 + m];
            uint32_t v = 0;
#pragma unroll
            for (int q = 0; q < 12; ++q)
                if (q == e) v = idx[q];
            o[m] = (int32_t)v;
        }
    }
}

inline int64_t n_cta(const Dims& d) { return (d.n_cells + kThreads - 1) / kThreads; }

struct Scratch {
    uint8_t* cases;
    uint32_t* vert_offset;
    uint2* block_sums;
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

inline size_t scratch_bytes(const Dims& d) {
    return align_up((size_t)d.n_cells, 256) + align_up((size_t)d.n_cells * 4, 256) + align_up((size_t)n_cta(d) * 8, 256);
}

inline Scratch carve(void* base, const Dims& d) {
    Scratch s;
    uint8_t* p = (uint8_t*)base;
    s.cases = p;
    p += align_up((size_t)d.n_cells, 256);
    s.vert_offset = (uint32_t*)p;
    p += align_up((size_t)d.n_cells * 4, 256);
    s.block_sums = (uint2*)p;
    return s;
}

int check_dims(int nx, int ny, int nz, const char* who) {
    NA_REQUIRE(nx >= 2 && ny >= 2 && nz >= 2, "%s: volume must be at least 2x2x2 samples (got %dx%dx%d)", who, nx, ny,
               nz);
    NA_REQUIRE((int64_t)(nx - 1) * (ny - 1) * (nz - 1) <= 0x7fffffffLL / 6, "%s: volume too large (vertex ids are 32-bit)",
               who);
    return NA_OK;
}

}  // namespace

extern "C" int64_t na_mc_scratch_bytes(int nx, int ny, int nz) {
    if (nx < 2 || ny < 2 || nz < 2) return 0;
    return (int64_t)scratch_bytes(make_dims(nx, ny, nz));
}

extern "C" int na_mc_count(const float* volume, int nx, int ny, int nz, double iso, void* scratch,
                           int64_t scratch_size, uint32_t* counts, void* stream) {
    na::clear_error();
    if (int rc = check_dims(nx, ny, nz, "mc_count")) return rc;
    NA_REQUIRE(volume && scratch && counts, "mc_count: null pointer");
    const Dims d = make_dims(nx, ny, nz);
    if (scratch_size < (int64_t)scratch_bytes(d)) {
        na::set_error("mc_count: scratch has %lld bytes, %lld needed", (long long)scratch_size,
                      (long long)scratch_bytes(d));
        return NA_ERR_CAPACITY;
    }
    NA_REQUIRE((uintptr_t)scratch % 256 == 0, "mc_count: scratch must be 256-byte aligned");
    const Scratch s = carve(scratch, d);
    cudaStream_t st = (cudaStream_t)stream;
    mc_classify_kernel<<<(unsigned)n_cta(d), kThreads, 0, st>>>(volume, d, iso, s.cases, s.block_sums);
    NA_CHECK_LAUNCH("mc_classify_kernel");
    mc_scan_blocks_kernel<<<1, kScanThreads, 0, st>>>(s.block_sums, n_cta(d), counts);
    NA_CHECK_LAUNCH("mc_scan_blocks_kernel");
    return NA_OK;
}

extern "C" int na_mc_emit(const float* volume, int nx, int ny, int nz, double iso, void* scratch,
                          int64_t scratch_size, double scale, const double* origin, double* verts, int64_t cap_verts,
                          int32_t* faces, int64_t cap_faces, void* stream) {
    na::clear_error();
    if (int rc = check_dims(nx, ny, nz, "mc_emit")) return rc;
    NA_REQUIRE(volume && scratch, "mc_emit: null pointer");
    NA_REQUIRE(cap_verts >= 0 && cap_faces >= 0 && (cap_verts == 0 || verts) && (cap_faces == 0 || faces),
               "mc_emit: null output with non-zero capacity");
    const Dims d = make_dims(nx, ny, nz);
    if (scratch_size < (int64_t)scratch_bytes(d)) {
        na::set_error("mc_emit: scratch has %lld bytes, %lld needed", (long long)scratch_size,
                      (long long)scratch_bytes(d));
        return NA_ERR_CAPACITY;
    }
    const Scratch s = carve(scratch, d);
    Xform xf;
    xf.scale = scale;
    for (int a = 0; a < 3; ++a) xf.origin[a] = origin ? origin[a] : 0.0;
    cudaStream_t st = (cudaStream_t)stream;
    mc_vertices_kernel<<<(unsigned)n_cta(d), kThreads, 0, st>>>(volume, d, iso, s.cases, s.block_sums, s.vert_offset,
                                                               verts, cap_verts, xf);
    NA_CHECK_LAUNCH("mc_vertices_kernel");
    mc_faces_kernel<<<(unsigned)n_cta(d), kThreads, 0, st>>>(d, s.cases, s.block_sums, s.vert_offset, faces, cap_faces);
    NA_CHECK_LAUNCH("mc_faces_kernel");
    return NA_OK;
}

// ---------------------------------------------------------------------------------------------------
// na_mc_run_static: marching cubes for callers that are not allowed to allocate device memory.
//
// Who needs this: code that runs where cudaMalloc is off limits -- a real-time or serving loop whose
// memory budget was fixed at start-up, a plugin called from inside somebody else's allocator or memory
// pool, a host thread that must not take the driver's allocation lock (cudaMalloc synchronises the
// whole device), or a process whose remaining device memory is already spoken for.  The regular pair
// na_mc_count / na_mc_emit asks such a caller for a scratch buffer it has no way to obtain.  Here the
// scratch lives inside the library image instead: a fixed array in device global memory that every GPU
// loading libna_b200.so reserves (about 1.3 MB per GPU) from load until the process exits.
//
// Limit: at most kStaticMaxCells = 64^3 cells per call (e.g. a 65^3-sample block, block_res = 64, the
// default of the reference's extract_mesh signature).  Larger volumes are rejected with NA_ERR_CAPACITY.
//
// Sharing: there is ONE array per GPU and every call on that GPU uses it.  It still holds whatever the
// previous call left behind; that is harmless because each call writes every scratch word (case byte,
// vertex offset, per-CTA sum) before it reads it, so nothing has to start at zero and nothing is
// cleared.  Two calls whose kernels can overlap on the same GPU (different streams, different host
// threads) would overwrite each other's scratch: callers must serialise them (same stream, or an event).
//
// Set-up: the array's name is only a host-side handle, not a device address; the address is looked up
// with cudaGetSymbolAddress the first time the entry point runs on a given device (that device being
// current) and remembered per device.  Make that first call outside of any stream capture.
// ---------------------------------------------------------------------------------------------------
#include <mutex>

namespace {

constexpr int64_t kStaticMaxCells = 64LL * 64 * 64;
constexpr size_t kStaticScratchBytes =
    ((size_t)kStaticMaxCells + 255) / 256 * 256 +                       // case index, 1 B per cell
    ((size_t)kStaticMaxCells * 4 + 255) / 256 * 256 +                   // vertex offset, 4 B per cell
    (((size_t)kStaticMaxCells + kThreads - 1) / kThreads * 8 + 255) / 256 * 256;  // (vertex, triangle) sums per CTA

static __device__ __align__(256) unsigned char g_mc_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
void* g_static_scratch_addr[kMaxDevices] = {nullptr};

// Device address of the static scratch on the CURRENT device (looked up once per device).
int static_scratch_address(void** out) {
    int dev = -1;
    if (int rc = na::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return rc;
    NA_REQUIRE(dev >= 0 && dev < kMaxDevices, "mc_run_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        if (int rc = na::check_cuda(cudaGetSymbolAddress(&p, g_mc_static_scratch), "cudaGetSymbolAddress")) return rc;
        NA_REQUIRE(p != nullptr, "mc_run_static: static scratch has no device address");
        g_static_scratch_addr[dev] = p;
    }
    *out = g_static_scratch_addr[dev];
    return NA_OK;
}

}  // namespace

extern "C" int64_t na_mc_static_max_cells(void) { return kStaticMaxCells; }

extern "C" int na_mc_run_static(const float* volume, int nx, int ny, int nz, double iso, double scale,
                                const double* origin, double* verts, int64_t cap_verts, int32_t* faces,
                                int64_t cap_faces, uint32_t* counts, void* stream) {
    na::clear_error();
    if (int rc = check_dims(nx, ny, nz, "mc_run_static")) return rc;
    NA_REQUIRE(volume && counts, "mc_run_static: null pointer");
    NA_REQUIRE(cap_verts >= 0 && cap_faces >= 0 && (cap_verts == 0 || verts) && (cap_faces == 0 || faces),
               "mc_run_static: null output with non-zero capacity");
    const Dims d = make_dims(nx, ny, nz);
    if (d.n_cells > kStaticMaxCells) {
        na::set_error("mc_run_static: %dx%dx%d samples = %lld cells exceed the built-in scratch (max %lld cells, "
                      "e.g. 65^3 samples); use na_mc_count/na_mc_emit with caller-owned scratch",
                      nx, ny, nz, (long long)d.n_cells, (long long)kStaticMaxCells);
        return NA_ERR_CAPACITY;
    }
    void* base = nullptr;
    if (int rc = static_scratch_address(&base)) return rc;
    const Scratch s = carve(base, d);
    Xform xf;
    xf.scale = scale;
    for (int a = 0; a < 3; ++a) xf.origin[a] = origin ? origin[a] : 0.0;
    cudaStream_t st = (cudaStream_t)stream;
    const unsigned ctas = (unsigned)n_cta(d);
    mc_classify_kernel<<<ctas, kThreads, 0, st>>>(volume, d, iso, s.cases, s.block_sums);
    NA_CHECK_LAUNCH("mc_classify_kernel");
    mc_scan_blocks_kernel<<<1, kScanThreads, 0, st>>>(s.block_sums, n_cta(d), counts);
    NA_CHECK_LAUNCH("mc_scan_blocks_kernel");
    mc_vertices_kernel<<<ctas, kThreads, 0, st>>>(volume, d, iso, s.cases, s.block_sums, s.vert_offset, verts,
                                                  cap_verts, xf);
    NA_CHECK_LAUNCH("mc_vertices_kernel");
    mc_faces_kernel<<<ctas, kThreads, 0, st>>>(d, s.cases, s.block_sums, s.vert_offset, faces, cap_faces);
    NA_CHECK_LAUNCH("mc_faces_kernel");
    return NA_OK;
}
