// (a3, a4, a5) batched strided block copies: the data movement of permute_legs
// (abelian.cpp:988-1006), combine_legs (abelian.cpp:1196-1217) and split_legs
// (abelian.cpp:3418-3433), for ALL blocks of a tensor in one launch.
//
// A job is a rank-r box copied between two strided views.  A tiny prep kernel canonicalises every
// job (drops extent-1 axes, orders axes by destination stride, merges axes contiguous on both
// sides) and classifies it:
//   mode 0  "rows"      source and destination share the unit-stride axis -> coalesced row copies
//   mode 1  "transpose" unit-stride axes differ -> 32x32 tiles staged through shared memory so
//                        that both the global reads and the global writes are coalesced
//   mode 2  "generic"   no unit-stride axis on one side (extent-1 slices): element gather
// Work is cut into units of <= 4096 elements; a persistent grid (multiple of the SM count) walks a
// contiguous range of units each.  HBM-bound: algorithmic bytes = 2 * sum(block bytes).
#include "common.cuh"

namespace cyb {

constexpr int CP_THREADS = 256;
constexpr int CP_TILE = 32;
constexpr int CP_CTAS_PER_SM = 6;   // resident CTAs per SM

struct CanonJob {
    int64_t src_off, dst_off;
    int64_t sstride[CYB_MAX_RANK];
    int64_t dstride[CYB_MAX_RANK];
    uint32_t shape[CYB_MAX_RANK];
    int32_t rank;      // canonical rank (>= 1)
    int32_t mode;      // 0 rows, 1 transpose, 2 generic
    int32_t sA;        // mode 1: axis with unit source stride
    int32_t tx;        // mode 0: threads along the row (power of two <= 256)
    uint32_t rows;     // mode 0: product of outer extents; mode 1: product of the other extents
    uint32_t n_a, n_b; // mode 0: (row groups, col chunks); mode 1: (tiles along sA, tiles along last)
    uint32_t pad_;
    uint64_t total;    // number of elements
};

// Canonical form + number of work units of one job (shared by the device prep kernel and the host-side
// workspace sizing, so both agree exactly).
constexpr int CP_UNIT_ELEMS = 2048;   // elements per mode-0 unit (8 per thread)

__host__ __device__ inline int64_t canon_job(const cyb_copy_job& J, CanonJob& C)
{
    int64_t shp[CYB_MAX_RANK], ss[CYB_MAX_RANK], ds[CYB_MAX_RANK];
    int r = 0;
    uint64_t total = 1;
    for (int a = 0; a < (int)J.rank; ++a) {
        total *= (uint64_t)J.shape[a];
        if (J.shape[a] != 1) { shp[r] = J.shape[a]; ss[r] = J.src_stride[a]; ds[r] = J.dst_stride[a]; ++r; }
    }
    C.src_off = J.src_off; C.dst_off = J.dst_off; C.total = total; C.pad_ = 0;
    if (total == 0) {
        C.rank = 1; C.mode = 0; C.sA = 0; C.tx = 1; C.rows = 0; C.n_a = 0; C.n_b = 0;
        for (int a = 0; a < CYB_MAX_RANK; ++a) { C.shape[a] = 1; C.sstride[a] = 0; C.dstride[a] = 0; }
        return 0;
    }
    // order by destination stride, largest first (insertion sort, r <= 8)
    for (int a = 1; a < r; ++a) {
        int64_t s0 = shp[a], s1 = ss[a], s2 = ds[a];
        int b = a - 1;
        while (b >= 0 && ds[b] < s2) { shp[b + 1] = shp[b]; ss[b + 1] = ss[b]; ds[b + 1] = ds[b]; --b; }
        shp[b + 1] = s0; ss[b + 1] = s1; ds[b + 1] = s2;
    }
    // merge neighbours that are contiguous in both views
    int m = 0;
    for (int a = 0; a < r; ++a) {
        if (m > 0 && ss[m - 1] == ss[a] * shp[a] && ds[m - 1] == ds[a] * shp[a]) {
            shp[m - 1] *= shp[a]; ss[m - 1] = ss[a]; ds[m - 1] = ds[a];
        } else { shp[m] = shp[a]; ss[m] = ss[a]; ds[m] = ds[a]; ++m; }
    }
    if (m == 0) { shp[0] = 1; ss[0] = 1; ds[0] = 1; m = 1; }
    r = m;
    for (int a = 0; a < CYB_MAX_RANK; ++a) {
        C.shape[a] = a < r ? (uint32_t)shp[a] : 1u;
        C.sstride[a] = a < r ? ss[a] : 0;
        C.dstride[a] = a < r ? ds[a] : 0;
    }
    C.rank = r;
    const int last = r - 1;
    int sA = -1;
    for (int a = 0; a < r; ++a) if (ss[a] == 1) sA = a;
    int64_t nunits;
    if (ds[last] == 1 && ss[last] == 1) {
        // rows: a unit is <= CP_UNIT_ELEMS elements = a chunk of one row, or several consecutive rows along
        // the second-innermost axis.  rows = rows per unit, n_a = units along that axis, n_b = chunks per row
        C.mode = 0; C.sA = last; C.tx = 1;
        const uint32_t L = (uint32_t)shp[last];
        const uint32_t E2 = r >= 2 ? (uint32_t)shp[r - 2] : 1u;
        uint32_t rpu = L >= (uint32_t)CP_UNIT_ELEMS ? 1u : (uint32_t)CP_UNIT_ELEMS / L;
        if (rpu > E2) rpu = E2;
        C.rows = rpu;
        C.n_a = (E2 + rpu - 1) / rpu;
        C.n_b = (L + CP_UNIT_ELEMS - 1) / CP_UNIT_ELEMS;
        nunits = (int64_t)(total / ((uint64_t)L * E2)) * C.n_a * C.n_b;
    } else if (ds[last] == 1 && sA >= 0) {
        C.mode = 1; C.sA = sA; C.tx = 32;
        const uint32_t S = (uint32_t)shp[sA], D = (uint32_t)shp[last];
        C.rows = (uint32_t)(total / ((uint64_t)S * D));
        C.n_a = (S + CP_TILE - 1) / CP_TILE;
        C.n_b = (D + CP_TILE - 1) / CP_TILE;
        nunits = (int64_t)C.rows * C.n_a * C.n_b;
    } else {
        C.mode = 2; C.sA = 0; C.tx = 1; C.rows = 0; C.n_a = 0; C.n_b = 0;
        nunits = (int64_t)((total + 1023) / 1024);
    }
    return nunits;
}

__global__ void canon_jobs_kernel(const cyb_copy_job* __restrict__ jobs, int64_t njobs,
                                  CanonJob* __restrict__ out, int64_t* __restrict__ units)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= njobs) return;
    CanonJob C;
    units[j] = canon_job(jobs[j], C);
    out[j] = C;
}

// exclusive scan over njobs unit counts -> unit_start[njobs + 1]; single CTA
__global__ void __launch_bounds__(1024) scan_units_kernel(const int64_t* __restrict__ units,
                                                          int64_t njobs,
                                                          int64_t* __restrict__ unit_start)
{
    __shared__ int64_t s_scan[33];
    int64_t carry = 0;
    for (int64_t base = 0; base < njobs; base += 1024) {
        const int64_t j = base + threadIdx.x;
        int64_t v = j < njobs ? units[j] : 0;
        int64_t total;
        int64_t pos = block_exclusive_scan<int64_t>(v, s_scan, &total);
        if (j < njobs) unit_start[j] = carry + pos;
        carry += total;
    }
    if (threadIdx.x == 0) unit_start[njobs] = carry;
}

template <class T, bool CONJ> __device__ __forceinline__ T maybe_conj(T v) { return v; }
template <> __device__ __forceinline__ double2 maybe_conj<double2, true>(double2 v) { v.y = -v.y; return v; }

// decompose `idx` over the axes of C except `skip0` and `skip1` (row-major over the canonical
// order) and return the source / destination element offsets
__device__ __forceinline__ void outer_offsets(const CanonJob& C, uint32_t idx, int skip0, int skip1,
                                              int64_t& so, int64_t& d_o)
{
    so = 0; d_o = 0;
    for (int a = C.rank - 1; a >= 0; --a) {
        if (a == skip0 || a == skip1) continue;
        const uint32_t e = C.shape[a];
        const uint32_t q = idx / e, rem = idx - q * e;
        so += (int64_t)rem * C.sstride[a];
        d_o += (int64_t)rem * C.dstride[a];
        idx = q;
    }
}

// One work unit, fully resolved at plan time (no index arithmetic left for the copy kernel):
//   mode 0: n0 rows x n1 contiguous elements, row strides s1 (source) / d1 (destination)
//   mode 1: a (n0 x n1) tile, n0 along the source-fast axis, n1 along the destination-fast axis;
//           s1 = source stride of the destination-fast axis, d1 = destination stride of the source-fast axis
//   mode 2: 1024 elements of job `job` starting at element `src` (gather through the CanonJob)
struct CopyUnit {
    int64_t src, dst;
    int64_t s1, d1;
    int32_t n0, n1;
    int32_t mode, job;
};

__global__ void fill_units_kernel(const CanonJob* __restrict__ cjobs, const int64_t* __restrict__ unit_start,
                                  int64_t njobs, CopyUnit* __restrict__ units)
{
    const int64_t total = unit_start[njobs];
    for (int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; u < total; u += (int64_t)gridDim.x * blockDim.x) {
        int64_t lo = 0, hi = njobs;
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (unit_start[mid] <= u) lo = mid; else hi = mid;
        }
        const CanonJob& C = cjobs[lo];
        uint32_t lu = (uint32_t)(u - unit_start[lo]);
        CopyUnit U;
        U.mode = C.mode; U.job = (int32_t)lo;
        const int last = C.rank - 1;
        if (C.mode == 0) {
            const uint32_t L = C.shape[last];
            const uint32_t E2 = C.rank >= 2 ? C.shape[C.rank - 2] : 1u;
            const uint32_t cb = lu % C.n_b; lu /= C.n_b;
            const uint32_t ra = lu % C.n_a; const uint32_t outer = lu / C.n_a;
            int64_t so, d_o;
            outer_offsets(C, outer, last, C.rank >= 2 ? C.rank - 2 : -1, so, d_o);
            const uint32_t row0 = ra * C.rows, col0 = cb * (uint32_t)CP_UNIT_ELEMS;
            U.s1 = C.rank >= 2 ? C.sstride[C.rank - 2] : 0;
            U.d1 = C.rank >= 2 ? C.dstride[C.rank - 2] : 0;
            U.src = C.src_off + so + (int64_t)row0 * U.s1 + col0;
            U.dst = C.dst_off + d_o + (int64_t)row0 * U.d1 + col0;
            U.n0 = (int32_t)min(C.rows, E2 - row0);
            U.n1 = (int32_t)min((uint32_t)CP_UNIT_ELEMS, L - col0);
        } else if (C.mode == 1) {
            const int sA = C.sA;
            const uint32_t S = C.shape[sA], D = C.shape[last];
            const uint32_t tiles = C.n_a * C.n_b;
            const uint32_t o = lu / tiles, t = lu - o * tiles;
            const uint32_t ta = t / C.n_b, tb = t - ta * C.n_b;
            int64_t so, d_o;
            outer_offsets(C, o, sA, last, so, d_o);
            const uint32_t s0 = ta * CP_TILE, d0 = tb * CP_TILE;
            U.s1 = C.sstride[last];
            U.d1 = C.dstride[sA];
            U.src = C.src_off + so + (int64_t)d0 * U.s1 + s0;
            U.dst = C.dst_off + d_o + (int64_t)s0 * U.d1 + d0;
            U.n0 = (int32_t)min((uint32_t)CP_TILE, S - s0);
            U.n1 = (int32_t)min((uint32_t)CP_TILE, D - d0);
        } else {
            U.src = (int64_t)lu * 1024; U.dst = 0; U.s1 = U.d1 = 0; U.n0 = U.n1 = 0;
        }
        units[u] = U;
    }
}

template <class T, bool CONJ>
__global__ void __launch_bounds__(CP_THREADS, CP_CTAS_PER_SM) copy_blocks_kernel(
    const T* __restrict__ src, T* __restrict__ dst, const CanonJob* __restrict__ cjobs,
    const CopyUnit* __restrict__ units, const int64_t* __restrict__ unit_total)
{
    __shared__ T tile[CP_TILE][CP_TILE + 1];
    const int64_t total_units = *unit_total;
    const int tid = threadIdx.x;
    for (int64_t u = blockIdx.x; u < total_units; u += gridDim.x) {
        const CopyUnit U = units[u];
        if (U.mode == 0) {
            const T* __restrict__ sp = src + U.src;
            T* __restrict__ dp = dst + U.dst;
            const int n = U.n0 * U.n1;
            T v[CP_UNIT_ELEMS / CP_THREADS];
            if (U.n0 == 1) {
#pragma unroll
                for (int k = 0; k < CP_UNIT_ELEMS / CP_THREADS; ++k) {
                    const int e = tid + k * CP_THREADS;
                    if (e < n) v[k] = __ldcs(sp + e);
                }
#pragma unroll
                for (int k = 0; k < CP_UNIT_ELEMS / CP_THREADS; ++k) {
                    const int e = tid + k * CP_THREADS;
                    if (e < n) __stcs(dp + e, maybe_conj<T, CONJ>(v[k]));
                }
            } else {
#pragma unroll
                for (int k = 0; k < CP_UNIT_ELEMS / CP_THREADS; ++k) {
                    const int e = tid + k * CP_THREADS;
                    if (e < n) { const int r = e / U.n1, c = e - r * U.n1; v[k] = __ldcs(sp + (int64_t)r * U.s1 + c); }
                }
#pragma unroll
                for (int k = 0; k < CP_UNIT_ELEMS / CP_THREADS; ++k) {
                    const int e = tid + k * CP_THREADS;
                    if (e < n) { const int r = e / U.n1, c = e - r * U.n1; __stcs(dp + (int64_t)r * U.d1 + c, maybe_conj<T, CONJ>(v[k])); }
                }
            }
        } else if (U.mode == 1) {
            const T* __restrict__ sp = src + U.src;
            T* __restrict__ dp = dst + U.dst;
            const int tx = tid & 31, ty = tid >> 5;  // 32 x 8
            __syncthreads();
#pragma unroll
            for (int k = 0; k < CP_TILE; k += 8) {
                const int di = ty + k;
                if (di < U.n1 && tx < U.n0) tile[di][tx] = __ldcs(sp + (int64_t)di * U.s1 + tx);
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < CP_TILE; k += 8) {
                const int si = ty + k;
                if (si < U.n0 && tx < U.n1) __stcs(dp + (int64_t)si * U.d1 + tx, maybe_conj<T, CONJ>(tile[tx][si]));
            }
        } else {
            const CanonJob& C = cjobs[U.job];
            const T* __restrict__ sp = src + C.src_off;
            T* __restrict__ dp = dst + C.dst_off;
            const uint64_t e0 = (uint64_t)U.src;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint64_t e = e0 + tid + (uint64_t)k * CP_THREADS;
                if (e < C.total) {
                    int64_t so, d_o;
                    outer_offsets(C, (uint32_t)e, -1, -1, so, d_o);
                    dp[d_o] = maybe_conj<T, CONJ>(sp[so]);
                }
            }
        }
    }
}

}  // namespace cyb

using namespace cyb;

extern "C" {

// workspace layout: CanonJob[njobs] | units-per-job[njobs] | unit_start[njobs + 1] | CopyUnit[total units]
static size_t ws_tables_bytes(int64_t njobs)
{
    return align_up(sizeof(CanonJob) * (size_t)njobs, 256) + align_up(sizeof(int64_t) * (2 * (size_t)njobs + 2), 256);
}

int64_t cyb_copy_blocks_workspace(const cyb_copy_job* jobs_host, int64_t njobs)
{
    if (njobs <= 0) return 0;
    int64_t units = 0;
    CanonJob C;
    for (int64_t j = 0; j < njobs; ++j) units += canon_job(jobs_host[j], C);
    return (int64_t)(ws_tables_bytes(njobs) + sizeof(CopyUnit) * (size_t)(units + 1));
}

int cyb_copy_blocks_prepare(const cyb_copy_job* jobs, int64_t njobs, void* workspace,
                            int64_t workspace_bytes, cyb_stream_t stream_)
{
    CYB_REQUIRE(njobs >= 0, "copy_blocks_prepare: njobs < 0");
    if (njobs == 0) return CYB_OK;
    CYB_REQUIRE(workspace && workspace_bytes >= (int64_t)(ws_tables_bytes(njobs) + sizeof(CopyUnit)), "copy_blocks_prepare: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    char* ws = static_cast<char*>(workspace);
    CanonJob* cj = reinterpret_cast<CanonJob*>(ws);
    int64_t* units = reinterpret_cast<int64_t*>(ws + align_up(sizeof(CanonJob) * (size_t)njobs, 256));
    int64_t* unit_start = units + njobs;
    CopyUnit* cu = reinterpret_cast<CopyUnit*>(ws + ws_tables_bytes(njobs));
    canon_jobs_kernel<<<(unsigned)((njobs + 255) / 256), 256, 0, stream>>>(jobs, njobs, cj, units);
    CYB_LAUNCH_CHECK();
    scan_units_kernel<<<1, 1024, 0, stream>>>(units, njobs, unit_start);
    CYB_LAUNCH_CHECK();
    // the caller sized the workspace with cyb_copy_blocks_workspace(host copy of the same jobs)
    fill_units_kernel<<<sm_count() * 4, 256, 0, stream>>>(cj, unit_start, njobs, cu);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_copy_blocks_run(const void* src, void* dst, const void* workspace, int64_t njobs, int dtype,
                        int conj, cyb_stream_t stream_)
{
    CYB_REQUIRE(njobs >= 0, "copy_blocks_run: njobs < 0");
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "copy_blocks_run: bad dtype %d", dtype);
    if (njobs == 0) return CYB_OK;
    CYB_REQUIRE(workspace != nullptr, "copy_blocks_run: no prepared workspace");
    cudaStream_t stream = (cudaStream_t)stream_;
    const char* ws = static_cast<const char*>(workspace);
    const CanonJob* cj = reinterpret_cast<const CanonJob*>(ws);
    const int64_t* unit_total = reinterpret_cast<const int64_t*>(ws + align_up(sizeof(CanonJob) * (size_t)njobs, 256)) + 2 * njobs;
    const CopyUnit* cu = reinterpret_cast<const CopyUnit*>(ws + ws_tables_bytes(njobs));
    const int grid = sm_count() * CP_CTAS_PER_SM;
    if (dtype == CYB_F64) {
        copy_blocks_kernel<double, false><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double*>(src), static_cast<double*>(dst), cj, cu, unit_total);
    } else if (conj) {
        copy_blocks_kernel<double2, true><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double2*>(src), static_cast<double2*>(dst), cj, cu, unit_total);
    } else {
        copy_blocks_kernel<double2, false><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double2*>(src), static_cast<double2*>(dst), cj, cu, unit_total);
    }
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

/* One-shot form: `jobs` is a DEVICE array and `jobs_host` its host copy (used to size the schedule). */
int cyb_copy_blocks(const void* src, void* dst, const cyb_copy_job* jobs, const cyb_copy_job* jobs_host,
                    int64_t njobs, int dtype, int conj, cyb_stream_t stream_)
{
    CYB_REQUIRE(njobs >= 0, "copy_blocks: njobs < 0");
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "copy_blocks: bad dtype %d", dtype);
    if (njobs == 0) return CYB_OK;
    CYB_REQUIRE(jobs_host != nullptr, "copy_blocks: need the host copy of the jobs");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t need = cyb_copy_blocks_workspace(jobs_host, njobs);
    Scratch own((size_t)need, stream);
    if (!own.ptr) return CYB_ERR_CUDA;
    const int rc = cyb_copy_blocks_prepare(jobs, njobs, own.ptr, need, stream_);
    if (rc != CYB_OK) return rc;
    return cyb_copy_blocks_run(src, dst, own.ptr, njobs, dtype, conj, stream_);
}

}  // extern "C"
