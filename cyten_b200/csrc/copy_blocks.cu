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

__global__ void canon_jobs_kernel(const cyb_copy_job* __restrict__ jobs, int64_t njobs,
                                  CanonJob* __restrict__ out, int64_t* __restrict__ units)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= njobs) return;
    const cyb_copy_job J = jobs[j];
    int64_t shp[CYB_MAX_RANK], ss[CYB_MAX_RANK], ds[CYB_MAX_RANK];
    int r = 0;
    uint64_t total = 1;
    for (int a = 0; a < (int)J.rank; ++a) {
        total *= (uint64_t)J.shape[a];
        if (J.shape[a] != 1) { shp[r] = J.shape[a]; ss[r] = J.src_stride[a]; ds[r] = J.dst_stride[a]; ++r; }
    }
    CanonJob C;
    C.src_off = J.src_off; C.dst_off = J.dst_off; C.total = total; C.pad_ = 0;
    if (total == 0) {
        C.rank = 1; C.mode = 0; C.sA = 0; C.tx = 1; C.rows = 0; C.n_a = 0; C.n_b = 0;
        for (int a = 0; a < CYB_MAX_RANK; ++a) { C.shape[a] = 1; C.sstride[a] = 0; C.dstride[a] = 0; }
        out[j] = C; units[j] = 0;
        return;
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
        C.mode = 0; C.sA = last;
        const uint32_t L = (uint32_t)shp[last];
        int tx = 1;
        while (tx < 256 && (uint32_t)tx < L) tx <<= 1;
        C.tx = tx;
        const uint32_t ty = CP_THREADS / tx;
        uint64_t rows = total / L;
        C.rows = (uint32_t)rows;
        C.n_a = (uint32_t)((rows + 4ull * ty - 1) / (4ull * ty));
        C.n_b = (L + 4u * tx - 1) / (4u * tx);
        nunits = (int64_t)C.n_a * C.n_b;
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
    out[j] = C;
    units[j] = nunits;
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

template <class T, bool CONJ>
__global__ void __launch_bounds__(CP_THREADS) copy_blocks_kernel(
    const T* __restrict__ src, T* __restrict__ dst, const CanonJob* __restrict__ cjobs,
    const int64_t* __restrict__ unit_start, int64_t njobs)
{
    __shared__ CanonJob sC;
    __shared__ T tile[CP_TILE][CP_TILE + 1];
    const int64_t total_units = unit_start[njobs];
    const int64_t per = (total_units + gridDim.x - 1) / gridDim.x;
    int64_t u = (int64_t)blockIdx.x * per;
    const int64_t u_end = min(total_units, u + per);
    if (u >= u_end) return;
    // locate the job of the first unit: largest j with unit_start[j] <= u
    int64_t lo = 0, hi = njobs;
    while (hi - lo > 1) {
        int64_t mid = (lo + hi) >> 1;
        if (unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    int64_t job = lo;
    int64_t loaded = -1;
    const int tid = threadIdx.x;
    for (; u < u_end; ++u) {
        while (unit_start[job + 1] <= u) ++job;
        if (job != loaded) {
            __syncthreads();
            const int nwords = sizeof(CanonJob) / 4;
            const uint32_t* g = reinterpret_cast<const uint32_t*>(&cjobs[job]);
            uint32_t* s = reinterpret_cast<uint32_t*>(&sC);
            for (int w = tid; w < nwords; w += CP_THREADS) s[w] = g[w];
            loaded = job;
            __syncthreads();
        }
        const uint32_t lu = (uint32_t)(u - unit_start[job]);
        const T* __restrict__ sp = src + sC.src_off;
        T* __restrict__ dp = dst + sC.dst_off;
        if (sC.mode == 0) {
            const int last = sC.rank - 1;
            const uint32_t L = sC.shape[last];
            const uint32_t txn = (uint32_t)sC.tx, tyn = CP_THREADS / txn;
            const uint32_t tx = tid & (txn - 1), ty = tid / txn;
            const uint32_t ua = lu / sC.n_b, ub = lu - ua * sC.n_b;
            const uint32_t row0 = ua * 4u * tyn, col0 = ub * 4u * txn;
            int64_t so[4], d_o[4];
            bool rok[4];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr) {
                const uint32_t row = row0 + ty + rr * tyn;
                rok[rr] = row < sC.rows;
                so[rr] = 0; d_o[rr] = 0;
                if (rok[rr]) outer_offsets(sC, row, last, -1, so[rr], d_o[rr]);
            }
            T v[4][4];
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    const uint32_t col = col0 + tx + cc * txn;
                    if (rok[rr] && col < L) v[rr][cc] = sp[so[rr] + col];
                }
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    const uint32_t col = col0 + tx + cc * txn;
                    if (rok[rr] && col < L) dp[d_o[rr] + col] = maybe_conj<T, CONJ>(v[rr][cc]);
                }
        } else if (sC.mode == 1) {
            const int last = sC.rank - 1, sA = sC.sA;
            const uint32_t S = sC.shape[sA], D = sC.shape[last];
            const uint32_t tiles = sC.n_a * sC.n_b;
            const uint32_t o = lu / tiles, t = lu - o * tiles;
            const uint32_t ta = t / sC.n_b, tb = t - ta * sC.n_b;
            int64_t so, d_o;
            outer_offsets(sC, o, sA, last, so, d_o);
            const uint32_t s0 = ta * CP_TILE, d0 = tb * CP_TILE;
            const uint32_t tx = tid & 31, ty = tid >> 5;  // 32 x 8
            const int64_t s_last = sC.sstride[last];      // source stride of the destination-fast axis
            const int64_t d_sA = sC.dstride[sA];          // destination stride of the source-fast axis
            __syncthreads();
#pragma unroll
            for (int k = 0; k < CP_TILE; k += 8) {
                const uint32_t di = d0 + ty + k, si = s0 + tx;
                if (di < D && si < S) tile[ty + k][tx] = sp[so + (int64_t)di * s_last + si];
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < CP_TILE; k += 8) {
                const uint32_t si = s0 + ty + k, di = d0 + tx;
                if (si < S && di < D) dp[d_o + (int64_t)si * d_sA + di] = maybe_conj<T, CONJ>(tile[tx][ty + k]);
            }
        } else {
            const uint64_t e0 = (uint64_t)lu * 1024u;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint64_t e = e0 + tid + (uint64_t)k * CP_THREADS;
                if (e < sC.total) {
                    int64_t so, d_o;
                    outer_offsets(sC, (uint32_t)e, -1, -1, so, d_o);
                    dp[d_o] = maybe_conj<T, CONJ>(sp[so]);
                }
            }
        }
    }
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int64_t cyb_copy_blocks_workspace(int64_t njobs)
{
    return (int64_t)(align_up(sizeof(CanonJob) * (size_t)njobs, 256) + sizeof(int64_t) * (2 * (size_t)njobs + 2));
}

int cyb_copy_blocks_prepare(const cyb_copy_job* jobs, int64_t njobs, void* workspace,
                            int64_t workspace_bytes, cyb_stream_t stream_)
{
    CYB_REQUIRE(njobs >= 0, "copy_blocks_prepare: njobs < 0");
    if (njobs == 0) return CYB_OK;
    CYB_REQUIRE(workspace && workspace_bytes >= cyb_copy_blocks_workspace(njobs), "copy_blocks_prepare: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    char* ws = static_cast<char*>(workspace);
    CanonJob* cj = reinterpret_cast<CanonJob*>(ws);
    int64_t* units = reinterpret_cast<int64_t*>(ws + align_up(sizeof(CanonJob) * (size_t)njobs, 256));
    int64_t* unit_start = units + njobs;
    canon_jobs_kernel<<<(unsigned)((njobs + 255) / 256), 256, 0, stream>>>(jobs, njobs, cj, units);
    CYB_LAUNCH_CHECK();
    scan_units_kernel<<<1, 1024, 0, stream>>>(units, njobs, unit_start);
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
    const int64_t* unit_start = reinterpret_cast<const int64_t*>(ws + align_up(sizeof(CanonJob) * (size_t)njobs, 256)) + njobs;
    const int grid = sm_count() * 4;
    if (dtype == CYB_F64) {
        copy_blocks_kernel<double, false><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double*>(src), static_cast<double*>(dst), cj, unit_start, njobs);
    } else if (conj) {
        copy_blocks_kernel<double2, true><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double2*>(src), static_cast<double2*>(dst), cj, unit_start, njobs);
    } else {
        copy_blocks_kernel<double2, false><<<grid, CP_THREADS, 0, stream>>>(
            static_cast<const double2*>(src), static_cast<double2*>(dst), cj, unit_start, njobs);
    }
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_copy_blocks(const void* src, void* dst, const cyb_copy_job* jobs, int64_t njobs, int dtype,
                    int conj, void* workspace, int64_t workspace_bytes, cyb_stream_t stream_)
{
    CYB_REQUIRE(njobs >= 0, "copy_blocks: njobs < 0");
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "copy_blocks: bad dtype %d", dtype);
    if (njobs == 0) return CYB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t need = cyb_copy_blocks_workspace(njobs);
    Scratch own(workspace ? 0 : (size_t)need, stream);
    void* ws = workspace;
    if (workspace) {
        CYB_REQUIRE(workspace_bytes >= need, "copy_blocks: workspace too small");
    } else {
        if (!own.ptr) return CYB_ERR_CUDA;
        ws = own.ptr;
    }
    const int rc = cyb_copy_blocks_prepare(jobs, njobs, ws, need, stream_);
    if (rc != CYB_OK) return rc;
    return cyb_copy_blocks_run(src, dst, ws, njobs, dtype, conj, stream_);
}

}  // extern "C"
