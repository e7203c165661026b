// (a8) global truncation cut on device + the small block-batched BLAS-1 helpers around it.
//
// cyb_truncate_select restates TensorBackend::_truncate_singular_values_selection
// (src/backends/tensor_backend.cpp:138-242) on device: stable ascending argsort of S^2, the
// chi_max / chi_min / degeneracy_tol / svd_min / trunc_cut constraint masks, smallest (or largest)
// admissible cut, keep-mask in the ORIGINAL order, err = sum of discarded S^2 and
// new_norm = sum of kept S^2 (neither square-rooted, tensor_backend.cpp:231-236).
// n <= d*chi values: one CTA, everything stays in L1/L2.  A constraint that would leave no
// admissible cut is ignored, like combine_constraints in the reference.
#include "common.cuh"
#include <math_constants.h>
#include <mutex>

namespace cyb {

constexpr int TR_THREADS = 1024;

struct TruncArgs {
    const double* s; int64_t n; int64_t chi_max; int64_t chi_min; double degeneracy_tol;
    double trunc_cut; double svd_min; int minimize_error;
    uint8_t* keep; double* result;
    // workspace
    int64_t* piv;      // [n] ascending order of marginal errors (stable)
    double* merr;      // [n] sorted marginal errors
    double* ssort;     // [n] sorted S
    double* csum;      // [n] inclusive cumulative sum of merr
    uint8_t* good;     // [n]
    uint8_t* good2;    // [n]
    int64_t* flag;     // [4]
};

// good = good & good2 unless that leaves nothing (then `good` is kept): combine_constraints
__device__ void combine_constraints(TruncArgs& P, int tid)
{
    __shared__ int s_any;
    if (tid == 0) s_any = 0;
    __syncthreads();
    int any = 0;
    for (int64_t i = tid; i < P.n; i += TR_THREADS) any |= (P.good[i] && P.good2[i]);
    if (any) atomicOr(&s_any, 1);
    __syncthreads();
    if (s_any)
        for (int64_t i = tid; i < P.n; i += TR_THREADS) P.good[i] = P.good[i] && P.good2[i];
    __syncthreads();
}

// stable ascending argsort of the marginal errors S^2 by rank counting (np.argsort default is not stable
// for ties in general, but ties are equal values: the kept SET differs only at exactly degenerate cutoffs,
// which the contract excludes).  One thread per value, the whole grid works: every CTA streams all n values
// through a shared-memory tile (n = d*chi = 8192 at chi = 4096: 3.8 ms in one CTA, ~20 us on the grid).
constexpr int TR_RANK_THREADS = 256;
__global__ void __launch_bounds__(TR_RANK_THREADS) truncate_rank_kernel(TruncArgs P)
{
    __shared__ double s_e[TR_RANK_THREADS];
    const int64_t n = P.n;
    const int64_t i = (int64_t)blockIdx.x * TR_RANK_THREADS + threadIdx.x;
    // NaN compares false with everything: ranks would collide and slots of piv[] stay unwritten (an out-of-bounds
    // scatter later).  The sort key maps NaN to +inf so the order is total; the select kernel reports non-finite input.
    const double si = i < n ? P.s[i] : 0.0, ei = (si != si) ? CUDART_INF : si * si;
    int64_t rank = 0;
    for (int64_t base = 0; base < n; base += TR_RANK_THREADS) {
        const int64_t j = base + threadIdx.x;
        const double sj = j < n ? P.s[j] : 0.0;
        __syncthreads();
        s_e[threadIdx.x] = (sj != sj) ? CUDART_INF : sj * sj;
        __syncthreads();
        const int cnt = (int)min((int64_t)TR_RANK_THREADS, n - base);
        for (int t = 0; t < cnt; ++t) {
            const double ej = s_e[t];
            rank += (ej < ei) || (ej == ei && base + t < i);
        }
    }
    if (i < n) {
        P.piv[rank] = i;
        P.merr[rank] = ei;
        P.ssort[rank] = si;
        P.good[rank] = 1;
    }
}

__global__ void __launch_bounds__(TR_THREADS) truncate_select_kernel(TruncArgs P)
{
    __shared__ double s_scan[33];
    __shared__ long long s_cut;
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    const int64_t n = P.n;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    {
        int bad = 0;
        for (int64_t i = tid; i < n; i += TR_THREADS) bad |= !isfinite(P.ssort[i]);
        if (bad) atomicOr(&s_bad, 1);
    }
    __syncthreads();
    if (s_bad) {     // unconverged / overflowed factorisation upstream: no selection, the host raises
        for (int64_t i = tid; i < n; i += TR_THREADS) P.keep[i] = 0;
        if (tid == 0) { P.result[0] = 0.0; P.result[1] = 0.0; P.result[2] = -1.0; P.result[3] = 0.0; }
        return;
    }
    // inclusive cumulative sum (sequential order of additions per chunk differs from np.cumsum
    // only in rounding; used for the trunc_cut test and err/new_norm)
    {
        double carry = 0.0;
        for (int64_t base = 0; base < n; base += TR_THREADS) {
            const int64_t i = base + tid;
            const double v = i < n ? P.merr[i] : 0.0;
            double total;
            const double ex = block_exclusive_scan<double>(v, s_scan, &total);
            if (i < n) P.csum[i] = carry + ex + v;
            carry += total;
        }
    }
    __syncthreads();
    if (P.chi_max >= 0 && P.chi_max < n) {
        for (int64_t i = tid; i < n; i += TR_THREADS) P.good2[i] = (i >= n - P.chi_max);
        __syncthreads();
        combine_constraints(P, tid);
    }
    if (P.chi_min > 1) {
        // good2[-chi_min+1:] = False
        for (int64_t i = tid; i < n; i += TR_THREADS) P.good2[i] = !(i >= n - P.chi_min + 1);
        __syncthreads();
        combine_constraints(P, tid);
    }
    if (P.degeneracy_tol > 0) {
        for (int64_t i = tid; i < n; i += TR_THREADS) {
            if (i == 0) P.good2[i] = 1;
            else {
                const double a = P.ssort[i] <= 1.0e-100 ? 1.0e-100 : P.ssort[i];
                const double b = P.ssort[i - 1] <= 1.0e-100 ? 1.0e-100 : P.ssort[i - 1];
                P.good2[i] = (log(a) - log(b)) >= P.degeneracy_tol;
            }
        }
        __syncthreads();
        combine_constraints(P, tid);
    }
    if (P.svd_min >= 0) {
        for (int64_t i = tid; i < n; i += TR_THREADS) P.good2[i] = P.ssort[i] >= P.svd_min;
        __syncthreads();
        combine_constraints(P, tid);
    }
    {
        const double t2 = P.trunc_cut * P.trunc_cut;
        for (int64_t i = tid; i < n; i += TR_THREADS) P.good2[i] = P.csum[i] > t2;
        __syncthreads();
        combine_constraints(P, tid);
    }
    // smallest / largest good cut
    if (tid == 0) s_cut = P.minimize_error ? (long long)n : -1ll;
    __syncthreads();
    long long mine = P.minimize_error ? (long long)n : -1ll;
    for (int64_t i = tid; i < n; i += TR_THREADS)
        if (P.good[i]) mine = P.minimize_error ? min(mine, (long long)i) : max(mine, (long long)i);
    if (P.minimize_error) atomicMin(&s_cut, mine); else atomicMax(&s_cut, mine);
    __syncthreads();
    long long cut = s_cut;
    if (cut < 0) cut = 0;
    if (cut > (long long)n) cut = n;
    for (int64_t i = tid; i < n; i += TR_THREADS) P.keep[P.piv[i]] = (i >= cut) ? 1 : 0;
    if (tid == 0) {
        const double total = n > 0 ? P.csum[n - 1] : 0.0;
        const double err = cut > 0 ? P.csum[cut - 1] : 0.0;
        // new_norm = sum(merr[cut:]) evaluated as a separate sum like the reference
        double nn = 0.0;
        for (int64_t i = cut; i < n; ++i) nn += P.merr[i];
        P.result[0] = err;
        P.result[1] = nn;
        P.result[2] = (double)(n - cut);
        P.result[3] = total;
    }
}

template <bool CX>
__global__ void sumsq_kernel(const double* __restrict__ x, int64_t ndoubles, double* __restrict__ out)
{
    __shared__ double s_red[32];
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < ndoubles; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = x[i];
        acc += v * v;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < (blockDim.x >> 5) ? s_red[threadIdx.x] : 0.0;
        v = warp_sum(v);
        if (threadIdx.x == 0) atomicAdd(out, v);
    }
}

__global__ void cutoff_inverse_kernel(const double* __restrict__ x, double* __restrict__ y, int64_t n, double cutoff)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = x[i];
        y[i] = fabs(v) < cutoff ? 0.0 : 1.0 / v;
    }
}

__global__ void axpby_kernel(const double* __restrict__ x, double* __restrict__ y, int64_t nd, double alpha, double beta)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nd; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = (beta == 0.0) ? alpha * x[i] : alpha * x[i] + beta * y[i];
}

// matrix i (m x n row-major, ES doubles per element): row r *= d[d_off + r] (axis 0) or col (axis 1)
__global__ void scale_rows_cols_kernel(double* __restrict__ x, const int64_t* __restrict__ dims,
                                       const int64_t* __restrict__ off, const double* __restrict__ d,
                                       const int64_t* __restrict__ d_off, int axis, int es)
{
    const int64_t mat = blockIdx.y;
    const int64_t m = dims[2 * mat], n = dims[2 * mat + 1];
    double* p = x + off[mat] * es;
    const double* dd = d + d_off[mat];
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < m * n; e += (int64_t)gridDim.x * blockDim.x) {
        const double f = axis == 0 ? dd[e / n] : dd[e % n];
        for (int k = 0; k < es; ++k) p[e * es + k] *= f;
    }
}

// keep-mask compaction of rows (axis 0) or columns (axis 1) of a batch of matrices:
// out_i = in_i[idx, :] or in_i[:, idx]; idx lists are concatenated in `idx` with idx_ptr
__global__ void gather_rows_cols_kernel(const double* __restrict__ src, double* __restrict__ dst,
                                        const int64_t* __restrict__ desc, const int64_t* __restrict__ idx,
                                        int axis, int es)
{
    // desc row: (m_in, n_in, src_off, dst_off, idx_begin, idx_count)
    const int64_t* D = desc + 6 * blockIdx.y;
    const int64_t m = D[0], n = D[1], k = D[5];
    const double* s = src + D[2] * es;
    double* o = dst + D[3] * es;
    const int64_t* ix = idx + D[4];
    const int64_t mo = axis == 0 ? k : m, no = axis == 0 ? n : k;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < mo * no; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / no, c = e % no;
        const int64_t sr = axis == 0 ? ix[r] : r, sc = axis == 0 ? c : ix[c];
        for (int q = 0; q < es; ++q) o[e * es + q] = s[(sr * n + sc) * es + q];
    }
}

// ---- scratch of cyb_truncate_select_static: no allocation at call time -------------------------------------------
// Same layout as the cudaMallocAsync scratch of cyb_truncate_select, sized for TR_STATIC_MAX_N values; every GPU that
// loads the library holds these TR_STATIC_BYTES (2.1 MiB) until the process exits.
constexpr int64_t TR_STATIC_MAX_N = 65536;          // == CYB_TRUNCATE_STATIC_MAX_N in include/cyten_b200.h
constexpr size_t tr_align256(size_t x) { return (x + 255) / 256 * 256; }
constexpr size_t TR_STATIC_BYTES = tr_align256(8 * (size_t)TR_STATIC_MAX_N) * 4 + tr_align256((size_t)TR_STATIC_MAX_N) * 2 + 256;
static __device__ __align__(256) unsigned char tr_static_scratch[TR_STATIC_BYTES];

// In host code `tr_static_scratch` is only a stand-in for the symbol, not a device address: the address a kernel may use
// comes from cudaGetSymbolAddress, asked once per device with that device current.
static int tr_static_address(unsigned char** out)
{
    static std::mutex mu;
    static unsigned char* addr[64] = {};
    int dev = 0;
    CYB_CHECK_CUDA(cudaGetDevice(&dev));
    CYB_REQUIRE(dev >= 0 && dev < 64, "truncate_select_static: device index %d outside [0, 64)", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        unsigned char* p = nullptr;
        CYB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, tr_static_scratch));
        CYB_REQUIRE(p != nullptr, "truncate_select_static: cudaGetSymbolAddress returned a null address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return CYB_OK;
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int cyb_truncate_select(const double* s, int64_t n, int64_t chi_max, int64_t chi_min,
                        double degeneracy_tol, double trunc_cut, double svd_min,
                        int minimize_error, uint8_t* keep, double* result, cyb_stream_t stream_)
{
    CYB_REQUIRE(n >= 1, "truncate_select: need at least one singular value");
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t N = (size_t)n;
    const size_t bytes = align_up(8 * N, 256) * 4 + align_up(N, 256) * 2 + 256;
    Scratch ws(bytes, stream);
    if (!ws.ptr) return CYB_ERR_CUDA;
    TruncArgs P;
    P.s = s; P.n = n; P.chi_max = chi_max; P.chi_min = chi_min; P.degeneracy_tol = degeneracy_tol;
    P.trunc_cut = trunc_cut; P.svd_min = svd_min; P.minimize_error = minimize_error;
    P.keep = keep; P.result = result;
    size_t o = 0;
    P.piv = ws.as<int64_t>(o); o += align_up(8 * N, 256);
    P.merr = ws.as<double>(o); o += align_up(8 * N, 256);
    P.ssort = ws.as<double>(o); o += align_up(8 * N, 256);
    P.csum = ws.as<double>(o); o += align_up(8 * N, 256);
    P.good = ws.as<uint8_t>(o); o += align_up(N, 256);
    P.good2 = ws.as<uint8_t>(o); o += align_up(N, 256);
    P.flag = ws.as<int64_t>(o);
    truncate_rank_kernel<<<(unsigned)((n + TR_RANK_THREADS - 1) / TR_RANK_THREADS), TR_RANK_THREADS, 0, stream>>>(P);
    truncate_select_kernel<<<1, TR_THREADS, 0, stream>>>(P);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_truncate_select_static(const double* s, int64_t n, int64_t chi_max, int64_t chi_min,
                               double degeneracy_tol, double trunc_cut, double svd_min,
                               int minimize_error, uint8_t* keep, double* result, cyb_stream_t stream_)
{
    CYB_REQUIRE(n >= 1, "truncate_select_static: need at least one singular value");
    if (n > TR_STATIC_MAX_N) {
        set_error("truncate_select_static: n = %lld exceeds the static scratch (CYB_TRUNCATE_STATIC_MAX_N = %lld values); "
                  "use cyb_truncate_select", (long long)n, (long long)TR_STATIC_MAX_N);
        return CYB_ERR_LIMIT;
    }
    CYB_REQUIRE(s != nullptr && keep != nullptr && result != nullptr, "truncate_select_static: s / keep / result must not be null");
    cudaStream_t stream = (cudaStream_t)stream_;
    unsigned char* base = nullptr;
    const int rc = tr_static_address(&base);
    if (rc != CYB_OK) return rc;
    const size_t N = (size_t)n;
    TruncArgs P;
    P.s = s; P.n = n; P.chi_max = chi_max; P.chi_min = chi_min; P.degeneracy_tol = degeneracy_tol;
    P.trunc_cut = trunc_cut; P.svd_min = svd_min; P.minimize_error = minimize_error;
    P.keep = keep; P.result = result;
    // The kernels write every scratch slot they later read (ranks are a permutation of 0 .. n-1), so nothing has to start
    // at zero and the previous call's contents are harmless.  Two calls that overlap on the GPU (different streams) race.
    size_t o = 0;
    P.piv = reinterpret_cast<int64_t*>(base + o); o += tr_align256(8 * N);
    P.merr = reinterpret_cast<double*>(base + o); o += tr_align256(8 * N);
    P.ssort = reinterpret_cast<double*>(base + o); o += tr_align256(8 * N);
    P.csum = reinterpret_cast<double*>(base + o); o += tr_align256(8 * N);
    P.good = reinterpret_cast<uint8_t*>(base + o); o += tr_align256(N);
    P.good2 = reinterpret_cast<uint8_t*>(base + o); o += tr_align256(N);
    P.flag = reinterpret_cast<int64_t*>(base + o);
    truncate_rank_kernel<<<(unsigned)((n + TR_RANK_THREADS - 1) / TR_RANK_THREADS), TR_RANK_THREADS, 0, stream>>>(P);
    truncate_select_kernel<<<1, TR_THREADS, 0, stream>>>(P);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_sumsq(const void* x, int64_t n, int dtype, double* out, cyb_stream_t stream_)
{
    cudaStream_t stream = (cudaStream_t)stream_;
    CYB_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(double), stream));
    const int64_t nd = n * (dtype == CYB_C128 ? 2 : 1);
    if (nd == 0) return CYB_OK;
    int64_t grid = (nd + 1023) / 1024;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    sumsq_kernel<false><<<(unsigned)grid, 256, 0, stream>>>(static_cast<const double*>(x), nd, out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_cutoff_inverse(const double* x, double* y, int64_t n, double cutoff, cyb_stream_t stream_)
{
    if (n == 0) return CYB_OK;
    int64_t grid = (n + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    cutoff_inverse_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream_>>>(x, y, n, cutoff);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_axpby(const void* x, void* y, int64_t n, int dtype, double alpha, double beta,
              cyb_stream_t stream_)
{
    const int64_t nd = n * (dtype == CYB_C128 ? 2 : 1);
    if (nd == 0) return CYB_OK;
    int64_t grid = (nd + 1023) / 1024;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    axpby_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream_>>>(static_cast<const double*>(x), static_cast<double*>(y), nd, alpha, beta);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_scale_rows_cols(void* x, const int64_t* dims, const int64_t* off, const double* d,
                        const int64_t* d_off, int64_t nmat, int axis, int dtype,
                        cyb_stream_t stream_)
{
    if (nmat <= 0) return CYB_OK;
    CYB_REQUIRE(nmat <= 65535, "scale_rows_cols: too many blocks");
    CYB_REQUIRE(axis == 0 || axis == 1, "scale_rows_cols: axis must be 0 or 1");
    dim3 grid(64, (unsigned)nmat);
    scale_rows_cols_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(static_cast<double*>(x), dims, off, d, d_off, axis, dtype == CYB_C128 ? 2 : 1);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_gather_rows_cols(const void* src, void* dst, const int64_t* desc, const int64_t* idx,
                         int64_t nmat, int axis, int dtype, cyb_stream_t stream_)
{
    if (nmat <= 0) return CYB_OK;
    CYB_REQUIRE(nmat <= 65535, "gather_rows_cols: too many blocks");
    CYB_REQUIRE(axis == 0 || axis == 1, "gather_rows_cols: axis must be 0 or 1");
    dim3 grid(64, (unsigned)nmat);
    gather_rows_cols_kernel<<<grid, 256, 0, (cudaStream_t)stream_>>>(static_cast<const double*>(src), static_cast<double*>(dst), desc, idx, axis, dtype == CYB_C128 ? 2 : 1);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // extern "C"
