// Block-batched BLAS-1 over tensor arenas: the Krylov/Lanczos vector operations that sit between two H_eff
// matvecs (reference: src/backends/abelian.cpp:2159-2211 inner, 2253-2301
// linear_combination, 2759-2778 mul, 2780-2791 norm, 3177-3232 scale_axis; loop: src/tensors/krylov_based.cpp:844-882).
//
// The reference walks the block lists on the host and makes one numpy call per block.  Here one launch covers
// all blocks of the tensors: a segment table (one row per pair of matching blocks: the element offsets of the
// block in every operand arena + its size) is cut into chunks of CHUNK double2 "pairs"; a warp takes chunks
// grid-stride, finds the chunk's segment by binary search in the (L1-resident) chunk prefix array and streams
// the chunk with 16-byte loads.  All kernels are HBM-bound; the algorithmic bytes are (operands read + result
// written) x block bytes.  Reductions are deterministic: one partial per CTA in a fixed chunk assignment, the
// last CTA to finish sums the partials in index order (no floating-point atomics).
#include "common.cuh"
#include <algorithm>

namespace cyb {

constexpr int B1_THREADS = 256;
constexpr int B1_CHUNK = 256;        // double2 pairs per chunk: 4 KB per operand, 8 pairs per lane (one batch of loads)
constexpr int B1_MAX_GRID = 148 * 8;
constexpr int B1_BATCH = 8;           // loads issued back to back per operand before the first use

struct B1Seg {                        // device view of the segment table
    const int64_t* seg;               // [nseg][ncol]: operand offsets (elements, -1 = block absent), then size
    const int64_t* chunk0;            // [nseg + 1]: first chunk of every segment
    int nseg, ncol;
    int64_t nchunk;
    int staged;                       // tables fit the dynamic shared memory of the launch
};

__device__ __forceinline__ int b1_find(const int64_t* chunk0, int nseg, int64_t c)
{
    int lo = 0, hi = nseg;            // chunk0[lo] <= c < chunk0[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (chunk0[mid] <= c) lo = mid; else hi = mid;
    }
    return lo;
}

// The tables are read by every warp for every chunk and a binary search is a chain of dependent loads: from global
// memory that is ~10 L2 round trips (5-8 us on a 20 us kernel).  They are staged in shared memory when they fit.
constexpr size_t B1_SMEM_MAX = 64 * 1024;
extern __shared__ int64_t b1_smem[];

__device__ __forceinline__ void b1_stage(const int64_t*& prefix, const int64_t*& rows, int nseg, int ncol, int staged)
{
    if (!staged) return;
    for (int i = threadIdx.x; i <= nseg; i += blockDim.x) b1_smem[i] = prefix[i];
    for (int i = threadIdx.x; i < nseg * ncol; i += blockDim.x) b1_smem[nseg + 1 + i] = rows[i];
    __syncthreads();
    prefix = b1_smem;
    rows = b1_smem + nseg + 1;
}

static size_t b1_smem_bytes(int64_t nseg, int ncol)
{
    const size_t need = (size_t)(nseg + 1 + nseg * ncol) * sizeof(int64_t);
    return need <= B1_SMEM_MAX ? need : 0;
}

__device__ __forceinline__ double2 b1_ld(const double* p, int64_t d, int64_t nd)
{
    if (d + 1 < nd) return *reinterpret_cast<const double2*>(p + d);
    return make_double2(p[d], 0.0);
}

// predicated variant: zero beyond the end of the block (a zero contributes nothing to a dot product or a sum)
__device__ __forceinline__ double2 b1_ldz(const double* p, int64_t d, int64_t nd)
{
    double2 r = make_double2(0.0, 0.0);
    if (d + 1 < nd) r = *reinterpret_cast<const double2*>(p + d);
    else if (d < nd) r.x = p[d];
    return r;
}

__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// CTA-wide deterministic sum of NV values per thread into partial[blockIdx.x * NV + v]; the last CTA adds the
// partials in order, applies nothing else and writes out[v]; it re-arms the counter for the next launch.
template <int NV>
__device__ __forceinline__ void b1_finish(double (&acc)[NV], double* __restrict__ partial, unsigned* __restrict__ counter,
                                          double* __restrict__ out)
{
    __shared__ double s_red[NV][B1_THREADS / 32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        double a = acc[v];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
        if (lane == 0) s_red[v][warp] = a;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double a = 0.0;
            for (int w = 0; w < B1_THREADS / 32; ++w) a += s_red[v][w];
            partial[(size_t)blockIdx.x * NV + v] = a;
        }
        __threadfence();
        s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < 32) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double a = 0.0;                                   // lane l sums blocks l, l+32, ... in order
            for (unsigned b = threadIdx.x; b < gridDim.x; b += 32) a += __ldcg(partial + (size_t)b * NV + v);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
            if (threadIdx.x == 0) out[v] = a;
        }
        if (threadIdx.x == 0) *counter = 0u;
    }
}

// out[0..1] = sum over segments of conj?(x) * y  (real dtype: out[1] = 0)
template <bool CX>
__global__ void __launch_bounds__(B1_THREADS) block_dot_kernel(const double* __restrict__ x, const double* __restrict__ y,
                                                               B1Seg T, int conj_x, double* __restrict__ partial,
                                                               unsigned* __restrict__ counter, double* __restrict__ out)
{
    constexpr int ES = CX ? 2 : 1;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * (B1_THREADS / 32) + (threadIdx.x >> 5);
    const int64_t nwarp = (int64_t)gridDim.x * (B1_THREADS / 32);
    double acc[2] = {0.0, 0.0};
    const int64_t *chunk0 = T.chunk0, *seg = T.seg;
    b1_stage(chunk0, seg, T.nseg, T.ncol, T.staged);
    for (int64_t c = warp0; c < T.nchunk; c += nwarp) {
        const int s = b1_find(chunk0, T.nseg, c);
        const int64_t* row = seg + (size_t)s * T.ncol;
        const int64_t ox = row[0], oy = row[1], nd = row[2] * ES;
        const double* px = x + ox * ES;
        const double* py = y + oy * ES;
        const int64_t d0 = (c - chunk0[s]) * (2 * B1_CHUNK);
        // all loads of a batch are issued before the first use: 8 x 2 x 16 B in flight per thread
#pragma unroll
        for (int j0 = 0; j0 < B1_CHUNK / 32; j0 += B1_BATCH) {
            if (d0 + (int64_t)(j0 * 32) * 2 >= nd) break;             // warp-uniform
            double2 a[B1_BATCH], b[B1_BATCH];
#pragma unroll
            for (int j = 0; j < B1_BATCH; ++j) {
                const int64_t d = d0 + (int64_t)((j0 + j) * 32 + lane) * 2;
                a[j] = b1_ldz(px, d, nd);
                b[j] = b1_ldz(py, d, nd);
            }
#pragma unroll
            for (int j = 0; j < B1_BATCH; ++j) {
                if (CX) {
                    const double ay = (conj_x & 1) ? -a[j].y : a[j].y;
                    acc[0] += a[j].x * b[j].x - ay * b[j].y;
                    acc[1] += a[j].x * b[j].y + ay * b[j].x;
                } else {
                    acc[0] += a[j].x * b[j].x + a[j].y * b[j].y;
                }
            }
        }
    }
    if (conj_x & 2) acc[1] = 0.0;     // caller wants Re <x, y> only (the Lanczos alpha of a Hermitian operator)
    b1_finish<2>(acc, partial, counter, out);
}

struct B1Coef {
    double c[3][2];                   // host factors (re, im) of x, y, z
    const double* dev[3];             // optional device factors (re, im): effective = host * device
};

// out = a x + b y + c z per segment (absent operands contribute 0); optional fused ||out||^2
template <bool CX, bool NORM>
__global__ void __launch_bounds__(B1_THREADS) block_lincomb_kernel(const double* x, const double* y, const double* z, double* o,   // o may alias an operand
                                                                   B1Seg T, B1Coef K, double* __restrict__ partial,
                                                                   unsigned* __restrict__ counter, double* __restrict__ sumsq)
{
    constexpr int ES = CX ? 2 : 1;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * (B1_THREADS / 32) + (threadIdx.x >> 5);
    const int64_t nwarp = (int64_t)gridDim.x * (B1_THREADS / 32);
    double2 k[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        k[i] = make_double2(K.c[i][0], K.c[i][1]);
        if (K.dev[i]) {
            const double2 d = make_double2(__ldg(K.dev[i]), CX ? __ldg(K.dev[i] + 1) : 0.0);
            k[i] = CX ? cmul(k[i], d) : make_double2(k[i].x * d.x, 0.0);
        }
    }
    const double* ops[3] = {x, y, z};
    double acc[1] = {0.0};
    const int64_t *chunk0 = T.chunk0, *seg = T.seg;
    b1_stage(chunk0, seg, T.nseg, T.ncol, T.staged);
    for (int64_t c = warp0; c < T.nchunk; c += nwarp) {
        const int s = b1_find(chunk0, T.nseg, c);
        const int64_t* row = seg + (size_t)s * T.ncol;
        const double* p[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const int64_t off = row[i];
            p[i] = (ops[i] != nullptr && off >= 0) ? ops[i] + off * ES : nullptr;
        }
        double* po = o + row[3] * ES;
        const int64_t nd = row[4] * ES;
        const int64_t d0 = (c - chunk0[s]) * (2 * B1_CHUNK);
#pragma unroll
        for (int j0 = 0; j0 < B1_CHUNK / 32; j0 += B1_BATCH / 2) {
            if (d0 + (int64_t)(j0 * 32) * 2 >= nd) break;             // warp-uniform
            double2 v[3][B1_BATCH / 2];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                if (p[i] == nullptr) continue;                        // warp-uniform
#pragma unroll
                for (int j = 0; j < B1_BATCH / 2; ++j)
                    v[i][j] = b1_ldz(p[i], d0 + (int64_t)((j0 + j) * 32 + lane) * 2, nd);
            }
#pragma unroll
            for (int j = 0; j < B1_BATCH / 2; ++j) {
                const int64_t d = d0 + (int64_t)((j0 + j) * 32 + lane) * 2;
                if (d >= nd) break;
                double2 r = make_double2(0.0, 0.0);
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    if (p[i] == nullptr) continue;
                    if (CX) { const double2 t = cmul(k[i], v[i][j]); r.x += t.x; r.y += t.y; }
                    else    { r.x += k[i].x * v[i][j].x; r.y += k[i].x * v[i][j].y; }
                }
                if (d + 1 < nd) *reinterpret_cast<double2*>(po + d) = r; else po[d] = r.x;
                if (NORM) acc[0] += r.x * r.x + (d + 1 < nd ? r.y * r.y : 0.0);
            }
        }
    }
    if (NORM) b1_finish<1>(acc, partial, counter, sumsq);
}

// out block = in block scaled along one axis: block viewed as (outer, dleg, inner); element (a, r, b) *= d[r]
// desc int64[nblk][6] = (off_in, off_out, outer, dleg, inner, d_off)
// row index of element e of a block viewed as (outer, dleg, inner); 32-bit arithmetic whenever the block allows it
// (a 64-bit div + mod per element costs more than the 16 bytes it moves), no division at all for the two common cases
// (scaled axis last: inner == 1; scaled axis first: outer == 1)
__device__ __forceinline__ int64_t b1_row(int64_t e, int64_t dleg, int64_t inner, bool small, bool first_axis)
{
    if (small) {
        const uint32_t q = inner == 1 ? (uint32_t)e : (uint32_t)e / (uint32_t)inner;
        return first_axis ? q : q % (uint32_t)dleg;
    }
    const int64_t q = e / inner;
    return first_axis ? q : q % dleg;
}

template <bool CX, bool DCX>
__global__ void __launch_bounds__(B1_THREADS) scale_axis_kernel(const double* __restrict__ in, double* __restrict__ out,
                                                                const int64_t* desc, const double* __restrict__ dg,
                                                                const int64_t* chunk0, int nblk, int64_t nchunk, int staged)
{
    constexpr int ES = CX ? 2 : 1;
    b1_stage(chunk0, desc, nblk, 6, staged);
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * (B1_THREADS / 32) + (threadIdx.x >> 5);
    const int64_t nwarp = (int64_t)gridDim.x * (B1_THREADS / 32);
    for (int64_t c = warp0; c < nchunk; c += nwarp) {            // same chunking as the dot / lincomb kernels
        const int b = b1_find(chunk0, nblk, c);
        const int64_t* D = desc + (size_t)b * 6;
        const int64_t outer = D[2], dleg = D[3], inner = D[4], n = outer * dleg * inner, nd = n * ES;
        const double* pi = in + D[0] * ES;
        double* po = out + D[1] * ES;
        const double* pd = dg + D[5] * (DCX ? 2 : 1);
        const int64_t d0 = (c - chunk0[b]) * (2 * B1_CHUNK);
        const bool small = n <= 0x7fffffffLL, first_axis = outer == 1;
        double2 v[B1_BATCH];
#pragma unroll
        for (int j = 0; j < B1_BATCH; ++j) v[j] = b1_ldz(pi, d0 + (int64_t)(j * 32 + lane) * 2, nd);
#pragma unroll
        for (int j = 0; j < B1_BATCH; ++j) {
            const int64_t d = d0 + (int64_t)(j * 32 + lane) * 2;
            if (d >= nd) break;
            if (CX) {
                const int64_t r = b1_row(d >> 1, dleg, inner, small, first_axis);
                const double2 s = DCX ? *reinterpret_cast<const double2*>(pd + 2 * r) : make_double2(__ldg(pd + r), 0.0);
                *reinterpret_cast<double2*>(po + d) = cmul(s, v[j]);
            } else {
                const double s0 = __ldg(pd + b1_row(d, dleg, inner, small, first_axis));
                if (d + 1 < nd) {
                    const double s1 = __ldg(pd + b1_row(d + 1, dleg, inner, small, first_axis));
                    *reinterpret_cast<double2*>(po + d) = make_double2(v[j].x * s0, v[j].y * s1);
                } else {
                    po[d] = v[j].x * s0;
                }
            }
        }
    }
}

// Persistent grid: as many CTAs as are resident at once (occupancy x SMs), every warp walks the chunks grid-stride,
// so there is no partial last wave of CTAs; small inputs get one warp per chunk.
template <class K>
static int b1_grid(K kernel, size_t smem, int64_t nchunk)
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, B1_THREADS, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int64_t want = (nchunk + B1_THREADS / 32 - 1) / (B1_THREADS / 32);
    const int64_t cap = std::min<int64_t>((int64_t)sm_count() * per_sm, B1_MAX_GRID);
    return (int)std::max<int64_t>(1, std::min<int64_t>(want, cap));
}

// dynamic shared memory above 48 KB is opt-in per kernel and per DEVICE
static cudaError_t b1_attr()
{
    static PerDeviceOnce configured;
    if (!configured.first()) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    auto set = [&](const void* f) {
        if (e == cudaSuccess) e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)B1_SMEM_MAX);
    };
    set((const void*)block_dot_kernel<true>); set((const void*)block_dot_kernel<false>);
    set((const void*)block_lincomb_kernel<true, true>); set((const void*)block_lincomb_kernel<true, false>);
    set((const void*)block_lincomb_kernel<false, true>); set((const void*)block_lincomb_kernel<false, false>);
    set((const void*)scale_axis_kernel<true, true>); set((const void*)scale_axis_kernel<true, false>);
    set((const void*)scale_axis_kernel<false, false>);
    return e;
}

}  // namespace cyb

using namespace cyb;

extern "C" {

int64_t cyb_blas1_workspace_bytes(void) { return (int64_t)(B1_MAX_GRID * 2 * sizeof(double) + 64); }

int64_t cyb_blas1_chunk_elems(int dtype) { return dtype == CYB_C128 ? B1_CHUNK : 2 * B1_CHUNK; }

int cyb_block_dot(const void* x, const void* y, const int64_t* seg, const int64_t* chunk0, int64_t nseg, int64_t nchunk,
                  int dtype, int conj_x, double* out, void* work, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "block_dot: dtype must be CYB_F64 or CYB_C128");
    CYB_REQUIRE(out != nullptr && work != nullptr, "block_dot: out / work must not be null");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (nseg == 0 || nchunk == 0) { CYB_CHECK_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), stream)); return CYB_OK; }
    const size_t smem = b1_smem_bytes(nseg, 3);
    B1Seg T{seg, chunk0, (int)nseg, 3, nchunk, smem != 0};
    double* partial = static_cast<double*>(work);
    unsigned* counter = reinterpret_cast<unsigned*>(static_cast<char*>(work) + B1_MAX_GRID * 2 * sizeof(double));
    CYB_CHECK_CUDA(b1_attr());
    if (dtype == CYB_C128)
        block_dot_kernel<true><<<b1_grid(block_dot_kernel<true>, smem, nchunk), B1_THREADS, smem, stream>>>(
            (const double*)x, (const double*)y, T, conj_x, partial, counter, out);
    else
        block_dot_kernel<false><<<b1_grid(block_dot_kernel<false>, smem, nchunk), B1_THREADS, smem, stream>>>(
            (const double*)x, (const double*)y, T, conj_x, partial, counter, out);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_block_lincomb(const void* x, const void* y, const void* z, void* out, const int64_t* seg, const int64_t* chunk0,
                      int64_t nseg, int64_t nchunk, int dtype, const double* coef, const double* const* coef_dev,
                      double* sumsq, void* work, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "block_lincomb: dtype must be CYB_F64 or CYB_C128");
    CYB_REQUIRE(out != nullptr && coef != nullptr, "block_lincomb: out / coef must not be null");
    CYB_REQUIRE(sumsq == nullptr || work != nullptr, "block_lincomb: the fused norm needs the workspace");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (nseg == 0 || nchunk == 0) {
        if (sumsq) CYB_CHECK_CUDA(cudaMemsetAsync(sumsq, 0, sizeof(double), stream));
        return CYB_OK;
    }
    const size_t smem = b1_smem_bytes(nseg, 5);
    B1Seg T{seg, chunk0, (int)nseg, 5, nchunk, smem != 0};
    CYB_CHECK_CUDA(b1_attr());
    B1Coef K;
    for (int i = 0; i < 3; ++i) {
        K.c[i][0] = coef[2 * i]; K.c[i][1] = coef[2 * i + 1];
        K.dev[i] = coef_dev ? coef_dev[i] : nullptr;
    }
    double* partial = static_cast<double*>(work);
    unsigned* counter = work ? reinterpret_cast<unsigned*>(static_cast<char*>(work) + B1_MAX_GRID * 2 * sizeof(double)) : nullptr;
    const double *px = (const double*)x, *py = (const double*)y, *pz = (const double*)z;
    double* po = (double*)out;
    const bool cx = dtype == CYB_C128;
    if (cx && sumsq)        block_lincomb_kernel<true, true><<<b1_grid(block_lincomb_kernel<true, true>, smem, nchunk), B1_THREADS, smem, stream>>>(px, py, pz, po, T, K, partial, counter, sumsq);
    else if (cx)            block_lincomb_kernel<true, false><<<b1_grid(block_lincomb_kernel<true, false>, smem, nchunk), B1_THREADS, smem, stream>>>(px, py, pz, po, T, K, partial, counter, sumsq);
    else if (sumsq)         block_lincomb_kernel<false, true><<<b1_grid(block_lincomb_kernel<false, true>, smem, nchunk), B1_THREADS, smem, stream>>>(px, py, pz, po, T, K, partial, counter, sumsq);
    else                    block_lincomb_kernel<false, false><<<b1_grid(block_lincomb_kernel<false, false>, smem, nchunk), B1_THREADS, smem, stream>>>(px, py, pz, po, T, K, partial, counter, sumsq);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

int cyb_scale_axis(const void* in, void* out, const int64_t* desc, const int64_t* work0, int64_t nblk, int64_t nwork,
                   const void* diag, int dtype, int diag_dtype, cyb_stream_t stream_)
{
    CYB_REQUIRE(dtype == CYB_F64 || dtype == CYB_C128, "scale_axis: dtype must be CYB_F64 or CYB_C128");
    CYB_REQUIRE(!(dtype == CYB_F64 && diag_dtype == CYB_C128), "scale_axis: promote the tensor to complex128 first");
    if (nblk == 0 || nwork == 0) return CYB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t smem = b1_smem_bytes(nblk, 6);
    const int staged = smem != 0;
    CYB_CHECK_CUDA(b1_attr());
    const double *pi = (const double*)in, *pd = (const double*)diag;
    double* po = (double*)out;
    if (dtype == CYB_C128 && diag_dtype == CYB_C128)
        scale_axis_kernel<true, true><<<b1_grid(scale_axis_kernel<true, true>, smem, nwork), B1_THREADS, smem, stream>>>(
            pi, po, desc, pd, work0, (int)nblk, nwork, staged);
    else if (dtype == CYB_C128)
        scale_axis_kernel<true, false><<<b1_grid(scale_axis_kernel<true, false>, smem, nwork), B1_THREADS, smem, stream>>>(
            pi, po, desc, pd, work0, (int)nblk, nwork, staged);
    else
        scale_axis_kernel<false, false><<<b1_grid(scale_axis_kernel<false, false>, smem, nwork), B1_THREADS, smem, stream>>>(
            pi, po, desc, pd, work0, (int)nblk, nwork, staged);
    CYB_LAUNCH_CHECK();
    return CYB_OK;
}

}  // extern "C"
